"""Placeholder: the op::Net backend binding (NetB200) lands here."""
