// empty stand-in: the reference sources compiled for the harness include but never use it
