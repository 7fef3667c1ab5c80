"""Seeded people arrays for the post-step parity tests (keepTopNPeople / KeypointScaler / hand & face detectors)."""
import numpy as np

from openpose_b200._lib import POSE_BODY_25, POSE_COCO_18

PEOPLE_CASES = [("body25_12", POSE_BODY_25, 25, 12, 1), ("coco_7", POSE_COCO_18, 18, 7, 2),
                ("body25_1", POSE_BODY_25, 25, 1, 3), ("body25_40_ties", POSE_BODY_25, 25, 40, 4)]
SCALE_CASES = [(2, 1.7066667, 1.959, (1280, 720)), (1, 1.0, 1.959, (1280, 720)), (3, 1.0, 2.0, (1280, 720)),
               (4, 1.0, 2.0, (640, 480)), (5, 1.0, 2.0, (1921, 1080)), (6, 1.0, 2.0, (333, 777)),
               (0, 3.0, 2.0, (100, 100)), (2, 1.0, 1.0, (10, 10))]      # (ScaleMode, in->out, net->out, producer)
TOP_N = [1, 3, 5, 50, -1]


def make_people(parts: int, people: int, seed: int):
    """-> keypoints (people, parts, 3), scores (people). Scores straddle both detector thresholds (0.03 / 0.25);
    the `ties` case repeats whole people so keepTopPeople's on-threshold rule is exercised."""
    rng = np.random.default_rng(4000 + seed)
    kp = np.zeros((people, parts, 3), np.float32)
    kp[..., 0] = rng.uniform(0, 1280, (people, parts))
    kp[..., 1] = rng.uniform(0, 720, (people, parts))
    kp[..., 2] = rng.choice([0.0, 0.02, 0.04, 0.2, 0.3, 0.6, 0.9], size=(people, parts)) * rng.uniform(0.9, 1.1, (people, parts))
    missing = kp[..., 2] == 0
    kp[missing] = 0                                    # undetected parts are (0, 0, 0) in the reference
    scores = rng.uniform(0.1, 1.0, people).astype(np.float32)
    if people >= 40:
        for dst, src in ((5, 2), (9, 2), (17, 2), (30, 11), (31, 11)):
            kp[dst], scores[dst] = kp[src], scores[src]
    return kp.astype(np.float32), scores
