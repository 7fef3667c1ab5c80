"""Seeded cases for the output-serialisation tests (tests/test_filestream_cpu.py, tests/golden/make_json_golden.py)."""
import numpy as np

NAMES9 = ["person_id", "pose_keypoints_2d", "face_keypoints_2d", "hand_left_keypoints_2d", "hand_right_keypoints_2d",
          "pose_keypoints_3d", "face_keypoints_3d", "hand_left_keypoints_3d", "hand_right_keypoints_3d"]


def _keypoints(rng, people, parts, dims=3, missing=0.3):
    kp = np.zeros((people, parts, dims), np.float32)
    kp[..., 0] = rng.uniform(0, 1920, (people, parts))
    kp[..., 1] = rng.uniform(0, 1080, (people, parts))
    if dims == 4:
        kp[..., 2] = rng.uniform(-3, 3, (people, parts))
    kp[..., -1] = rng.uniform(0.05, 1, (people, parts))
    kp[rng.uniform(size=(people, parts)) < missing] = 0
    return kp


def people_cases():
    """name -> (keypointVector, candidates)"""
    rng = np.random.default_rng(77)
    cases = {}
    cases["empty"] = ([(None, n) for n in NAMES9], None)
    body = _keypoints(rng, 3, 25)
    cases["body_only"] = ([(np.full((3,), -1, np.float32), NAMES9[0]), (body, NAMES9[1])] +
                          [(None, n) for n in NAMES9[2:]], None)
    face, hl, hr = _keypoints(rng, 3, 70), _keypoints(rng, 3, 21), _keypoints(rng, 3, 21)
    cases["body_face_hands"] = ([(np.arange(3, dtype=np.float32), NAMES9[0]), (body, NAMES9[1]), (face, NAMES9[2]),
                                 (hl, NAMES9[3]), (hr, NAMES9[4]), (_keypoints(rng, 3, 25, 4), NAMES9[5]),
                                 (None, NAMES9[6]), (None, NAMES9[7]), (None, NAMES9[8])], None)
    cand = [np.stack([rng.uniform(0, 1920, n), rng.uniform(0, 1080, n), rng.uniform(0.05, 1, n)], 1).astype(np.float32)
            for n in rng.integers(0, 6, 25)]
    cases["with_candidates"] = ([(body, NAMES9[1])], cand)
    cases["candidates_no_people"] = ([(None, NAMES9[1])], [np.zeros((0, 3), np.float32)] * 18)
    # ragged: the second array has fewer people than the first (the reference reads the rows it has: only use equal
    # or larger per the max rule with rows present) -> same people count, different part counts
    cases["coco18_single"] = ([(_keypoints(rng, 1, 18), NAMES9[1])], None)
    # number formats: tiny, huge, negative, integers, ties
    odd = np.array([[[1e-7, -1e-5, 0.0001], [0.00012345678, 123456.5, 999999.4], [999999.5, 1e6, 1.5e7],
                     [-0.0, 0.1, 100000.5], [2.5, 0.30000001, 1234.5625], [float(np.float32(9.9999995)), 99999.95, 5e-5]]],
                   np.float32)
    cases["number_formats"] = ([(odd, NAMES9[1])], None)
    return cases


def coco_cases():
    """name -> dict(frames=[(kp, scores)], names=[...], parts, pose_model, variants, fmt)"""
    rng = np.random.default_rng(78)
    cases = {}
    frames = []
    for people in (2, 0, 4, 1):
        kp = _keypoints(rng, people, 25, missing=0.4)
        if people == 4:
            kp[1] = 0          # a person without any valid keypoint is skipped
            kp[2, 19:] = 0     # a person without foot keypoints: skipped in the foot file only
        frames.append((kp, rng.uniform(0, 1, people).astype(np.float32)))
    names = ["COCO_val2014_000000000139", "000000000285", "img_00632", "x/y/000000001000"]
    cases["body25_body_foot"] = dict(frames=frames, names=names, parts=25, pose_model=0, variants=3, fmt=0)
    cases["body25_body"] = dict(frames=frames, names=names, parts=25, pose_model=0, variants=1, fmt=0)
    frames18 = [(_keypoints(rng, p, 18), rng.uniform(0, 1, p).astype(np.float32)) for p in (3, 1)]
    cases["coco18"] = dict(frames=frames18, names=["000000000001", "000000000002"], parts=18, pose_model=1,
                           variants=1, fmt=0)
    return cases
