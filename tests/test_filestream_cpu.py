"""Output serialisation: the product's JSON writers (libopb200, csrc/host/people_json.cpp,
through the C ABI) against (1) the committed bytes the reference's own object code wrote (tests/golden/
json_golden.npz), (2) that object code live where oracle/_ref exists, (3) Python's correctly rounded '%g' for the
number formatter, plus the batch call, the file writers and the error rules. Host only: no GPU needed."""
import json
import os

import numpy as np
import pytest

from oracle.post import Ref
from openpose_b200 import filestream as fs
from openpose_b200._lib import Opb200Error
from tests import cases_json

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "json_golden.npz")
PEOPLE = cases_json.people_cases()
COCO = cases_json.coco_cases()


def _coco_files(c, hr, tmp_path):
    path = str(tmp_path / "coco.json")
    s = fs.CocoJsonSaver(path, c["pose_model"], bool(hr), c["variants"], c["fmt"])
    for f, ((kp, sc), name) in enumerate(zip(c["frames"], c["names"])):
        s.record(kp, sc, name, f)
    s.close()
    return {fn: open(str(tmp_path / fn), "rb").read() for fn in sorted(os.listdir(str(tmp_path)))}


@pytest.mark.parametrize("hr", [0, 1])
@pytest.mark.parametrize("case", sorted(PEOPLE))
def test_people_json_equals_reference_bytes(case, hr):
    kv, cand = PEOPLE[case]
    got = fs.peopleJsonBytes(kv, cand, bool(hr))
    assert got == np.load(GOLDEN)["people/%s/%d" % (case, hr)].tobytes()
    doc = json.loads(got.replace(b"-nan", b"NaN").replace(b"nan", b"NaN"))
    assert doc["version"] == 1.3 and isinstance(doc["people"], list)


@pytest.mark.parametrize("hr", [0, 1])
@pytest.mark.parametrize("case", sorted(COCO))
def test_coco_json_equals_reference_bytes(case, hr, tmp_path):
    g = np.load(GOLDEN)
    files = _coco_files(COCO[case], hr, tmp_path)
    want = {k.split("/", 3)[3]: g[k].tobytes() for k in g.files if k.startswith("coco/%s/%d/" % (case, hr))}
    assert files == want
    for data in files.values():
        assert isinstance(json.loads(data), list)


@pytest.mark.skipif(not Ref.available(), reason="reference object code not built")
def test_fixture_still_equals_reference_object_code(tmp_path):
    ref, g = Ref(), np.load(GOLDEN)
    for name, (kv, cand) in PEOPLE.items():
        for hr in (0, 1):
            assert ref.save_people_json(kv, cand, bool(hr)) == g["people/%s/%d" % (name, hr)].tobytes()
    for name, c in COCO.items():
        files = ref.coco_json(c["frames"], c["names"], c["parts"], c["pose_model"], True, c["variants"], c["fmt"])
        for fn, data in files.items():
            assert data == g["coco/%s/1/%s" % (name, fn)].tobytes()


@pytest.mark.skipif(not Ref.available(), reason="reference object code not built")
def test_random_arrays_against_reference_object_code():
    ref = Ref()
    rng = np.random.default_rng(5)
    for trial in range(25):
        people = int(rng.integers(0, 5))
        kv = []
        for name in cases_json.NAMES9[:int(rng.integers(1, 9))]:
            kind = int(rng.integers(0, 3))
            if kind == 0 or people == 0:
                kv.append((None, name))
            elif kind == 1:
                kv.append((rng.normal(0, 10.0 ** rng.integers(-6, 7), (people,)).astype(np.float32), name))
            else:
                parts, dims = int(rng.integers(1, 30)), int(rng.integers(1, 5))
                kv.append((rng.normal(0, 10.0 ** rng.integers(-6, 7), (people, parts, dims)).astype(np.float32), name))
        cand = None
        if trial % 3 == 0:
            cand = [rng.normal(0, 300, (int(rng.integers(0, 4)), 3)).astype(np.float32) for _ in range(int(rng.integers(1, 26)))]
        for hr in (False, True):
            assert fs.peopleJsonBytes(kv, cand, hr) == ref.save_people_json(kv, cand, hr), trial


def test_number_format_is_ostream_float():
    """`ostream << float` = printf('%g') of the float's exact value; Python's % operator is correctly rounded too."""
    rng = np.random.default_rng(11)
    vals = [rng.uniform(-2000, 2000, 60000), rng.normal(0, 1, 20000) * 10.0 ** rng.integers(-9, 9, 20000),
            rng.integers(-1000000, 1000000, 20000).astype(np.float64) / 2.0 ** rng.integers(0, 12, 20000),
            np.round(rng.uniform(0, 1000, 20000), 1), np.round(rng.uniform(0, 1, 20000), 5)]
    # six-digit ties that are exactly representable, carries at every power of ten, range boundaries
    ties = [(2 * n + 1) * 5.0 * 10.0 ** e for n in range(49990, 50010) for e in range(-7, 2)]
    edges = [10.0 ** e * (1 + s * 2.0 ** -k) for e in range(-6, 8) for s in (-1, 0, 1) for k in (20, 22, 23, 24)]
    near = [9.9999949 * 10.0 ** e for e in range(-5, 6)] + [9.9999951 * 10.0 ** e for e in range(-5, 6)]
    v = np.concatenate([np.asarray(x, np.float64) for x in vals + [ties, edges, near]]).astype(np.float32)
    v = np.concatenate([v, np.array([0.0, -0.0, np.inf, -np.inf, 1e-45, 3.4e38, 1e-4, 999999.44, 999999.5], np.float32)])
    got = fs.peopleJsonBytes([(v, "x")], None, False)
    body = got[len(b'{"version":1.3,"people":['):-2]
    rows = body.split(b"},{")
    assert len(rows) == v.size
    for x, row in zip(v, rows):
        txt = row.split(b"[", 1)[1].split(b"]", 1)[0].decode()
        assert txt == "%g" % float(x), (float(x).hex(), txt)
    assert fs.peopleJsonBytes([(np.array([np.nan], np.float32), "x")], None, False).count(b"nan") == 1


def test_batch_equals_frame_by_frame():
    rng = np.random.default_rng(3)
    frames, cap, parts = 6, 8, 25
    kp = rng.uniform(0, 700, (frames, cap, parts, 3)).astype(np.float32)
    counts = np.array([0, 1, 8, 3, 0, 5], np.int32)
    for hr in (False, True):
        docs = fs.peopleJsonBatch(kp, counts, hr)
        assert len(docs) == frames
        for f in range(frames):
            n = int(counts[f])
            kv = [(np.full((n,), -1, np.float32), "person_id"), (kp[f, :n], "pose_keypoints_2d")] + \
                 [(None, name) for name in cases_json.NAMES9[2:]]
            assert docs[f] == fs.peopleJsonBytes(kv, None, hr)
    with pytest.raises(Opb200Error):
        fs.peopleJsonBatch(kp, np.array([0, 1, 9, 3, 0, 5], np.int32))


def test_savers_write_the_same_bytes_to_disk(tmp_path):
    kv, cand = PEOPLE["with_candidates"]
    saver = fs.PeopleJsonSaver(str(tmp_path / "out" / "json"))
    saver.save(kv, cand, "000000000012_keypoints", False)
    data = open(str(tmp_path / "out" / "json" / "000000000012_keypoints.json"), "rb").read()
    assert data == fs.peopleJsonBytes(kv, cand, False)
    with pytest.raises(Opb200Error, match="could not be opened"):
        fs.savePeopleJson(kv, cand, str(tmp_path / "missing_dir" / "a.json"), True)


def test_error_rules(tmp_path):
    with pytest.raises(Opb200Error, match="getNumberDimensions"):            # fileStream.cpp:314-317
        fs.peopleJsonBytes([(np.zeros((2, 3), np.float32), "x")], None, True)
    with pytest.raises(Opb200Error, match="Empty path"):                      # cocoJsonSaver.cpp:37-39
        fs.CocoJsonSaver("")
    with pytest.raises(Opb200Error, match="Unknown value for cocoJsonFormat"):  # :40-42
        fs.CocoJsonSaver(str(tmp_path / "c.json"), 0, True, 32)
    s = fs.CocoJsonSaver(str(tmp_path / "c.json"))
    kp = np.ones((1, 25, 3), np.float32)
    with pytest.raises(Opb200Error, match="Dimension mismatch"):              # :101-102
        s.record(kp, np.ones((2,), np.float32), "000001")
    with pytest.raises(Opb200Error, match="cannot obtain the ID"):            # :9-26
        s.record(kp, np.ones((1,), np.float32), "no_number_here_")
    with pytest.raises(Opb200Error, match="Invalid number of body parts"):    # :196-198
        s.record(np.ones((1, 15, 3), np.float32), np.ones((1,), np.float32), "000001")
    s.close()
