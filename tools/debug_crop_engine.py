import numpy as np, torch
from openpose_b200 import modelgen, synth
from openpose_b200.crop_extractors import _CropNet, crop_frames_gpu, crop_matrix
from openpose_b200.engine import CropEngine
cuda = torch.device("cuda:0")
proto, cm, _, _ = modelgen.build_files("hand", 4321)
w, h, side = 320, 240, 96
frame = synth.frames(1, h, w, 70)[0]
m = np.stack([crop_matrix((30.0, 40.0, 120.0, 120.0), side, False), crop_matrix((100.0, 50.0, 90.0, 90.0), side, True)])
for batch in (2, 4):
    eng = CropEngine(proto, cm, 22, gpus=[0], frame_size=(w, h), net_side=side, batch=batch, lanes=2)
    eng.submit(1, frame, m)
    fid, p1 = eng.next()
    eng.submit(2, frame, m)
    fid, p2 = eng.next()
    eng.close()
    ref = _CropNet(proto, cm, 0, 8.0); ref.init()
    x = crop_frames_gpu(torch.from_numpy(frame).to(cuda), list(m), (side, side))
    w1 = ref.run(x).copy(); w2 = ref.run(x).copy()
    xx = torch.zeros((batch, 3, side, side), device=cuda); xx[:2] = x
    w3 = ref.run(xx)[:2].copy()
    ref.net.close()
    print("batch", batch, "engine repeat equal", np.array_equal(p1, p2), "ref repeat equal", np.array_equal(w1, w2),
          "engine==ref", np.array_equal(p1, w1), "engine==ref(padded batch)", np.array_equal(p1, w3), "ref==ref(padded)", np.array_equal(w1, w3))
    print(p1[0, :3], w1[0, :3], w3[0, :3])
