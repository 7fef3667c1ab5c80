"""Debug/evidence tool (GPU): per-layer error of NetB200 against the FP32 oracle and its bf16-rounding twin.

    python -m tests.net_layerwise [model] [H] [W] [N]

Prints, for every conv/pool/concat top in network order, max-abs error vs (a) the FP32 restatement and (b) the
same graph with bf16-rounded conv inputs/weights (what the B200 path computes), so a bug (large error vs b)
is told apart from rounding (small vs b, growing slowly vs a).
"""
import sys

import numpy as np


def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "body_25"
    H = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    W = int(sys.argv[3]) if len(sys.argv) > 3 else 96
    N = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    import torch
    from openpose_b200 import synth
    from openpose_b200.net import NetB200
    from oracle import ref_net
    proto, cm, ref = ref_net.build(model, 1234)
    x = synth.frames_to_net_input(synth.frames(N, H, W, 7))
    want32 = ref.forward(x)
    want16 = ref.forward(x, bf16_activations=True)
    net = NetB200(proto, cm)
    net.initializationOnThread()
    net.forwardPass(x)
    torch.cuda.synchronize()
    worst = 0.0
    for L in ref.netdef.layers:
        if L.type in ("ReLU", "PReLU"):
            continue
        top = L.tops[0]
        got = net.readBlob(top) if top != "net_output" or L.type != "Concat" else net.getOutputBlobArray().cpu().numpy()
        e32 = float(np.abs(got - want32[top]).max())
        e16 = float(np.abs(got - want16[top]).max())
        rms = float(np.sqrt((want32[top] ** 2).mean()))
        flag = "  <-- BAD" if not np.isfinite(e16) or e16 > 0.05 * max(rms, 1e-3) + 0.02 else ""
        worst = max(worst, e16)
        print("%-28s %-12s %-18s rms %.4f  err_vs_fp32 %.5f  err_vs_bf16sim %.5f%s"
              % (top, L.type, str(got.shape), rms, e32, e16, flag))
    out = net.getOutputBlobArray().cpu().numpy()
    print("net_output: max-abs vs fp32 oracle %.5f (tolerance 2e-2), vs bf16 twin %.5f, out rms %.4f"
          % (np.abs(out - want32["net_output"]).max(), np.abs(out - want16["net_output"]).max(),
             np.sqrt((want32["net_output"] ** 2).mean())))
    inf = net.info()
    print("convs %d launches %d GFLOP %.2f arena MB %.1f" % (inf.conv_layers, inf.kernel_launches_per_forward,
                                                             inf.flops_per_forward / 1e9, inf.arena_bytes / 2 ** 20))


if __name__ == "__main__":
    main()
