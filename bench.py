#!/usr/bin/env python
"""bench.py -- BODY_25 656x368 hot-path throughput (frames/s) on N B200s, frame-sharded (no collective).

A "step" = one batch of B synthetic frames through the hot path on every rank:
  [net forward (NetB200, bf16 tcgen05 convs)] -> resizeAndMerge x8 -> nms -> bodyPartConnector
`value`     : frames/s with the step's inputs already resident in HBM (CUDA events, max over ranks). A random-init net
              emits noise (no trained weights offline), so by default the post-processing of the step runs on planted
              8-person net outputs of the same shape while the net runs on the frames (--post-input net: on the net's own
              output); rank 0 checks the people it assembles against the reference's CPU code (`parity_check`).
`e2e`       : frames/s through the library's frame-level API (opb200_engine_*): u8 BGR 1280x720 frames in pinned host
              memory -> H2D -> device pre-processing -> net -> resize -> nms -> connector (on the net's own output) -> D2H
              of keypoints/scores/counts, every copy inside the timed region, two batches in flight.
`ref_gpu_ms`: the reference's OWN CUDA kernels (oracle/_ref/libopref_cuda.so) for the post-processing stages on one frame
              of the same inputs, same box: the GPU-vs-GPU baseline next to `stages_ms`.
`latency`   : one frame (batch 1) through the same path, device-resident input.
`roofline`  : the dominant kernel of the step (conv family when the net is on the path: tensor bound;
              resizeAndMerge otherwise: hbm bound), timed live with CUDA events on the launching stream.
`cpu_baseline` / `--impl reference`: the reference's CPU code for the path on the host cores
              (oracle/_ref = the reference's own nms/connector object code; oracle port for resize;
              torch-CPU restatement of Caffe for the net, labelled).
Inputs per step are larger than L2 (B x 75 MB heat-maps), so no explicit L2 flush is needed; see config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

NET_H, NET_W = 368, 656          # BODY_25 net resolution 656x368 (BASELINE.json metric)
FRAME_H, FRAME_W = 720, 1280     # producer resolution of the end-to-end leg (BASELINE.json configs[0])
OUT_H, OUT_W = NET_H // 8, NET_W // 8
PARTS, CHANNELS = 25, 78
MAXP = 127
SCALE_NET_TO_OUTPUT = 1.95913    # 1280x720 frame -> 656x368 net
NMS_OFFSET = float(0.5 / SCALE_NET_TO_OUTPUT)
CONV_GFLOP_PER_FRAME = 287.26
RESIZE_BYTES = 4 * (CHANNELS * OUT_H * OUT_W + CHANNELS * NET_H * NET_W)      # 76.50 MB
NMS_BYTES = 4 * PARTS * NET_H * NET_W + 4 * PARTS * (MAXP + 1) * 3            # 24.18 MB


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


def conv_traffic(batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum over the conv launches of one step, from the committed ncu capture
    of this command at the same batch (profiles/r02_conv_traffic.json); None when there is no capture for this batch."""
    p = os.path.join(ROOT, "profiles", "r02_conv_traffic.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return float(d["dram_bytes_per_step"]) if int(d["batch"]) == batch else None
    except (OSError, ValueError, KeyError):
        return None


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for nme, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------------
def planted_net_outputs(batch: int, people: int, seed: int) -> np.ndarray:
    """(batch, 78, 46, 82) planted net-output-resolution maps (SURVEY.md 8(d) C3/C5 generator)."""
    from openpose_b200 import synth
    outs = []
    for b in range(batch):
        m, _ = synth.planted_maps(0, OUT_H, OUT_W, people, seed=seed + b, sigma=1.2, paf_halfwidth=0.8,
                                  noise=0.005, limb_len=(3.0, 7.0))
        outs.append(m)
    return np.stack(outs).astype(np.float32)


class PostPath:
    """resizeAndMerge -> nms -> bodyPartConnector on the device, through the C ABI (openpose_b200.post)."""

    def __init__(self, batch: int, device, capacity: int = 256):
        import torch
        from openpose_b200 import post
        self.torch, self.post, self.batch, self.device = torch, post, batch, device
        self.heat = torch.empty((batch, CHANNELS, NET_H, NET_W), device=device)
        self.peaks = torch.zeros((batch, PARTS, MAXP + 1, 3), device=device)
        self.params = post.make_connect_params(0, scale_factor=SCALE_NET_TO_OUTPUT)
        self.kp = torch.empty((batch, capacity, PARTS, 3), device=device)
        self.sc = torch.empty((batch, capacity), device=device)
        self.cnt = torch.empty((batch,), dtype=torch.int32, device=device)
        self.capacity = capacity
        # kernels + memsets per step: resize 1, nms (ticket memset + kernel) 2, connector (memset + candidates + match +
        # assemble) 4 -- csrc/post/{resize_merge,nms,paf_connect}.cu
        self.launches_per_step = 7

    def run(self, net_out, ev=None):
        """net_out: (B,78,46,82) device tensor. ev: optional list collecting (name, start, end) events."""
        t, post = self.torch, self.post

        def stage(name, fn):
            if ev is None:
                fn()
                return
            a, b = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            ev.append((name, a, b))
        stage("resize", lambda: post.resize_and_merge_gpu(self.heat, [net_out]))
        stage("nms", lambda: post.nms_gpu(self.peaks, self.heat, 0.05, (NMS_OFFSET, NMS_OFFSET)))
        stage("connect", lambda: post.connect_body_parts_gpu(self.heat, self.peaks, self.params,
                                                             capacity=self.capacity,
                                                             out=(self.kp, self.sc, self.cnt)))
        return self.kp, self.sc, self.cnt


def cpu_reference_post(net_out: np.ndarray):
    """The reference's CPU post-processing for ONE frame: oracle port resize (cuda.hu math, the judged
    semantics) + the reference's own nmsCpu / connectBodyPartsCpu object code when oracle/_ref exists."""
    from oracle.post import Port, Ref
    port = Port()
    use_ref = Ref.available()
    impl = Ref() if use_ref else port
    t0 = time.perf_counter()
    try:  # what resizeAndMergeCpu itself calls (resizeAndMergeBase.cpp:51): cv::resize INTER_CUBIC per channel
        import cv2
        cv2.setNumThreads(1)
        heat_cv = np.empty((CHANNELS, NET_H, NET_W), np.float32)
        for c in range(CHANNELS):
            cv2.resize(net_out[c], (NET_W, NET_H), dst=heat_cv[c], interpolation=cv2.INTER_CUBIC)
        resize_impl = "cv2.resize INTER_CUBIC"
    except ImportError:
        heat_cv = None
        resize_impl = "oracle port (scalar C)"
    t1 = time.perf_counter()
    heat = port.resize_merge([net_out], NET_H, NET_W)   # untimed when cv2 exists: cuda.hu semantics for nms input
    if heat_cv is None:
        t1 = time.perf_counter()
    t1b = time.perf_counter()
    peaks = impl.nms(heat, 0.05, PARTS, off=(NMS_OFFSET, NMS_OFFSET))
    t2 = time.perf_counter()
    kp, sc = impl.connect(heat, peaks, 0, scale=SCALE_NET_TO_OUTPUT)[:2]
    t3 = time.perf_counter()
    return dict(keypoints=kp, scores=sc, resize_ms=(t1 - t0) * 1e3, nms_ms=(t2 - t1b) * 1e3, connect_ms=(t3 - t2) * 1e3,
                total_s=(t1 - t0) + (t3 - t1b), resize_impl=resize_impl, people=int(kp.shape[0]), kind="reference" if use_ref else "port")


# ----------------------------------------------------------------------------------------------------
# distributed helpers (frame sharding: no data-path collective; gloo/nccl only for barrier + max)
# ----------------------------------------------------------------------------------------------------
def dist_init(n_gpus: int):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        backend = os.environ.get("OPB200_DIST_BACKEND", "nccl")   # "gloo": the CPU tests of this plumbing
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def whole_job_rate(units_local: float, seconds_local: float, world: int, device) -> float:
    """Units all ranks processed / the slowest rank's time (frame sharding: ranks never exchange data)."""
    return sum_over_ranks(units_local, world, device) / max_over_ranks(seconds_local, world, device)


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ----------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    rank, world, local = dist_init(args.gpus)
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback on this path)"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    from openpose_b200 import _lib
    _lib.lib()  # fail loudly if libopb200.so is missing
    peaks_tbl = load_peaks()
    B = args.batch
    have_net = False
    net = None
    launches_net = 0
    if args.stages in ("full", "net"):
        from openpose_b200 import net as netmod
        net = netmod.bench_net(B, NET_H, NET_W, dev)      # NetB200 with seeded random-init BODY_25 weights
        have_net = True
    post = PostPath(B, dev) if args.stages in ("full", "post") else None

    # synthetic inputs (per rank, different seeds: weak scaling, fixed per-GPU work)
    planted_dev = None
    if have_net:
        from openpose_b200 import synth
        host_in = torch.from_numpy(synth.frames_to_net_input(synth.frames(B, NET_H, NET_W, 100 + rank)))
        if post is not None and args.post_input == "planted":
            planted_host = planted_net_outputs(B, args.people, 1000 * (rank + 1))
            planted_dev = torch.from_numpy(planted_host).to(dev)
    else:
        host_in = torch.from_numpy(planted_net_outputs(B, args.people, 1000 * (rank + 1)))
    host_in = host_in.pin_memory()
    dev_in = host_in.to(dev)
    h_kp = torch.empty((B, 256, PARTS, 3)).pin_memory()
    h_sc = torch.empty((B, 256)).pin_memory()
    h_cnt = torch.empty((B,), dtype=torch.int32).pin_memory()

    def step_device(ev=None):
        x = dev_in
        if have_net:
            x = net.forward_device(x, ev)
        if post is not None:
            return post.run(planted_dev if planted_dev is not None else x, ev)
        return x

    def step_host():
        if have_net:   # the public call takes the HOST array (as op::Net::forwardPass does); H2D happens inside it
            x = net.forward_device(host_in)
            out = post.run(x) if post is not None else x
        else:
            dev_in.copy_(host_in, non_blocking=True)
            out = step_device()
        if post is not None:
            h_kp.copy_(out[0], non_blocking=True)
            h_sc.copy_(out[1], non_blocking=True)
            h_cnt.copy_(out[2], non_blocking=True)
        else:
            net.copy_output_to_host()
        torch.cuda.current_stream().synchronize()

    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()

    # ---- timed region 1: device-resident inputs -----------------------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier(world)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_events = []
    e0.record()
    for _ in range(args.steps):
        step_device(stage_events)
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    ms_total = max_over_ranks(e0.elapsed_time(e1), world, dev)
    clocks = sampler.stop() if rank == 0 else None
    frames_all = sum_over_ranks(float(B * args.steps), world, dev)
    value = frames_all / (ms_total * 1e-3)

    if have_net:
        launches_net = net.launches_per_step() + 1        # + the input copy node
    stage_ms = {}
    for name, a, b in stage_events:
        stage_ms.setdefault(name, []).append(a.elapsed_time(b))
    stage_avg = {k: float(np.mean(v)) for k, v in stage_ms.items()}

    # ---- timed region 2: end to end through host buffers -------------------------------------------
    parity = None
    if post is not None and planted_dev is not None and rank == 0:
        parity = (post.kp[0].cpu().numpy(), post.sc[0].cpu().numpy(), int(post.cnt[0].item()), planted_host[0])
    engine_mode = have_net and post is not None and not args.no_engine
    if engine_mode:
        # The library's frame-level API: u8 frames in pinned staging -> people, two batches in flight per device. The
        # staging buffers are filled once (the frames stay in pinned host memory); every timed step copies its batch H2D
        # and its results D2H. Steady state: each iteration takes the oldest batch and submits the next.
        from openpose_b200 import modelgen, synth
        from openpose_b200.engine import PoseEngine
        net.net.close()            # the engine owns its own net; free this one's arena first
        del dev_in
        torch.cuda.empty_cache()
        proto, cm, _, _ = modelgen.build_files("body_25", 1234)
        eng = PoseEngine(proto, cm, gpus=[local], net_size=(NET_W, NET_H), frame_size=(FRAME_W, FRAME_H), batch=B,
                         lanes=2, capacity=256)
        frames = synth.frames(B, FRAME_H, FRAME_W, 300 + rank)
        fid = 0
        for _ in range(2):                       # fill both lanes' staging buffers
            eng.acquire()[:] = frames
            eng.commit(fid, B)
            fid += B
        for _ in range(3):                       # warm-up in steady state
            eng.next(copy=False)
            eng.acquire()
            eng.commit(fid, B)
            fid += B
        barrier(world)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            eng.next(copy=False)
            eng.acquire()
            eng.commit(fid, B)
            fid += B
        wall = time.perf_counter() - t0
        last_id = -1
        while eng.in_flight():
            last_id, kp_e, _, cnt_e = eng.next(copy=False)
        assert last_id == fid - B and kp_e.shape[0] == B and cnt_e.shape[0] == B, "engine lost a batch"
        barrier(world)
        eng.close()
        ms_e2e = max_over_ranks(wall * 1e3, world, dev)
        h2d = B * FRAME_H * FRAME_W * 3
        d2h = B * (256 * PARTS * 3 + 256 + 1) * 4
        e2e_mode = ("opb200_engine_*: u8 BGR %dx%d frames in pinned host staging -> H2D -> device pre-processing -> net -> "
                    "resize -> nms -> connector (on the net's own output) -> D2H; 2 batches in flight; host wall clock over "
                    "%d steady-state steps" % (FRAME_W, FRAME_H, args.steps))
    else:
        run_steps = lambda n: [step_host() for _ in range(n)]   # noqa: E731
        run_steps(3)
        barrier(world)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        run_steps(args.steps)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier(world)
        ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), wall * 1e3), world, dev)
        h2d = host_in.numel() * host_in.element_size()
        d2h = (h_kp.numel() + h_sc.numel() + h_cnt.numel()) * 4 if post is not None else net.output_bytes()
        e2e_mode = "one call per step through the Python mirror of the operator API, every step synchronous"
    e2e_value = frames_all / (ms_e2e * 1e-3)

    if rank != 0:
        return
    # ---- roofline of the dominant kernel --------------------------------------------------------------
    if have_net:
        conv_ms = stage_avg["net_convs"]
        achieved = CONV_GFLOP_PER_FRAME * B / conv_ms          # GFLOP/ms = TFLOP/s
        roof = {"bound": "tensor", "kernel": "whole net forward of one step: input copy + first conv + 113 conv_igemm_tcgen05 launches",
                "achieved": achieved, "peak": peaks_tbl["tf_sustained"], "unit": "TFLOP/s",
                "frac": achieved / peaks_tbl["tf_sustained"], "frac_of_burst_peak": achieved / peaks_tbl["tf_burst"],
                "traffic": conv_traffic(B), "traffic_source": "committed ncu capture profiles/r02_conv_traffic.json "
                "(not measured in this run)" if conv_traffic(B) is not None else None,
                "peak_source": peaks_tbl["source"] + " sustained bf16 (burst %.0f)" % peaks_tbl["tf_burst"],
                "launch_ms": conv_ms}
    else:
        rms = stage_avg["resize"]
        achieved = RESIZE_BYTES * B / (rms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "resize8Kernel", "achieved": achieved, "peak": peaks_tbl["hbm"],
                "unit": "GB/s", "frac": achieved / peaks_tbl["hbm"], "traffic": args.traffic,
                "peak_source": peaks_tbl["source"] + " copy bandwidth", "launch_ms": rms}
    stages = {k: round(v, 4) for k, v in stage_avg.items()}
    if "nms" in stage_avg:
        stages["nms_GBps"] = round(NMS_BYTES * B / (stage_avg["nms"] * 1e-3) / 1e9, 1)
        stages["resize_GBps"] = round(RESIZE_BYTES * B / (stage_avg["resize"] * 1e-3) / 1e9, 1)

    # ---- CPU baseline (rank 0, N=1 only): bounded sample = 1 frame of the same workload ----------
    cpu = None
    if world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args, host_in, parity[3] if parity is not None else None)
    parity_check = None
    if parity is not None and cpu is not None and "_keypoints" in cpu:
        kp_cpu, sc_cpu = cpu.pop("_keypoints"), cpu.pop("_scores")
        n = parity[2]
        parity_check = {"frame": 0, "people_gpu": n, "people_cpu": int(kp_cpu.shape[0]),
                        "keypoints_bit_equal": bool(n == kp_cpu.shape[0] and np.array_equal(parity[0][:n], kp_cpu)
                                                    and np.array_equal(parity[1][:n], sc_cpu))}
        assert parity_check["people_gpu"] == parity_check["people_cpu"] > 0 and parity_check["keypoints_bit_equal"], parity_check
    ref_gpu = reference_gpu_stages(parity[3], dev) if (parity is not None and world == 1 and not args.no_cpu) else None
    latency = measure_latency(args, dev) if (have_net and post is not None and world == 1 and not args.no_latency) else None

    launches = (launches_net if have_net else 0) + (post.launches_per_step if post is not None else 0)
    line = {
        "metric": "BODY_25 656x368 frames/s", "value": value, "unit": "frames/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 (convs, fp32 accumulate) + f32 (post-processing)" if have_net else "f32",
        "data": "synthetic",
        "config": {"workload": ("BODY_25 net 656x368, scale_number 1, batch of %d synthetic frames per GPU, "
                                "random-init caffemodel; stages=%s; post-processing input: %s" % (
                                    B, args.stages, ("planted %d-person net outputs (78x46x82)" % args.people)
                                    if planted_dev is not None else "the net's own output")) if have_net else
                               ("BODY_25 post-processing only (net not on the path yet): %d planted %d-person "
                                "net outputs (78x46x82) per GPU -> resizeAndMerge x8 -> nms -> connector"
                                % (B, args.people)),
                   "batch_per_gpu": B, "stages": args.stages,
                   "l2": "inputs larger than L2: %d MB of heat-maps per step vs 126 MB L2"
                         % (B * CHANNELS * NET_H * NET_W * 4 // 2 ** 20),
                   "sharding": "frame-sharded, no collective",
                   "timed_region": "%d steps = %.2f s per region: at the default 100 steps the board sits at its power cap "
                                   "for the whole region, so `value` and `e2e` are sustained rates (20-step regions read "
                                   "about 5 %% higher)" % (args.steps, ms_total / 1e3)},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps, "mode": e2e_mode},
        "gpu_launches": launches * args.steps,
        "roofline": roof, "stages_ms": stages, "clocks": clocks,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if parity_check is not None:
        line["parity_check"] = parity_check
    if ref_gpu is not None:
        line["ref_gpu_ms"] = ref_gpu
    if latency is not None:
        line["latency"] = latency
    print(json.dumps(line))


def reference_gpu_stages(net_out: np.ndarray, dev, reps: int = 5):
    """The reference's own CUDA kernels (oracle/_ref/libopref_cuda.so, compiled unmodified from /root/reference) on ONE
    frame of the step's post-processing input: ms per frame and stage, CUDA events on the legacy default stream they launch
    on (connectBodyPartsGpu's host half -- D2H of the score cube, sort, assembly -- by host clock). None when the library
    was not built."""
    import torch
    from oracle.ref_cuda import RefCuda, REFCUDA_SO
    if not os.path.exists(REFCUDA_SO):
        return None
    rc = RefCuda()
    src = torch.from_numpy(np.ascontiguousarray(net_out[None])).to(dev)
    heat = torch.empty((1, CHANNELS, NET_H, NET_W), device=dev)
    peaks = torch.zeros((1, PARTS, MAXP + 1, 3), device=dev)
    kernel = torch.empty(heat.numel() + 16 * (NET_W + 16), dtype=torch.int32, device=dev)
    cube = torch.empty((26, MAXP, MAXP), device=dev)
    t = {"resize": [], "nms": [], "connect": []}
    people = 0
    for r in range(reps + 1):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        torch.cuda.synchronize()
        ev[0].record()
        rc.resize_merge(heat, [src])
        ev[1].record()
        rc.nms(peaks, heat, 0.05, (NMS_OFFSET, NMS_OFFSET), kernel=kernel)
        ev[2].record()
        torch.cuda.synchronize()
        h0 = time.perf_counter()
        kp, _, _ = rc.connect(heat[0], peaks[0], 0, scale=SCALE_NET_TO_OUTPUT, cube_dev=cube, want_cube=False)
        torch.cuda.synchronize()
        h1 = time.perf_counter()
        people = int(kp.shape[0])
        if r:   # first repetition = warm-up
            t["resize"].append(ev[0].elapsed_time(ev[1]))
            t["nms"].append(ev[1].elapsed_time(ev[2]))
            t["connect"].append((h1 - h0) * 1e3)
    return {"per": "frame", "resize": round(float(np.median(t["resize"])), 4), "nms": round(float(np.median(t["nms"])), 4),
            "connect": round(float(np.median(t["connect"])), 4), "people": people,
            "what": "resizeAndMergeGpu, nmsGpu, connectBodyPartsGpu of the reference (CUDA-build semantics) on frame 0"}


def measure_latency(args, dev, steps: int = 100):
    """One frame at a time (batch 1) through net -> resize -> nms -> connector, device-resident input."""
    import torch
    from openpose_b200 import net as netmod
    from openpose_b200 import synth
    net1 = netmod.bench_net(1, NET_H, NET_W, dev)
    post1 = PostPath(1, dev)
    x = torch.from_numpy(synth.frames_to_net_input(synth.frames(1, NET_H, NET_W, 7))).to(dev)
    planted = torch.from_numpy(planted_net_outputs(1, args.people, 77)).to(dev) if args.post_input == "planted" else None
    for _ in range(10):
        out = net1.forward_device(x)
        post1.run(planted if planted is not None else out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = net1.forward_device(x)
        post1.run(planted if planted is not None else out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    net1.net.close()
    out = {"batch": 1, "ms_per_frame": round(ms, 4), "frames_per_s": round(1e3 / ms, 1), "steps": steps,
           "what": "device-resident input, frames issued back to back on one stream"}
    # BASELINE config 1 end to end: ONE u8 1280x720 frame in host memory -> people in host memory through the engine,
    # nothing else in flight (host wall clock per frame: H2D, pre-processing, net, resize, nms, connector, D2H, the
    # worker thread's hand-over)
    from openpose_b200 import modelgen
    from openpose_b200.engine import PoseEngine
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    eng = PoseEngine(proto, cm, gpus=[dev.index or 0], net_size=(NET_W, NET_H), frame_size=(FRAME_W, FRAME_H), batch=1,
                     lanes=1, capacity=256)
    frame = synth.frames(1, FRAME_H, FRAME_W, 9)
    for _ in range(10):
        eng.submit(0, frame)
        eng.next(copy=False)
    t0 = time.perf_counter()
    for i in range(steps):
        eng.submit(i, frame)
        eng.next(copy=False)
    out["e2e_ms_per_frame"] = round((time.perf_counter() - t0) * 1e3 / steps, 4)
    out["e2e_what"] = ("one u8 %dx%d frame at a time through opb200_engine_* (submit -> next): copy from a pageable host "
                       "buffer into the pinned staging lane, H2D, device chain, D2H" % (FRAME_W, FRAME_H))
    # the producer writes the frame straight into the staging lane (acquire / commit): no host copy in the timed region
    eng.acquire()[:] = frame
    eng.commit(0, 1)
    eng.next(copy=False)
    t0 = time.perf_counter()
    for i in range(steps):
        eng.acquire()                     # (the lane still holds the frame)
        eng.commit(i, 1)
        eng.next(copy=False)
    out["e2e_in_place_ms_per_frame"] = round((time.perf_counter() - t0) * 1e3 / steps, 4)
    eng.close()
    return out


# ----------------------------------------------------------------------------------------------------
# CPU baseline + reference arm (host cores; the reference's CPU code where it compiled, see oracle/)
# ----------------------------------------------------------------------------------------------------
def cpu_baseline(args, host_in, post_input=None):
    """One frame of the same workload on the host cores. Post: single thread (the reference has no
    OpenMP in src/openpose/net). Net (when on the path): torch-CPU FP32 restatement of Caffe, all cores."""
    cores = os.cpu_count() or 1
    parts = {}
    total = 0.0
    kind = "port"
    if args.stages in ("full", "net"):
        from oracle import ref_net
        net_out, net_s = ref_net.timed_forward_body25(host_in[:1].numpy(), threads=cores)
        parts["net_ms (torch-CPU restatement of Caffe, %d threads)" % cores] = round(net_s * 1e3, 1)
        total += net_s
    else:
        net_out = host_in[:1].numpy()
    if args.stages in ("full", "post"):
        # the post-processing input of the GPU arm: planted net outputs (default) or the net's own output
        r = cpu_reference_post(np.ascontiguousarray(post_input if post_input is not None else net_out[0]))
        # "reference" only when every timed stage is the reference's own object code; with the net on the
        # path 95 % of the time is the torch-CPU restatement of Caffe, so the line as a whole is a port
        kind = r["kind"] if args.stages == "post" else "port"
        parts["post_impl"] = ("reference object code (oracle/_ref)" if r["kind"] == "reference"
                              else "C restatement (oracle/)")
        parts.update({"resize_ms": round(r["resize_ms"], 2), "nms_ms": round(r["nms_ms"], 2),
                      "connect_ms": round(r["connect_ms"], 2), "people": r["people"],
                      "resize_impl": r["resize_impl"]})
        total += r["total_s"]
        out_kp, out_sc = r["keypoints"], r["scores"]
    else:
        out_kp = out_sc = None
    extra = {"_keypoints": out_kp, "_scores": out_sc} if (out_kp is not None and post_input is not None) else {}
    return {**extra, "value": 1.0 / total, "unit": "frames/s", "cores": cores if args.stages != "post" else 1,
            "kind": kind, "sample": "1 frame of the step's batch (frame 0), all stages once", "stages": parts}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    B = 1
    if args.stages in ("full", "net"):
        from openpose_b200 import synth
        import torch
        host_in = torch.from_numpy(synth.frames_to_net_input(synth.frames(B, NET_H, NET_W, 100)))
    else:
        import torch
        host_in = torch.from_numpy(planted_net_outputs(B, args.people, 1000))
    steps = max(1, min(args.steps, 3))
    warm = 1 if args.warmup > 0 else 0
    last = None
    post_in = planted_net_outputs(1, args.people, 1000)[0] if (args.stages == "full" and args.post_input == "planted") else None
    for _ in range(warm):
        last = cpu_baseline(args, host_in, post_in)
    dt = 0.0   # sum of the timed stage regions (excludes the oracle's own bookkeeping between stages)
    for _ in range(steps):
        last = cpu_baseline(args, host_in, post_in)
        dt += 1.0 / last["value"]
    for k in ("_keypoints", "_scores"):
        last.pop(k, None)
    v = steps * B / dt
    line = {"impl": "reference", "metric": "BODY_25 656x368 frames/s", "value": v, "unit": "frames/s",
            "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": steps, "warmup": warm,
            "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "same as the GPU arm, bounded sample: 1 frame per step on the host cores",
                       "stages": args.stages},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": last["cores"], "kind": last["kind"],
                             "sample": "1 frame per step, %d steps" % steps, "stages": last["stages"]},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)   # 1.2 s per timed region: the sustained (power-capped) rate
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=47,
                    help="frames per step per GPU (47 frames: the 46x82 levels are 1464 tiles = 366 quads = 4.95 waves of the 74 "
                         "CTA pairs; 44 frames, the default until late in round 2, are 4.65 waves)")
    ap.add_argument("--people", type=int, default=8, help="planted people per frame (post-only workload)")
    ap.add_argument("--stages", default=DEFAULT_STAGES, choices=["full", "post", "net"])
    ap.add_argument("--post-input", default="planted", choices=["planted", "net"],
                    help="what the post-processing of a full step runs on: planted --people-person net outputs (default; a "
                         "random-init net emits noise that assembles nobody) or the net's own output")
    ap.add_argument("--no-engine", action="store_true",
                    help="end-to-end region through the Python mirror of the operator API instead of opb200_engine_*")
    ap.add_argument("--no-latency", action="store_true", help="skip the batch-1 latency leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--traffic", type=float, default=None, help="ncu dram bytes per launch, if known")
    args = ap.parse_args()
    try:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_gpu(args)
    finally:
        # leave the process group cleanly: otherwise NCCL's exit warning is the last line a log shows, not the JSON line
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.destroy_process_group()
        except Exception:
            pass


DEFAULT_STAGES = "full"

if __name__ == "__main__":
    main()
