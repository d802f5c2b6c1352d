#!/usr/bin/env python3
"""bench.py -- tensor_transform hot path on B200: 1080p frames/s, GB/s, HBM-roofline fraction.

Workload (BASELINE.json configs[1], the configuration the metric is quoted on):
  one tensor 3:1920:1080:32 uint8 (199 065 600 B, uniform 0..255, seed 2) ->
  tensor_transform mode=arithmetic
     option=typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,
            div:58.395@0,div:57.12@1,div:57.375@2                (ImageNet mean/std)
  -> 3:1920:1080:32 float32 (796 262 400 B).  A "step" = one pass over that 32-frame batch.

  value  : device-resident frames/s (inputs already in HBM), CUDA events, max over ranks.
  e2e    : same metric through the host-buffer path: pinned host input -> H2D -> kernel ->
           D2H into pinned host output, every step, copies inside the timed region.
  roofline: algorithmic bytes per launch (5 B/element) / measured launch time vs the measured
           HBM copy bandwidth of this pool (MEASURED_PEAKS.json).
  cpu_baseline: the reference's own C code (oracle/_ref) or the C port timed on the host cores.

Multi-GPU (--gpus N under torchrun): streams are independent, so each rank runs its own
32-frame batch on its own GPU ("weak" scaling); NCCL is used only for the barrier and the
max-over-ranks of the timings -- the data path has no collective.

`--impl reference` times the reference CPU implementation of the same step on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W, H, C, B = 1920, 1080, 3, 32
DIM = [C, W, H, 1]
FRAME_ELEMS = C * W * H
OPTION = ("typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
          "div:58.395@0,div:57.12@1,div:57.375@2")
ORACLE_OPS = [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1),
              ("add", -103.53, 2), ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)]
WORKLOAD = "3:1920:1080:32 uint8 -> float32, arithmetic " + OPTION
ALGO_BYTES_PER_ELEM = 5           # sizeof(u8) + sizeof(f32), SURVEY.md 8d
NCU_TRAFFIC_BYTES_PER_LAUNCH = 938.4e6   # measured once with ncu, see profiles/
METRIC = "tensor_transform_1080p_frames_per_sec"
UNIT = "frames/s"


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------
# CPU side (oracle): cpu_baseline leg and --impl reference
# ---------------------------------------------------------------------------------------
def cpu_checker():
    from oracle import oracle as O
    if O.ref_available():
        return O.ref(), "reference"
    return O.port(), "port"


def cpu_checker_port():
    from oracle import oracle as O
    return O.port()


def cpu_run_frames(checker, frames, n_threads):
    """Run the reference CPU arithmetic on `frames` (list of u8 arrays), one frame per task,
    n_threads worker threads (ctypes releases the GIL).  Returns wall seconds."""
    idx = [0]
    lock = threading.Lock()

    def worker():
        while True:
            with lock:
                i = idx[0]
                idx[0] += 1
            if i >= len(frames):
                return
            checker.arithmetic(frames[i], DIM, ORACLE_OPS, True, 0)

    ths = [threading.Thread(target=worker) for _ in range(n_threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    return time.perf_counter() - t0


def cpu_baseline(sample_frames=None):
    checker, kind = cpu_checker()
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(2)
    frame = rng.integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
    t1 = cpu_run_frames(checker, [frame], 1)            # warm-up + per-frame cost
    if sample_frames is None:
        # ~15 s of CPU work in total, at least one frame per core
        sample_frames = int(max(cores, min(4 * cores, 15.0 / max(t1, 1e-3))))
    frames = [frame] * sample_frames
    dt = cpu_run_frames(checker, frames, cores)
    return {"value": sample_frames / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%d 1080p frames of the bench workload, one frame per task on %d threads "
                      "(%.3f s single-thread per frame)" % (sample_frames, cores, t1)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    checker, kind = cpu_checker()
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(2)
    frame = rng.integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
    frames = [frame] * B
    for _ in range(args.warmup):
        cpu_run_frames(checker, frames[:cores], cores)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_run_frames(checker, frames, cores)
    value = args.steps * B / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_frames": B},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "each step = the full 32-frame batch, one frame per task on "
                                   "%d host threads" % cores},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------
def time_launches(dev, fn, steps, warmup):
    for _ in range(warmup):
        fn()
    dev.sync()
    e0, e1 = dev.event(), dev.event()
    dev.record(e0)
    for _ in range(steps):
        fn()
    dev.record(e1)
    ms = dev.elapsed_ms(e0, e1)
    dev.event_destroy(e0)
    dev.event_destroy(e1)
    return ms


def extra_kernels(dev, peak, with_cpu=True):
    """Device-resident GB/s of the other modes at their BASELINE configs, each with the
    reference CPU function timed beside it on ONE frame, one host thread (informational)."""
    out = {}
    rng = np.random.default_rng(3)
    checker = cpu_checker()[0] if with_cpu else None

    def run(name, mode, option, itype, dim, batch, bytes_per_elem, make, cpu=None):
        plan = dev.plan(mode, option, itype, dim)
        frame_n = int(np.prod(dim))
        n = frame_n * batch
        host = make(n)
        d_in = dev.upload(host)
        d_out = dev.alloc(plan.out_size * batch)
        ms = time_launches(dev, lambda: plan.launch(d_in, d_out, batch), 10, 3) / 10
        gbs = n * bytes_per_elem / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "GB/s": gbs, "frac_of_measured_peak": gbs / peak,
                     "launches": plan.last_launch_count}
        dev.free(d_in)
        dev.free(d_out)
        plan.close()
        if cpu is not None and checker is not None:
            # bound the CPU sample to one 1080p-sized frame (the reference takes 0.06-0.4 s for it)
            sample = min(frame_n, FRAME_ELEMS)
            sdim = dim if sample == frame_n else [C, W, H, 1]
            xs = host[:sample]
            t0 = time.perf_counter()
            cpu(checker, xs, sdim)
            dt = time.perf_counter() - t0
            out[name]["cpu_reference_1_thread_GB/s"] = sample * bytes_per_elem / dt / 1e9
            out[name]["cpu_sample_elements"] = sample

    f32 = lambda n: rng.standard_normal(n, dtype=np.float32)
    u8 = lambda n: rng.integers(0, 256, n, dtype=np.uint8)
    scalar_ops = [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]
    run("typecast_u8_f32_1080p_b32", "typecast", "float32", "uint8", DIM, B, 5, u8,
        lambda c, x, d: c.typecast(x, "float32"))
    run("arith_scalar_u8_f32_1080p_b32", "arithmetic", "typecast:float32,add:-127.5,div:127.5",
        "uint8", DIM, B, 5, u8, lambda c, x, d: c.arithmetic(x, d, scalar_ops))
    # configs[0]: MobileNet pre-processing, 224x224 RGB, 256 frames per launch
    run("arith_mobilenet_224_b256", "arithmetic", "typecast:float32,add:-127.5,div:127.5",
        "uint8", [3, 224, 224, 1], 256, 5, u8, lambda c, x, d: c.arithmetic(x, d, scalar_ops))
    run("transpose_hwc_chw_4k_f32", "transpose", "1:2:0:3", "float32", [3, 3840, 2160, 1], 1, 8, f32,
        lambda c, x, d: c.transpose(x, d, (1, 2, 0, 3)))
    run("transpose_hwc_chw_4k_f32_b8", "transpose", "1:2:0:3", "float32", [3, 3840, 2160, 1], 8, 8, f32)
    run("transpose_2_0_1_4k_f32", "transpose", "2:0:1:3", "float32", [3, 3840, 2160, 1], 1, 8, f32,
        lambda c, x, d: c.transpose(x, d, (2, 0, 1, 3)))
    run("transpose_1_0_2_4k_f32", "transpose", "1:0:2:3", "float32", [3, 3840, 2160, 1], 1, 8, f32)
    run("transpose_0_2_1_4k_f32", "transpose", "0:2:1:3", "float32", [3, 3840, 2160, 1], 1, 8, f32)
    run("transpose_2_1_0_4k_f32", "transpose", "2:1:0:3", "float32", [3, 3840, 2160, 1], 1, 8, f32)
    for o in ("2:0:1:3", "1:0:2:3", "0:2:1:3", "2:1:0:3"):
        run("transpose_%s_4k_f32_b8" % o[:5].replace(":", "_"), "transpose", o, "float32",
            [3, 3840, 2160, 1], 8, 8, f32)
    run("dimchg_0_2_4k_f32", "dimchg", "0:2", "float32", [3, 3840, 2160, 1], 1, 8, f32,
        lambda c, x, d: c.dimchg(x, d, 0, 2))
    run("clamp_f32_1080p_b64", "clamp", "-33.3:77.7", "float32", DIM, 64, 8, f32,
        lambda c, x, d: c.clamp(x, -33.3, 77.7))
    run("stand_default_f32_1080p_b64_per_frame", "stand", "default", "float32", DIM, 64, 8, f32,
        lambda c, x, d: c.stand(x, d, "default"))
    run("stand_dc_perch_f32_1080p_b64_per_frame", "stand", "dc-average,per-channel:true",
        "float32", DIM, 64, 8, f32, lambda c, x, d: c.stand(x, d, "dc-average", None, True))
    run("stand_default_f32_1080p_x64_one_tensor", "stand", "default", "float32",
        [3, 1920, 1080, 64], 1, 8, f32)
    # converter stride strip (K8): 1918x1080 RGB frames, rows padded to a multiple of 4 bytes
    w = 1918
    row, pitch, rows = 3 * w, (3 * w + 3) // 4 * 4, 1080 * 32
    src = u8(pitch * rows)
    d_src = dev.upload(src)
    d_dst = dev.alloc(row * rows)
    ms = time_launches(dev, lambda: dev.strip(d_src, d_dst, row, pitch, rows), 10, 3) / 10
    gbs = 2 * row * rows / (ms * 1e-3) / 1e9
    out["stride_strip_1918x1080_rgb_b32"] = {"ms": ms, "GB/s": gbs, "frac_of_measured_peak": gbs / peak,
                                             "launches": 1}
    if checker is not None:
        t0 = time.perf_counter()
        cpu_checker_port().stride_strip(src[:pitch * 1080], 3, w, 1080)
        dt = time.perf_counter() - t0
        out["stride_strip_1918x1080_rgb_b32"]["cpu_reference_1_thread_GB/s"] = 2 * row * 1080 / dt / 1e9
    # tensor_decoder direct_video (K9): the inverse, dense rows -> rows padded to 4 bytes
    d_pad = dev.alloc(pitch * rows)
    ms = time_launches(dev, lambda: dev.pad(d_dst, d_pad, row, pitch, rows), 10, 3) / 10
    gbs = 2 * row * rows / (ms * 1e-3) / 1e9
    out["direct_video_pad_1918x1080_rgb_b32"] = {"ms": ms, "GB/s": gbs,
                                                 "frac_of_measured_peak": gbs / peak, "launches": 1}
    if checker is not None:
        dense = src[:row * 1080]
        t0 = time.perf_counter()
        cpu_checker_port().direct_video(dense, 3, w, 1080)
        dt = time.perf_counter() - t0
        out["direct_video_pad_1918x1080_rgb_b32"]["cpu_reference_1_thread_GB/s"] = 2 * row * 1080 / dt / 1e9
    dev.free(d_pad)
    dev.free(d_src)
    dev.free(d_dst)
    out.update(aggregator_extras(dev, peak, checker is not None))
    out.update(merge_extras(dev, peak, checker is not None))
    return out


def merge_extras(dev, peak, with_cpu):
    """tensor_merge mode=linear option=0 (K10): a 1080p RGB frame and an alpha plane -> RGBA, 32
    frames per launch through nnsb_merge_launch.  Algorithmic bytes: 2 x output bytes."""
    import ctypes as C
    from nnstreamer_b200 import _lib as L
    rng = np.random.default_rng(12)
    frames = 32
    px = W * H * frames
    rgb = rng.integers(0, 256, 3 * px, dtype=np.uint8)
    alpha = rng.integers(0, 256, px, dtype=np.uint8)
    d_rgb, d_a = dev.upload(rgb), dev.upload(alpha)
    d_out = dev.alloc(4 * px)
    ptrs = (C.c_void_p * 2)(d_rgb, d_a)
    blk = (C.c_uint64 * 2)(3, 1)
    lib = L.lib()

    def launch():
        L.check(lib.nnsb_merge_launch(dev.stream, 2, ptrs, blk, C.c_void_p(d_out), px))

    ms = time_launches(dev, launch, 10, 3) / 10
    gbs = 2 * 4 * px / (ms * 1e-3) / 1e9
    out = {"merge_rgb_alpha_1080p_b32": {"ms": ms, "GB/s": gbs, "frac_of_measured_peak": gbs / peak,
                                          "launches": 1}}
    if with_cpu:
        n1 = W * H
        t0 = time.perf_counter()
        cpu_checker_port().merge_linear([rgb[:3 * n1], alpha[:n1]], [[3, W, H, 1], [1, W, H, 1]], 0)
        dt = time.perf_counter() - t0
        out["merge_rgb_alpha_1080p_b32"]["cpu_reference_1_thread_GB/s"] = 2 * 4 * n1 / dt / 1e9
    for p in (d_rgb, d_a, d_out):
        dev.free(p)
    # feature maps: 64 + 32 float32 channels per position (256- and 128-byte blocks)
    pos = 1 << 21
    fa = rng.standard_normal(64 * pos, dtype=np.float32)
    fb = rng.standard_normal(32 * pos, dtype=np.float32)
    d_fa, d_fb = dev.upload(fa), dev.upload(fb)
    d_out = dev.alloc(96 * 4 * pos)
    ptrs2 = (C.c_void_p * 2)(d_fa, d_fb)
    blk2 = (C.c_uint64 * 2)(256, 128)

    def launch2():
        L.check(lib.nnsb_merge_launch(dev.stream, 2, ptrs2, blk2, C.c_void_p(d_out), pos))

    ms = time_launches(dev, launch2, 10, 3) / 10
    gbs = 2 * 96 * 4 * pos / (ms * 1e-3) / 1e9
    out["merge_f32_64ch_32ch_2M_positions"] = {"ms": ms, "GB/s": gbs, "frac_of_measured_peak": gbs / peak,
                                               "launches": 1}
    for p in (d_fa, d_fb, d_out):
        dev.free(p)
    return out


def aggregator_extras(dev, peak, with_cpu):
    """tensor_aggregator: 1080p RGB uint8 frames already in HBM, frames-out=32.
    frames-dim=3 batches by appending (one D2D copy per frame + one per output); frames-dim=1
    concatenates along the width (adapter copy + one permutation launch per output).  Algorithmic
    bytes: 2 x frame bytes per frame (read once, written once)."""
    from nnstreamer_b200 import TensorAggregator, device_memory
    out = {}
    rng = np.random.default_rng(11)
    frame = rng.integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
    d_frame = dev.upload(frame)
    mem = device_memory(d_frame, frame.nbytes)
    caps = ("other/tensor, dimension=(string)3:1920:1080:1, type=(string)uint8, "
            "framerate=(fraction)30/1")
    for name, fdim in (("aggregator_1080p_u8_out32_append", 3), ("aggregator_1080p_u8_out32_concat_w", 1)):
        agg = TensorAggregator(frames_out=32, frames_dim=fdim)
        agg.set_caps(caps)
        n_frames = 32 * 6
        for i in range(32):                         # warm-up: one output
            _, mems, _ = agg.chain(mem)
            for m in mems:
                m.release()
        t0 = time.perf_counter()
        produced = 0
        for i in range(n_frames):
            _, mems, _ = agg.chain(mem)
            for m in mems:
                produced += 1
                m.release()
        dt = time.perf_counter() - t0
        out[name] = {"frames/s": n_frames / dt, "GB/s": 2 * frame.nbytes * n_frames / dt / 1e9,
                     "outputs": produced, "kernel_launches": agg.kernel_launches,
                     "timing": "wall clock around synchronous chain calls, device-resident frames"}
        agg.close()
    if with_cpu:
        port = cpu_checker_port()
        frames4 = np.tile(frame, 4)
        t0 = time.perf_counter()
        port.aggregator_concat(frames4, frame.nbytes, 4, 3 * 1920)
        dt = time.perf_counter() - t0
        out["aggregator_1080p_u8_out32_concat_w"]["cpu_reference_1_thread_GB/s"] = \
            2 * frames4.nbytes / dt / 1e9
    dev.free(d_frame)
    return out


def multi_stream_chain(device_id, n_streams=8, frames_per_stream=40):
    """BASELINE configs[4] on ONE GPU: `n_streams` concurrent 1080p pipelines, each
    tensor_converter -> tensor_transform typecast:float32 -> arithmetic add:-127.5,div:127.5 ->
    transpose 1:2:0:3, one host thread and one set of elements (own CUDA streams) per pipeline.
    Returns frames/s device-resident (frames pre-staged in HBM) and with the H2D of every
    frame inside the timed region (the float CHW result stays on the device for the consumer)."""
    from nnstreamer_b200 import Device, TensorConverter, TensorTransform, device_memory
    caps = ("video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, framerate=(fraction)30/1" % (W, H))
    out = {}
    for mode in ("device_resident", "h2d_per_frame"):
        dev = Device(device_id)
        pipes = []
        for sidx in range(n_streams):
            conv = TensorConverter(device_id=device_id)
            c1 = conv.set_caps(caps)
            t1 = TensorTransform(device_id=device_id, mode="typecast", option="float32")
            c2 = t1.negotiate(c1)
            t2 = TensorTransform(device_id=device_id, mode="arithmetic", option="add:-127.5,div:127.5")
            c3 = t2.negotiate(c2)
            t3 = TensorTransform(device_id=device_id, mode="transpose", option="1:2:0:3")
            t3.negotiate(c3)
            rng = np.random.default_rng(sidx)
            if mode == "device_resident":
                frame = device_memory(dev.upload(rng.integers(0, 256, FRAME_ELEMS, dtype=np.uint8)), FRAME_ELEMS)
            else:
                arr, _ptr = dev.pinned(FRAME_ELEMS, np.uint8)
                arr[:] = rng.integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
                frame = arr
            pipes.append((conv, t1, t2, t3, frame))

        def run(pipe, n):
            conv, t1, t2, t3, frame = pipe
            for _ in range(n):
                flow, m0 = conv.chain(frame)
                flow, (m1,) = t1.transform([m0])
                flow, (m2,) = t2.transform([m1])
                flow, (m3,) = t3.transform([m2])
                for m in (m0, m1, m2, m3):
                    m.release()

        for p in pipes:
            run(p, 3)
        ths = [threading.Thread(target=run, args=(p, frames_per_stream)) for p in pipes]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        fps = n_streams * frames_per_stream / dt
        # algorithmic bytes per element along the chain: converter 0 (+1 when uploaded),
        # typecast 5, arithmetic 8, transpose 8
        out[mode] = {"frames/s": fps, "GB/s": fps * FRAME_ELEMS * 21 / 1e9,
                     "streams": n_streams, "frames_per_stream": frames_per_stream}
        for conv, t1, t2, t3, frame in pipes:
            for e in (conv, t1, t2, t3):
                e.close()
        dev.close()
    return out


def native_stream_chain(device_id, n_streams=8, frames_per_stream=300):
    """The same chain driven by tools/stream_bench (native threads through the C-ABI): what the
    element mirrors cost without the Python interpreter in the per-frame loop."""
    exe = os.path.join(ROOT, "tools", "stream_bench")
    if not os.path.isfile(exe):
        return {"unavailable": "tools/stream_bench not built (python -c 'import __graft_entry__ as g; g.build()')"}
    out = {}
    for mode, key in (("resident", "device_resident"), ("h2d", "h2d_per_frame")):
        try:
            r = subprocess.run([exe, str(device_id), str(n_streams), str(frames_per_stream), mode],
                               capture_output=True, text=True, timeout=300)
            out[key] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else \
                {"error": r.stderr.strip()[-300:]}
        except Exception as exc:                    # noqa: BLE001
            out[key] = {"error": str(exc)}
    return out


def bind_to_gpu_numa_node(gpu_index):
    """Pin this process to the CPUs NVML reports as local to its GPU before any pinned host memory
    is allocated, so the staging buffers of the end-to-end path sit on the GPU's NUMA node (the
    host-side limiter of multi-GPU scaling).  Best effort: returns what was done."""
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(int(gpu_index))
        pynvml.nvmlDeviceSetCpuAffinity(handle)
        return "bound to %d cpus" % len(os.sched_getaffinity(0))
    except Exception as exc:                        # noqa: BLE001
        return "not bound (%s)" % (str(exc)[:60],)


def run_b200(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    numa = bind_to_gpu_numa_node(local_rank)

    from nnstreamer_b200 import Device
    dev = Device(local_rank)
    peak, peak_src = measured_peak()

    rng = np.random.default_rng(2 + rank)
    n = FRAME_ELEMS * B
    h_in, h_in_ptr = dev.pinned(n, np.uint8)
    h_in[:] = rng.integers(0, 256, n, dtype=np.uint8)
    d_in = dev.alloc(n)
    d_out = dev.alloc(4 * n)
    dev.h2d(d_in, h_in)
    dev.sync()
    plan = dev.plan("arithmetic", OPTION, "uint8", DIM)

    from nnstreamer_b200 import sharding

    def barrier():
        dev.sync()
        sharding.barrier(dist)

    def max_over_ranks(x):
        return sharding.max_over_ranks(x, dist, "cuda" if dist is not None else "cpu")

    # ---- device-resident: K launches, CUDA events on the launching stream ----
    step = lambda: plan.launch(d_in, d_out, B)
    # clocks / throttle reasons are sampled from before the warm-up until after the end-to-end
    # region, so that the (milliseconds-long) device-resident region is covered too
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    for _ in range(args.warmup):
        step()
    barrier()
    e0, e1 = dev.event(), dev.event()
    dev.record(e0)
    for _ in range(args.steps):
        step()
    dev.record(e1)
    ms_dev = dev.elapsed_ms(e0, e1)
    barrier()
    ms_dev = max_over_ranks(ms_dev)
    launches = plan.last_launch_count * args.steps

    # ---- end to end: the element's transform() with HOST memories, every step ----
    # tensor_transform mirror (nnstreamer_b200.TensorTransform): same properties / caps as the
    # reference element; the input lives in pinned host memory, transform() uploads it,
    # launches, and downloads into a pinned output memory it allocates from its pool.  The
    # copies are inside the timed region (wall clock around synchronous transform() calls,
    # cross-checked with CUDA events on the element's streams being idle at both ends).
    import ctypes as C

    from nnstreamer_b200 import TensorTransform
    from nnstreamer_b200 import _lib as L
    elem = TensorTransform(device_id=local_rank, mode="arithmetic", option=OPTION)
    elem.negotiate("other/tensors, format=(string)static, num_tensors=(int)1, "
                   "dimensions=(string)3:1920:1080:%d, types=(string)uint8, framerate=(fraction)30/1" % B)
    e_in = elem.alloc(n, on_device=False)               # pinned host GstMemory stand-in
    C.memmove(e_in.data, h_in.ctypes.data, n)
    ins = (L.Memory * 1)(e_in)
    outs = (L.Memory * L.SIZE_LIMIT)()
    n_out = C.c_uint32(0)
    e2e_first = None

    def e2e_step(keep=False):
        flow = elem.transform_raw(ins, 1, outs, n_out)
        if flow != 0 or n_out.value != 1:
            raise RuntimeError("transform failed: %s" % L.lib().nnsb_last_error())
        if keep:
            buf = (C.c_uint8 * (4 * FRAME_ELEMS)).from_address(outs[0].data)
            res = np.frombuffer(buf, dtype=np.float32).copy()
        else:
            res = None
        L.lib().nnsb_memory_release(C.byref(outs[0]))
        return res

    for _ in range(max(1, min(args.warmup, 3))):
        e2e_first = e2e_step(keep=True)
    barrier()
    launches_before = elem.kernel_launches
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    ms_e2e = (time.perf_counter() - t0) * 1e3
    e2e_launches = elem.kernel_launches - launches_before
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e = max_over_ranks(ms_e2e)

    # ---- CPU baseline (with a parity spot check of the bench data) and the other kernels ----
    parity = None
    cpu = None
    extras = None
    if rank == 0:
        if world == 1:
            if not args.no_cpu_baseline:
                # the cpu_baseline leg: the CPU code is timed, and the frame it computes anyway is
                # compared with the first frame of the GPU's end-to-end result
                cpu = cpu_baseline()
                want = cpu_checker_port().arithmetic(h_in[:FRAME_ELEMS], DIM, ORACLE_OPS, True, 0)
                parity = bool(e2e_first.tobytes() == want.tobytes())
            if not args.no_extras:
                dev.free(d_in)
                dev.free(d_out)
                d_in = d_out = None
                extras = extra_kernels(dev, peak, with_cpu=not args.no_cpu_baseline)
                extras["chain_8_streams_1080p_converter_typecast_arith_transpose"] = \
                    multi_stream_chain(local_rank)
                extras["chain_8_streams_1080p_native_driver"] = native_stream_chain(local_rank)

    if rank == 0:
        ms_step = ms_dev / args.steps
        algo_bytes = n * ALGO_BYTES_PER_ELEM
        achieved = algo_bytes / (ms_step * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": world * B * args.steps / (ms_dev * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_frames": B,
                       "l2": "inputs larger than L2 (199 MB in + 796 MB out per step vs 126 MB L2)",
                       "sharding": "one independent 32-frame stream per GPU, no collective"},
            "GB/s": world * achieved,
            "e2e": {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": n, "d2h_bytes_per_step": 4 * n,
                    "ms_per_step": ms_e2e / args.steps,
                    "api": "nnsb_transform_transform (element mirror) on pinned host memories, "
                           "chunks pipelined on two streams; wall clock around synchronous calls"},
            "gpu_launches": launches + e2e_launches,
            "gpu_launches_detail": {"device_resident": launches, "e2e": e2e_launches},
            "host_affinity": numa,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_nominal_8TBs": achieved / 8000.0,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full
                         # (profiles/r1_lut_chain_u8_f32_summary.md): 199.1 MB + 739.3 MB
                         "traffic": NCU_TRAFFIC_BYTES_PER_LAUNCH, "peak_source": peak_src,
                         "kernel": "lut_chain_kernel<uint8,float,false,4> (k_chain.cuh)",
                         "algorithmic_bytes_per_launch": algo_bytes},
            "cpu_baseline": cpu,
            "clocks": clocks,
            "parity_vs_oracle_first_frame": parity,
        }
        if extras is not None:
            line["kernels"] = extras
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    dev.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the informational per-kernel table")
    ap.add_argument("--no-cpu-baseline", action="store_true",
                    help="skip the host-core timing of the reference code (profiling runs)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
