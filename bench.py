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
           D2H into pinned host output, every step, copies inside the timed region; with
           `host_link` = bare cudaMemcpyAsync of the same bytes by all ranks at once (the ceiling
           the host side sets) and the per-rank times.
  e2e_consumer_on_device : the same batch when the consumer is CUDA-aware (converter uploads the
           32 host frames, transform, tensor_filter hand-off on the device): H2D only.
  roofline: algorithmic bytes per launch (5 B/element) / measured launch time vs the measured
           HBM copy bandwidth of this pool (MEASURED_PEAKS.json); traffic from the committed ncu CSV.
  cpu_baseline: the reference's own C code (oracle/_ref) or the C port timed on the host cores.
  kernels: configs[4] (`chain_64_streams_*`: 8 streams per GPU through converter -> typecast ->
           arithmetic -> transpose, resident / H2D / H2D+D2H, 1920 and 1918 wide) on EVERY N;
           at N=1 also the other modes at their BASELINE configs, the narrow-row permutes (with
           `yardstick_*` = an element-wise kernel on the same small batches), the aggregator
           (Python calls and `*_native_thread` = tools/agg_bench), the 8-stream chain from
           Python threads (one call per element, and `*_linked` = one nnsb_branch_push per frame).

Multi-GPU (--gpus N under torchrun): streams are independent, so each rank runs its own
32-frame batch (and its own 8 streams) on its own GPU ("weak" scaling); NCCL is used only for the
barrier and the reductions of the timings -- the data path has no collective.

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
C_ = C                              # (`C` is rebound to ctypes inside some functions)
DIM = [C, W, H, 1]
FRAME_ELEMS = C * W * H
OPTION = ("typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
          "div:58.395@0,div:57.12@1,div:57.375@2")
ORACLE_OPS = [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1),
              ("add", -103.53, 2), ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)]
WORKLOAD = "3:1920:1080:32 uint8 -> float32, arithmetic " + OPTION
ALGO_BYTES_PER_ELEM = 5           # sizeof(u8) + sizeof(f32), SURVEY.md 8d
# ncu --set full capture of the headline kernel (one launch of this very workload), committed
NCU_RAW_CSV = os.path.join(ROOT, "profiles", "r2_lut_chain_u8_f32_raw.csv")
METRIC = "tensor_transform_1080p_frames_per_sec"
UNIT = "frames/s"
# the same dict in both arms (`same_config`): what the step is and how L2 is kept out of it
CONFIG = {"workload": WORKLOAD, "batch_frames": B,
          "l2": "inputs larger than L2 (199 MB in + 796 MB out per step vs 126 MB L2)",
          "sharding": "one independent 32-frame stream per GPU, no collective"}


def ncu_traffic_bytes(path=NCU_RAW_CSV, launch="0"):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch, read from the committed ncu
    CSV (profiles/summarize.py format: launch,metric,unit,value).  None if the file is absent."""
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    try:
        with open(path) as fh:
            for line in fh:
                parts = line.rstrip("\n").split(",")
                if len(parts) == 4 and parts[0] == launch and parts[1] in (
                        "dram__bytes_read.sum", "dram__bytes_write.sum"):
                    total += float(parts[3]) * scale[parts[2]]
                    seen += 1
    except Exception:
        return None
    return total if seen == 2 else None


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


def cpu_baseline(sample_frames=None, budget_s=15.0):
    checker, kind = cpu_checker()
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(2)
    frame = rng.integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
    t1 = cpu_run_frames(checker, [frame], 1)            # warm-up + per-frame cost
    if sample_frames is None:
        # ~budget_s of CPU work in total, at least one frame per core
        sample_frames = int(max(cores, min(4 * cores, budget_s / max(t1, 1e-3))))
    distinct = min(sample_frames, B)
    pool = rng.integers(0, 256, FRAME_ELEMS * distinct, dtype=np.uint8)
    frames = [pool[(i % distinct) * FRAME_ELEMS:(i % distinct + 1) * FRAME_ELEMS]
              for i in range(sample_frames)]
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
    # 32 different frames (199 MB): the batch does not sit in the last-level cache
    batch = rng.integers(0, 256, FRAME_ELEMS * B, dtype=np.uint8)
    frames = [batch[i * FRAME_ELEMS:(i + 1) * FRAME_ELEMS] for i in range(B)]
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
        "config": CONFIG,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "each step = the full batch of 32 distinct frames, one frame "
                                   "per task on %d host threads" % cores},
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
        # launches rotate over enough buffer pairs that their footprint is >= 3 x L2 (126 MB): a
        # single 4K frame (99.5 MB in, 99.5 MB out) would otherwise be served partly from L2
        pair = plan.in_size * batch + plan.out_size * batch
        rot = int(max(1, min(8, -(-3 * 126_000_000 // pair))))
        d_ins = [dev.upload(host) for _ in range(rot)]
        d_outs = [dev.alloc(plan.out_size * batch) for _ in range(rot)]
        it = [0]

        def launch():
            k = it[0] % rot
            it[0] += 1
            plan.launch(d_ins[k], d_outs[k], batch)

        steps = 20
        ms = time_launches(dev, launch, steps, 3) / steps
        gbs = n * bytes_per_elem / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "GB/s": gbs, "frac_of_measured_peak": gbs / peak,
                     "launches": plan.last_launch_count, "rotating_buffers": rot}
        for p in d_ins + d_outs:
            dev.free(p)
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
    # real video rows whose byte counts are not multiples of 16 (narrow-vector / odd-row paths)
    u16 = lambda n: rng.integers(0, 65536, n, dtype=np.uint16)
    run("transpose_0_2_1_1080p_rgb_u8_b32", "transpose", "0:2:1:3", "uint8", DIM, 32, 2, u8,
        lambda c, x, d: c.transpose(x, d, (0, 2, 1, 3)))
    run("transpose_2_1_0_227x227x3_f32_b64", "transpose", "2:1:0:3", "float32", [3, 227, 227, 1], 64, 8, f32)
    run("transpose_1_2_0_225x113x3_u16_b256", "transpose", "1:2:0:3", "uint16", [3, 225, 113, 1], 256, 4, u16)
    run("transpose_0_2_1_225x113x3_u16_b256", "transpose", "0:2:1:3", "uint16", [3, 225, 113, 1], 256, 4, u16)
    # the same shapes in launches long enough to leave the launch / ramp-up cost behind (the
    # batches above move 79 MB: 17 us at 4.6 TB/s), and an element-wise kernel on those small
    # batches as a yardstick for what a launch of that size can reach at all
    run("transpose_2_1_0_227x227x3_f32_b512", "transpose", "2:1:0:3", "float32", [3, 227, 227, 1], 512, 8, f32)
    run("transpose_1_2_0_225x113x3_u16_b2048", "transpose", "1:2:0:3", "uint16", [3, 225, 113, 1], 2048, 4, u16)
    run("transpose_0_2_1_225x113x3_u16_b2048", "transpose", "0:2:1:3", "uint16", [3, 225, 113, 1], 2048, 4, u16)
    run("yardstick_add0_227x227x3_f32_b64", "arithmetic", "add:0", "float32", [3, 227, 227, 1], 64, 8, f32)
    run("yardstick_add0_225x113x3_u16_b256", "arithmetic", "add:0", "uint16", [3, 225, 113, 1], 256, 4, u16)
    run("transpose_1_0_2_1080p_gray_u8_b32", "transpose", "1:0:2:3", "uint8", [1, 1920, 1080, 1], 32, 2, u8)
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
    """tensor_aggregator on 1080p RGB uint8 frames, frames-out=32.  The producer
    writes each frame into the memory the aggregator hands out (nnsb_aggregator_acquire: the tail
    of its ring), so pushing moves nothing.  frames-dim=3 (append): an output is a share of the
    ring -- zero bytes moved, the figure is the rate of the host calls.  frames-dim=1 (concatenate
    along the width): one permutation launch per output; algorithmic bytes = 2 x frame bytes per
    frame.  Wall clock around the asynchronous chain calls, one wait at the end."""
    from nnstreamer_b200 import TensorAggregator
    out = {}
    nbytes = FRAME_ELEMS
    caps = ("other/tensor, dimension=(string)3:1920:1080:1, type=(string)uint8, "
            "framerate=(fraction)30/1")
    for name, fdim in (("aggregator_1080p_u8_out32_append", 3), ("aggregator_1080p_u8_out32_concat_w", 1)):
        agg = TensorAggregator(frames_out=32, frames_dim=fdim)
        agg.set_caps(caps)
        n_frames = 32 * 12

        def feed(count):
            last, produced = None, 0
            for _ in range(count):
                slot = agg.acquire(nbytes)              # where the producer would write the frame
                _, mems, _ = agg.chain(slot)
                slot.release()
                for m in mems:
                    produced += 1
                    if last is not None:
                        last.release()
                    last = m
            if last is not None:
                last.sync()
                last.release()
            return produced

        feed(64)                                        # warm-up: two outputs
        l0 = agg.kernel_launches
        t0 = time.perf_counter()
        produced = feed(n_frames)
        dt = time.perf_counter() - t0
        cin, cout = agg.bytes_copied
        out[name] = {"frames/s": n_frames / dt, "outputs": produced,
                     "kernel_launches": agg.kernel_launches - l0,
                     "bytes_copied_into_the_adapter": cin, "bytes_moved_by_concatenation": cout,
                     "timing": "wall clock around asynchronous chain calls, frames written in place "
                               "(nnsb_aggregator_acquire), one wait at the end"}
        if fdim == 1:
            out[name]["GB/s"] = 2 * nbytes * n_frames / dt / 1e9
            out[name]["frac_of_measured_peak"] = out[name]["GB/s"] / peak
        agg.close()
    # the same calls from one native streaming thread (tools/agg_bench.cpp)
    exe = os.path.join(ROOT, "tools", "agg_bench")
    for name, fdim in (("aggregator_1080p_u8_out32_append", 3), ("aggregator_1080p_u8_out32_concat_w", 1)):
        try:
            r = subprocess.run([exe, str(dev.device_id), "32", str(fdim), "12"], capture_output=True,
                               text=True, timeout=300)
            res = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else \
                {"error": r.stderr.strip()[-200:]}
        except Exception as exc:                    # noqa: BLE001
            res = {"error": str(exc)[:200]}
        if fdim == 1 and "GB/s" in res:
            res["frac_of_measured_peak"] = res["GB/s"] / peak
        out[name + "_native_thread"] = res
    if with_cpu:
        port = cpu_checker_port()
        frame = np.random.default_rng(11).integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
        frames4 = np.tile(frame, 4)
        t0 = time.perf_counter()
        port.aggregator_concat(frames4, frame.nbytes, 4, 3 * 1920)
        dt = time.perf_counter() - t0
        out["aggregator_1080p_u8_out32_concat_w"]["cpu_reference_1_thread_GB/s"] = \
            2 * frames4.nbytes / dt / 1e9
    return out


def python_stream_chain(device_id, n_streams=8, frames_per_stream=300, linked=False):
    """BASELINE configs[4] on ONE GPU driven from Python threads: the same C-ABI calls as
    tools/stream_bench (nnsb_converter_chain, nnsb_transform_transform x3, nnsb_memory_release),
    made through ctypes with pre-built argument arrays -- what the stream-ordered memories buy a
    caller that is NOT native code.  Device-resident frames, one thread and one set of elements
    per pipeline; each thread waits once, for its last output.  linked=True: the four elements
    are linked into an nnsb_branch and a frame is ONE call (GStreamer's push model), so the
    interpreter lock is taken once per frame instead of eight times."""
    from nnstreamer_b200 import Branch, DeviceMemory, TensorConverter, TensorTransform
    from nnstreamer_b200 import _lib as L
    import ctypes as C
    lib = L.lib()
    caps = ("video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, framerate=(fraction)30/1" % (W, H))
    pipes = []
    for sidx in range(n_streams):
        conv = TensorConverter(device_id=device_id)
        c1 = conv.set_caps(caps)
        t1 = TensorTransform(device_id=device_id, mode="typecast", option="float32")
        t2 = TensorTransform(device_id=device_id, mode="arithmetic", option="add:-127.5,div:127.5")
        t3 = TensorTransform(device_id=device_id, mode="transpose", option="1:2:0:3")
        t3.negotiate(t2.negotiate(t1.negotiate(c1)))
        x = np.random.default_rng(sidx).integers(0, 256, FRAME_ELEMS, dtype=np.uint8)
        m = t1.alloc(FRAME_ELEMS, on_device=True)
        assert lib.nnsb_transform_upload(t1.handle, C.byref(m), x.ctypes.data_as(C.c_void_p)) == 0
        pipes.append((conv, t1, t2, t3, DeviceMemory(m)))
    branches = [Branch(p[0], list(p[1:4])) for p in pipes] if linked else []

    def run_linked(branch, frame, n):
        push, release = lib.nnsb_branch_push, lib.nnsb_memory_release
        outs = (L.Memory * L.SIZE_LIMIT)()
        n_out = C.c_uint32(0)
        fr, p_n, handle = C.byref(frame.mem), C.byref(n_out), branch.handle
        last = L.Memory()
        p_last = C.byref(last)
        size = C.sizeof(L.Memory)
        for _ in range(n):
            if push(handle, fr, 1, outs, p_n) != 0:
                raise RuntimeError("branch failed: %s" % lib.nnsb_last_error())
            release(p_last)
            C.memmove(p_last, outs, size)
        L.check(lib.nnsb_memory_sync(p_last))
        release(p_last)

    def run(pipe, n):
        if linked:
            return run_linked(branches[pipes.index(pipe)], pipe[4], n)
        conv, t1, t2, t3, frame = pipe
        chain, transform, release = lib.nnsb_converter_chain, lib.nnsb_transform_transform, lib.nnsb_memory_release
        m = [(L.Memory * L.SIZE_LIMIT)() for _ in range(4)]
        produced, n_out = C.c_int(0), C.c_uint32(0)
        fr = C.byref(frame.mem)
        p_produced, p_n = C.byref(produced), C.byref(n_out)
        last = L.Memory()
        p_last = C.byref(last)
        handles = (t1.handle, t2.handle, t3.handle)
        for _ in range(n):
            if chain(conv.handle, fr, m[0], p_produced) != 0:
                raise RuntimeError("converter failed")
            for k in range(3):
                if transform(handles[k], m[k], 1, m[k + 1], p_n) != 0:
                    raise RuntimeError("transform failed: %s" % lib.nnsb_last_error())
            release(m[0])
            release(m[1])
            release(m[2])
            release(p_last)
            C.memmove(p_last, m[3], C.sizeof(L.Memory))
        L.check(lib.nnsb_memory_sync(p_last))
        release(p_last)

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
    for b in branches:
        b.close()
    for conv, t1, t2, t3, frame in pipes:
        frame.release()
        for e in (conv, t1, t2, t3):
            e.close()
    return {"frames/s": fps, "GB/s": fps * FRAME_ELEMS * 21 / 1e9, "streams": n_streams,
            "frames_per_stream": frames_per_stream,
            "driver": ("Python threads, one nnsb_branch_push per frame (elements linked), "
                       "device-resident frames") if linked else
                      "Python threads, raw ctypes calls of the C-ABI element mirrors, device-resident frames"}


STREAMS_PER_GPU = 8
TOTAL_STREAMS = 64                  # BASELINE configs[4]


def sharded_stream_chain(local_rank, rank, world, dist, reduce_device, frames_per_stream, check):
    """BASELINE configs[4]: 64 concurrent 1080p streams through converter -> typecast -> arithmetic
    -> transpose, 8 per GPU (nnstreamer_b200.sharding: stream s -> GPU s mod G), so N GPUs carry
    8 N of them.  Every rank runs tools/stream_bench (native threads through the C-ABI, one per
    stream) on ITS GPU; all ranks start their timed region together (a file barrier after warm-up), the
    frames are summed over the ranks and divided by the slowest rank's time.  Three modes
    (device-resident / H2D per frame / H2D + D2H per frame) x two widths (1920, and 1918 whose
    padded rows the converter strips).  `check` (rank 0, inside the cpu_baseline leg) compares the
    last output of stream 0 with the CPU reference chain."""
    from nnstreamer_b200 import sharding
    exe = os.path.join(ROOT, "tools", "stream_bench")
    if not os.path.isfile(exe):
        return {"unavailable": "tools/stream_bench not built (python -c 'import __graft_entry__ as g; g.build()')"}
    n_local = len(sharding.streams_of_rank(min(TOTAL_STREAMS, STREAMS_PER_GPU * world), world, rank))
    out = {}
    import tempfile
    tmp = tempfile.mkdtemp(prefix="nnsb_chain_")
    for width in (1920, 1918):
        entry = {"streams": n_local * world, "streams_per_gpu": n_local, "gpus": world,
                 "frames_per_stream": frames_per_stream,
                 "chain": "tensor_converter -> typecast:float32 -> arithmetic add:-127.5,div:127.5 "
                          "-> transpose 1:2:0:3", "driver": "tools/stream_bench, one process per GPU, "
                          "one native thread per stream"}
        for mode in ("resident", "h2d", "h2d_d2h"):
            dump = os.path.join(tmp, "w%d" % width) if (mode == "resident" and rank == 0
                                                        and check is not None) else "-"
            # same start instant on every rank: a file barrier after each process's own warm-up
            # (the ranks of one node share /tmp)
            bdir = "/tmp/nnsb_chain_barrier_%s_%d_%s" % (os.environ.get("MASTER_PORT", "0"), width, mode)
            os.makedirs(bdir, exist_ok=True)
            try:
                os.remove(os.path.join(bdir, "ready.%d" % rank))
            except OSError:
                pass
            sharding.max_over_ranks(0.0, dist, reduce_device)       # nobody starts before all cleaned up
            res = None
            try:
                r = subprocess.run([exe, str(local_rank), str(n_local), str(frames_per_stream), mode,
                                    str(width), dump, "barrier:%s:%d:%d" % (bdir, rank, world)],
                                   capture_output=True, text=True, timeout=600)
                if r.returncode == 0:
                    res = json.loads(r.stdout.strip().splitlines()[-1])
                else:
                    entry[mode] = {"error": r.stderr.strip()[-300:]}
            except Exception as exc:                # noqa: BLE001
                entry[mode] = {"error": str(exc)}
            ok = sharding.sum_over_ranks(1.0 if res else 0.0, dist, reduce_device)
            try:
                os.remove(os.path.join(bdir, "ready.%d" % rank))
                os.rmdir(bdir)
            except OSError:
                pass
            if ok < world:
                entry.setdefault(mode, {"error": "a rank failed"})
                continue
            frames = sharding.sum_over_ranks(n_local * frames_per_stream, dist, reduce_device)
            slowest = sharding.max_over_ranks(res["seconds"], dist, reduce_device)
            skew = sharding.max_over_ranks(res["start_skew_s"], dist, reduce_device)
            h = int(res["stream0_last_output_fnv1a"], 16)
            hashes = sharding.gather_over_ranks([float(h >> 32), float(h & 0xffffffff)], dist, reduce_device)
            fps = frames / slowest
            elems = 3 * width * H
            bpe = 21 + (2 if (width % 4 and mode == "resident") else 0)
            entry[mode] = {"frames/s": fps, "GB/s": fps * elems * bpe / 1e9,
                           "slowest_rank_seconds": slowest, "this_rank_frames/s": res["frames/s"],
                           "h2d_bytes_per_frame": res["h2d_bytes_per_frame"],
                           "d2h_bytes_per_frame": res["d2h_bytes_per_frame"],
                           "start_barrier_wait_s_max": skew,
                           "all_ranks_same_output_as_rank0": all(x == hashes[0] for x in hashes)}
            if dump != "-":
                entry["parity_vs_oracle_stream0_last_frame"] = check(dump, width)
                for ext in (".in", ".out"):
                    try:
                        os.remove(dump + ext)
                    except OSError:
                        pass
        out["chain_64_streams_%dx1080" % width] = entry
    try:
        os.rmdir(tmp)
    except OSError:
        pass
    return out


def chain_parity_check(dump, width):
    """cpu_baseline leg only: last output of stream 0 vs the CPU reference chain on the same frame."""
    try:
        port = cpu_checker_port()
        frame = np.fromfile(dump + ".in", dtype=np.uint8)
        got = np.fromfile(dump + ".out", dtype=np.float32)
        dense = port.stride_strip(frame, 3, width, H) if width % 4 else frame
        f = port.arithmetic(dense, [3, width, H, 1], [("typecast", "float32"), ("add", -127.5), ("div", 127.5)])
        want = port.transpose(f, [3, width, H, 1], (1, 2, 0, 3))
        return bool(got.tobytes() == want.tobytes())
    except Exception as exc:                        # noqa: BLE001
        return "check failed: %s" % (str(exc)[:80],)


def host_link_ceiling(dev, h_in_ptr, h_out_ptr, d_in, d_out, n, steps, barrier):
    """What the host link of THIS rank delivers while every other rank does the same: bare
    cudaMemcpyAsync of the step's upload (n bytes) on one stream and of its download (4 n bytes)
    on another, from / to the same pinned buffers, no kernels.  ms per step."""
    import ctypes as C
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    s2 = C.c_void_p()
    L.check(lib.nnsb_stream_create(dev.ctx, C.byref(s2)))

    def step():
        L.check(lib.nnsb_h2d_async(dev.stream, C.c_void_p(d_in), C.c_void_p(h_in_ptr), n))
        L.check(lib.nnsb_d2h_async(s2, C.c_void_p(h_out_ptr), C.c_void_p(d_out), 4 * n))
        L.check(lib.nnsb_stream_sync(dev.stream))
        L.check(lib.nnsb_stream_sync(s2))

    step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    barrier()
    lib.nnsb_stream_destroy(s2)
    return ms


def consumer_on_device_e2e(local_rank, h_batch, steps, barrier):
    """The same 32-frame batch end to end when the consumer is CUDA-aware:
    tensor_converter frames-per-tensor=32 takes the 32 HOST frames (H2D inside the timed region),
    tensor_transform runs the bench chain on the device, and a tensor_filter whose sub-plugin is a
    CUDA kernel (this library's clamp, enqueued on the stream the filter hands it) consumes the
    float32 tensor in HBM.  Nothing is downloaded.  Returns (ms per step, bytes up, bytes down)."""
    import ctypes as C
    from nnstreamer_b200 import Device, TensorConverter, TensorFilter, TensorTransform
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    conv = TensorConverter(device_id=local_rank, frames_per_tensor=B)
    c1 = conv.set_caps("video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, "
                       "framerate=(fraction)30/1" % (W, H))
    t = TensorTransform(device_id=local_rank, mode="arithmetic", option=OPTION)
    t.negotiate(c1)
    aux = Device(local_rank)
    out_dim = [C_, W, H, B]
    plan = aux.plan("clamp", "-1.5:1.5", "float32", out_dim)
    wrapped = {}

    def invoke(inputs, outputs, cuda_stream):
        if cuda_stream not in wrapped:
            sw = C.c_void_p()
            L.check(lib.nnsb_stream_wrap(aux.ctx, C.c_void_p(cuda_stream), C.byref(sw)))
            wrapped[cuda_stream] = sw
        plan.launch(inputs[0][0], outputs[0][0], 1, stream=wrapped[cuda_stream])
        return 0

    cfg = TensorFilter.make_config([("float32", out_dim)], rate=(30, 1))
    filt = TensorFilter(invoke, cfg, cfg, device_id=local_rank)
    frames = [h_batch[i * FRAME_ELEMS:(i + 1) * FRAME_ELEMS] for i in range(B)]

    def step(prev):
        tensor = None
        for f in frames:
            flow, tensor = conv.chain(f)
        flow, (y,) = t.transform([tensor])
        flow, (z,) = filt.transform([y])
        if flow != 0:
            raise RuntimeError("consumer-on-device step failed: %s" % lib.nnsb_last_error())
        tensor.release()
        y.release()
        if prev is not None:
            prev.release()
        return z

    z = step(None)
    z = step(z)
    z.sync()
    barrier()
    launches0 = t.kernel_launches
    t0 = time.perf_counter()
    for _ in range(steps):
        z = step(z)
    z.sync()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    barrier()
    launches = (t.kernel_launches - launches0) + steps          # + one clamp launch per step
    up, down, invokes = filt.counters
    z.release()
    for sw in wrapped.values():
        lib.nnsb_stream_destroy(sw)
    plan.close()
    for e in (filt, t, conv):
        e.close()
    aux.close()
    return ms, FRAME_ELEMS * B, up + down, launches


def per_frame_mobilenet(dev, local_rank, with_cpu):
    """BASELINE configs[0] as the reference runs it: frames-per-tensor=1, one 224x224 RGB frame per
    buffer through the element (host memory in, host memory out).  Launch- and copy-latency
    bound on a GPU; the one configuration where the CPU element is a fair opponent."""
    from nnstreamer_b200 import TensorTransform
    t = TensorTransform(device_id=local_rank, mode="arithmetic", option="typecast:float32,add:-127.5,div:127.5")
    t.negotiate("other/tensor, dimension=(string)3:224:224:1, type=(string)uint8, framerate=(fraction)30/1")
    n = 3 * 224 * 224
    arr, ptr = dev.pinned(n, np.uint8)
    arr[:] = np.random.default_rng(8).integers(0, 256, n, dtype=np.uint8)
    import ctypes as C
    from nnstreamer_b200 import _lib as L
    ins = (L.Memory * 1)()
    ins[0].data = ptr
    ins[0].size = n
    outs = (L.Memory * L.SIZE_LIMIT)()
    n_out = C.c_uint32(0)
    lib = L.lib()

    def one():
        if t.transform_raw(ins, 1, outs, n_out) != 0:
            raise RuntimeError("per-frame transform failed")
        lib.nnsb_memory_release(C.byref(outs[0]))

    for _ in range(50):
        one()
    frames = 2000
    t0 = time.perf_counter()
    for _ in range(frames):
        one()
    dt = time.perf_counter() - t0
    res = {"frames/s": frames / dt, "us_per_frame": dt / frames * 1e6, "frames": frames,
           "api": "nnsb_transform_transform, one pinned host frame in, one pinned host frame out per call"}
    if with_cpu:
        checker = cpu_checker()[0]
        ops = [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]
        x = np.array(arr)
        checker.arithmetic(x, [3, 224, 224, 1], ops)
        t0 = time.perf_counter()
        for _ in range(200):
            checker.arithmetic(x, [3, 224, 224, 1], ops)
        dtc = time.perf_counter() - t0
        res["cpu_reference_1_thread_frames/s"] = 200 / dtc
    dev.pinned_free(ptr)
    t.close()
    return res


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
    ms_e2e_rank = (time.perf_counter() - t0) * 1e3
    e2e_launches = elem.kernel_launches - launches_before
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    reduce_device = "cuda" if dist is not None else "cpu"
    ms_e2e = max_over_ranks(ms_e2e_rank)
    e2e_per_rank = [v[0] / args.steps for v in
                    sharding.gather_over_ranks([ms_e2e_rank], dist, reduce_device)]

    # ---- the same step with a PAGEABLE source (an upstream GstMemory is malloc'd memory): the
    #      element stages it through its pinned ring with helper threads ----
    pageable = None
    if rank == 0 and world == 1 and not args.no_extras:
        x_page = np.array(h_in)                         # ordinary numpy allocation
        ins_p = (L.Memory * 1)()
        ins_p[0].data = x_page.ctypes.data
        ins_p[0].size = n
        outs_p = (L.Memory * L.SIZE_LIMIT)()

        def page_step():
            if elem.transform_raw(ins_p, 1, outs_p, n_out) != 0:
                raise RuntimeError("transform (pageable) failed: %s" % L.lib().nnsb_last_error())
            L.lib().nnsb_memory_release(C.byref(outs_p[0]))

        page_step()
        psteps = max(3, min(args.steps, 10))
        t0 = time.perf_counter()
        for _ in range(psteps):
            page_step()
        ms_page = (time.perf_counter() - t0) * 1e3 / psteps
        pageable = {"frames/s": B / (ms_page * 1e-3), "ms_per_step": ms_page,
                    "frac_of_pinned_source_e2e": (ms_e2e_rank / args.steps) / ms_page,
                    "staging": "4 pinned slots, 2 helper threads (NNSB_STAGING_THREADS)"}

    # ---- what the host link gives every rank at the same time: the ceiling of e2e ----
    h_out, h_out_ptr = dev.pinned(4 * n, np.uint8)
    link_steps = max(3, min(args.steps, 10))
    ms_link_rank = host_link_ceiling(dev, h_in_ptr, h_out_ptr, d_in, d_out, n, link_steps, barrier)
    ms_link = max_over_ranks(ms_link_rank)
    link_per_rank = [v[0] for v in sharding.gather_over_ranks([ms_link_rank], dist, reduce_device)]
    dev.pinned_free(h_out_ptr)

    # ---- same batch, consumer on the device: H2D only, nothing downloaded ----
    cod_steps = max(3, min(args.steps, 10))
    ms_cod_rank, cod_up, cod_filter_copies, cod_launches = consumer_on_device_e2e(
        local_rank, h_in, cod_steps, barrier)
    ms_cod = max_over_ranks(ms_cod_rank)

    # ---- CPU baseline (with a parity spot check of the bench data) and the other kernels ----
    parity = None
    cpu = None
    extras = None
    with_cpu = rank == 0 and not args.no_cpu_baseline
    if with_cpu:
        # the cpu_baseline leg: the CPU code is timed (shorter sample when other ranks wait), and
        # the frame it computes anyway is compared with the first frame of the GPU's result
        cpu = cpu_baseline(budget_s=15.0 if world == 1 else 6.0)
        want = cpu_checker_port().arithmetic(h_in[:FRAME_ELEMS], DIM, ORACLE_OPS, True, 0)
        parity = bool(e2e_first.tobytes() == want.tobytes())
    if not args.no_extras:
        dev.free(d_in)
        dev.free(d_out)
        d_in = d_out = None
        elem.close()
        # configs[4] on every N: all ranks take part (each drives its own GPU)
        barrier()
        chains = sharded_stream_chain(local_rank, rank, world, dist, reduce_device,
                                      args.chain_frames, chain_parity_check if with_cpu else None)
        if rank == 0:
            extras = dict(chains)
            if world == 1:
                extras.update(extra_kernels(dev, peak, with_cpu=with_cpu))
                extras["chain_8_streams_1080p_python_threads"] = python_stream_chain(local_rank)
                extras["chain_8_streams_1080p_python_threads_linked"] = python_stream_chain(local_rank, linked=True)
                extras["mobilenet_224_per_frame_element_e2e"] = per_frame_mobilenet(dev, local_rank, with_cpu)
                extras["e2e_pageable_source"] = pageable

    if rank == 0:
        ms_step = ms_dev / args.steps
        algo_bytes = n * ALGO_BYTES_PER_ELEM
        achieved = algo_bytes / (ms_step * 1e-3) / 1e9
        e2e_ms_step = ms_e2e / args.steps
        line = {
            "metric": METRIC, "value": world * B * args.steps / (ms_dev * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": CONFIG,
            "GB/s": world * achieved,
            "e2e": {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": n, "d2h_bytes_per_step": 4 * n,
                    "ms_per_step": e2e_ms_step,
                    "ms_per_step_per_rank": e2e_per_rank,
                    # bare cudaMemcpyAsync of the same bytes from the same pinned buffers, all
                    # ranks at once, no kernels: the ceiling the host side sets for this step
                    "host_link": {"ms_per_step": ms_link, "ms_per_step_per_rank": link_per_rank,
                                  "GB/s_whole_box": world * 5 * n / (ms_link * 1e-3) / 1e9,
                                  "frames/s_ceiling": world * B / (ms_link * 1e-3)},
                    "frac_of_host_link": ms_link / e2e_ms_step,
                    "api": "nnsb_transform_transform (element mirror) on pinned host memories, "
                           "chunks pipelined on two streams; wall clock around synchronous calls"},
            # the same batch when the consumer is CUDA-aware (tensor_filter hand-off on the device):
            # converter (H2D of 32 host frames) -> transform -> filter kernel, zero bytes downloaded
            "e2e_consumer_on_device": {"value": world * B / (ms_cod * 1e-3), "unit": UNIT,
                                       "ms_per_step": ms_cod, "h2d_bytes_per_step": cod_up,
                                       "d2h_bytes_per_step": 0,
                                       "bytes_copied_by_the_filter_element": cod_filter_copies,
                                       "api": "nnsb_converter_chain x32 (host frames) -> "
                                              "nnsb_transform_transform -> nnsb_filter_transform "
                                              "(sub-plugin = clamp kernel on the filter's stream)"},
            "gpu_launches": launches + e2e_launches + cod_launches,
            "gpu_launches_detail": {"device_resident": launches, "e2e": e2e_launches,
                                    "e2e_consumer_on_device": cod_launches},
            "host_affinity": numa,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_nominal_8TBs": achieved / 8000.0,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one launch, read from the
                         # committed ncu --set full capture of this kernel on this workload
                         "traffic": ncu_traffic_bytes(),
                         "traffic_source": os.path.relpath(NCU_RAW_CSV, ROOT),
                         "peak_source": peak_src,
                         "kernel": "lut_chain_kernel<uint8,float,false,4> (k_chain.cuh)",
                         "algorithmic_bytes_per_launch": algo_bytes},
            "cpu_baseline": cpu,
            "clocks": clocks,
            "parity_vs_oracle_first_frame": parity,
        }
        if extras is not None:
            line["kernels"] = extras
        print(json.dumps(line))
    if d_in is not None:
        elem.close()
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
    ap.add_argument("--chain-frames", type=int, default=300,
                    help="frames per stream of the configs[4] multi-stream chain")
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
