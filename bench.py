#!/usr/bin/env python
"""bench.py -- headline benchmark of the NJUCloud hot path on B200 (contract: see DESIGN.md section 6).

Main line (one JSON object on stdout, rank 0):
  workload C2 = BASELINE.json configs[1]: batch preprocess of 4096 synthetic 1024x768x3 uploads
  -> 28x28 normalised float (train-side chain: cubic resize -> gray -> *1/255), per GPU.
  value  = whole-job images/s with the uploads already resident in HBM (CUDA events, max over ranks)
  e2e    = the same through the public API with HOST buffers (pinned), H2D + kernel + D2H inside
  roofline / cpu_baseline / clocks / gpu_launches as the contract asks.
Extra objects in the same line: "predict_chain" (the streaming, HBM-bound kernel of the preprocess
half), "train" (LeNet train step, data parallel over the ranks), "predict" (LeNet inference).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
Under torchrun (N > 1) one rank per GPU; preprocess = independent replicas (no collective),
training = NCCL all-reduce of the flat gradient buffer.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

H, W, C = 768, 1024, 3
ALG_BYTES_TRAIN_CHAIN = 28 * 28 * 16 * 3 + 784 + 3136          # 41 552 B / picture
ALG_BYTES_PREDICT_CHAIN = H * W * C + 3136                      # 2 362 432 B / picture


def _emit(line):
    print(json.dumps(line), flush=True)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def traffic_table():
    p = os.path.join(ROOT, "profiles", "traffic.json")
    return json.load(open(p)) if os.path.exists(p) else {}


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU through NVML while a timed region runs."""

    def __init__(self, index, period=0.02):
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            self.nv = None
            self.err = repr(e)

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8)),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40)),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20)),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)),
        }
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(get_reasons(self.h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.nv:
            self._stop.clear()
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()

    def stop(self):
        if self._t:
            self._stop.set()
            self._t.join()
            self._t = None

    def summary(self):
        if not self.nv or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the same cv2 calls the reference makes, on the host cores
# ------------------------------------------------------------------------------------------
_CPU_SAMPLE = None


def _cpu_init(sample_n):
    """worker initialiser: every process builds the same bounded sample once"""
    global _CPU_SAMPLE
    import cv2
    cv2.setNumThreads(1)   # parallelism comes from one picture per worker process
    _CPU_SAMPLE = np.random.default_rng(0).integers(0, 256, (sample_n, H, W, C), np.uint8)


def _cpu_work(i):
    """One picture through the reference's own library calls, in memory: cv2.resize INTER_CUBIC ->
    BGR2GRAY -> *1/255 (preprocess.py:29,40; construct_distribute.py:503-504) and PIL convert('L') ->
    resize((20,20)) -> threshold/canvas -> /255 (construct_inference.py:312-330)."""
    from oracle import cvref
    img = _CPU_SAMPLE[i % len(_CPU_SAMPLE)]
    cvref.train_chain(img)
    cvref.predict_chain(img[:, :, ::-1], compat=False)
    return 1


def _cpu_pool(procs, sample_n):
    import multiprocessing as mp
    return mp.get_context("fork").Pool(procs, initializer=_cpu_init, initargs=(sample_n,))


def cpu_baseline_throughput(seconds, procs, sample_n=16):
    with _cpu_pool(procs, sample_n) as pool:
        pool.map(_cpu_work, range(procs))          # warm-up: sample built, libraries loaded
        done = 0
        t0 = time.perf_counter()
        while True:
            done += sum(pool.map(_cpu_work, range(done, done + 8 * procs), chunksize=2))
            dt = time.perf_counter() - t0
            if dt >= seconds:
                break
    return done / dt, done, dt


WORKLOAD = ("C2 batch preprocess: %d synthetic 1024x768x3 uint8 uploads per GPU through the whole preprocess half -> "
            "28x28 float32: train-side chain (cv2-exact cubic resize -> gray -> *1/255) and predict-side chain "
            "(PIL-exact L -> antialiased bicubic 20x20 -> threshold -> canvas -> /255)")


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per_step = 256
    with _cpu_pool(threads, 16) as pool:
        def step():
            return sum(pool.map(_cpu_work, range(per_step), chunksize=2))
        for _ in range(min(args.warmup, 1)):
            step()
        t0 = time.perf_counter()
        n = 0
        for _ in range(args.steps):
            n += step()
        dt = time.perf_counter() - t0
    val = n / dt
    line = {
        "impl": "reference", "metric": "images/sec", "value": val, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD % per_step, "pictures_per_step": per_step},
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": "16 distinct 1024x768x3 pictures per worker cycled, %d per step; the reference's own cv2 / "
                                   "Pillow calls for both chains, in memory (no JPEG), IPP off, one worker process per core"
                                   % per_step},
        "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _emit(line)


# ------------------------------------------------------------------------------------------
def time_steps(torch, fn, steps, warmup, dist_on, counter=None, count_out=None):
    """W warm-up, then exactly K steps between barrier+synchronize, CUDA events on the current
    stream; returns this rank's milliseconds for the K steps.  counter() (the library's own
    launch counter) is read on both sides of the timed loop into count_out[0]."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist_on:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0 = counter() if counter else 0
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if counter and count_out is not None:
        count_out[0] = counter() - c0
    if dist_on:
        torch.distributed.barrier()
    return e0.elapsed_time(e1)


def max_over_ranks(torch, ms, dist_on, device):
    if not dist_on:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


# LeNet step model: 82.05 MFLOP / picture; HBM bytes per step (fp32, every saved
# activation and activation gradient written once + read once, parameters read fwd + dgrad, gradient
# written, Adagrad reads g,w,acc and writes w,acc)
LENET_FLOP_PER_IMG = 82.05e6
LENET_PARAMS = 3274634


def lenet_step_bytes(b):
    return 4 * b * (2 * 48858 + 2 * 48074) + 4 * 8 * LENET_PARAMS


def run_net_bench(torch, ctx, args, rank, world, device, peaks, dist_on):
    """LeNet (BASELINE configs 2-3): train step images/s (data parallel over the ranks, NCCL
    all-reduce of the flat gradient buffer) and inference images/s.  Synthetic MNIST-shaped data
    resident in HBM (60 000 x 784 float32), random-init weights."""
    from cloud_server_b200.net import Net, DataParallel
    from cloud_server_b200.configs import LENET
    out = {}
    g = torch.Generator(device=device)
    g.manual_seed(99)        # same data on every rank; each rank takes its slice of the global batch
    n_img = 60000
    xs = torch.rand((n_img, 784), device=device, generator=g)
    ys = torch.nn.functional.one_hot(torch.randint(0, 10, (n_img,), device=device, generator=g), 10).float()
    net = Net(LENET, ctx, seed=2)
    net.set_math(args.math)
    # N > 1: the gradient exchange runs in the library (its own NCCL communicator, tail of the buffer overlapped with the
    # rest of backward) unless --dp torch asks for torch.distributed's all-reduce after backward
    dp = DataParallel(net, in_library=(args.dp == "library"))
    steps = max(10, args.steps)

    def bench_train(global_batch, key):
        local = global_batch // world
        nb = n_img // global_batch
        state = {"i": 0}

        def step():
            b = state["i"] % nb
            state["i"] += 1
            lo = b * global_batch + rank * local
            dp.train_step(xs[lo:lo + local], ys[lo:lo + local], "entropy", "compat")

        cnt = [0]
        ms = time_steps(torch, step, steps, args.warmup, dist_on, counter=lambda: ctx.launches, count_out=cnt)
        ms = max_over_ranks(torch, ms, dist_on, device)
        eager_per = ms / steps
        graph = None
        if world == 1 and not args.no_graph:
            # the same step replayed from a CUDA graph (one graph launch instead of ~30 kernel launches); each step still
            # copies its own batch into the graph's input buffers
            from cloud_server_b200.net import GraphedTrainStep
            graph = GraphedTrainStep(net, local, "entropy", "compat")

            def gstep():
                b = state["i"] % nb
                state["i"] += 1
                lo = b * global_batch
                graph(xs[lo:lo + local], ys[lo:lo + local])

            ms = time_steps(torch, gstep, steps, args.warmup, dist_on)
        per = ms / steps
        byts = lenet_step_bytes(local)
        out[key] = {
            "value": global_batch * steps / (ms * 1e-3), "unit": "images/s", "ms_per_step": per, "global_batch": global_batch,
            "per_gpu_batch": local, "optimizer": "Adagrad(1e-4) (what the reference runs)", "gpu_launches_per_step": cnt[0] / steps,
            "cuda_graph": graph is not None, "ms_per_step_eager": eager_per,
            "gradient_exchange": None if world == 1 else ("library NCCL communicator, dense-layer gradients overlapped with the "
                                                          "convolution backward" if dp.in_library else "torch.distributed all-reduce after backward"),
            "math": {0: "fp32 simt", 1: "tcgen05 3xTF32 (fp32-class accuracy)", 2: "tcgen05 1xTF32", 3: "tcgen05 3xTF32 forward + 1xTF32 gradients"}[args.math],
            "roofline": {"bound": "hbm", "achieved": byts / (per * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": byts / (per * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes_per_step": byts,
                         "tflops": LENET_FLOP_PER_IMG * local / (per * 1e-3) / 1e12},
        }

    if world == 1:
        bench_train(256, "train")                     # C3: batch 256
        bench_train(2048, "train_b2048")              # C4's single-device reference point
    else:
        bench_train(2048, "train")                    # C4: global batch 2048 over the ranks (strong)
        bench_train(256 * world, "train_weak")        # 256 per GPU (weak)

    # e2e train step: host (pinned) batch in, loss out, every step
    if world == 1:
        xb = xs[:256].cpu().pin_memory()
        yb = ys[:256].cpu().pin_memory()
        xd, yd = torch.empty_like(xs[:256]), torch.empty_like(ys[:256])
        lh = torch.zeros(1).pin_memory()

        gstep = None
        if not args.no_graph:
            from cloud_server_b200.net import GraphedTrainStep
            gstep = GraphedTrainStep(net, 256, "entropy", "compat")

        def e2e():
            if gstep is not None:      # pinned host batch -> the graph's input buffers -> replay -> loss back to the host
                gstep.x.copy_(xb, non_blocking=True)
                gstep.y.copy_(yb, non_blocking=True)
                gstep.graph.replay()
                lh.copy_(gstep.loss, non_blocking=True)
                return
            xd.copy_(xb, non_blocking=True)
            yd.copy_(yb, non_blocking=True)
            loss, _ = net.train_step(xd, yd, "entropy", "compat")
            lh.copy_(loss, non_blocking=True)

        ms = time_steps(torch, e2e, steps, 3, False)
        out["train"]["e2e"] = {"value": 256 * steps / (ms * 1e-3), "unit": "images/s",
                               "h2d_bytes_per_step": 256 * (784 + 10) * 4, "d2h_bytes_per_step": 4}

    # C5: the user-constructed deeper CNN (4 conv + 3 dense) on 32x32x3, global batch 512 over the ranks
    from cloud_server_b200.configs import DEEPNET
    n5 = 8192
    x5 = torch.rand((n5, 32 * 32 * 3), device=device, generator=g)
    y5 = torch.nn.functional.one_hot(torch.randint(0, 10, (n5,), device=device, generator=g), 10).float()
    net5 = Net(DEEPNET, ctx, 32, 32, 3, seed=3)
    net5.set_math(args.math)
    dp5 = DataParallel(net5, in_library=(args.dp == "library"))
    gb5 = 512
    loc5 = gb5 // world
    st5 = {"i": 0}

    def step5():
        b = st5["i"] % (n5 // gb5)
        st5["i"] += 1
        lo = b * gb5 + rank * loc5
        dp5.train_step(x5[lo:lo + loc5], y5[lo:lo + loc5], "entropy", "compat")

    ms5 = max_over_ranks(torch, time_steps(torch, step5, steps, args.warmup, dist_on), dist_on, device)
    byts5 = 4 * loc5 * (2 * 114442 + 2 * 111370) + 32 * 2297130          # SURVEY 8(d) C5 formula
    out["train_c5_deepnet"] = {
        "value": gb5 * steps / (ms5 * 1e-3), "unit": "images/s", "ms_per_step": ms5 / steps, "global_batch": gb5,
        "per_gpu_batch": loc5, "input": "32x32x3", "params": net5.param_count,
        "roofline": {"bound": "hbm", "achieved": byts5 / (ms5 / steps * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": byts5 / (ms5 / steps * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes_per_step": byts5,
                     "tflops": 158.5e6 * loc5 / (ms5 / steps * 1e-3) / 1e12}}

    # inference: sess.run(argmax) (construct_inference.py:335-342), batched
    for nb_, key in ((4096, "predict"), (99, "predict_c1_99_pictures"), (1, "predict_single")):
        xb = xs[:nb_]

        def pstep():
            net.predict(xb)

        ms = time_steps(torch, pstep, steps, 3, dist_on)
        ms = max_over_ranks(torch, ms, dist_on, device)
        out[key] = {"value": world * nb_ * steps / (ms * 1e-3), "unit": "images/s", "ms_per_batch": ms / steps, "batch": nb_}
    return out


def run_ours(args, rank, local_rank, world):
    import torch
    from cloud_server_b200.context import Context

    dist_on = world > 1
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if dist_on:
        torch.distributed.init_process_group("nccl", device_id=device)
    ctx = Context(local_rank)
    peaks, peak_src = measured_peaks()
    traffic = traffic_table()
    n = args.batch

    # ---- synthetic uploads, resident in HBM (9.66 GB at the default batch: far larger than L2)
    g = torch.Generator(device=device)
    g.manual_seed(1234 + rank)
    batch = torch.randint(0, 256, (n, H, W, C), dtype=torch.uint8, device=device, generator=g)
    outs = (torch.empty((n, 28, 28), dtype=torch.uint8, device=device),
            torch.empty((n, 784), dtype=torch.float32, device=device),
            torch.empty((n, 784), dtype=torch.float32, device=device))

    def step():            # one pass of the preprocess half over the batch: 2 kernel launches
        ctx.preprocess_uploads(batch, channel_order="BGR", out=outs)

    def step_train():
        ctx.resize_gray_norm(batch, out_u8=outs[0], out_f32=outs[1])

    def step_pred():
        ctx.predict_chain(batch, channel_order="BGR", out=outs[2])

    sampler = ClockSampler(local_rank)
    sampler.start()
    cnt = [0]
    ms = time_steps(torch, step, args.steps, args.warmup, dist_on, counter=lambda: ctx.launches, count_out=cnt)
    launches = cnt[0]
    # keep the same kernels running ~0.5 s more so the NVML sampler sees clocks under this load
    t_end = time.perf_counter() + 0.5
    while time.perf_counter() < t_end:
        for _ in range(5):
            step()
        torch.cuda.synchronize()
    sampler.stop()
    ms_max = max_over_ranks(torch, ms, dist_on, device)
    value = world * n * args.steps / (ms_max * 1e-3)
    # the opt-in single-pass kernel (both chains from one read of every upload), same step, for the record
    os.environ["NC_FUSED_CHAINS"] = "1"
    ms_fused = time_steps(torch, step, args.steps, 3, dist_on) / args.steps
    del os.environ["NC_FUSED_CHAINS"]
    # each kernel alone, same inputs (per-launch duration for the rooflines)
    ms_t = time_steps(torch, step_train, args.steps, 3, dist_on) / args.steps
    ms_p = time_steps(torch, step_pred, args.steps, 3, dist_on) / args.steps
    ach_t = ALG_BYTES_TRAIN_CHAIN * n / (ms_t * 1e-3) / 1e9
    ach_p = ALG_BYTES_PREDICT_CHAIN * n / (ms_p * 1e-3) / 1e9
    tr_t, tr_p = traffic.get("train_chain_rows_kernel"), traffic.get("predict_chain_mma_kernel")
    peak = peaks["hbm_gbs"]

    line = {
        "metric": "images/sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD % n, "pictures_per_gpu": n, "parallelism": "replicas x%d (no collective)" % world,
                   "l2": "inputs (%.2f GB) larger than L2" % (n * H * W * C / 1e9)},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": ach_p, "peak": peak, "unit": "GB/s", "frac": ach_p / peak,
                     "traffic": tr_p * n if tr_p else None, "kernel": "predict_chain_mma_kernel (dominant: %.0f%% of the step)"
                     % (100 * ms_p / (ms_p + ms_t)), "peak_source": peak_src + " (burst copy bandwidth)",
                     "algorithmic_bytes_per_launch": ALG_BYTES_PREDICT_CHAIN * n, "launch_ms": ms_p,
                     "note": "peak = copy bandwidth (half reads, half writes); this kernel only reads (99.7 %), and a "
                             "read-only stream can pass the copy figure: frac > 1 is possible.  CUDA events: 82-85 % of the 8.18 TB/s "
                             "pin rate (profiles/r01n_ncu_preprocess.md)"},
        "train_side_chain": {"value": world * n / (ms_t * 1e-3), "unit": "images/s", "launch_ms": ms_t,
                             "roofline": {"bound": "hbm", "achieved": ach_t, "peak": peak, "unit": "GB/s", "frac": ach_t / peak,
                                          "traffic": tr_t * n if tr_t else None, "kernel": "train_chain_rows_kernel",
                                          "algorithmic_bytes_per_launch": ALG_BYTES_TRAIN_CHAIN * n,
                                          "dram_frac": (tr_t * n / (ms_t * 1e-3) / 1e9 / peak) if tr_t else None,
                                          "note": "4x4-tap resize: DRAM moves the 112 source rows whole (64 B bursts, taps "
                                                  "110 B apart), 8.3x the algorithmic bytes; dram_frac = ncu traffic / time / peak"}},
        "predict_side_chain": {"value": world * n / (ms_p * 1e-3), "unit": "images/s", "launch_ms": ms_p},
        "single_pass_opt_in": {"value": world * n / (ms_fused * 1e-3), "unit": "images/s", "ms_per_step": ms_fused,
                               "note": "NC_FUSED_CHAINS=1: one kernel, every upload read once (13% less HBM traffic); "
                                       "not the default because it is no faster"},
        "clocks": sampler.summary(),
    }

    # ---- e2e: pinned host uploads through the public API (chunked H2D overlapped with the kernels),
    #      all three results read back to the host
    n_e = min(n, args.e2e_batch)
    host = torch.empty((n_e, H, W, C), dtype=torch.uint8).pin_memory()
    host.copy_(batch[:n_e])
    res_host = [torch.empty(o[:n_e].shape, dtype=o.dtype).pin_memory() for o in outs]
    outs_e = tuple(o[:n_e] for o in outs)

    def e2e_step():
        r = ctx.preprocess_uploads(host, channel_order="BGR", out=outs_e)
        for hbuf, dbuf in zip(res_host, r):
            hbuf.copy_(dbuf, non_blocking=True)

    ks = max(3, min(args.steps, 10))
    ms_e = max_over_ranks(torch, time_steps(torch, e2e_step, ks, 3, dist_on), dist_on, device)
    torch.cuda.synchronize()
    assert all(torch.equal(hb, db.cpu()) for hb, db in zip(res_host, outs_e))

    # train-side chain alone, zero copy: the kernel bulk-copies only the source rows it needs over PCIe
    def e2e_train_only():
        ctx.resize_gray_norm(host, out_u8=outs_e[0], out_f32=outs_e[1])
        res_host[1].copy_(outs_e[1], non_blocking=True)

    ms_z = max_over_ranks(torch, time_steps(torch, e2e_train_only, ks, 3, dist_on), dist_on, device)
    line["e2e"] = {
        "value": world * n_e * ks / (ms_e * 1e-3), "unit": "images/s", "h2d_bytes_per_step": int(n_e * H * W * C),
        "d2h_bytes_per_step": int(n_e * (784 + 3136 + 3136)), "pictures_per_step": n_e,
        "mode": "pinned host batch -> 256-picture chunks, H2D on a copy stream overlapped with both kernels -> D2H of all outputs",
    }
    line["train_side_chain"]["e2e"] = {
        "value": world * n_e * ks / (ms_z * 1e-3), "unit": "images/s", "h2d_bytes_per_step": int(n_e * 112 * W * C),
        "d2h_bytes_per_step": int(n_e * 3136), "mode": "zero copy: cp.async.bulk of the 112 needed rows straight from pinned host memory"}
    del host

    # ---- CNN train / predict sub-benchmarks (filled in when the net library is present)
    line.update(run_net_bench(torch, ctx, args, rank, world, device, peaks, dist_on))

    # ---- CPU baseline on this box's host cores (rank 0, N == 1 only), bounded sample
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, done, dt = cpu_baseline_throughput(args.cpu_seconds, threads)
        line["cpu_baseline"] = {
            "value": v, "unit": "images/s", "cores": threads, "kind": "port",
            "sample": "16 distinct 1024x768x3 pictures per worker cycled for %.1f s (%d pictures); the reference's own cv2 / "
                      "Pillow calls for both chains (preprocess.py:29,40; construct_inference.py:312-330) in memory, IPP off, "
                      "one worker process per core" % (dt, done)}
    if rank == 0:
        _emit(line)
    if dist_on:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="uploads per GPU (C2: 4096)")
    ap.add_argument("--e2e-batch", type=int, default=1024, help="uploads per e2e step (host pinned)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dp", default="library", choices=["library", "torch"], help="who sums the gradients when N > 1")
    ap.add_argument("--no-graph", action="store_true", help="time the train step eagerly only (default: eager and CUDA-graph replay)")
    ap.add_argument("--math", type=int, default=3, help="CNN math: 0 fp32 SIMT, 1 tcgen05 3xTF32, 2 tcgen05 1xTF32, 3 (default) 3xTF32 forward + 1xTF32 gradients")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    # The contract is ONE JSON line on stdout.  Libraries write banners to fd 1 behind Python's back (NCCL prints its
    # version line there on the first communicator): park the real stdout, point fd 1 at stderr for the whole run, and
    # print the result line to the parked descriptor.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    global _emit

    def _emit(line):
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        print("bench.py: --gpus %d needs torchrun with that many ranks" % args.gpus, file=sys.stderr)
        sys.exit(2)
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
