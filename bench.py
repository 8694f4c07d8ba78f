#!/usr/bin/env python
"""bench.py -- headline benchmark of the NJUCloud hot path on B200 (contract: see DESIGN.md section 6).

Main line (one JSON object on stdout, rank 0), the same workload at every N so that the driver's 1 -> 8 curve measures
the gradient exchange:
  workload C4 = BASELINE.json configs[3] (N = 1: its single-device reference point, configs[2]'s net at the same batch):
  LeNet (2 conv + 2 dense) train step -- forward, softmax-CE, backward, Adagrad(1e-4) -- at GLOBAL batch 2048 split over the
  ranks (strong scaling), gradients summed by NCCL inside the step, the step replayed from a CUDA graph.
  value     = whole-job images/s with the 60 000-picture set resident in HBM (CUDA events, max over ranks)
  e2e       = the same step with every batch coming from PINNED HOST memory (H2D of pictures and labels, loss read back)
  roofline  = the dominant kernel of the step (per-stage CUDA events inside the library), tensor-bound
  cpu_baseline / --impl reference = the same step in torch-CPU on all host cores (oracle/torchref.py, kind "port":
              TensorFlow 1.x, what the reference runs, cannot be installed here)
Sub-objects of the same line: "preprocess" (BASELINE configs[1], C2: the whole preprocess half over 4096 uploads, with its
own roofline / e2e / cpu_baseline -- replicas, no collective), "train_b256" (C3), "train_weak" (256 pictures per GPU),
"train_c5_deepnet" (C5), "predict*" (C1-style inference), "dp_parity" (N > 1: the data-parallel step against the
single-device step at the global batch, asserted before anything is timed).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
Under torchrun (N > 1) one rank per GPU.
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


_T0 = time.time()


def note(msg):
    """progress on stderr (stdout carries exactly one JSON line)"""
    if int(os.environ.get("RANK", "0")) == 0:
        print("[bench %6.1fs] %s" % (time.time() - _T0, msg), file=sys.stderr, flush=True)


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


def _cpu_init(sample_n, ipp=False):
    """worker initialiser: every process builds the same bounded sample once"""
    global _CPU_SAMPLE
    import cv2
    from oracle import cvref  # noqa: F401  (importing it switches IPP off: the canonical arithmetic)
    cv2.setNumThreads(1)   # parallelism comes from one picture per worker process
    cv2.ipp.setUseIPP(bool(ipp))
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


def _cpu_pool(procs, sample_n, ipp=False):
    import multiprocessing as mp
    # spawn, not fork: by now this process owns CUDA, NCCL and OpenCV / OpenMP worker threads, and a forked child that
    # touches any of their locks can hang
    return mp.get_context("spawn").Pool(procs, initializer=_cpu_init, initargs=(sample_n, ipp))


def cpu_baseline_throughput(seconds, procs, sample_n=16, ipp=False):
    with _cpu_pool(procs, sample_n, ipp) as pool:
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
TRAIN_WORKLOAD = ("C4 LeNet train step (conv5x5x32-relu-pool-conv5x5x64-relu-pool-fc1024-relu-fc10, softmax-CE, Adagrad 1e-4 = "
                  "what the reference runs), global batch %d over %d GPU(s), 60000 synthetic MNIST-shaped pictures")
GLOBAL_BATCH = 2048


def cpu_train_step_throughput(batch, steps, warmup, threads):
    """The reference arm of the train-step benchmark: the same graph, loss and optimizer in torch-CPU on all host cores
    (oracle.torchref.TorchStep; TensorFlow 1.x itself is not installable here).  Synthetic data of the benchmark's shape."""
    import torch
    from oracle.torchref import TorchStep
    from oracle import tfref
    torch.set_num_threads(threads)
    ts = TorchStep(tfref.LENET, seed=2, threads=threads, opt="adagrad", lr=1e-4)
    g = torch.Generator().manual_seed(99)
    n_img = max(batch * 4, 8192)
    xs = torch.rand((n_img, 784), generator=g)
    ys = torch.nn.functional.one_hot(torch.randint(0, 10, (n_img,), generator=g), 10).float()
    nb = n_img // batch
    for i in range(warmup):
        ts.step(xs[(i % nb) * batch:(i % nb + 1) * batch], ys[(i % nb) * batch:(i % nb + 1) * batch])
    t0 = time.perf_counter()
    for i in range(steps):
        b = (i + warmup) % nb
        ts.step(xs[b * batch:(b + 1) * batch], ys[b * batch:(b + 1) * batch])
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    val, dt = cpu_train_step_throughput(GLOBAL_BATCH, args.steps, min(args.warmup, 2), threads)
    sample = ("%d steps of the same LeNet step at batch %d in torch-CPU (oracle/torchref.py: explicit asymmetric SAME padding, "
              "mean softmax-CE, autograd, TF-1.x Adagrad acc0 = 0.1), %d threads" % (args.steps, GLOBAL_BATCH, threads))
    line = {
        "impl": "reference", "metric": "images/sec", "value": val, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": TRAIN_WORKLOAD % (GLOBAL_BATCH, args.gpus), "global_batch": GLOBAL_BATCH},
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
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
MATH_NAMES = {0: "fp32 simt", 1: "tcgen05 3xTF32 (fp32-class accuracy)", 2: "tcgen05 1xTF32",
              3: "tcgen05 3xTF32 forward + single rounded-TF32 products for the gradients"}


def lenet_step_bytes(b):
    return 4 * b * (2 * 48858 + 2 * 48074) + 4 * 8 * LENET_PARAMS


def lenet_stage_flops(name, b):
    """algorithmic FLOPs of one profiled stage of the LeNet step at per-GPU batch b (2 * MACs of the contraction it holds)"""
    if "conv 5x5x32->64" in name:
        return 2.0 * b * 196 * 800 * 64
    if "connect 3136->1024" in name:
        return 2.0 * b * 3136 * 1024
    if "conv 5x5x1->32" in name or "L2 pool" in name or "L0 conv" in name:      # the fused first block (conv1 + relu + pool)
        return 2.0 * b * 784 * 25 * 32
    return 0.0


def dp_parity_check(torch, ctx, world, rank, device, math):
    """N > 1, before anything is timed: two data-parallel steps (library communicator, exchange overlapped with backward)
    against two single-device steps at the global batch, from the same weights.  Returns the error relative to the update."""
    from cloud_server_b200.net import Net, DataParallel
    from cloud_server_b200.configs import LENET
    g = torch.Generator(device=device)
    g.manual_seed(7)
    B = 256 * world if 256 * world <= 2048 else 2048
    x = torch.rand((2, B, 784), device=device, generator=g)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (2, B), device=device, generator=g), 10).float()
    net = Net(LENET, ctx, seed=100 + rank)
    net.set_math(math)
    if rank == 0:
        # damped start (dense x0.1, filters x0.3): at the reference's sigma-0.1 initial values SGD 0.05 is chaotic and any
        # last-bit difference (the summation order of the W partial gradients) grows by orders of magnitude in two steps
        w = net.get_params()
        for off, size, shape in net.tensors:
            if len(shape) == 2:
                w[off:off + size] *= 0.1
            elif len(shape) == 4:
                w[off:off + size] *= 0.3
        net.set_params(w)
    dp = DataParallel(net, in_library=True)
    w0 = net.params.clone()
    single = Net(LENET, ctx, seed=5)
    single.set_math(math)
    single.params.copy_(w0)
    n = B // world
    for s_ in range(2):
        dp.train_step(x[s_, rank * n:(rank + 1) * n], y[s_, rank * n:(rank + 1) * n], "entropy", "sgd", 0.05)
        single.train_step(x[s_], y[s_], "entropy", "sgd", 0.05)
    upd = float((single.params - w0).abs().max().item())
    err = float((net.params - single.params).abs().max().item()) / max(upd, 1e-30)
    t = torch.tensor([err], dtype=torch.float64, device=device)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    same = net.params.clone()
    torch.distributed.all_reduce(same, op=torch.distributed.ReduceOp.MAX)
    identical = bool(torch.equal(same, net.params))
    net.detach_nccl()
    net.close(); single.close()
    err = float(t.item())
    assert err < 1e-3 and identical, "data-parallel step differs from the single-device step: %.3e, identical ranks %s" % (err, identical)
    return {"net": "LeNet", "global_batch": B, "steps": 2, "optimizer": "SGD 0.05", "error_relative_to_update": err,
            "ranks_bit_identical": identical, "tolerance": 1e-3,
            "target": "single-device step at the global batch (construct_distribute.py:413 with batch = global)"}


def run_train_bench(torch, ctx, args, rank, world, device, peaks, dist_on, sampler):
    """The main line (C4) and the other CNN sub-benchmarks.  Synthetic MNIST-shaped data resident in HBM
    (60 000 x 784 float32 = 188 MB, larger than L2; a different batch every step), random-init weights."""
    from cloud_server_b200.net import Net, DataParallel, GraphedTrainStep
    from cloud_server_b200.configs import LENET, DEEPNET
    main, out = {}, {}
    g = torch.Generator(device=device)
    g.manual_seed(99)        # same data on every rank; each rank takes its slice of the global batch
    n_img = 60000
    xs = torch.rand((n_img, 784), device=device, generator=g)
    ys = torch.nn.functional.one_hot(torch.randint(0, 10, (n_img,), device=device, generator=g), 10).float()
    net = Net(LENET, ctx, seed=2)
    net.set_math(args.math)
    dp = DataParallel(net, in_library=(args.dp == "library"))
    steps = args.steps
    traffic = traffic_table()

    def bench_train(global_batch, want_profile=False, want_e2e=False, clock=False):
        local = global_batch // world
        nb = n_img // global_batch
        state = {"i": 0}

        def batch_of():
            b = state["i"] % nb
            state["i"] += 1
            lo = b * global_batch + rank * local
            return xs[lo:lo + local], ys[lo:lo + local]

        def step():
            xb, yb = batch_of()
            dp.train_step(xb, yb, "entropy", "compat")

        cnt = [0]
        ms = time_steps(torch, step, max(3, min(steps, 10)), args.warmup, dist_on, counter=lambda: ctx.launches, count_out=cnt)
        eager_per = max_over_ranks(torch, ms, dist_on, device) / max(3, min(steps, 10))
        launches_per_step = cnt[0] / max(3, min(steps, 10))
        prof = None
        if want_profile:
            xb, yb = batch_of()
            prof = net.profile(lambda: dp.train_step(xb, yb, "entropy", "compat"))
        graph = None
        per = eager_per
        if clock and (args.no_graph or not (world == 1 or dp.in_library)):
            sampler.start()
            ms = time_steps(torch, step, steps, args.warmup, dist_on)
            sampler.stop()
            per = eager_per = max_over_ranks(torch, ms, dist_on, device) / steps
        if not args.no_graph and (world == 1 or dp.in_library):
            # the same step replayed from a CUDA graph (kernels and, for N > 1, the NCCL all-reduces captured once); each step
            # still copies its own batch into the graph's input buffers
            graph = GraphedTrainStep(net, local, "entropy", "compat")

            def gstep():
                xb, yb = batch_of()
                graph(xb, yb)

            if clock:
                sampler.start()
            ms = time_steps(torch, gstep, steps, args.warmup, dist_on)
            if clock:
                t_end = time.perf_counter() + 0.5     # keep the same kernels running so the NVML sampler sees clocks under load
                while time.perf_counter() < t_end:
                    for _ in range(20):
                        gstep()
                    torch.cuda.synchronize()
                sampler.stop()
            per = max_over_ranks(torch, ms, dist_on, device) / steps
        byts = lenet_step_bytes(local)
        res = {
            "value": global_batch / (per * 1e-3), "unit": "images/s", "ms_per_step": per, "global_batch": global_batch,
            "per_gpu_batch": local, "optimizer": "Adagrad(1e-4) (what the reference runs)", "gpu_launches_per_step": launches_per_step,
            "cuda_graph": graph is not None, "ms_per_step_eager": eager_per,
            "gradient_exchange": None if world == 1 else ("library NCCL communicator on a side stream, dense-layer gradients overlapped "
                                                          "with the convolution backward, captured in the graph" if dp.in_library
                                                          else "torch.distributed all-reduce after backward"),
            "math": MATH_NAMES[args.math],
            "step_roofline": {"bound": "hbm", "achieved": byts / (per * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                              "frac": byts / (per * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes_per_step": byts,
                              "tflops": LENET_FLOP_PER_IMG * local / (per * 1e-3) / 1e12,
                              "note": "layer-boundary byte model of SURVEY 8(d) for the whole step"},
        }
        if prof:
            total = sum(ms_ for _, ms_ in prof)
            res["stages_ms"] = [[nm, round(ms_, 5)] for nm, ms_ in prof]
            tens = [(nm, ms_) for nm, ms_ in prof if lenet_stage_flops(nm, local) > 0]
            nm, ms_ = max(tens, key=lambda t: t[1])
            fl = lenet_stage_flops(nm, local)
            peak = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
            res["roofline"] = {
                "bound": "tensor", "achieved": fl / (ms_ * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                "frac": fl / (ms_ * 1e-3) / 1e12 / peak, "traffic": traffic.get("train_dominant_b%d" % local),
                "kernel": nm + (" (dominant stage: %.0f%% of the eager step's %.3f ms)" % (100 * ms_ / total, total)),
                "algorithmic_flops_per_launch": fl, "launch_ms": ms_, "peak_source": "measured (cuBLAS bf16, sustained)",
                "note": "peak is the dense bf16 rate; this contraction runs kind::tf32 (half that rate) and, in the forward pass, as "
                        "3 TF32 products per fp32 product (hi.hi + lo.hi + hi.lo) to hold north_star's 1e-3 on logits: the ceiling of "
                        "a forward stage is peak / 6, of a gradient stage peak / 2",
            }
        if want_e2e:
            # end to end: the batch comes from PINNED HOST memory every step (pictures + one-hot labels), the loss goes back
            hx = [xs[i * local:(i + 1) * local].cpu().pin_memory() for i in range(4)]
            hy = [ys[i * local:(i + 1) * local].cpu().pin_memory() for i in range(4)]
            lh = torch.zeros(1).pin_memory()
            xd, yd = torch.empty_like(xs[:local]), torch.empty_like(ys[:local])
            k = {"i": 0}

            def e2e():
                i = k["i"] % 4
                k["i"] += 1
                if graph is not None:
                    # the public host-fed loop: step on the batch whose copy was started a step ago, start the next copy
                    loss, _ = graph.run_prefetched()
                    lh.copy_(loss, non_blocking=True)
                    graph.prefetch(hx[i], hy[i])
                else:
                    xd.copy_(hx[i], non_blocking=True)
                    yd.copy_(hy[i], non_blocking=True)
                    loss, _ = dp.train_step(xd, yd, "entropy", "compat")
                    lh.copy_(loss, non_blocking=True)

            if graph is not None:
                graph.prefetch(hx[3], hy[3])
            ms = max_over_ranks(torch, time_steps(torch, e2e, steps, 3, dist_on), dist_on, device)
            res["e2e"] = {"value": global_batch * steps / (ms * 1e-3), "unit": "images/s",
                          "h2d_bytes_per_step": world * local * (784 + 10) * 4, "d2h_bytes_per_step": 4 * world,
                          "mode": "per rank, every step: H2D of one batch (pictures + labels) from pinned host memory on a copy stream -- the "
                                  "batch of the NEXT step, in flight under this step (GraphedTrainStep.prefetch / run_prefetched) -> graph "
                                  "replay of the step on the batch copied a step ago -> loss D2H"}
        del graph
        return res

    # ---- main line: C4, global batch 2048 over the ranks (strong scaling); N = 1 is its single-device point
    m = bench_train(GLOBAL_BATCH, want_profile=True, want_e2e=True, clock=True)
    note("main line done; other train configurations")
    main.update(m)
    main["grad_exchange"] = "p2p" if getattr(dp, "p2p", False) else ("nccl" if world > 1 else None)
    if getattr(dp, "p2p", False) and net.p2p_timed_out():
        raise RuntimeError("the peer-memory gradient exchange gave up waiting for a rank: the timed steps are void")
    if world == 1:
        out["train_b256"] = bench_train(256, want_profile=True, want_e2e=True)       # C3
    else:
        out["train_weak"] = bench_train(256 * world, want_profile=True)            # 256 per GPU
    if dp.in_library and world > 1:
        net.detach_nccl()

    # C5: the user-constructed deeper CNN (4 conv + 3 dense) on 32x32x3, global batch 512 over the ranks
    n5 = 8192
    x5 = torch.rand((n5, 32 * 32 * 3), device=device, generator=g)
    y5 = torch.nn.functional.one_hot(torch.randint(0, 10, (n5,), device=device, generator=g), 10).float()
    net5 = Net(DEEPNET, ctx, 32, 32, 3, seed=3)
    net5.set_math(args.math)
    dp5 = DataParallel(net5, in_library=(args.dp == "library"))
    gb5 = 512
    loc5 = gb5 // world
    st5 = {"i": 0}
    graph5 = GraphedTrainStep(net5, loc5, "entropy", "compat") if (not args.no_graph and (world == 1 or dp5.in_library)) else None

    def step5():
        b = st5["i"] % (n5 // gb5)
        st5["i"] += 1
        lo = b * gb5 + rank * loc5
        if graph5 is not None:
            graph5(x5[lo:lo + loc5], y5[lo:lo + loc5])
        else:
            dp5.train_step(x5[lo:lo + loc5], y5[lo:lo + loc5], "entropy", "compat")

    ms5 = max_over_ranks(torch, time_steps(torch, step5, steps, args.warmup, dist_on), dist_on, device)
    byts5 = 4 * loc5 * (2 * 114442 + 2 * 111370) + 32 * 2297130          # SURVEY 8(d) C5 formula
    out["train_c5_deepnet"] = {
        "value": gb5 * steps / (ms5 * 1e-3), "unit": "images/s", "ms_per_step": ms5 / steps, "global_batch": gb5,
        "per_gpu_batch": loc5, "input": "32x32x3", "params": net5.param_count, "cuda_graph": graph5 is not None,
        "step_roofline": {"bound": "hbm", "achieved": byts5 / (ms5 / steps * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                          "frac": byts5 / (ms5 / steps * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes_per_step": byts5,
                          "tflops": 158.5e6 * loc5 / (ms5 / steps * 1e-3) / 1e12}}
    del graph5
    if dp5.in_library and world > 1:
        net5.detach_nccl()
    net5.close()

    # inference: sess.run(argmax) (construct_inference.py:335-342), batched -- per-rank replicas
    for nb_, key in ((4096, "predict"), (99, "predict_c1_99_pictures"), (1, "predict_single")):
        xb = xs[:nb_]

        def pstep():
            net.predict(xb)

        ms = time_steps(torch, pstep, steps, 3, dist_on)
        ms = max_over_ranks(torch, ms, dist_on, device)
        out[key] = {"value": world * nb_ * steps / (ms * 1e-3), "unit": "images/s", "ms_per_batch": ms / steps, "batch": nb_}
    net.close()
    return main, out


def run_preprocess_bench(torch, ctx, args, rank, world, device, peaks, peak_src, dist_on):
    """BASELINE configs[1] (C2): the whole preprocess half over n uploads of 1024x768x3 resident in HBM (9.66 GB at the
    default n = 4096: far larger than L2) -- independent per picture, so under torchrun these are replicas, no collective."""
    traffic = traffic_table()
    n = args.batch
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

    steps = args.steps
    cnt = [0]
    ms = time_steps(torch, step, steps, args.warmup, dist_on, counter=lambda: ctx.launches, count_out=cnt)
    ms_max = max_over_ranks(torch, ms, dist_on, device)
    value = world * n * steps / (ms_max * 1e-3)
    os.environ["NC_FUSED_CHAINS"] = "1"
    ms_fused = time_steps(torch, step, steps, 3, dist_on) / steps
    del os.environ["NC_FUSED_CHAINS"]
    ms_t = time_steps(torch, step_train, steps, 3, dist_on) / steps
    ms_p = time_steps(torch, step_pred, steps, 3, dist_on) / steps
    ach_t = ALG_BYTES_TRAIN_CHAIN * n / (ms_t * 1e-3) / 1e9
    ach_p = ALG_BYTES_PREDICT_CHAIN * n / (ms_p * 1e-3) / 1e9
    tr_t, tr_p = traffic.get("train_chain_rows_kernel"), traffic.get("predict_chain_mma_kernel")
    peak = peaks["hbm_gbs"]
    sub = {
        "metric": "images/sec", "value": value, "unit": "images/s", "ms_per_step": ms_max / steps, "scaling": "weak", "dtype": "u8",
        "config": {"workload": WORKLOAD % n, "pictures_per_gpu": n, "parallelism": "replicas x%d (no collective)" % world,
                   "l2": "inputs (%.2f GB) larger than L2" % (n * H * W * C / 1e9)},
        "gpu_launches": cnt[0],
        "roofline": {"bound": "hbm", "achieved": ach_p, "peak": peak, "unit": "GB/s", "frac": ach_p / peak,
                     "traffic": tr_p * n if tr_p else None, "kernel": "predict_chain_mma_kernel (dominant: %.0f%% of the step)"
                     % (100 * ms_p / (ms_p + ms_t)), "peak_source": peak_src + " (burst copy bandwidth)",
                     "algorithmic_bytes_per_launch": ALG_BYTES_PREDICT_CHAIN * n, "launch_ms": ms_p,
                     "note": "peak = copy bandwidth (half reads, half writes); this kernel only reads (99.7 %), and a "
                             "read-only stream can pass the copy figure: frac > 1 is possible"},
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
    }
    # ---- e2e: pinned host uploads through the public API (chunked H2D overlapped with the kernels), all results read back
    n_e = min(n, args.e2e_batch)
    host = torch.empty((n_e, H, W, C), dtype=torch.uint8).pin_memory()
    host.copy_(batch[:n_e])
    res_host = [torch.empty(o[:n_e].shape, dtype=o.dtype).pin_memory() for o in outs]
    outs_e = tuple(o[:n_e] for o in outs)

    def e2e_step():
        r = ctx.preprocess_uploads(host, channel_order="BGR", out=outs_e)
        for hbuf, dbuf in zip(res_host, r):
            hbuf.copy_(dbuf, non_blocking=True)

    ks = max(3, min(steps, 10))
    ms_e = max_over_ranks(torch, time_steps(torch, e2e_step, ks, 3, dist_on), dist_on, device)
    torch.cuda.synchronize()
    assert all(torch.equal(hb, db.cpu()) for hb, db in zip(res_host, outs_e))

    def e2e_train_only():
        ctx.resize_gray_norm(host, out_u8=outs_e[0], out_f32=outs_e[1])
        res_host[1].copy_(outs_e[1], non_blocking=True)

    ms_z = max_over_ranks(torch, time_steps(torch, e2e_train_only, ks, 3, dist_on), dist_on, device)
    sub["e2e"] = {
        "value": world * n_e * ks / (ms_e * 1e-3), "unit": "images/s", "h2d_bytes_per_step": int(n_e * H * W * C),
        "d2h_bytes_per_step": int(n_e * (784 + 3136 + 3136)), "pictures_per_step": n_e,
        "mode": "pinned host batch -> 256-picture chunks, H2D on a copy stream overlapped with both kernels -> D2H of all outputs",
    }
    sub["train_side_chain"]["e2e"] = {
        "value": world * n_e * ks / (ms_z * 1e-3), "unit": "images/s", "h2d_bytes_per_step": int(n_e * 112 * W * C),
        "d2h_bytes_per_step": int(n_e * 3136), "mode": "zero copy: cp.async.bulk of the 112 needed rows straight from pinned host memory"}
    del host
    note("preprocess: e2e from JPEG files")
    # ---- e2e from the FILES the platform really holds: JPEG uploads (cv2.imread, preprocess.py:28) -> both chains.
    #      Huffman decoding on the host cores, coefficients over PCIe, IDCT / upsampling / colour + both chains on the device.
    try:
        import cv2
        cv2.setNumThreads(1)
        rng = np.random.default_rng(5)
        distinct = 16
        files = []
        for i in range(distinct):
            img = cv2.GaussianBlur(rng.integers(0, 256, (H, W, C), np.uint8), (0, 0), 3)     # natural-ish spectrum
            ok, buf = cv2.imencode(".jpg", img)                                          # cv2.imwrite defaults: q95, 4:2:0
            files.append(buf.tobytes())
        n_j = 256
        batch_files = [files[i % distinct] for i in range(n_j)]
        dec = torch.empty((n_j, H, W, C), dtype=torch.uint8, device=device)
        outs_j = tuple(o[:n_j] for o in outs)
        res_j = [torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs_j]

        def jpeg_step():
            d = ctx.decode_jpegs(batch_files, channel_order="BGR", out=dec)
            r = ctx.preprocess_uploads(d, channel_order="BGR", out=outs_j)
            for hbuf, dbuf in zip(res_j, r):
                hbuf.copy_(dbuf, non_blocking=True)

        jpeg_step()
        torch.cuda.synchronize()
        want = cv2.imdecode(np.frombuffer(files[0], np.uint8), cv2.IMREAD_COLOR)
        assert np.array_equal(dec[0].cpu().numpy(), want), "device JPEG decode differs from cv2.imread"
        t0 = time.perf_counter()
        kj = 3
        for _ in range(kj):
            jpeg_step()
        torch.cuda.synchronize()
        dtj = time.perf_counter() - t0
        t1 = time.perf_counter()
        for f in files[:8]:
            cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR)
        cv_ms = (time.perf_counter() - t1) / 8 * 1e3
        sub["e2e_from_jpeg"] = {
            "value": n_j * kj / dtj, "unit": "images/s", "pictures_per_step": n_j, "jpeg_bytes_per_step": int(sum(len(f) for f in batch_files)),
            "h2d_bytes_per_step": int(ctx.last_jpeg_h2d_bytes), "decoded_bytes_per_step": int(n_j * H * W * C),
            "d2h_bytes_per_step": int(n_j * (784 + 3136 + 3136)), "host_threads": os.cpu_count(),
            "cv2_imdecode_ms_per_picture_1_thread": cv_ms,
            "mode": "q95 4:2:0 JPEG files in host memory -> Huffman decode on the host cores -> nonzero coefficients H2D -> device IDCT / "
                    "fancy upsampling / YCbCr->BGR (== cv2.imread, asserted) -> both chains -> D2H; wall clock, per rank",
        }
    except ImportError:
        pass
    del batch
    torch.cuda.empty_cache()
    return sub


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
    sampler = ClockSampler(local_rank)

    note("context up; data-parallel parity check" if dist_on else "context up")
    parity = dp_parity_check(torch, ctx, world, rank, device, args.math) if dist_on else None
    note("train benchmarks")
    main, subs = run_train_bench(torch, ctx, args, rank, world, device, peaks, dist_on, sampler)
    local = GLOBAL_BATCH // world
    line = {
        "metric": "images/sec", "value": main["value"], "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "tf32", "data": "synthetic",
        "config": {"workload": TRAIN_WORKLOAD % (GLOBAL_BATCH, world), "global_batch": GLOBAL_BATCH, "per_gpu_batch": local,
                   "parallelism": "single device" if world == 1 else "dp%d: %s all-reduce (sum) of the 13.1 MB flat gradient buffer inside the step" % (
                       world, "NVLink peer-memory (own kernel)" if main.get("grad_exchange") == "p2p" else "NCCL"),
                   "l2": "inputs: the 60000-picture set (188 MB float32) is larger than L2 and every step takes a different batch"},
        "gpu_launches": int(round(main["gpu_launches_per_step"] * args.steps)),
        "roofline": main.get("roofline"), "e2e": main.get("e2e"), "clocks": sampler.summary(),
    }
    for k in ("step_roofline", "stages_ms", "cuda_graph", "ms_per_step_eager", "gradient_exchange", "math", "optimizer", "gpu_launches_per_step"):
        line[k] = main.get(k)
    if parity:
        line["dp_parity"] = parity
    line.update(subs)

    # ---- C2: the preprocess half (sub-object with its own roofline / e2e / cpu_baseline)
    if not args.no_preprocess:
        note("preprocess benchmark (C2)")
        line["preprocess"] = run_preprocess_bench(torch, ctx, args, rank, world, device, peaks, peak_src, dist_on)

    # ---- CPU baselines on this box's host cores (rank 0, N == 1 only), bounded samples
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        ksteps = 8
        note("cpu baseline: torch-CPU train step, %d threads" % threads)
        v, dt = cpu_train_step_throughput(GLOBAL_BATCH, ksteps, 1, threads)
        line["cpu_baseline"] = {
            "value": v, "unit": "images/s", "cores": threads, "kind": "port",
            "sample": "%d steps of the same LeNet step at batch %d in torch-CPU (oracle/torchref.py; TensorFlow 1.x is not "
                      "installable here), %.1f s" % (ksteps, GLOBAL_BATCH, dt)}
        if "preprocess" in line:
            note("cpu baseline: cv2 / Pillow preprocess chains, %d worker processes" % threads)
            pv, done, pdt = cpu_baseline_throughput(args.cpu_seconds, threads)
            line["preprocess"]["cpu_baseline"] = {
                "value": pv, "unit": "images/s", "cores": threads, "kind": "port",
                "sample": "16 distinct 1024x768x3 pictures per worker cycled for %.1f s (%d pictures); the reference's own cv2 / "
                          "Pillow calls for both chains (preprocess.py:29,40; construct_inference.py:312-330) in memory, IPP off, "
                          "one worker process per core" % (pdt, done)}
            pv2, done2, pdt2 = cpu_baseline_throughput(min(args.cpu_seconds, 6.0), threads, ipp=True)
            line["preprocess"]["cpu_baseline_ipp_on"] = {"value": pv2, "unit": "images/s", "cores": threads, "kind": "port",
                                                         "sample": "same, cv2 with its default IPP arithmetic, %.1f s (%d pictures)" % (pdt2, done2)}
    note("done")
    if rank == 0:
        _emit(line)
    if dist_on:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="uploads per GPU (C2: 4096)")
    ap.add_argument("--e2e-batch", type=int, default=1024, help="uploads per e2e step (host pinned)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-preprocess", action="store_true", help="skip the C2 preprocess sub-benchmark")
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
