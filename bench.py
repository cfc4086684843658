#!/usr/bin/env python
"""bench.py — env-steps/sec (select-action + TD update) of the Q-policy hot path on B200.

    python bench.py --gpus 1 --steps K --warmup W            # our arm, one GPU
    torchrun ... bench.py --gpus N --steps K --warmup W      # our arm, N GPUs (one rank per GPU)
    python bench.py --impl reference --steps K --warmup W    # the CPU arm (oracle port, all host cores)

A bench "step" is one pass of the hot path over one batch of synthetic input: a full episode
(series_len - window env-steps, 799 at the default sizes) of every episode of the workload.
N = 1 workload: BASELINE.json configs[1] — 4096 parallel episodes, hist 200 (window 201), hidden 200,
fp32, one synthetic 1000-step quarter-tick random-walk series per episode, budget 1000, 0 shares.
N > 1: the same per GPU (weak scaling), episodes sharded by rank, one gradient all-reduce per env-step.

value  : whole-job env-steps/s with prices already resident in HBM (CUDA events on the library's stream).
e2e    : the same through qp_train with HOST buffers: upload of the price series from pinned memory,
         every env-step, download of the portfolios — all inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WINDOW, HIDDEN, SERIES_LEN, BUDGET0, SHARES0, SEED = 201, 200, 1000, 1000.0, 0, 42
D = WINDOW + 2
P = D * HIDDEN + 3 * HIDDEN
FLOP_PER_ENV_STEP = 6 * D * HIDDEN + 16 * HIDDEN          # 246 800
ROLLOUT = SERIES_LEN - WINDOW                             # env-steps per bench step
SCALING = "weak"                                          # per-GPU episodes fixed as N grows (configs[1], configs[2])
METRIC = "env-steps/sec (act+TD update)"
UNIT = "env-steps/s"
WORKLOAD_NAME = "configs[1]"
EPISODES_PER_SERIES = 1                                    # configs[4] (router fan-out): 256 episodes share a series


def set_workload(name: str, args):
    """configs[1] (default): 4096 episodes, window 201, hidden 200.  'wide' = configs[3]: hist 512
    (window 513), hidden 2048, batch 65 536 on one GPU; a bench step is then a 16-env-step slice."""
    global WINDOW, HIDDEN, D, P, FLOP_PER_ENV_STEP, ROLLOUT, WORKLOAD_NAME, SCALING, EPISODES_PER_SERIES
    if name == "fanout":
        # configs[4]: the router's fan-out (TrainerRouterActor.scala:60-66): 512 tickers x 256 episodes per GPU, the episodes
        # of a ticker on ONE series (series_of_env = repeat(tickers, 256))
        WORKLOAD_NAME, EPISODES_PER_SERIES = "configs[4] (router fan-out, 256 episodes per ticker)", 256
        if args.envs_per_gpu == 4096:
            args.envs_per_gpu = 131072
        if args.rollout_steps:
            ROLLOUT = args.rollout_steps
        return
    if name == "wide":
        WINDOW, HIDDEN, WORKLOAD_NAME = 513, 2048, "configs[3] (wide policy)"
        D = WINDOW + 2
        P = D * HIDDEN + 3 * HIDDEN
        FLOP_PER_ENV_STEP = 6 * D * HIDDEN + 16 * HIDDEN
        ROLLOUT = args.rollout_steps or 16
        if args.envs_per_gpu == 4096:
            args.envs_per_gpu = 65536 // max(1, args.gpus)     # the config names the total batch: strong scaling
            SCALING = "strong"
    elif args.rollout_steps:
        ROLLOUT = args.rollout_steps


def workload_inputs(O, first_env: int, n_envs: int):
    """Prices and episode -> series map of n_envs episodes starting at global episode first_env."""
    if EPISODES_PER_SERIES == 1:
        return O.gen_prices(SEED, first_env, n_envs, SERIES_LEN), np.arange(n_envs, dtype=np.int32)
    first_series = first_env // EPISODES_PER_SERIES
    n_series = -(-(first_env + n_envs) // EPISODES_PER_SERIES) - first_series
    soe = ((first_env + np.arange(n_envs)) // EPISODES_PER_SERIES - first_series).astype(np.int32)
    return O.gen_prices(SEED, first_series, n_series, SERIES_LEN), soe


def bytes_per_env_step(envs_per_gpu: int) -> float:
    """BASELINE.md §3 / SURVEY.md §8(d): 4*(Wn+1) + 40 + 24*P/N_g algorithmic bytes."""
    return 4 * (WINDOW + 1) + 40 + 24.0 * P / envs_per_gpu


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        m = json.load(open(path))
        return dict(hbm_gbs=float(m["hbm_gbs"]), bf16_tflops=float(m["bf16_tflops"]),
                    bf16_sustained=float(m.get("bf16_tflops_sustained", m["bf16_tflops"])), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md).  One
    nvidia-smi process (rank 0) watches every GPU of the job; it is started before the warm-up because its
    own start-up takes longer than a short timed region, and only the samples whose timestamp falls inside
    [mark_begin, mark_end] are kept."""
    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_indices):
        self.gpus = list(gpu_indices)
        self.proc = None
        self.path = None
        self.t0 = self.t1 = None

    def start(self):
        if not self.gpus:
            return
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", ",".join(str(g) for g in self.gpus)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        for _ in range(40):                 # nvidia-smi's own start-up can outlast a short region: wait for one sample
            try:
                if os.path.getsize(self.path) > 0:
                    break
            except OSError:
                break
            time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        try:
            lines = open(self.path).read().splitlines()
            os.unlink(self.path)
        except Exception:
            lines = []
        return self.summarise(lines, self.t0, self.t1, len(self.gpus))

    @staticmethod
    def summarise(lines, t0, t1, n_gpus):
        """nvidia-smi CSV lines (the query in Q) -> the `clocks` object of the bench line; only samples whose
        timestamp lies in [t0, t1] count, unless there are none."""
        import datetime
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        rows = []
        for line in lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(f[2]), float(f[3]), f[5:9]))
            except ValueError:
                continue
        inside = [r for r in rows if t0 is not None and t1 is not None and t0 <= r[0] <= t1 + 0.02]
        window = "timed region"
        if not inside:                      # region shorter than the sampling period / nvidia-smi's start-up
            inside, window = rows, "nearest samples (warm-up before / idle after a timed region shorter than the sampler's start-up)"
        reasons = set()
        for r in inside:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3]):
                if v == "Active":
                    reasons.add(name)
        if inside:
            sm = [r[1] for r in inside]
            out.update(sm_mhz=statistics.median(sm), sm_min_mhz=min(sm), sm_max_mhz=max(r[2] for r in inside),
                       samples=len(inside), gpus=n_gpus, window=window)
        out["reasons"] = sorted(reasons)
        return out


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# =================================================================================================
# reference arm: the CPU implementation of the path (oracle port), all host cores
# =================================================================================================
def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    from oracle import oracle as O
    n_envs = args.envs_per_gpu * max(1, args.gpus)
    threads = O.max_threads()
    cfg = O.Config(wn=WINDOW, hidden=HIDDEN)
    prices, soe = workload_inputs(O, 0, n_envs)
    o = O.Oracle(cfg)
    o.env_create(prices, soe, BUDGET0, SHARES0)
    # size one step as a bounded sample: the first S env-steps of the workload's episodes
    t0 = time.perf_counter()
    o.env_step(1, threads=threads)
    t1 = time.perf_counter() - t0
    budget_s = min(120.0 / max(1, args.steps + args.warmup), 20.0)
    S = int(max(1, min(40, budget_s / max(t1, 1e-6))))
    total_needed = S * (args.steps + args.warmup)
    if total_needed + 1 > o.n_steps_total:
        S = max(1, (o.n_steps_total - 1) // (args.steps + args.warmup))
    for _ in range(args.warmup):
        o.env_step(S, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        o.env_step(S, threads=threads)
    dt = time.perf_counter() - t0
    value = args.steps * S * n_envs / dt
    sample = f"{S} env-steps x {n_envs} episodes per bench step (prefix of the {SERIES_LEN - WINDOW}-step episode)"
    if S * (args.steps + args.warmup) + 1 > o.n_steps_total:
        raise SystemExit("reference arm: episode too short for the requested steps")
    line = dict(
        impl="reference", metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
        ms_per_step=dt / args.steps * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="fp32",
        data="synthetic", config=workload_config(args, n_envs),
        cpu_baseline=dict(value=value, unit=UNIT, cores=threads, kind="port", sample=sample),
        e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
        note="reference tensorflow_scala path cannot run here (no JVM; artefact is darwin-only): timed arm is the "
             "C restatement oracle/qp_oracle.c, batched-SUM, pthreads over all host cores",
    )
    print(json.dumps(line))
    return 0


def workload_config(args, n_envs):
    return dict(workload=f"{WORKLOAD_NAME}: {args.envs_per_gpu} parallel episodes per GPU, hist {WINDOW - 1} (window {WINDOW}), "
                         f"hidden {HIDDEN}, fp32, {SERIES_LEN}-step random-walk series per episode, budget 1000",
                episodes=n_envs, env_steps_per_bench_step=ROLLOUT, series_len=SERIES_LEN,
                bookkeeping="REFERENCE", l2="flushed between bench steps (256 MiB write inside the timed region)",
                episodes_per_series=EPISODES_PER_SERIES,
                parallelism=f"dp{max(1, args.gpus)} (episodes sharded, gradient all-reduce per env-step)" if args.gpus > 1
                else ("single GPU, GEMM-organised step (tcgen05 3xTF32 GEMMs)" if HIDDEN > 256
                      else ("single GPU, fan-out path (window part of layer 1 once per series; three kernels + AdaGrad per env-step)"
                            if EPISODES_PER_SERIES > 1 else "single GPU, persistent cooperative kernel")))


# =================================================================================================
# our arm
# =================================================================================================
def run_ours(args):
    import torch
    import scala_akka_tensorflow_sharetrade_helper_b200 as q
    from oracle import oracle as O

    rank, world, local = dist_env()
    if world != max(1, args.gpus):
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    n_local = args.envs_per_gpu
    n_total = n_local * world
    n_env_steps = ROLLOUT
    # synthetic inputs: this rank's shard of the global series / episodes
    prices_np, soe_np = workload_inputs(O, rank * n_local, n_local)
    prices_pin = torch.from_numpy(prices_np).pin_memory()
    soe_pin = torch.from_numpy(soe_np).pin_memory()
    ports_pin = torch.empty(n_local, dtype=torch.float64).pin_memory()

    pol = q.QDecisionPolicy(q.PolicyConfig(window=WINDOW, hidden=HIDDEN, device=local, seed=SEED))
    if world > 1:
        pol.set_option("rsag", args.rsag)
        pol.dist_init(dist, rank, world)
    stream = torch.cuda.ExternalStream(pol.stream, device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def resident_pass():
        with torch.cuda.stream(stream):
            flush.zero_()                           # L2 flush (126 MB L2 < 256 MiB)
        pol.env_reset()
        pol.env_step(n_env_steps, sync=False)

    def e2e_pass():
        with torch.cuda.stream(stream):
            flush.zero_()
        if n_env_steps == SERIES_LEN - WINDOW:
            pol.train(prices_pin.numpy(), soe_pin.numpy(), BUDGET0, SHARES0, env_id_base=rank * n_local)
        else:   # a slice of the episode: the same three calls qp_train makes, stopped early
            pol.env_create(prices_pin.numpy(), soe_pin.numpy(), BUDGET0, SHARES0, env_id_base=rank * n_local)
            pol.env_step(n_env_steps, sync=False)
            pol.portfolios()

    pol.env_create(prices_np, soe_np, BUDGET0, SHARES0, env_id_base=rank * n_local)

    def timed(fn, k, w):
        # rank 0 watches the clocks of every GPU of the job
        sampler = ClockSampler(([local] if world == 1 else range(world)) if rank == 0 else [])
        sampler.start()
        for _ in range(w):
            fn()
        barrier()
        l0 = pol.launch_count
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.mark_begin()
        t0 = time.perf_counter()
        with torch.cuda.stream(stream):
            ev0.record()
        for _ in range(k):
            fn()
        with torch.cuda.stream(stream):
            ev1.record()
        barrier()
        wall = time.perf_counter() - t0
        sampler.mark_end()
        clocks = sampler.stop()
        ms = ev0.elapsed_time(ev1)
        launches = pol.launch_count - l0
        if dist is not None:
            t = torch.tensor([ms, wall], dtype=torch.float64, device=f"cuda:{local}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1])
        return ms, wall, clocks, launches

    ms, wall, clocks, launches = timed(resident_pass, args.steps, max(args.warmup, 3))
    value = args.steps * n_env_steps * n_total / (ms * 1e-3)
    e2e_ms, e2e_wall, e2e_clocks, _ = timed(e2e_pass, args.steps, 1)
    e2e_value = args.steps * n_env_steps * n_total / (e2e_wall)

    # ---- result check, outside the timed regions: the last e2e pass left a finished run on every rank.
    # REFERENCE book-keeping from (1000, 0) never changes a portfolio's value, and every rank applied
    # the same all-reduced gradient sums, so the replicas' weights must be bit-identical.
    import hashlib
    ports_now = pol.portfolios()
    params_now, _ = pol.get_weights()
    digest = hashlib.blake2b(params_now.tobytes(), digest_size=8).hexdigest()
    digests = [digest]
    problems = []
    if dist is not None:
        digests = [None] * world
        dist.all_gather_object(digests, digest)
    if not (ports_now == BUDGET0).all():
        problems.append(f"rank {rank}: {(ports_now != BUDGET0).sum()} portfolios differ from {BUDGET0}")
    if not np.isfinite(params_now).all():
        problems.append(f"rank {rank}: non-finite weights")
    if len(set(digests)) != 1:
        problems.append(f"replicas' weights differ across ranks: {digests}")
    if dist is not None:
        allp = [None] * world
        dist.all_gather_object(allp, problems)
        problems = [x for pr in allp for x in pr]
    result_check = "ok" if not problems else "FAILED: " + "; ".join(sorted(set(problems)))
    data_plane = dict(fanout_active=int(pol.get_option("fanout_active")),
                      fused_active=int(pol.get_option("fused_active")), ep_tc_active=int(pol.get_option("ep_tc_active")),
                      ep_c2_active=int(pol.get_option("ep_c2_active")), nvls_active=int(pol.get_option("nvls_active")),
                      p2p_active=int(pol.get_option("p2p_active")), fused_step_active=int(pol.get_option("fused_step_active")),
                      tiles_per_sm=-(-((n_local + 31) // 32) // 148), params_digest=digest,
                      share_exchange=("none" if world == 1 else ("reduce-scatter + all-gather" if pol.get_option("rsag_active") else "all-to-all")),
                      gradient_exchange=("none (one GPU)" if world == 1 else
                                         (("in-kernel exchange: one tagged 64-bit multicast store per share (multimem.st through the NVSwitch), "
                                           "no NCCL call in the step") if pol.get_option("nvls_active") else
                                          ("in-kernel all-to-all over NVLink peer memory (tagged 64-bit stores), no NCCL call in the step"
                                           if pol.get_option("p2p_active") else "ncclAllReduce between kernels"))))

    # ---- roofline of the dominant kernel, measured live with CUDA events on the launching stream
    peaks = measured_peaks()
    fp32_uniform_peak = pol.measure_fp32_peak()       # dependency-free FFMA with kernel-uniform multiplicands
    fp32_peak = pol.measure_fp32_outer_peak()         # three-register outer-product FFMA: what a contraction can issue
    n_prof = (200 if n_local <= 32 * 148 else 24) if HIDDEN <= 256 else 8
    fused = bool(pol.get_option("fused_active"))
    wide = bool(pol.get_option("wide"))

    def split_path_us():
        pol.env_reset()
        pol.env_step(20)
        kms = pol.env_step_timed(n_prof)
        return (kms / n_prof * 1e3).tolist()

    fused_step = bool(pol.get_option("fused_step_active"))
    if pol.get_option("fanout_active"):
        # three kernels + AdaGrad per env-step; the row kernel dominates and is bound by fp32 instruction issue
        per_kernel_us = split_path_us()
        names = ["fan_zwin_kernel", "fan_row_kernel", "fan_grad_kernel + apply_kernel"]
        dur = per_kernel_us[1] * 1e-6
        step_us = sum(per_kernel_us)
        # what fan_row_kernel does per (episode, hidden column): four hidden evaluations (2 FMA + max) with their three Q
        # FMAs each (40 flop), and the backward sums (hidden again, h*dz2, three masked sums: 11 flop)
        row_flops = float(n_local) * HIDDEN * 51.0
        row_bytes = float(n_local) * 40.0 + (len(np.unique(soe_np)) * 2.0 + 6.0 * -(-n_local // 256)) * HIDDEN * 4.0
        roofline = dict(
            bound="hbm", kernel="fan_row_kernel (fan-out path: per-episode part of the step; the window part of layer 1 is computed once per series)",
            achieved=row_bytes / dur / 1e9, peak=peaks["hbm_gbs"], unit="GB/s", frac=row_bytes / dur / 1e9 / peaks["hbm_gbs"], traffic=None,
            peak_source=peaks["source"] + " (burst copy)", avg_launch_us=per_kernel_us[1], kernel_us=dict(zip(names, per_kernel_us)),
            binding=dict(bound="fp32-fma", achieved=row_flops / dur / 1e12, peak=fp32_peak, unit="TFLOP/s",
                         frac=row_flops / dur / 1e12 / fp32_peak if fp32_peak else None,
                         note="flops the kernel executes (51 per episode and hidden column), not the 246 800 per env-step of the unfactorised step",
                         peak_source="measured live: qp_measure_fp32_outer_peak (8x8 outer-product FFMA, three register sources)"),
            step=dict(us=step_us, unfactorised_flop_per_env_step=FLOP_PER_ENV_STEP,
                      unfactorised_equivalent_tflops=FLOP_PER_ENV_STEP * n_local / (step_us * 1e-6) / 1e12))
    elif fused_step:
        pol.env_reset()
        pol.env_step(20)
        ep_us = float(pol.env_step_timed(n_prof)[0]) / n_prof * 1e3
        dur = ep_us * 1e-6
        step_bytes = bytes_per_env_step(n_local) * n_local
        step_flops = float(FLOP_PER_ENV_STEP) * n_local
        roofline = dict(
            bound="hbm", kernel="episode_kernel (one fused step) + sum + NCCL all-reduce + AdaGrad, per GPU",
            achieved=step_bytes / dur / 1e9, peak=peaks["hbm_gbs"], unit="GB/s",
            frac=step_bytes / dur / 1e9 / peaks["hbm_gbs"], traffic=None, peak_source=peaks["source"] + " (burst copy)",
            us_per_env_step=ep_us, avg_launch_us=ep_us,
            binding=dict(bound="fp32-fma", achieved=step_flops / dur / 1e12, peak=fp32_peak, unit="TFLOP/s",
                         frac=step_flops / dur / 1e12 / fp32_peak if fp32_peak else None,
                         peak_source="measured live: qp_measure_fp32_outer_peak (8x8 outer-product FFMA, three register sources); the "
                                     f"uniform-operand FFMA stream reaches {fp32_uniform_peak:.1f} TFLOP/s on the same GPU"),
            bytes_per_env_step=bytes_per_env_step(n_local), flop_per_env_step=FLOP_PER_ENV_STEP)
    elif wide:
        pol.env_reset()
        pol.env_step(2)
        ep_us = float(pol.env_step_timed(n_prof)[0]) / n_prof * 1e3
        dur = ep_us * 1e-6
        step_flops = float(FLOP_PER_ENV_STEP) * n_local
        step_bytes = bytes_per_env_step(n_local) * n_local
        tc_on = bool(pol.get_option("tc_active"))
        passes = 3
        # TF32 tensor peak is not in MEASURED_PEAKS.json; it is half the measured bf16 figure by construction
        tf32_peak = peaks["bf16_tflops"] / 2.0
        if tc_on:
            roofline = dict(
                bound="tensor", kernel="tc_gemm_kernel (tcgen05 kind::tf32, 3xTF32) inside the GEMM-organised step",
                achieved=step_flops / dur / 1e12, peak=tf32_peak / passes, unit="TFLOP/s",
                frac=step_flops / dur / 1e12 / (tf32_peak / passes), traffic=None,
                peak_source=peaks["source"] + " bf16 burst / 2 (tf32) / 3 (three MMAs per product for fp32-grade accuracy)",
                us_per_env_step=ep_us, avg_launch_us=ep_us,
                note="achieved counts algorithmic FLOPs of the whole step (GEMMs + row kernel + transposes); the GEMMs "
                     "themselves are L2-bandwidth-bound at 128x128 tiles with fp32 hi/lo operands (DESIGN.md §4)",
                hbm=dict(achieved=step_bytes / dur / 1e9, peak=peaks["hbm_gbs"], unit="GB/s",
                         frac=step_bytes / dur / 1e9 / peaks["hbm_gbs"]),
                fp32_equivalent=dict(achieved=step_flops / dur / 1e12, peak=fp32_peak, unit="TFLOP/s",
                                     frac=step_flops / dur / 1e12 / fp32_peak if fp32_peak else None),
                a_lo_zero=bool(pol.get_option("tc_a_lo_zero")),
                bytes_per_env_step=bytes_per_env_step(n_local), flop_per_env_step=FLOP_PER_ENV_STEP)
        else:
            roofline = dict(
                bound="hbm", kernel="GEMM-organised step, fp32 FFMA GEMMs (G1, E1, E2, G2, apply)", achieved=step_bytes / dur / 1e9,
                peak=peaks["hbm_gbs"], unit="GB/s", frac=step_bytes / dur / 1e9 / peaks["hbm_gbs"], traffic=None,
                peak_source=peaks["source"] + " (burst copy)", avg_launch_us=ep_us, us_per_env_step=ep_us,
                binding=dict(bound="fp32-fma", achieved=step_flops / dur / 1e12, peak=fp32_peak, unit="TFLOP/s",
                             frac=step_flops / dur / 1e12 / fp32_peak if fp32_peak else None,
                             peak_source="measured live: qp_measure_fp32_outer_peak (8x8 outer-product FFMA, three register sources); the "
                                     f"uniform-operand FFMA stream reaches {fp32_uniform_peak:.1f} TFLOP/s on the same GPU"),
                bytes_per_env_step=bytes_per_env_step(n_local), flop_per_env_step=FLOP_PER_ENV_STEP)
    elif fused:
        # one persistent kernel runs every env-step of a launch: its duration / steps is the step time
        pol.env_reset()
        pol.env_step(20)
        ep_us = float(pol.env_step_timed(n_prof)[0]) / n_prof * 1e3
        pol.set_option("fused", 0)
        split_us = split_path_us()
        pol.set_option("fused", 1)
        names = ["step_kernel", "dw1_kernel", "apply_kernel"]
        dur = ep_us * 1e-6
        step_bytes = bytes_per_env_step(n_local) * n_local
        step_flops = float(FLOP_PER_ENV_STEP) * n_local
        ep_tc = bool(pol.get_option("ep_tc_active"))
        hbm = dict(achieved=step_bytes / dur / 1e9, peak=peaks["hbm_gbs"], unit="GB/s",
                   frac=step_bytes / dur / 1e9 / peaks["hbm_gbs"], peak_source=peaks["source"] + " (burst copy)")
        fma = dict(bound="fp32-fma", achieved=step_flops / dur / 1e12, peak=fp32_peak, unit="TFLOP/s",
                   frac=step_flops / dur / 1e12 / fp32_peak if fp32_peak else None,
                   peak_source="measured live: qp_measure_fp32_outer_peak (8x8 outer-product FFMA, three register sources); the "
                               f"uniform-operand FFMA stream reaches {fp32_uniform_peak:.1f} TFLOP/s on the same GPU")
        if ep_tc:
            # contractions on tcgen05 kind::tf32; prices are tf32-exact, W1 / dZ1 are hi+lo pairs: two MMAs per product
            ep_c2 = bool(pol.get_option("ep_c2_active"))
            tc_peak = peaks["bf16_tflops"] / 2.0 / 2.0
            kname = ("episode_c2_kernel (persistent, all env-steps of a launch; CTA pairs split the hidden dimension, 64 episodes per pair; "
                     "contractions on tcgen05, 2 x TF32 per product)" if ep_c2 else
                     "episode_tc_kernel (persistent, all env-steps of a launch; contractions on tcgen05, 2 x TF32 per product)")
            roofline = dict(
                bound="tensor", kernel=kname,
                achieved=step_flops / dur / 1e12, peak=tc_peak, unit="TFLOP/s", frac=step_flops / dur / 1e12 / tc_peak,
                traffic=None, peak_source=peaks["source"] + " bf16 burst / 2 (tf32) / 2 (hi and lo part of W1 / dZ1)",
                avg_launch_us=ep_us * n_prof, us_per_env_step=ep_us,
                note="achieved counts the algorithmic FLOPs of the whole env-step.  The step is a chain of dependent phases inside "
                     "one persistent kernel (operand build, W1^T stream + MMA, TMEM epilogues, element-wise TD work, partial "
                     "store, grid barrier, gradient sum + AdaGrad): see phase_cycles and DESIGN.md §4",
                hbm=hbm, fp32_fma_equivalent=fma,
                bytes_per_env_step=bytes_per_env_step(n_local), flop_per_env_step=FLOP_PER_ENV_STEP,
                three_kernel_path_us=dict(zip(names, split_us)))
            # DRAM and L2 traffic of one env-step: MEASURED, from this round's ncu --set full capture of a launch of the same
            # kernel at this workload (profiles/r2_ncu_traffic.json: dram__bytes_read/write.sum, lts__t_bytes.sum, per env-step)
            single_wave = n_local <= 32 * 148
            try:
                tr = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")))
                key = "episode_c2_kernel" if ep_c2 else "episode_tc_kernel"
                if n_local == 4096 and HIDDEN == 200 and key in tr:
                    roofline["traffic"] = tr[key]["dram_bytes_per_env_step"]
                    roofline["l2"] = dict(bytes_per_step=tr[key]["l2_bytes_per_env_step"], achieved=tr[key]["l2_bytes_per_env_step"] / dur / 1e9,
                                          unit="GB/s", note="lts__t_bytes.sum of the capture / env-steps of the captured launch")
                    roofline["traffic_note"] = tr[key]["source"]
            except Exception as exc:
                roofline["traffic_note"] = f"no ncu traffic summary: {exc}"
            try:
                if not single_wave:
                    raise RuntimeError("several tiles per CTA (waves): the per-phase stamps describe the last wave only; "
                                       "see profiles/ep_tc_phases.py")
                pol.set_option("phase_clocks", 1)
                pol.env_reset()
                pol.env_step(40)
                c = pol.phase_clocks(40)[5:]
                pol.set_option("phase_clocks", 0)
                if ep_c2:
                    order = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 12, 10]
                    labels = ["S0 wait for the last CTA's new weights (A tiles are built meanwhile)", "S1 W1^T stream + A-tile transpose + forward MMA",
                              "S2 z out of TMEM", "S3 hidden layer, Q partial sums, pair exchange, select", "S4 TD target, dZ1 hi/lo",
                              "S5 backward MMA (two halves), dW2 / budget / shares rows, first half of the partial out of TMEM",
                              "S6 rest of the partial out, next price", "store completion", "grid barrier (fence + counter)",
                              "phase B: partial segments summed (16-byte loads from L2)", "NVLink share exchange", "AdaGrad + W1^T publish"]
                else:
                    order = list(range(11))
                    labels = ["S0 A-operand build + tables", "S1 W1^T stream + forward MMA", "S2 z out of TMEM", "S3 Q(s), select, book-keeping, Q(s')",
                              "S4 TD target, dZ1 hi/lo", "S5 backward MMA, dW2, budget/shares rows", "S6 dW1 partial: TMEM -> TMA store, extra rows",
                              "S7 window row + store completion", "grid barrier", "phase B: bulk-copy sum, AdaGrad, W1^T publish"]
                d = np.array([(c[:, b_] - c[:, a_]).astype(np.float64).mean() for a_, b_ in zip(order[:-1], order[1:])])
                tot = float((c[1:, 0] - c[:-1, 0]).mean())
                roofline["phase_cycles"] = dict({k: float(v) for k, v in zip(labels, d)}, step_total=tot,
                                                note="clock64 stamps of CTA 0, mean over 35 env-steps")
            except Exception as exc:  # stamps are a tuning aid, never fatal for the bench line
                roofline["phase_cycles"] = dict(error=str(exc))
            if single_wave:
                pol.set_option("ep_tc", 0)
                pol.env_reset()
                pol.env_step(20)
                roofline["ffma_persistent_kernel_us_per_env_step"] = float(pol.env_step_timed(n_prof)[0]) / n_prof * 1e3
                pol.set_option("ep_tc", 1)
                if ep_c2:
                    pol.set_option("ep_c2", 0)
                    pol.env_reset()
                    pol.env_step(20)
                    roofline["one_tile_per_cta_kernel_us_per_env_step"] = float(pol.env_step_timed(n_prof)[0]) / n_prof * 1e3
                    pol.set_option("ep_c2", 1)
                if world == 1 and n_local == 4096 and HIDDEN == 200:
                    # what the reference's own kind of data gets (quotes in cents are not tf32 values; here the same series
                    # times 1.0371): the pair kernel takes x = tf32(x) + remainder as two episodes, 2 368 episodes per GPU;
                    # beyond that the three-kernel path serves them.  Outside every timed region; the policy is not used again.
                    cent = {}
                    for n_c in (2368, 4096):
                        pol.env_create(prices_np[:n_c] * np.float32(1.0371), soe_np[:n_c], BUDGET0, SHARES0)
                        pol.env_step(20)
                        us_c = float(np.sum(pol.env_step_timed(n_prof))) / n_prof * 1e3
                        cent[str(n_c)] = dict(us_per_env_step=us_c, env_steps_per_s=n_c / us_c * 1e6,
                                              path=("episode_c2_kernel, split form" if pol.get_option("ep_c2_split_active")
                                                    else ("persistent FFMA kernel" if pol.get_option("fused_active") else "three kernels per env-step")))
                    roofline["prices_not_tf32_exact"] = cent
            else:
                roofline["kernel"] += f"; {-(-((n_local + 31) // 32) // 148)} waves per step"
        else:
            roofline = dict(
                bound="hbm", kernel="episode_kernel (persistent, all env-steps of a launch)",
                achieved=hbm["achieved"], peak=hbm["peak"], unit="GB/s", frac=hbm["frac"], traffic=None,
                peak_source=hbm["peak_source"], avg_launch_us=ep_us * n_prof, us_per_env_step=ep_us,
                binding=dict(fma, note="compute-bound at fp32: 227 FLOP/B vs ridge ~11 (SURVEY.md §8d); "
                             "inside the SM the forward sweep is limited by shared-memory wavefronts (DESIGN.md §4)"),
                bytes_per_env_step=bytes_per_env_step(n_local), flop_per_env_step=FLOP_PER_ENV_STEP,
                three_kernel_path_us=dict(zip(names, split_us)))
        # DRAM traffic from the committed ncu --set full capture of a 30-step launch of this kernel
        # (profiles/r1_v2_ncu_episode_kernel_summary.csv: 9.40 MB read + 4.91 MB written), per env-step;
        # far below the algorithmic bytes because windows stay in shared memory and weights/partials in L2
        if n_local == 4096 and HIDDEN == 200 and not ep_tc:
            roofline["traffic"] = (9.397248e6 + 4.906752e6) / 30.0
            roofline["traffic_note"] = "bytes per env-step of 4096 episodes, from profiles/r1_v2_ncu_episode_kernel_summary.csv"
    else:
        per_kernel_us = split_path_us()
        names = ["step_kernel", "dw1_kernel", "apply_kernel(+allreduce)" if world > 1 else "apply_kernel"]
        dom = int(np.argmax(per_kernel_us))
        # algorithmic work of each kernel per launch
        k_bytes = [n_local * (4 * (WINDOW + 1) + 40) + 4 * P, n_local * 4 * (WINDOW + 2), 24.0 * P - 4 * P]
        k_flops = [n_local * (4 * D * HIDDEN + 16 * HIDDEN), n_local * 2 * D * HIDDEN, 6.0 * P]
        dur = per_kernel_us[dom] * 1e-6
        step_us = sum(per_kernel_us)
        roofline = dict(
            bound="hbm", kernel=names[dom], achieved=k_bytes[dom] / dur / 1e9, peak=peaks["hbm_gbs"], unit="GB/s",
            frac=k_bytes[dom] / dur / 1e9 / peaks["hbm_gbs"], traffic=None, peak_source=peaks["source"] + " (burst copy)",
            avg_launch_us=per_kernel_us[dom], kernel_us=dict(zip(names, per_kernel_us)),
            binding=dict(bound="fp32-fma", note="compute-bound at fp32: 227 FLOP/B vs ridge ~11 (SURVEY.md §8d)",
                         achieved=k_flops[dom] / dur / 1e12, peak=fp32_peak, unit="TFLOP/s",
                         frac=k_flops[dom] / dur / 1e12 / fp32_peak if fp32_peak else None,
                         peak_source="measured live: qp_measure_fp32_outer_peak (8x8 outer-product FFMA, three register sources); the "
                                     f"uniform-operand FFMA stream reaches {fp32_uniform_peak:.1f} TFLOP/s on the same GPU"),
            step=dict(us=step_us, hbm_frac=bytes_per_env_step(n_local) * n_local / (step_us * 1e-6) / 1e9 / peaks["hbm_gbs"],
                      fp32_frac=FLOP_PER_ENV_STEP * n_local / (step_us * 1e-6) / 1e12 / fp32_peak if fp32_peak else None,
                      bytes_per_env_step=bytes_per_env_step(n_local), flop_per_env_step=FLOP_PER_ENV_STEP),
        )

    if data_plane["fanout_active"]:
        dtype = "fp32 (FFMA; window part of layer 1 once per series; book-keeping fp64)"
    elif data_plane["ep_tc_active"]:
        dtype = "fp32 (contractions on tcgen05 as 2xTF32: tf32-exact prices x tf32 hi/lo weight pairs, fp32 accumulate in TMEM; element-wise fp32, book-keeping fp64)"
    elif wide and pol.get_option("tc_active"):
        dtype = "fp32 (GEMMs on tcgen05 as 3xTF32 hi/lo splits, fp32 accumulate in TMEM; element-wise fp32, book-keeping fp64)"
    else:
        dtype = "fp32 (FFMA contractions; book-keeping fp64)"
    line = dict(
        metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
        ms_per_step=ms / args.steps, higher_is_better=True, scaling=SCALING, vs_baseline=None, dtype=dtype,
        data="synthetic", config=workload_config(args, n_total), clocks=clocks,
        e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=int(prices_np.nbytes + soe_np.nbytes),
                 d2h_bytes_per_step=int(n_local * 8), ms_per_step=e2e_wall / args.steps * 1e3, clocks=e2e_clocks,
                 api="qp_train (TrainerChildActor.train): pinned host prices in, portfolios out"),
        gpu_launches=int(launches), roofline=roofline, wall_s=wall, result_check=result_check, data_plane=data_plane,
    )

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # bounded sample (~10-30 s of one core): a prefix of the episodes of a subset of the workload,
        # sized by FLOPs so that the wide workload does not take hours
        cfg = O.Config(wn=WINDOW, hidden=HIDDEN)
        budget_flop = 3.0e10
        n_s_envs = int(min(n_local, 4096))
        n_s = int(max(1, min(args.cpu_sample_steps, budget_flop / (FLOP_PER_ENV_STEP * n_s_envs))))
        o = O.Oracle(cfg)
        o.env_create(prices_np[:n_s_envs], soe_np[:n_s_envs], BUDGET0, SHARES0)
        t0 = time.perf_counter()
        o.env_step(n_s, mode=1)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = dict(
            value=n_s * n_s_envs / dt, unit=UNIT, cores=1, kind="port",
            sample=f"first {n_s} env-steps of the first {n_s_envs} episodes of the workload, sequential per-sample "
                   f"updates (the reference's one-mailbox order), oracle/qp_oracle.c, {dt:.1f} s",
            host_cores_available=O.max_threads())
    pol.close()
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs-per-gpu", type=int, default=4096)
    ap.add_argument("--cpu-sample-steps", type=int, default=30)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="config2", choices=["config2", "wide", "fanout"])
    ap.add_argument("--rollout-steps", type=int, default=0, help="env-steps per bench step (default: the whole episode)")
    ap.add_argument("--rsag", type=int, default=-1, help="share exchange: 1 reduce-scatter + all-gather, 0 all-to-all, -1 by world size")
    args = ap.parse_args()
    set_workload(args.workload, args)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
