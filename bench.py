#!/usr/bin/env python
"""bench.py -- path-steps/sec of the Monte Carlo hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c4|c5|c3|ref] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one whole simulation job (all paths x all time steps + every reduction the
reference's result dict holds) on synthetic parameters.  Default workload is BASELINE.json
configs[1] ("c2": single-asset GBM, 10M paths x 252 steps fp32, terminal stats + percentiles)
per GPU -- weak scaling: with N ranks the job is N x 10M paths, sharded by global path id and
merged with one small all-gather per stage.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEED = 0x123456789ABCDEF0
S0, MU, SIGMA, DT = 100.0, 0.05, 0.20, 1.0 / 252

# paths per GPU, steps, description
WORKLOADS = {
    "c2": dict(paths=10_000_000, steps=252, desc="single-asset GBM 10M paths x 252 steps fp32, terminal stats + "
                                                 "9 percentiles + CVaR + 2048-bin histogram (BASELINE configs[1])"),
    "c2full": dict(paths=10_000_000, steps=252, desc="config-2 shape with full-resolution normals (MC_FLAG_FULL_RES_NORMALS: 4 per "
                                                     "Philox call from 32-bit uniforms, MUFU sin/cos -- SURVEY 8d's own work definition)"),
    "c4": dict(paths=12_500_000, steps=2520, desc="single-asset GBM 2520 steps fp32, 12.5M paths per GPU "
                                                  "(= 100M over 8 GPUs, BASELINE configs[3])"),
    "c5": dict(paths=4_194_304, steps=1024, desc="GBM 4M paths x 1024 steps fp32, full paths stored to HBM + "
                                                 "per-step fan chart (BASELINE configs[4])"),
    "c3": dict(paths=1_000_000, steps=252, desc="16-asset Cholesky portfolio 1M paths x 252 steps fp32, "
                                                "VaR/CVaR (BASELINE configs[2])"),
    "c4fan": dict(paths=12_500_000, steps=2520, desc="config-4 shape WITH the per-step fan chart (app.py:643-645): 126 GB "
                                                     "of paths per GPU never exist; 8 GB chunks are simulated three times "
                                                     "(SURVEY 8f N1)"),
    "ref": dict(paths=10_000_000, steps=252, desc="reference model (bootstrap + GARCH(1,1), app.py:618-633) "
                                                  "10M paths x 252 steps fp32"),
    "ref64": dict(paths=10_000_000, steps=252, desc="reference model in the reference's own precision (float64, app.py:612-633) "
                                                    "10M paths x 252 steps, terminal statistics, no paths kept"),
    "c1": dict(paths=10_000, steps=252, desc="BASELINE configs[0]: the reference's whole routine (app.py:586-657) at 10k paths x "
                                             "252 steps, float64, paths matrix + per-step fan chart + terminal statistics"),
}
FP64_WORKLOADS = ("ref64", "c1")
# algorithmic work per path-step of the fused kernels -> path-steps per SM-clock
# (fused GBM: 16 ALU lane-ops / 128 lanes = 2 MUFU / 16 lanes = 1/8 clk per path-step).  Kept as `survey_8d`.
STEPS_PER_SM_CLK = {"c2": 8.0, "c2full": 8.0, "c4": 8.0, "ref": 7.5, "c3": 0.33, "ref64": 4.8, "c1": 4.8}
# The headline roofline: what the BINDING resource of the kernel AS BUILT allows, from the instruction mix of its
# hot loop (tools/sass_loops.py -> profiles/r02/sass_hot_loops.txt) at the issue rates MEASURED on sm_100a
# (tools/microbench*.cu, per SM per clock: IMAD.WIDE / IMAD.HI 31.6, IMAD 63, FFMA2 / FMUL2 62.3 -- all on the heavy
# FMA pipe, which they hold one after the other --, DFMA 62.7, 128 thread-instructions issued).
#   fused GBM      per normal: 2.12 IMAD.WIDE + 0.94 IMAD + 0.75 FFMA2 = 0.0940 clk of heavy pipe; 9.8 instr = 0.0766 clk of issue
#   Mode R fp32    per step:   3.40 IMAD.WIDE + 1.20 IMAD + 0.40 FFMA2 = 0.1330 clk;              17.6 instr = 0.1375 clk of issue
#   portfolio d=16 per step:   76 FFMA2 + 8 FMUL2 + 36 IMAD.WIDE + 5 IMAD = 2.566 clk;             288 instr = 2.25 clk of issue
#   Mode R fp64    per step:   13 DFMA/DMUL/DADD = 0.207 clk, 7.25 IMAD.WIDE/HI = 0.229 clk;      50.5 instr = 0.395 clk of issue
#   GBM full-res   per normal: 2 MUFU + 0.5 I2F = 0.156 clk of the 16-lane XU pipe; 4.25 IMAD.WIDE = 0.134 clk; 19 instr = 0.148 clk
BOUND_PER_SM_CLK = {"c2": 1 / 0.0940, "c2full": 1 / 0.156, "c4": 1 / 0.0940, "ref": 1 / 0.1375, "c3": 1 / 2.566, "ref64": 1 / 0.395, "c1": 1 / 0.395}
BOUND_PIPE = {"c2": "heavy FMA pipe: 2.12 IMAD.WIDE + 0.94 IMAD + 0.75 FFMA2 per normal",
              "c2full": "XU pipe: 2 MUFU (lg2 + sqrt per pair, sin, cos) + 0.5 I2F per normal at 16 lanes (issue and IMAD.WIDE within 15 %)",
              "c4": "heavy FMA pipe: 2.12 IMAD.WIDE + 0.94 IMAD + 0.75 FFMA2 per normal",
              "ref": "issue: 17.6 instructions per path-step (heavy FMA pipe close behind: 3.4 IMAD.WIDE + 1.2 IMAD)",
              "c3": "heavy FMA pipe: 76 FFMA2 + 8 FMUL2 + 36 IMAD.WIDE per 16-asset step",
              "ref64": "issue: 50.5 instructions per path-step (13 on the FP64 pipe, 7.25 IMAD.WIDE/HI)",
              "c1": "issue: 50.5 instructions per path-step; at 10k paths the job is launch-latency bound"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def wait_first(self, timeout_s):
        t_end = time.perf_counter() + timeout_s
        while self.proc is not None and not self.rows and time.perf_counter() < t_end:
            time.sleep(0.02)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, pw, reasons = [], [], [], set()
        for ts, line in self.rows:
            if t0 is not None and not (t0 <= ts <= t1 + 0.1):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


# --------------------------------------------------------------------------- this repo's arm
def hist_range(t):
    w = 6.0 * SIGMA * math.sqrt(t * DT)
    return (S0 * math.exp(-w), S0 * math.exp(w))


_HISTORY = {}


def synthetic_history():
    """The caller's inputs as they sit in host memory before the call: returns, last rvol_20, and the DataFrame
    app.py:1240 passes (built once: making it is the caller's business, not the path's)."""
    if not _HISTORY:
        import numpy as np
        rs = np.random.RandomState(0)
        _HISTORY["returns"], _HISTORY["rvol"] = rs.normal(0.0, 0.02, 1250), 0.02 * math.sqrt(252)
    return _HISTORY["returns"], _HISTORY["rvol"]


def synthetic_frame():
    if "df" not in _HISTORY:
        import pandas as pd
        returns, rvol = synthetic_history()
        _HISTORY["df"] = pd.DataFrame({"log_ret": returns, "rvol_20": rvol})
    return _HISTORY["df"]


def make_spec(workload, world):
    import numpy as np
    from montecarlo_b200 import PortfolioSpec, SimSpec
    from montecarlo_b200.api import bootstrap_pool
    w = WORKLOADS[workload]
    n, t = w["paths"] * world, w["steps"]
    pool = None
    if workload in ("c2", "c4", "c2full"):
        spec = SimSpec.gbm(n, t, S0, MU, SIGMA, DT, seed=SEED, hist_bins=2048, hist_range=hist_range(t),
                           full_res_normals=workload == "c2full")
    elif workload == "c5":
        spec = SimSpec.gbm(n, t, S0, MU, SIGMA, DT, seed=SEED, hist_bins=2048, hist_range=hist_range(t),
                           store_paths=True, fan_chart=True)
    elif workload == "c4fan":
        from montecarlo_b200.api import _fan_plan
        _, _, chunk = _fan_plan(False, True, n, t, "float32")
        spec = SimSpec.gbm(n, t, S0, MU, SIGMA, DT, seed=SEED, hist_bins=2048, hist_range=hist_range(t),
                           store_paths=True, fan_chart=True, fan_chunk_paths=chunk)
    elif workload in FP64_WORKLOADS:
        returns, rvol = synthetic_history()
        pool = bootstrap_pool(returns)
        full = workload == "c1"
        spec = SimSpec(n_paths=n, n_steps=t, s0=S0, var0=rvol ** 2 / 252, seed=SEED, dtype="float64", store_paths=full,
                       fan_chart=full)
    elif workload == "c3":
        d = 16
        corr = np.full((d, d), 0.3) + 0.7 * np.eye(d)
        pf = PortfolioSpec(s0=np.full(d, S0), mu=MU, sigma=0.15 + 0.01 * np.arange(d), weights=np.full(d, 1 / d),
                           corr=corr, dt=DT)
        spec = SimSpec(n_paths=n, n_steps=t, s0=S0, shock="normal", var_model="const", seed=SEED, portfolio=pf)
    else:
        returns, rvol = synthetic_history()
        pool = bootstrap_pool(returns)
        spec = SimSpec(n_paths=n, n_steps=t, s0=S0, var0=rvol ** 2 / 252, seed=SEED, hist_bins=2048, hist_range=hist_range(t))
    return spec, pool


def api_call(workload, world, group):
    """The same job through the public API with HOST inputs (what a user of the reference calls)."""
    import numpy as np
    import montecarlo_b200 as mc
    w = WORKLOADS[workload]
    n, t = w["paths"] * world, w["steps"]
    if workload in ("c2", "c4", "c2full"):
        return mc.simulate_gbm(n, t, S0, MU, SIGMA, dt=DT, seed=SEED, hist_bins=2048, group=group,
                               full_resolution=workload == "c2full")
    if workload == "c5":
        return mc.simulate_gbm(n, t, S0, MU, SIGMA, dt=DT, seed=SEED, hist_bins=2048, return_paths="device", group=group)
    if workload == "c4fan":
        return mc.simulate_gbm(n, t, S0, MU, SIGMA, dt=DT, seed=SEED, hist_bins=2048, return_paths=False, fan_chart=True,
                               group=group)
    if workload == "c3":
        d = 16
        corr = np.full((d, d), 0.3) + 0.7 * np.eye(d)
        return mc.simulate_portfolio(n, t, np.full(d, S0), MU, 0.15 + 0.01 * np.arange(d), np.full(d, 1 / d), corr,
                                     dt=DT, seed=SEED, group=group)
    df = synthetic_frame()
    if workload == "c1":        # the drop-in exactly as app.py:1240 calls it (host NumPy paths, fan chart, float64)
        return mc.garch_monte_carlo(df, n, t, S0, seed=SEED, group=group)
    if workload == "ref64":
        return mc.garch_monte_carlo(df, n, t, S0, seed=SEED, return_paths=False, fan_chart=False, group=group)
    return mc.garch_monte_carlo(df, n, t, S0, seed=SEED, dtype="float32", return_paths=False, fan_chart=False,
                                hist_bins=2048, hist_range=hist_range(t), group=group)


def io_bytes(workload, spec, pool):
    """Bytes copied per step: H2D = parameter struct + select plans (+ bootstrap pool);
    D2H = result block + histogram (+ fan chart)."""
    import ctypes
    from montecarlo_b200 import _lib
    plan_bytes = 592                      # sizeof(mc::SelectPlan)
    h2d = ctypes.sizeof(_lib.McParams) + plan_bytes * (2 if spec.fan_chart else 1)
    itemsize = 8 if spec.dtype == "float64" else 4
    if pool is not None:
        h2d += pool.size * itemsize
    if spec.portfolio is not None:
        d = len(spec.portfolio.s0)
        h2d += 4 * (d * (d + 1) // 2 + 2 * d)
    d2h = 504 + 8 * spec.hist_bins + (8 * len(spec.path_pcts) * (spec.n_steps + 1) if spec.fan_chart else 0)
    if workload == "c1":        # the drop-in hands the paths matrix back as host NumPy, like the reference
        d2h += 8 * spec.n_paths * (spec.n_steps + 1)
    return h2d, d2h


def run_workload(workload, steps, warmup, world, rank, group, device, sample_clocks=True):
    import torch
    import torch.distributed as dist
    from montecarlo_b200.engine import get_engine
    eng = get_engine(device)
    if os.environ.get("MC_BENCH_PILOT"):          # tuning sweeps only (tools/gpu_sweep.sh)
        eng.lib.mc_debug_set_pilot(int(os.environ["MC_BENCH_PILOT"]))
    if os.environ.get("MC_BENCH_WAVES") or os.environ.get("MC_BENCH_NBLOCK"):
        eng.debug_set_tuning(waves=int(os.environ.get("MC_BENCH_WAVES", 0)), nblock=int(os.environ.get("MC_BENCH_NBLOCK", 0)))
    spec, pool = make_spec(workload, world)
    pool_dev = eng.to_device(pool, torch.float64 if spec.dtype == "float64" else torch.float32) if pool is not None else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=eng.device)       # > 126 MB L2
    n_local = WORKLOADS[workload]["paths"]
    t = spec.n_steps

    def barrier():
        torch.cuda.synchronize(eng.device)
        if world > 1:
            dist.barrier(group)
            torch.cuda.synchronize(eng.device)

    def one_step():
        out = eng.simulate(spec, pool=pool_dev, group=group)
        return out

    # nvidia-smi takes a while to come up (NVML init stalls CUDA calls meanwhile): start it and wait for its first
    # sample BEFORE the warm-up, so that the GPU goes from the warm-up straight into the timed region
    sampler = ClockSampler(device) if (rank == 0 and sample_clocks and not os.environ.get("MC_BENCH_NO_CLOCKS")) else None
    if sampler:
        sampler.start()
        sampler.wait_first(10.0)
    last = None
    for _ in range(warmup):
        # `last` stays referenced across the call exactly as in the timed loop, so the caching allocator already
        # owns both result buffers that then alternate (else the second timed step pays a cudaMalloc)
        last = one_step()
        api_call(workload, world, group)
    barrier()
    t0 = time.perf_counter()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    sim_ms, launches = [], 0
    for a, b in ev:
        flush.zero_()                     # L2 flush between timed iterations (untimed)
        torch.cuda.synchronize(eng.device)
        if world > 1 and os.environ.get("MC_BENCH_STEP_BARRIER", "1") != "0":
            dist.barrier(group)           # ranks enter the step together: a step ends at its (device-side) exchanges,
            torch.cuda.synchronize(eng.device)      # so a rank that started late would bill its lateness to the others
        a.record()
        last = one_step()
        b.record()
        sim_ms.append(last.sim_ms)
        launches += last.launches
    barrier()
    t1 = time.perf_counter()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    if os.environ.get("MC_BENCH_DEBUG") and rank == 0:
        print("per-step ms:", [round(a.elapsed_time(b), 3) for a, b in ev], "sim:", [round(v, 3) for v in sim_ms], file=sys.stderr)
    # end to end: public API, host inputs, wall clock around the call (it synchronises)
    barrier()
    e2e_s = 0.0
    for _ in range(steps):
        flush.zero_()
        torch.cuda.synchronize(eng.device)
        if world > 1:
            dist.barrier(group)
        ta = time.perf_counter()
        api_call(workload, world, group)
        e2e_s += time.perf_counter() - ta
    clocks = sampler.stop(t0, time.perf_counter()) if sampler else None      # device-timed + end-to-end loops
    tm = torch.tensor([dev_ms, e2e_s * 1e3, sum(sim_ms)], dtype=torch.float64, device=eng.device)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX, group=group)
    dev_ms, e2e_ms, sim_total = [float(v) for v in tm.tolist()]
    total_steps = n_local * world * t * steps
    h2d, d2h = io_bytes(workload, spec, pool)
    return dict(value=total_steps / (dev_ms * 1e-3), ms_per_step=dev_ms / steps,
                e2e=dict(value=total_steps / (e2e_ms * 1e-3), unit="path-steps/s", h2d_bytes_per_step=h2d,
                         d2h_bytes_per_step=d2h, ms_per_step=e2e_ms / steps),
                sim_ms=sim_total / steps, gpu_launches=launches, clocks=clocks, n_local=n_local, n_steps=t,
                summary={k: float(v) for k, v in last.summary().items() if not isinstance(v, dict)},
                p5=last.pct.get("p5"))


def measured_traffic(workload):
    """DRAM bytes per launch of the workload's dominant kernel from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            return json.load(fh).get(workload, {}).get("dram_bytes")
    except (OSError, ValueError):
        return None


def roofline_for(workload, r, n_sm, clocks, peaks, peaks_kind):
    out = _roofline_for(workload, r, n_sm, clocks, peaks, peaks_kind)
    out["traffic"] = measured_traffic(workload)
    return out


def _roofline_for(workload, r, n_sm, clocks, peaks, peaks_kind):
    n, t, sim_s = r["n_local"], r["n_steps"], r["sim_ms"] * 1e-3
    if workload == "c4fan":
        # whole job: every stored price is written twice (pilot stage, first narrowing pass) and read once (the
        # later passes work from the ~10 % of values pass 1 set aside; the pilot reads 32768 columns)
        bytes_ = 3 * 4.0 * n * (t + 1)
        ach = bytes_ / (r["ms_per_step"] * 1e-3) / 1e9
        peak = float(peaks["hbm_gbs"])
        return {"bound": "hbm", "kernel": "whole job: 2 x sim32_kernel<STORE,V=4> + select_pass1_lut_f32 per 8 GB chunk, pass 2 from side lists",
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                "peak_source": f"{peaks_kind} copy bandwidth", "algorithmic_bytes_per_job": bytes_}
    if workload == "c5":
        bytes_ = 4.0 * n * (t + 1)            # 4 B written per stored price
        ach = bytes_ / sim_s / 1e9
        peak = float(peaks["hbm_gbs"])
        return {"bound": "hbm", "kernel": "sim32_kernel<NORMAL,CONST,STORE,V=4> (st.global.v4, time-major)", "achieved": ach, "peak": peak,
                "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": f"{peaks_kind} copy bandwidth",
                "algorithmic_bytes_per_launch": bytes_}
    mhz_max = float(peaks.get("sm_max_mhz", 1965.0))
    mhz = (clocks or {}).get("sm_mhz") or mhz_max
    ach = n * t / sim_s
    peak = BOUND_PER_SM_CLK[workload] * n_sm * mhz * 1e6
    survey = STEPS_PER_SM_CLK[workload] * n_sm * mhz * 1e6
    frac = ach / peak
    out = {"bound": "fp32 issue / heavy FMA pipe (not hbm, not tensor)" if workload not in FP64_WORKLOADS else "issue / FP64 pipe",
           "kernel": {"c3": "portfolio_kernel<float,16>", "ref64": "sim_kernel<double,BOOTSTRAP,GARCH>",
                      "c1": "sim_kernel<double,BOOTSTRAP,GARCH,STORE>"}.get(workload, "sim32_kernel<...,SINK> (fused RNG + step + histograms)"),
           "achieved": ach / 1e9, "peak": peak / 1e9, "unit": "Gpath-steps/s", "frac": min(frac, 1.0), "traffic": None,
           "peak_source": f"{BOUND_PIPE[workload]}: {BOUND_PER_SM_CLK[workload]:.2f} path-steps/SM-clk x {n_sm} SMs x {mhz:.0f} MHz "
                          "(instruction mix of the hot loop from the shipped SASS, issue rates measured on sm_100a; median SM "
                          "clock sampled during the timed region)",
           "survey_8d": {"peak": survey / 1e9, "frac": ach / survey,
                         "note": f"SURVEY 8d's model of the algorithm ({STEPS_PER_SM_CLK[workload]} path-steps/SM-clk): it assumed 4 "
                                 "normals per Philox call and 2 MUFU per normal; the kernels draw 8 per call and read sin/cos "
                                 "from a shared-memory table, so this fraction can exceed 1 and is not the headline"}}
    if frac > 1.0:
        out["frac_unclamped"] = frac        # the bound is wrong if this ever shows up
    return out


def callers_leg():
    """The callers either side of the path, at the reference's own sizes: the dashboard's default call
    through the drop-in, and 100 tickers from raw close prices in one batched call.  Wall clock, median of 7."""
    import numpy as np
    import pandas as pd
    import torch
    from montecarlo_b200 import garch_monte_carlo, garch_monte_carlo_from_close

    def med(f, reps=7):
        f()
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = f()
            ts.append((time.perf_counter() - t0) * 1e3)
        return statistics.median(ts), out
    rs = np.random.RandomState(0)
    df = pd.DataFrame({"log_ret": rs.normal(0.0, 0.02, 1250), "rvol_20": 0.02 * math.sqrt(252)})
    t_call, res = med(lambda: garch_monte_carlo(df, 1000, 20, 100.0, seed=1))
    close = 50.0 * np.exp(np.cumsum(rs.normal(3e-4, 0.017, (100, 1260)), axis=1))
    t_batch, many = med(lambda: garch_monte_carlo_from_close(close, 1000, 20, seed=1, return_paths="device"))
    return {"drop_in_1000x20": {"ms_per_call": t_call, "device_ms": res["timing"]["total_ms"],
                                "launches": res["timing"]["launches"], "note": "app.py:1240 with the sliders' defaults"},
            "from_close_100_tickers_1000x20": {"ms_per_call": t_batch, "device_ms": many[0]["timing"]["total_ms"],
                                               "launches": many[0]["timing"]["launches"],
                                               "note": "mc_prepare_inputs + mc_run_batch, all tickers in one simulation launch"}}


# --------------------------------------------------------------------------- CPU arms
# kind of loop the oracle port times for a workload that the reference itself cannot run (it has no GBM / portfolio)
CPU_KIND = {"c2": "gbm", "c2full": "gbm", "c4": "gbm", "c5": "gbm", "c3": "gbm", "ref": "garch", "c4fan": "gbm"}
# workloads that ARE the reference's routine: timed with the unmodified function when it travelled (oracle/_ref)
REFERENCE_WORKLOADS = ("c1", "ref64")


def pipelined_leg(workload, steps, warmup, device):
    """NOT the headline: the same job submitted through `montecarlo_b200.Pipeline(2)` -- two engines on two streams taking
    jobs in turn, so the host side of a call and the closing kernels of one job run under the simulation of the next.
    Wall clock between two device synchronisations around all `steps` jobs; jobs overlap, so there is no L2 flush
    between them (the job reads no input: its shocks come from counters)."""
    import torch
    from montecarlo_b200.engine import Pipeline
    spec, pool = make_spec(workload, 1)
    pipe = Pipeline(2, device)
    try:
        for _ in pipe.map([spec] * warmup, pool=pool):
            pass
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        last = None
        for last in pipe.map([spec] * steps, pool=pool):
            pass
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    finally:
        pipe.close()
    w = WORKLOADS[workload]
    return {"workload": w["desc"], "in_flight": 2, "steps": steps, "ms_per_step": dt * 1e3 / steps,
            "value": w["paths"] * w["steps"] * steps / dt, "unit": "path-steps/s", "timing": "wall clock, all jobs",
            "p50": last.pct.get("p50") if last is not None else None,
            "note": "serving mode (montecarlo_b200.Pipeline): every job's result is fetched by the host; not the headline"}


def _reference_function():
    """The unmodified `garch_monte_carlo` (app.py:586-657): from /root/reference in the build container, from the
    AST-extracted copy under oracle/_ref/ (git-ignored, written by __graft_entry__.build) on the GPU box."""
    from oracle import ref_loader
    path = ref_loader.reference_path()
    return (ref_loader.load_reference_function(path), path) if path else (None, None)


def _reference_chunk(args):
    n, t, seed = args
    import numpy as np
    import pandas as pd
    fn, _ = _reference_function()
    df = synthetic_frame()
    np.random.seed(seed)
    return float(fn(df, n, t, S0)["expected_return"])


def time_reference(n, t, processes=1, repeats=1):
    """(best seconds, path-steps/s) of `processes` concurrent calls of the unmodified reference, n paths each."""
    jobs = [(n, t, 1000 + i) for i in range(processes)]
    best = float("inf")
    if processes == 1:
        for _ in range(repeats):
            t0 = time.perf_counter()
            _reference_chunk(jobs[0])
            best = min(best, time.perf_counter() - t0)
    else:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(processes) as pool:
            pool.map(_reference_chunk, [(200, 2, 0)] * processes)
            for _ in range(repeats):
                t0 = time.perf_counter()
                pool.map(_reference_chunk, jobs)
                best = min(best, time.perf_counter() - t0)
    return best, n * t * processes / best


def cpu_baseline_leg(workload, target_seconds=8.0):
    """The CPU beside the GPU number, one core (NumPy's elementwise ops, choice and percentile are single-threaded).
    c1 / ref64: the UNMODIFIED reference function at BASELINE configs[0]'s size (10k x 252, all of app.py:586-657
    incl. the paths matrix and the axis=1 percentiles that are 61-72 % of its time).  Other workloads: the oracle
    port's loop for the model the reference does not have (GBM), terminal statistics only -- no paths matrix, no
    fan chart, so ~10x faster per path-step than the reference's own routine."""
    from oracle import cpu_baseline as cb
    if workload in REFERENCE_WORKLOADS and _reference_function()[0] is not None:
        n, t = 10_000, 252
        sec, rate = time_reference(n, t, 1, repeats=3)
        total, runs = sec * 3, 3
        while total < target_seconds and runs < 40:
            s2, r2 = time_reference(n, t, 1)
            total += s2; runs += 1
            if s2 < sec:
                sec, rate = s2, r2
        return {"value": rate, "unit": "path-steps/s", "cores": 1, "kind": "reference",
                "sample": f"unmodified garch_monte_carlo (app.py:586-657, AST-extracted), {n} paths x {t} steps per call "
                          f"(BASELINE configs[0]), best of {runs} calls, {sec:.3f} s each"}
    kind = "reference" if workload in REFERENCE_WORKLOADS else CPU_KIND[workload]
    t = min(WORKLOADS[workload]["steps"], 252)
    n = 200_000 if kind != "reference" else 10_000
    sec, rate, cores = cb.time_cpu(kind, n, t, processes=1)
    if sec < target_seconds / 4 and kind != "reference":     # grow the sample towards ~target_seconds
        n = int(min(2_000_000, n * target_seconds / max(sec, 1e-3) / 2))
        sec, rate, cores = cb.time_cpu(kind, n, t, processes=1)
    what = ("full routine of app.py:612-657 incl. paths matrix and fan chart" if kind == "reference" else
            f"{kind} loop of app.py:618-633 + terminal statistics ONLY (no paths matrix, no axis=1 percentiles: the "
            "reference's own routine is ~10x slower per path-step)")
    return {"value": rate, "unit": "path-steps/s", "cores": cores, "kind": "port",
            "sample": f"oracle NumPy port ({what}), {n} paths x {t} steps, {sec:.2f} s, single process"}


def reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation of the path on all host cores.  c1 / ref64: the
    unmodified function (oracle/_ref) in one process per core; the GBM workloads: the oracle port's GBM loop."""
    if rank != 0:
        return
    from oracle import cpu_baseline as cb
    workload = args.workload
    cores = cb.host_cores()
    procs = max(1, min(cores, 64))
    real = workload in REFERENCE_WORKLOADS and _reference_function()[0] is not None
    t = min(WORKLOADS[workload]["steps"], 252)
    total_s, total_steps = 0.0, 0
    if real:
        per_proc = 10_000
        for _ in range(args.warmup):
            time_reference(500, 8, procs)
        for _ in range(args.steps):
            sec, _ = time_reference(per_proc, t, procs)
            total_s += sec
            total_steps += per_proc * procs * t
        kind, what = "reference", "unmodified garch_monte_carlo (app.py:586-657, AST-extracted)"
    else:
        kind = "reference" if workload in REFERENCE_WORKLOADS else CPU_KIND[workload]
        per_proc = 50_000 if kind != "reference" else 10_000
        for _ in range(args.warmup):
            cb.time_cpu(kind, 2_000 * procs, 8, processes=procs)
        for _ in range(args.steps):
            sec, rate, _ = cb.time_cpu(kind, per_proc * procs, t, processes=procs)
            total_s += sec
            total_steps += per_proc * procs * t
        what = (f"oracle NumPy port ({kind}" + (": loop + terminal statistics only, no paths matrix / fan chart -- ~10x faster "
                                                "per path-step than the reference's own routine)" if kind != "reference" else ")"))
        kind = "port"
    value = total_steps / total_s
    line = {"impl": "reference", "metric": "path-steps/sec", "value": value, "unit": "path-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{workload}: {WORKLOADS[workload]['desc']}",
                       "sample": f"{per_proc * procs} paths x {t} steps per step"},
            "cpu_baseline": {"value": value, "unit": "path-steps/s", "cores": procs, "kind": kind,
                             "sample": f"{what}, {per_proc} paths x {t} steps in each of {procs} processes per step"},
            "e2e": {"value": value, "unit": "path-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the other BASELINE configs / CPU leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: montecarlo_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        group = dist.group.WORLD
    warmup = max(3, args.warmup)
    r = run_workload(args.workload, args.steps, warmup, world, rank, group, local_rank)
    peaks, peaks_kind = measured_peaks()
    n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
    line = {"metric": "path-steps/sec", "value": r["value"], "unit": "path-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {WORKLOADS[args.workload]['desc']}",
                       "paths_per_gpu": r["n_local"], "n_steps": r["n_steps"], "sharding": f"paths/{world} ranks",
                       "l2": "256 MB buffer written between timed iterations", "seed": hex(SEED)},
            "e2e": r["e2e"], "gpu_launches": r["gpu_launches"], "clocks": r["clocks"],
            "roofline": roofline_for(args.workload, r, n_sm, r["clocks"], peaks, peaks_kind),
            "sim_kernel_ms": r["sim_ms"], "result": r["summary"]}
    if args.workload in FP64_WORKLOADS:
        line["dtype"] = "f64"
    if not args.no_extras:
        # the other BASELINE configs ride along.  One GPU: all of them + the CPU leg + the callers.  Several ranks: the
        # configuration BASELINE quotes for 2/4/8 GPUs (config 4: 12.5M paths x 2520 steps per rank = 100M at 8) with
        # and without the per-step fan chart -- every rank runs them, rank 0 reports.
        if world == 1 and rank == 0:
            line["cpu_baseline"] = cpu_baseline_leg(args.workload)
        names = ("c4", "c5", "c3", "ref", "ref64", "c1", "c2full", "c4fan", "c2") if world == 1 else ("c4", "c4fan")
        extras = {}
        for wl in names:
            if wl == args.workload:
                continue
            try:
                x = run_workload(wl, 3 if wl != "c4fan" else 2, 3 if wl != "c4fan" else 1, world, rank, group, local_rank,
                                 sample_clocks=True)
                extras[wl] = {"workload": WORKLOADS[wl]["desc"], "value": x["value"], "ms_per_step": x["ms_per_step"],
                              "n_gpus": world, "e2e": x["e2e"]["value"], "sim_kernel_ms": x["sim_ms"], "clocks": x["clocks"],
                              "gpu_launches_per_step": x["gpu_launches"] / (3 if wl != "c4fan" else 2),
                              "roofline": roofline_for(wl, x, n_sm, x["clocks"], peaks, peaks_kind)}
                if world == 1 and wl in REFERENCE_WORKLOADS:
                    extras[wl]["cpu_baseline"] = cpu_baseline_leg(wl, target_seconds=4.0)
            except Exception as exc:      # an extra must never lose the headline line
                extras[wl] = {"error": repr(exc)[:200]}
            torch.cuda.empty_cache()
        if world == 1:
            try:
                extras["callers"] = callers_leg()
            except Exception as exc:
                extras["callers"] = {"error": repr(exc)[:200]}
            try:
                extras["c2_two_in_flight"] = pipelined_leg("c2", 40, 6, local_rank)
            except Exception as exc:
                extras["c2_two_in_flight"] = {"error": repr(exc)[:200]}
        line["extras"] = extras
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
