"""Timed CPU arm for bench.py (TEST/BENCH INFRASTRUCTURE, see oracle/__init__).

The reference is pure Python + NumPy and cannot travel to the GPU box, so the CPU
baseline is the oracle PORT of its loop (oracle/mc_numpy.py; kind = "port"), timed with
NumPy's own generator exactly as the reference draws (one [N] vector per step, app.py:620).
`workload` names follow BASELINE.json's configs."""
from __future__ import annotations

import os
import time

import numpy as np

from . import mc_numpy as port


def _gbm_chunk(args):
    n, t, seed = args
    rs = np.random.RandomState(seed)
    drift, var = port.gbm_step_params(0.05, 0.20, 1 / 252)
    vol = np.sqrt(var)
    s = np.full(n, 100.0)
    for _ in range(t):                      # loop shape of app.py:618-633, constant variance
        z = rs.standard_normal(n)
        s = s * np.exp(z * vol + drift)
    st = port.terminal_statistics(s, 100.0)
    return st["expected_return"]


def _garch_chunk(args):
    n, t, seed = args
    rs = np.random.RandomState(seed)
    returns = np.random.RandomState(0).normal(0, 0.02, 1250)
    # terminal statistics only (no paths matrix, no fan chart): the same work the GPU arm does
    pool = port.bootstrap_pool(returns)
    var = np.full(n, port.initial_variance(0.02 * np.sqrt(252)))
    s = np.full(n, 100.0)
    for _ in range(t):
        r = rs.choice(pool, size=n) * np.sqrt(var)
        s = s * np.exp(r)
        var = port.REF_OMEGA + port.REF_ALPHA * (r * r) + port.REF_BETA * var
    return port.terminal_statistics(s, 100.0)["expected_return"]


def _reference_shape_chunk(args):
    """The reference's full routine incl. paths matrix and fan chart (app.py:612-657)."""
    n, t, seed = args
    returns = np.random.RandomState(0).normal(0, 0.02, 1250)
    res = port.garch_monte_carlo_port(returns, 0.02 * np.sqrt(252), n, t, 100.0, rng=np.random.RandomState(seed))
    return res["expected_return"]


WORKLOADS = {"gbm": _gbm_chunk, "garch": _garch_chunk, "reference": _reference_shape_chunk}


def time_cpu(workload: str, n_paths: int, n_steps: int, processes: int = 1, repeats: int = 1):
    """Returns (best seconds, path-steps/s, cores used).  With processes > 1 the paths are split
    over a process pool (NumPy's elementwise ops are single-threaded)."""
    fn = WORKLOADS[workload]
    processes = max(1, int(processes))
    per = max(1, n_paths // processes)
    jobs = [(per, n_steps, 1000 + i) for i in range(processes)]
    best = float("inf")
    if processes == 1:
        for _ in range(repeats):
            t0 = time.perf_counter()
            fn(jobs[0])
            best = min(best, time.perf_counter() - t0)
    else:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(processes) as pool:
            pool.map(_gbm_chunk, [(1000, 2, 0)] * processes)      # spin the workers up
            for _ in range(repeats):
                t0 = time.perf_counter()
                pool.map(fn, jobs)
                best = min(best, time.perf_counter() - t0)
    total = per * processes * n_steps
    return best, total / best, processes


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1
