"""Host-side description of one simulation request and its translation to the
C ABI's `mc_params` (include/mc_b200.h).  Pure Python/NumPy: testable without a GPU.

Defaults are the reference's hard-coded locals (app.py:601-603, app.py:639, app.py:644,
app.py:656)."""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib as L

REF_OMEGA, REF_ALPHA, REF_BETA = 0.00001, 0.10, 0.85          # app.py:601-603
REF_TERMINAL_PCTS = (1, 5, 10, 25, 50, 75, 90, 95, 99)         # app.py:639
REF_PATH_PCTS = (5, 25, 50, 75, 95)                            # app.py:644
REF_TAIL_PCT = 5.0                                             # app.py:655-656
MIN_HISTORY = 100                                              # app.py:594

_SHOCKS = {"bootstrap": L.SHOCK_BOOTSTRAP, "normal": L.SHOCK_NORMAL, "injected": L.SHOCK_INJECTED}
_VARS = {"garch": L.VAR_GARCH, "const": L.VAR_CONST}
_DTYPES = {"float32": L.F32, "float64": L.F64}


@dataclass
class PortfolioSpec:
    """d correlated assets (north_star extension; no counterpart in the reference)."""
    s0: Sequence[float]
    mu: Sequence[float]
    sigma: Sequence[float]
    weights: Sequence[float]
    corr: Optional[np.ndarray] = None      # correlation matrix, or
    chol: Optional[np.ndarray] = None      # its lower Cholesky factor
    dt: float = 1.0 / 252

    def arrays(self):
        s0 = np.ascontiguousarray(self.s0, dtype=np.float64)
        d = s0.shape[0]
        mu = np.ascontiguousarray(np.broadcast_to(np.asarray(self.mu, dtype=np.float64), (d,)))
        sigma = np.ascontiguousarray(np.broadcast_to(np.asarray(self.sigma, dtype=np.float64), (d,)))
        w = np.ascontiguousarray(np.broadcast_to(np.asarray(self.weights, dtype=np.float64), (d,)))
        if self.chol is not None:
            chol = np.ascontiguousarray(np.tril(np.asarray(self.chol, dtype=np.float64)))
        elif self.corr is not None:
            chol = np.ascontiguousarray(np.linalg.cholesky(np.asarray(self.corr, dtype=np.float64)))
        else:
            chol = np.eye(d)
        if chol.shape != (d, d):
            raise ValueError("correlation / Cholesky factor must be d x d")
        if not (1 < d <= L.MC_MAX_ASSETS):
            raise ValueError(f"portfolio needs 2..{L.MC_MAX_ASSETS} assets")
        sigma_step = np.ascontiguousarray(sigma * math.sqrt(self.dt))
        drift_step = np.ascontiguousarray((mu - 0.5 * sigma * sigma) * self.dt)
        return dict(d=d, chol=chol, sigma_step=sigma_step, drift_step=drift_step, s0=s0, weights=w,
                    v0=float(np.dot(w, s0)))


@dataclass
class SimSpec:
    n_paths: int                    # GLOBAL number of paths (n_sims)
    n_steps: int                    # n_days
    s0: float                       # current_price
    shock: str = "bootstrap"        # bootstrap | normal | injected
    var_model: str = "garch"        # garch | const
    var0: float = 0.0               # initial per-step variance (app.py:606)
    omega: float = REF_OMEGA
    alpha: float = REF_ALPHA
    beta: float = REF_BETA
    drift: float = 0.0              # log-space drift per step; the reference has none
    dtype: str = "float32"
    seed: int = 0
    store_paths: bool = False
    fan_chart: bool = False
    fan_chunk_paths: int = 0        # > 0: fan chart by regeneration, this many paths share one buffer at a time
    pcts: Tuple[float, ...] = REF_TERMINAL_PCTS
    path_pcts: Tuple[float, ...] = REF_PATH_PCTS
    tail_pct: Optional[float] = REF_TAIL_PCT
    hist_bins: int = 0
    hist_range: Optional[Tuple[float, float]] = None    # None -> (min, max) like np.histogram
    portfolio: Optional[PortfolioSpec] = None
    full_res_normals: bool = False  # fp32 normal shocks from full 32-bit uniforms (4 per Philox call, |z| <= 6.66) instead of
                                    # the default 8 per call (|z| <= 5.27, P(|z| > 5) 27 % low): MC_FLAG_FULL_RES_NORMALS
    pool_len: int = 0
    _pf_arrays: dict = field(default=None, repr=False, compare=False)

    @staticmethod
    def gbm(n_paths, n_steps, s0, mu, sigma, dt=1.0 / 252, **kw):
        """Mode G: x += (mu - sigma^2/2) dt + sigma sqrt(dt) z."""
        return SimSpec(n_paths=n_paths, n_steps=n_steps, s0=s0, shock=kw.pop("shock", "normal"),
                       var_model="const", var0=sigma * sigma * dt,
                       drift=(mu - 0.5 * sigma * sigma) * dt, **kw)

    def validate(self):
        if self.shock not in _SHOCKS:
            raise ValueError(f"shock must be one of {sorted(_SHOCKS)}")
        if self.var_model not in _VARS:
            raise ValueError(f"var_model must be one of {sorted(_VARS)}")
        if self.dtype not in _DTYPES:
            raise ValueError("dtype must be 'float32' or 'float64'")
        if self.n_paths < 1 or self.n_steps < 1:
            raise ValueError("n_paths and n_steps must be positive")
        if self.fan_chart and not self.store_paths:
            raise ValueError("fan_chart needs store_paths=True")
        if len(self.pcts) > L.MC_MAX_PCTS or len(self.path_pcts) > L.MC_MAX_PATH_PCTS:
            raise ValueError("too many percentiles")
        if list(self.pcts) != sorted(self.pcts) or list(self.path_pcts) != sorted(self.path_pcts):
            raise ValueError("percentiles must be ascending")
        if self.tail_pct is not None and self.pcts and float(self.tail_pct) not in [float(p) for p in self.pcts]:
            raise ValueError("tail_pct must be one of pcts")
        if self.portfolio is not None and (self.shock == "bootstrap" or self.var_model != "const"):
            raise ValueError("portfolio mode needs normal/injected shocks and var_model='const'")
        if self.full_res_normals and self.portfolio is not None and self.dtype == "float32":
            raise ValueError("full_res_normals: single-asset kernels only (float64 portfolios already draw from 32-bit uniforms)")


def shard_range(n_global: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous global path-id range of `rank`: (offset, count).  The Philox counter is the
    global id, so every path's value is independent of `world`."""
    base, rem = divmod(int(n_global), int(world))
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def make_params(spec: SimSpec, path_offset: int = 0, n_local: Optional[int] = None,
                paths_ld: int = 0, n_ranks: int = 1, paths_volatile: bool = False, pass_group: int = 0) -> L.McParams:
    # a dashboard (and a bench loop) asks for the same job again and again: the struct is kept under everything it was
    # built from and handed out as a copy (16 -> 3 us per call).  Portfolio specs carry arrays and are always rebuilt.
    if spec.portfolio is not None:
        return _build_params(spec, path_offset, n_local, paths_ld, n_ranks, paths_volatile, pass_group)
    hr = spec.hist_range
    key = (spec.n_paths, spec.n_steps, spec.s0, spec.shock, spec.var_model, spec.var0, spec.omega, spec.alpha, spec.beta,
           spec.drift, spec.dtype, spec.seed, spec.store_paths, spec.fan_chart, tuple(spec.pcts), tuple(spec.path_pcts),
           spec.tail_pct, spec.hist_bins, None if hr is None else (hr[0], hr[1]), spec.full_res_normals, spec.pool_len,
           path_offset, n_local, paths_ld, n_ranks, paths_volatile, pass_group)
    proto = _PARAMS_CACHE.get(key)
    if proto is None:
        proto = _build_params(spec, path_offset, n_local, paths_ld, n_ranks, paths_volatile, pass_group)
        if len(_PARAMS_CACHE) >= 64:
            _PARAMS_CACHE.clear()
        _PARAMS_CACHE[key] = proto
    return L.McParams.from_buffer_copy(proto)


_PARAMS_CACHE: dict = {}


def _build_params(spec, path_offset, n_local, paths_ld, n_ranks, paths_volatile, pass_group) -> L.McParams:
    spec.validate()
    p = L.McParams()
    p.struct_size = C.sizeof(L.McParams)
    p.version = L.MC_ABI_VERSION
    p.n_paths = spec.n_paths if n_local is None else n_local
    p.path_offset = path_offset
    p.n_paths_global = spec.n_paths
    p.n_ranks = int(n_ranks)
    p.pass_group = int(pass_group)
    p.paths_ld = paths_ld
    p.n_steps = spec.n_steps
    p.n_assets = 1
    p.shock = _SHOCKS[spec.shock]
    p.var_model = _VARS[spec.var_model]
    p.dtype = _DTYPES[spec.dtype]
    flags = 0
    if spec.store_paths:
        flags |= L.FLAG_STORE_PATHS
    if spec.fan_chart and spec.path_pcts:
        flags |= L.FLAG_FAN_CHART
    if not spec.pcts:
        flags |= L.FLAG_NO_TERMINAL_SELECT
    if paths_volatile:
        flags |= L.FLAG_PATHS_VOLATILE
    if spec.full_res_normals and spec.shock == "normal":
        flags |= L.FLAG_FULL_RES_NORMALS
    p.flags = flags
    p.seed = spec.seed & 0xFFFFFFFFFFFFFFFF
    p.s0 = spec.s0
    p.var0, p.omega, p.alpha, p.beta, p.drift = spec.var0, spec.omega, spec.alpha, spec.beta, spec.drift
    p.pool_len = spec.pool_len
    p.hist_bins = spec.hist_bins
    if spec.hist_range is not None:
        p.hist_lo, p.hist_hi = float(spec.hist_range[0]), float(spec.hist_range[1])
        if not p.hist_lo < p.hist_hi:
            raise ValueError("hist_range must satisfy lo < hi")
    else:
        p.hist_lo, p.hist_hi = 0.0, 0.0
    p.n_pcts = len(spec.pcts)
    p.n_path_pcts = len(spec.path_pcts) if spec.fan_chart else 0
    for i, v in enumerate(spec.pcts):
        p.pcts[i] = float(v)
    for i, v in enumerate(spec.path_pcts):
        p.path_pcts[i] = float(v)
    p.tail_pct = float(spec.tail_pct) if (spec.tail_pct is not None and spec.pcts) else -1.0
    if spec.portfolio is not None:
        arr = spec._pf_arrays or spec.portfolio.arrays()
        spec._pf_arrays = arr
        p.n_assets = arr["d"]
        p.s0 = arr["v0"]
    return p


def make_portfolio(spec: SimSpec) -> Optional[L.McPortfolio]:
    if spec.portfolio is None:
        return None
    arr = spec._pf_arrays or spec.portfolio.arrays()
    spec._pf_arrays = arr            # keeps the arrays alive while the struct points into them
    pf = L.McPortfolio()
    dp = C.POINTER(C.c_double)
    pf.h_chol = arr["chol"].ctypes.data_as(dp)
    pf.h_sigma_step = arr["sigma_step"].ctypes.data_as(dp)
    pf.h_drift_step = arr["drift_step"].ctypes.data_as(dp)
    pf.h_s0 = arr["s0"].ctypes.data_as(dp)
    pf.h_weights = arr["weights"].ctypes.data_as(dp)
    return pf
