"""montecarlo_b200 -- B200-native Monte Carlo path simulation behind the call signature of
raasawin/montecarlo's `garch_monte_carlo` (reference app.py:586-657).

Host code is Python (PyTorch tensors for I/O); compute is hand-written sm_100a CUDA in
`libmc_b200.so`, reached through the C ABI declared in include/mc_b200.h.  There is no
CPU fallback: the compute entry points raise if the library or a CUDA device is missing.
"""
from .api import (garch_monte_carlo, garch_monte_carlo_from_close, calibrate_garch, simulate_gbm,
                  simulate_portfolio, bootstrap_pool)
from .spec import PortfolioSpec, SimSpec, make_params, shard_range

__all__ = ["garch_monte_carlo", "garch_monte_carlo_from_close", "calibrate_garch", "simulate_gbm", "simulate_portfolio", "bootstrap_pool",
           "SimSpec", "PortfolioSpec", "make_params", "shard_range", "Engine", "get_engine", "Pipeline"]


def __getattr__(name):
    if name in ("Engine", "get_engine", "run_stages", "RunOutput", "Pipeline"):
        from . import engine
        return getattr(engine, name)
    raise AttributeError(name)
