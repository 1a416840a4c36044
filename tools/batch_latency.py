"""Many price histories per call: 100 tickers x the dashboard's default 1000 paths x 20 days, from
raw close prices.  Times (a) the reference's way on the host (pandas producers + NumPy port of the routine, one
ticker after another), (b) the CUDA drop-in called once per ticker, (c) mc_prepare_inputs + mc_run_batch."""
import os, sys, time, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import mc_numpy as port, ref_loader
from montecarlo_b200 import garch_monte_carlo_from_close

def med(f, reps):
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); f(); torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    return statistics.median(ts)

for n_tickers, n_sims, n_days in ((100, 1000, 20), (100, 5000, 60), (500, 1000, 20), (100, 100_000, 60)):
    close = np.stack([ref_loader.synthetic_ohlcv(1260, 500 + b, 20.0 + b)["Close"].values for b in range(n_tickers)])
    keep = dict(return_paths="device") if n_sims <= 5000 else dict(return_paths=False, fan_chart=False)
    def host():
        np.random.seed(1)
        for b in range(min(n_tickers, 10)):
            inp = port.history_inputs(close[b])
            port.garch_monte_carlo_port(inp["returns"], inp["rvol_last"], min(n_sims, 5000), n_days, inp["s0"])
    t_host = med(host, 2) * n_tickers / min(n_tickers, 10) * (n_sims / min(n_sims, 5000))
    loop = lambda: garch_monte_carlo_from_close(close, n_sims, n_days, seed=3, batched=False, **keep)
    batch = lambda: garch_monte_carlo_from_close(close, n_sims, n_days, seed=3, batched=True, **keep)
    loop(); batch()
    t_loop, t_batch = med(loop, 5), med(batch, 5)
    r = batch()[0]["timing"]
    print(f"{n_tickers} tickers x {n_sims} x {n_days}: host port ~{t_host:.0f} ms (extrapolated from 10 tickers"
          f"{', 5000 paths' if n_sims > 5000 else ''}), one run per ticker {t_loop:.2f} ms, "
          f"batched {t_batch:.2f} ms (device {r['total_ms']:.3f} ms, sim {r['sim_ms']:.3f} ms, {r['launches']} launches)")
