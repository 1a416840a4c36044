"""GARCH(1,1) calibration.  Nothing like it exists in the reference (omega/alpha/beta are the
constants of app.py:601-603): parity is unpinned by the reference.  The CPU restatement is anchored on an
independent optimiser and on parameter recovery; the CUDA kernel is compared with the restatement."""
import os

import numpy as np
import pytest

from oracle import garch_fit, mc_numpy as port, ref_loader

# fp64 on both sides; the kernel contracts omega + alpha x + beta h into FMAs and sums z^2 in another order
OBJ_RTOL = 1e-9
PARAM_ATOL = 2e-4         # the objective is flat near its minimum: neighbouring fine-grid candidates can swap


def test_grid_refinement_is_not_beaten_by_nelder_mead():
    from scipy.optimize import minimize
    r = garch_fit.simulate_garch(1500, 2e-6, 0.08, 0.90, seed=4)
    fit = garch_fit.calibrate(r)
    z2 = (r / fit["scale"]) ** 2

    def f(x):
        a, b = x
        if a < 0 or b < 0 or a + b > garch_fit.PERSIST_MAX or a + b < garch_fit.PERSIST_MIN or a > garch_fit.ALPHA_MAX:
            return 1e300
        return float(garch_fit.objective(z2, a, b))
    starts = [(0.05, 0.90), (0.10, 0.85), (0.2, 0.6), (fit["alpha"], fit["beta"])]
    best = min(minimize(f, s, method="Nelder-Mead", options=dict(xatol=1e-7, fatol=1e-10, maxiter=800)).fun for s in starts)
    assert fit["objective"] <= best + 1e-7 * abs(best)
    assert abs(float(garch_fit.objective(z2, fit["alpha"], fit["beta"])) - fit["objective"]) < 1e-9 * abs(best)


def test_recovers_the_parameters_of_simulated_garch_data():
    r = garch_fit.simulate_garch(20000, 1e-5, 0.10, 0.85, seed=1)      # the reference's constants, app.py:601-603
    fit = garch_fit.calibrate(r)
    assert abs(fit["alpha"] - 0.10) < 0.02 and abs(fit["beta"] - 0.85) < 0.03
    assert abs(fit["omega"] / 1e-5 - 1) < 0.35
    long_run = fit["omega"] / (1 - fit["alpha"] - fit["beta"])
    assert abs(long_run / np.mean(r ** 2) - 1) < 1e-9                  # variance targeting


def _histories():
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_inputs_from_close.npz"))
    close, lengths = g["close"].copy(), g["lengths"].copy()
    # give two histories real volatility clustering: prices from simulated GARCH returns
    for b, (a, be) in ((0, (0.10, 0.85)), (5, (0.05, 0.93))):
        n = int(lengths[b])
        close[b, :n] = 50.0 * np.exp(np.cumsum(garch_fit.simulate_garch(n, 2e-4 * (1 - a - be), a, be, seed=40 + b)))
    return close, lengths


@pytest.mark.gpu
def test_cuda_calibration_against_the_restatement(engine):
    close, lengths = _histories()
    pool, info = engine.prepare_inputs(close, lengths, 199, "float64")
    _, fit = engine.calibrate_garch(pool, info, levels=8)
    assert list(fit["valid"]) == [1, 1, 1, 0, 0, 1]                    # < 100 returns: skipped, like app.py:594-595
    for b in (0, 1, 2, 5):
        exp = garch_fit.calibrate(port.history_inputs(close[b, :lengths[b]])["returns"])
        assert abs(fit["objective"][b] - exp["objective"]) <= OBJ_RTOL * abs(exp["objective"])
        assert abs(fit["alpha"][b] - exp["alpha"]) <= PARAM_ATOL and abs(fit["beta"][b] - exp["beta"]) <= PARAM_ATOL
        assert abs(fit["omega"][b] - exp["omega"]) <= 5e-3 * exp["omega"] + 1e-12
        assert 0 <= fit["alpha"][b] <= 0.5 and fit["beta"][b] >= 0 and fit["alpha"][b] + fit["beta"][b] <= 0.9995 + 1e-9
    # the clustered histories are told apart from the iid ones
    assert fit["alpha"][0] > 0.04 and fit["alpha"][5] > 0.02


@pytest.mark.gpu
def test_calibrated_run_batched_equals_one_run_per_history():
    from montecarlo_b200 import calibrate_garch, garch_monte_carlo_from_close
    close, lengths = _histories()
    fits = calibrate_garch(close, lengths=lengths)
    loop = garch_monte_carlo_from_close(close, 1500, 25, lengths=lengths, seed=8, calibrate=True, batched=False)
    batch = garch_monte_carlo_from_close(close, 1500, 25, lengths=lengths, seed=8, calibrate=True, batched=True)
    plain = garch_monte_carlo_from_close(close, 1500, 25, lengths=lengths, seed=8)
    for b, (a, c) in enumerate(zip(loop, batch)):
        assert (a is None) == (c is None) == (fits[b] is None)
        if a is None:
            continue
        assert a["garch"] == c["garch"] == fits[b]
        assert np.array_equal(a["paths"], c["paths"]) and a["percentiles"] == c["percentiles"]
        for k in a["path_percentiles"]:
            assert np.array_equal(a["path_percentiles"][k], c["path_percentiles"][k])
        assert "garch" not in plain[b] and not np.array_equal(plain[b]["paths"], c["paths"])   # opt-in
    one = garch_monte_carlo_from_close(close[0], 1500, 25, seed=8, calibrate=True)
    assert np.array_equal(one["paths"], batch[0]["paths"]) and one["garch"] == batch[0]["garch"]
