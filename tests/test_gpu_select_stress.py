"""Randomised sweep of the exact selection (app.py:638-645 semantics: np.percentile, method='linear') through
`mc_percentiles`: sizes from 1 to 400k, one to forty rows, value families that stress the order-preserving keys
(ties, signed zeros, denormals, infinities, huge dynamic range), random percentile lists.  Bit-exact or fail."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _values(rs, family, shape):
    if family == "lognormal":
        return 100 * np.exp(0.3 * rs.standard_normal(shape))
    if family == "heavy":
        return rs.standard_t(1.5, shape) * 1e3
    if family == "few_distinct":
        return rs.choice(rs.normal(size=5), size=shape)
    if family == "integers":
        return rs.randint(-50, 50, size=shape).astype(np.float64)
    if family == "zeros_and_denormals":
        x = rs.choice([0.0, -0.0, 1e-45, -1e-45, 1e-38, 5e-324, 1.0], size=shape)
        return x
    if family == "wide_range":
        return np.ldexp(rs.uniform(1, 2, shape), rs.randint(-120, 120, size=shape)) * rs.choice([-1, 1], size=shape)
    if family == "with_inf":
        x = rs.normal(size=shape)
        x[rs.uniform(size=shape) < 0.01] = np.inf
        x[rs.uniform(size=shape) < 0.01] = -np.inf
        return x
    if family == "sorted":
        return np.sort(rs.normal(size=shape), axis=-1)
    raise AssertionError(family)


FAMILIES = ["lognormal", "heavy", "few_distinct", "integers", "zeros_and_denormals", "wide_range", "with_inf", "sorted"]


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("seed", range(6))
def test_random_selection_problems(engine, dtype, seed):
    rs = np.random.RandomState(1000 + seed)
    for family in FAMILIES:
        rows = int(rs.choice([1, 1, 2, 7, 40]))
        n = int(rs.choice([1, 2, 3, 33, 1000, 4097, 50_000, 400_000 // rows]))
        x = _values(rs, family, (rows, n))
        xt = torch.tensor(x, dtype=dtype, device=engine.device)
        k = int(rs.randint(1, 12))
        pcts = sorted(set(np.round(rs.uniform(0, 100, k), int(rs.choice([0, 1, 3]))).tolist()) | ({0.0, 100.0} if seed % 2 else set()))
        got = engine.percentiles(xt, pcts)
        with np.errstate(all="ignore"):
            ref = np.percentile(xt.cpu().numpy().astype(np.float64), pcts, axis=1).T
        same = (got == ref) | (np.isnan(got) & np.isnan(ref))
        assert same.all(), (family, rows, n, pcts, got[~same][:3], ref[~same][:3])


@pytest.mark.parametrize("seed", range(4))
def test_table_pass_edge_cases(engine, seed):
    """<= 5 percentiles in fp32 take the table-driven pass 1 (select_pass1_lut_f32_kernel): rows long enough for many
    batches per thread, unaligned rows (scalar loads), lengths that are no multiple of 4, constant rows (no usable
    table scale: every value takes the exact path), rows whose outer brackets are infinite, heavy ties at a bracket
    bound, and a row that is one huge outlier away from constant."""
    rs = np.random.RandomState(7000 + seed)
    n = int(rs.choice([262_147, 1_000_003, 2_097_152 + 5]))
    rows = []
    rows.append(100 * np.exp(0.02 * rs.standard_normal(n)))                       # narrow lognormal (early fan-chart rows)
    rows.append(np.full(n, 100.0))                                                 # constant (row 0 of the paths)
    x = rs.normal(size=n); x[rs.uniform(size=n) < 0.08] = np.inf; x[rs.uniform(size=n) < 0.08] = -np.inf
    rows.append(x)                                                                 # p5 / p95 are infinite
    rows.append(np.round(rs.normal(size=n), 1))                                    # ~60 distinct values: ties everywhere
    x = np.full(n, 3.25); x[rs.randint(n)] = 1e30; x[rs.randint(n)] = -1e30
    rows.append(x)
    rows.append(rs.standard_t(2.0, n) * 1e-3)                                      # heavy tails around zero, both signs
    full = torch.tensor(np.stack(rows), dtype=torch.float32, device=engine.device)
    for view in (full, full[:, 1:], full[:, 3:n - 2]):                             # aligned / unaligned starts and ends
        for pcts in ([5, 25, 50, 75, 95], [50.0], [0.0, 100.0], [1, 99], [33.3, 66.6, 99.99]):
            got = engine.percentiles(view, pcts)
            with np.errstate(all="ignore"):
                ref = np.percentile(view.cpu().numpy().astype(np.float64), pcts, axis=1).T
            same = (got == ref) | (np.isnan(got) & np.isnan(ref))
            assert same.all(), (seed, n, tuple(view.shape), pcts, np.argwhere(~same)[:4].tolist(), got[~same][:3], ref[~same][:3])
