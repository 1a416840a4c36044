"""`mc_histogram_static` (include/mc_b200.h; csrc/mc_static.cu): the fixed-bin histogram for callers that must not
allocate -- its scratch is a file-scope __device__ array, not cudaMalloc'd handle buffers.  No counterpart in the
reference.  Checked through the C ABI against the existing entry point
`mc_histogram` (whose counts the suite pins to np.histogram), exactly, on calls made back to back: the shared array
keeps the previous call's contents, so the second call must not see the first one's counts or edges.

Collected last on purpose: a failure here hides nothing behind it."""
import ctypes as C

import numpy as np
import pytest
import torch

from montecarlo_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _static_hist(lib, x, bins, lo, hi, device=0):
    """One call through the C ABI.  The caller (this test) owns input and output; the entry point allocates nothing."""
    counts = torch.full((bins,), -1, dtype=torch.int64, device=x.device)        # poisoned: every bin must be written
    err = C.create_string_buffer(256)
    dtype = L.F64 if x.dtype == torch.float64 else L.F32
    stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    rc = lib.mc_histogram_static(device, C.c_void_p(x.data_ptr()), dtype, x.numel(), bins, lo, hi,
                                 C.c_void_p(counts.data_ptr()), stream, err, len(err))
    return rc, counts, err.value.decode()


def _data(seed, n, dtype, lo, hi, bins):
    rs = np.random.RandomState(seed)
    x = 100 * np.exp(rs.normal(size=n) * 0.3)
    edges = np.linspace(lo, hi, bins + 1)
    k = min(bins + 1, n // 4)
    x[:k] = edges[:k]                                   # values exactly on edges
    x[k:2 * k - 1] = np.nextafter(edges[1:k], 0)        # and one ulp below them
    x[-6:] = [lo, hi, lo - 1e-9, hi + 1e-9, -5.0, 1e9]  # the ends, and values outside the range
    return torch.tensor(x, dtype=dtype, device="cuda:0")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_static_histogram_equals_mc_histogram_on_calls_in_a_row(engine, dtype):
    lib = engine.lib
    # three calls back to back on one stream: many bins, then FEWER bins over another range (stale counts / edges of the
    # first call lie behind them in the shared array), then the maximum size
    cases = [(1, 300_007, 2048, 40.0, 250.0), (2, 50_001, 77, 10.0, 500.0), (3, 1_000_003, L.MC_STATIC_HIST_MAX_BINS, 55.5, 180.25)]
    xs = [_data(seed, n, dtype, lo, hi, bins) for seed, n, bins, lo, hi in cases]
    got = []
    for x, (_, _, bins, lo, hi) in zip(xs, cases):
        rc, counts, msg = _static_hist(lib, x, bins, lo, hi)
        assert rc == 0, msg
        got.append(counts)
    torch.cuda.synchronize()
    for x, counts, (_, n, bins, lo, hi) in zip(xs, got, cases):
        ref = engine.histogram(x, bins, lo, hi)                               # the existing entry point (mc_histogram)
        assert np.array_equal(counts.cpu().numpy(), ref), (bins, lo, hi)      # integer counts: exact
        assert np.array_equal(ref, np.histogram(x.cpu().numpy().astype(np.float64), bins=bins, range=(lo, hi))[0])
    # the same call again gives the same answer (nothing accumulates across calls)
    rc, again, msg = _static_hist(lib, xs[0], cases[0][2], cases[0][3], cases[0][4])
    assert rc == 0 and torch.equal(again, got[0])


def test_static_histogram_empty_input_and_limits(engine):
    lib = engine.lib
    x = torch.empty(0, dtype=torch.float32, device="cuda:0")
    rc, counts, msg = _static_hist(lib, x, 16, 0.0, 1.0)
    assert rc == 0 and int(counts.abs().sum()) == 0
    x = torch.rand(1000, dtype=torch.float32, device="cuda:0")
    rc, _, msg = _static_hist(lib, x, L.MC_STATIC_HIST_MAX_BINS + 1, 0.0, 1.0)
    assert rc == -1 and "at most 4096" in msg and "mc_histogram" in msg       # beyond the static array: refused, clearly
    rc, _, msg = _static_hist(lib, x, 0, 0.0, 1.0)
    assert rc == -1 and msg
    rc, _, msg = _static_hist(lib, x, 8, 1.0, 1.0)
    assert rc == -1 and "lo < hi" in msg
    rc, _, msg = _static_hist(lib, x, 8, 0.0, 1.0, device=63)
    assert rc != 0 and msg
    assert torch.cuda.current_device() == 0
    rc, counts, msg = _static_hist(lib, x, 8, 0.0, 1.0)                          # and it still works afterwards
    assert rc == 0 and int(counts.sum()) == 1000
