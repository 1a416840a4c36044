"""The multi-rank reduction protocol on CPU: montecarlo_b200.engine.run_stages (the
product's orchestration loop) driven over gloo, world_size 2 and 3, with the NumPy
stage model standing in for the device kernels.  Checked against NumPy on the
concatenated data (the reductions of app.py:636-656)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.select_model import NumpyStageBackend, keys_of, values_of

PCTS = (1, 5, 10, 25, 50, 75, 90, 95, 99)


def _data(dtype, n=20011, seed=5):
    rs = np.random.RandomState(seed)
    x = (100 * np.exp(rs.standard_t(4, n) * 0.1)).astype(dtype)
    x[:50] = x[50]            # ties
    x[100:110] *= -1          # negative values (portfolio with short legs)
    return x


def _expect(x, s0):
    xd = x.astype(np.float64)
    pct = {p: np.percentile(xd, p) for p in PCTS}
    tail = xd[xd <= pct[5]]
    return dict(pct=pct, mean=xd.mean(), std=xd.std(), n_above=int((xd > s0).sum()), tail_count=tail.size,
                tail_sum=tail.sum(), hist=np.histogram(xd, bins=64, range=(50.0, 200.0))[0])


def _check(res, exp):
    for p in PCTS:
        assert res["pct"][p] == exp["pct"][p], p          # exact order statistics + NumPy lerp
    assert res["n_above"] == exp["n_above"] and res["tail_count"] == exp["tail_count"]
    assert abs(res["mean"] - exp["mean"]) < 1e-12 * abs(exp["mean"])
    assert abs(res["std"] - exp["std"]) < 1e-10 * exp["std"]
    assert abs(res["tail_sum"] - exp["tail_sum"]) < 1e-12 * abs(exp["tail_sum"])
    assert np.array_equal(np.asarray(res["hist"]), exp["hist"])


def test_key_roundtrip_is_order_preserving():
    for dtype in (np.float32, np.float64):
        x = np.array([-np.inf, -3.5, -0.0, 0.0, 1e-30, 1.0, 2.5, np.inf], dtype=dtype)
        k = keys_of(x)
        assert np.all(np.diff(k.astype(np.float64)) >= 0) and np.all(k[[0, 1]] < k[3])
        assert np.array_equal(values_of(k, dtype), x.astype(np.float64))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_single_rank(dtype):
    from montecarlo_b200.engine import run_stages
    x = _data(dtype)
    b = NumpyStageBackend(x, x.size, 100.0, PCTS, 5, 64, (50.0, 200.0))
    run_stages(b, None)
    _check(b.result, _expect(x, 100.0))


def _worker(rank, world, port, dtype_name, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from montecarlo_b200.engine import run_stages
        from montecarlo_b200.spec import shard_range
        x = _data(np.dtype(dtype_name).type)
        off, cnt = shard_range(x.size, rank, world)
        b = NumpyStageBackend(x[off:off + cnt], x.size, 100.0, PCTS, 5, 64, (50.0, 200.0))
        run_stages(b, dist.group.WORLD)
        q.put((rank, b.result))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,dtype_name", [(2, "float32"), (3, "float32"), (2, "float64")])
def test_sharded_over_gloo(world, dtype_name):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, dtype_name, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x = _data(np.dtype(dtype_name).type)
    exp = _expect(x, 100.0)
    for _, res in results:          # every rank holds the same merged result
        _check(res, exp)
