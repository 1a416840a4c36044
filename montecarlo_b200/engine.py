"""Engine: owns one `mc_handle` (one GPU) and runs simulation requests on it.

Single GPU: one `mc_run` call.  Path-sharded multi-GPU (one process per GPU,
`torch.distributed` over NCCL): the staged C entry points with ONE small
all-gather per stage (`run_stages`) -- no path data ever crosses GPUs.
The orchestration is backend-agnostic so the same loop is exercised on CPU with the
gloo backend in tests."""
from __future__ import annotations

import contextlib
import ctypes as C
import functools
import threading
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib as L
from .spec import SimSpec, make_params, make_portfolio, shard_range


@dataclass
class RunOutput:
    spec: SimSpec
    n: int
    n_above: int
    mean: float
    std: float
    min: float
    max: float
    pct: dict                       # {'p5': value, ...}
    pct_bounds: dict                # {'p5': (lo order stat, hi order stat)}
    tail_count: int
    tail_sum: float
    hist: Optional[np.ndarray]      # int64[bins]
    hist_range: Optional[tuple]
    fan: Optional[dict]             # {'p5': float64[T+1], ...}
    terminal: Optional[torch.Tensor]   # this rank's final prices (device)
    paths: Optional[torch.Tensor]      # this rank's [(T+1), n_local] (device) or None
    sim_ms: float
    total_ms: float
    launches: int
    path_offset: int = 0

    @property
    def center(self) -> float:
        return self.spec._pf_arrays["v0"] if self.spec.portfolio is not None else self.spec.s0

    def summary(self) -> dict:
        """The scalar keys of the reference's result dict (app.py:652-656)."""
        s0 = self.center
        out = {
            "percentiles": dict(self.pct),
            "expected_return": (self.mean / s0 - 1) * 100,
            "expected_vol": self.std / s0 * 100,
            "prob_profit": self.n_above / self.n * 100,
        }
        tp = self.spec.tail_pct
        if tp is not None and self.pct:
            key = f"p{tp:g}"
            out[f"var_{100 - tp:g}"] = (s0 - self.pct[key]) / s0 * 100
            tail_mean = self.tail_sum / self.tail_count if self.tail_count else float("nan")
            out[f"cvar_{100 - tp:g}"] = (s0 - tail_mean) / s0 * 100
        return out


def run_stages(backend, group=None):
    """Drive a stage backend to completion.  `backend` offers begin(), stage_words(),
    stage_view(words), gather_view(total_words), next(gathered, n_ranks) -> done."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if group is not None else 1
    backend.begin()
    done = False
    while not done:
        words = backend.stage_words()
        local = backend.stage_view(words)
        if world > 1:
            gathered = backend.gather_view(world * words)
            dist.all_gather_into_tensor(gathered, local, group=group)
        else:
            gathered = local
        done = backend.next(gathered, world)
    return world


class _CudaStageBackend:

    def __init__(self, engine, params, pf, pool, shocks, paths, terminal, world):
        self.e, self.params, self.pf = engine, params, pf
        self.pool, self.shocks, self.paths, self.terminal = pool, shocks, paths, terminal
        cap = int(engine.lib.mc_stage_max_words(C.byref(params)))
        self.stage = torch.empty(cap, dtype=torch.int64, device=engine.device)
        self.gathered = torch.empty(cap * world, dtype=torch.int64, device=engine.device) if world > 1 else None
        self.cap = cap

    @staticmethod
    def _ptr(t):
        return C.c_void_p(t.data_ptr()) if t is not None and t.numel() else C.c_void_p(0)

    def begin(self):
        e = self.e
        e.check(e.lib.mc_stage_begin(e.handle, C.byref(self.params), C.byref(self.pf) if self.pf else None,
                                     self._ptr(self.pool), self._ptr(self.shocks), self._ptr(self.paths),
                                     self._ptr(self.terminal), self._ptr(self.stage), self.cap, e.stream_ptr()))

    def stage_words(self):
        return int(self.e.lib.mc_stage_words(self.e.handle))

    def stage_view(self, words):
        return self.stage[:words]

    def gather_view(self, total):
        return self.gathered[:total]

    def next(self, gathered, n_ranks, leader=None):
        """`leader`: another backend of this GPU that has already taken this step on the same gathered buffer -- its
        selection state is copied instead of being derived again (`mc_stage_adopt`)."""
        done = C.c_int(0)
        e = self.e
        if leader is not None and leader is not self:
            e.check(e.lib.mc_stage_adopt(e.handle, leader.e.handle, C.c_void_p(gathered.data_ptr()), n_ranks,
                                         self._ptr(self.stage), C.byref(done), e.stream_ptr()))
        else:
            e.check(e.lib.mc_stage_next(e.handle, C.c_void_p(gathered.data_ptr()), n_ranks, self._ptr(self.stage),
                                        C.byref(done), e.stream_ptr()))
        return bool(done.value)


@functools.lru_cache(maxsize=64)
def _pct_names(pcts) -> tuple:
    return tuple(f"p{p:g}" for p in pcts)


class Engine:
    """One per (thread, device).  Thread-safe through the handle's own mutex, but a
    handle serialises its calls: use one Engine per concurrent caller."""

    def __init__(self, device: Optional[int] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("montecarlo_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = L.load_library()
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.device = torch.device("cuda", self.device_index)
        h = C.c_void_p()
        with torch.cuda.device(self.device):       # the C entry points select their handle's device: restore the caller's
            rc = self.lib.mc_create(self.device_index, C.byref(h))
        if rc != 0:
            raise RuntimeError(f"mc_create failed: {self.lib.mc_last_error(None).decode()}")
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            with torch.cuda.device(self.device):
                self.lib.mc_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != 0:
            raise RuntimeError(f"montecarlo_b200 error {rc}: {self.lib.mc_last_error(self.handle).decode()}")

    def stream_ptr(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ helpers
    def _torch_dtype(self, spec):
        return torch.float64 if spec.dtype == "float64" else torch.float32

    def to_device(self, x, dtype):
        """Host array -> device through pinned memory (or pass a device tensor through)."""
        if x is None:
            return None
        if isinstance(x, torch.Tensor):
            if x.device == self.device and x.dtype == dtype and x.is_contiguous():
                return x
            return x.to(device=self.device, dtype=dtype).contiguous()
        host = torch.from_numpy(np.ascontiguousarray(x)).to(dtype)
        if host.numel() * host.element_size() <= (1 << 16):
            return host.to(self.device)                  # small (a bootstrap pool): pinning it costs more than the copy
        return host.pin_memory().to(self.device, non_blocking=True)

    POOL_STAGE_ELEMS = 1 << 16

    def _stage_pool(self, pool, dtype):
        """The bootstrap pool of ONE simulate() call on the device.  A host array (the drop-in's case: ~10 KB) goes through
        a page-locked buffer into a device buffer that both live as long as the engine -- no allocation, no pageable
        staging copy (27 -> 10 us).  The view is only valid until the next call, which is why this is not `to_device`."""
        if isinstance(pool, torch.Tensor) or np.size(pool) > self.POOL_STAGE_ELEMS:
            return self.to_device(pool, dtype)
        bufs = getattr(self, "_pool_stage", None)
        if bufs is None:
            bufs = self._pool_stage = {}
        if dtype not in bufs:
            bufs[dtype] = (torch.empty(self.POOL_STAGE_ELEMS, dtype=dtype, pin_memory=True),
                           torch.empty(self.POOL_STAGE_ELEMS, dtype=dtype, device=self.device))
        pin, dev = bufs[dtype]
        n = int(np.size(pool))
        # the previous call's copy has long finished: every simulate() ends with a synchronising fetch
        pin[:n].copy_(torch.from_numpy(np.ascontiguousarray(pool).reshape(-1)))
        dev[:n].copy_(pin[:n], non_blocking=True)
        return dev[:n]

    # ------------------------------------------------------------------ main entry
    def simulate(self, spec: SimSpec, pool=None, shocks=None, group=None, keep_paths=True) -> RunOutput:
        """Run `spec`.  With `group` (a torch.distributed process group, or True for the
        default group) the spec's n_paths are sharded over the group's ranks."""
        import torch.distributed as dist
        dt = self._torch_dtype(spec)
        world, rank = 1, 0
        if group is True:
            group = dist.group.WORLD
        if group is not None:
            world, rank = dist.get_world_size(group), dist.get_rank(group)
        offset, n_local = shard_range(spec.n_paths, rank, world)
        if spec.shock == "bootstrap":
            if pool is None:
                raise ValueError("bootstrap shocks need a pool")
            pool_t = self._stage_pool(pool, dt)
            spec.pool_len = int(pool_t.numel())
        else:
            pool_t = None
        shocks_t = self.to_device(shocks, dt) if spec.shock == "injected" else None
        if shocks_t is not None:
            rows = spec.n_steps * (spec.portfolio.arrays()["d"] if spec.portfolio is not None else 1)
            if shocks_t.numel() != rows * n_local:
                raise ValueError(f"injected shocks must have shape [{rows}, {n_local}] on this rank")
        if spec.fan_chart and spec.fan_chunk_paths and n_local > spec.fan_chunk_paths:
            if shocks_t is not None:
                raise ValueError("fan chart by regeneration needs device-generated shocks")
            return self._simulate_regenerating(spec, pool_t, group, world, rank, offset, n_local, dt)
        # every buffer below names its device and the C entry points guard their own: the context manager (~8 us) is only
        # entered when the calling thread sits on another GPU
        with (torch.cuda.device(self.device) if torch.cuda.current_device() != self.device_index else contextlib.nullcontext()):
            terminal = torch.empty(n_local, dtype=dt, device=self.device)
            paths = torch.empty((spec.n_steps + 1, n_local), dtype=dt, device=self.device) if spec.store_paths else None
            params = make_params(spec, path_offset=offset, n_local=n_local, n_ranks=world)
            pf = make_portfolio(spec)
            res = L.McResult()
            hist = np.zeros(spec.hist_bins, dtype=np.int64) if spec.hist_bins else None     # the library writes uint64 counts
            fan = np.zeros((len(spec.path_pcts), spec.n_steps + 1)) if (spec.fan_chart and spec.path_pcts) else None
            hist_p = C.c_void_p(hist.ctypes.data) if hist is not None else C.c_void_p(0)
            fan_p = C.c_void_p(fan.ctypes.data) if fan is not None else C.c_void_p(0)
            ptr = _CudaStageBackend._ptr
            if world == 1:
                self.check(self.lib.mc_run(self.handle, C.byref(params), C.byref(pf) if pf else None, ptr(pool_t),
                                           ptr(shocks_t), ptr(paths), ptr(terminal), C.byref(res), hist_p, fan_p,
                                           self.stream_ptr()))
            else:
                staged = True
                if self.fused_request(spec, world) and self.attach_peers(group):
                    # two exchanges, both carried out by the kernels through peer-mapped memory: no stage calls
                    self.check(self.lib.mc_run_sharded(self.handle, C.byref(params), C.byref(pf) if pf else None,
                                                       ptr(pool_t), ptr(shocks_t), ptr(paths), ptr(terminal),
                                                       self.stream_ptr()))
                    self.check(self.lib.mc_fetch_result(self.handle, C.byref(res), hist_p, fan_p, self.stream_ptr()))
                    staged = res.status != 0       # identical on every rank: redo the reductions the general way
                    if staged:
                        params.flags = (params.flags | L.FLAG_TERMINAL_GIVEN) & ~L.FLAG_STORE_PATHS
                if staged:
                    backend = _CudaStageBackend(self, params, pf, pool_t, shocks_t, paths, terminal, world)
                    run_stages(backend, group)
                    self.check(self.lib.mc_fetch_result(self.handle, C.byref(res), hist_p, fan_p, self.stream_ptr()))
        return self._pack(spec, res, hist, fan, terminal, paths if keep_paths else None, offset)

    # ------------------------------------------------------------------ enqueue now, read later
    def submit(self, spec: SimSpec, pool=None, shocks=None):
        """Enqueue a single-GPU job on the current stream and return at once; `collect(ticket)` waits for it and returns
        what `simulate` would have.  For requests whose reductions the fused closing takes (`fused_request`: no fan chart,
        no stored paths, an explicit histogram range) -- `mc_run_sharded` + `mc_fetch_result` with one rank
        (include/mc_b200.h).  One job in flight per engine: alternate between engines (`Pipeline`) to overlap the host
        side and the closing of one job with the simulation of the next."""
        if getattr(self, "_in_flight", False):
            raise RuntimeError("this engine already has a job in flight: collect() it first, or use another engine")
        if spec.store_paths or not self.fused_request(spec, 1):
            raise ValueError("submit() takes jobs without stored paths / fan chart and with an explicit histogram range")
        dt = self._torch_dtype(spec)
        n = spec.n_paths
        pool_t = None
        if spec.shock == "bootstrap":
            if pool is None:
                raise ValueError("bootstrap shocks need a pool")
            pool_t = self._stage_pool(pool, dt)
            spec.pool_len = int(pool_t.numel())
        shocks_t = self.to_device(shocks, dt) if spec.shock == "injected" else None
        stream = torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            terminal = torch.empty(n, dtype=dt, device=self.device)
            params = make_params(spec, path_offset=0, n_local=n, n_ranks=1)
            pf = make_portfolio(spec)
            ptr = _CudaStageBackend._ptr
            self.check(self.lib.mc_run_sharded(self.handle, C.byref(params), C.byref(pf) if pf else None, ptr(pool_t),
                                               ptr(shocks_t), None, ptr(terminal), C.c_void_p(stream.cuda_stream)))
        self._in_flight = True
        return dict(spec=spec, params=params, pf=pf, pool=pool_t, shocks=shocks_t, terminal=terminal, stream=stream)

    def collect(self, ticket) -> RunOutput:
        """Wait for the job of `submit` and return its RunOutput."""
        spec, params, pf = ticket["spec"], ticket["params"], ticket["pf"]
        res = L.McResult()
        hist = np.zeros(spec.hist_bins, dtype=np.int64) if spec.hist_bins else None
        hist_p = C.c_void_p(hist.ctypes.data) if hist is not None else C.c_void_p(0)
        sp = C.c_void_p(ticket["stream"].cuda_stream)
        ptr = _CudaStageBackend._ptr
        with torch.cuda.device(self.device):
            try:
                self.check(self.lib.mc_fetch_result(self.handle, C.byref(res), hist_p, None, sp))
            finally:
                self._in_flight = False
            if res.status != 0:          # the key window missed: the terminal values are there, reduce them the general way
                params.flags = (params.flags | L.FLAG_TERMINAL_GIVEN) & ~L.FLAG_STORE_PATHS
                self.check(self.lib.mc_run(self.handle, C.byref(params), C.byref(pf) if pf else None, ptr(ticket["pool"]),
                                           ptr(ticket["shocks"]), None, ptr(ticket["terminal"]), C.byref(res), hist_p,
                                           None, sp))
        return self._pack(spec, res, hist, None, ticket["terminal"], None, 0)

    # ------------------------------------------------------------------ peers (runs sharded without host round trips)
    @staticmethod
    def fused_request(spec: SimSpec, world: int) -> bool:
        """What `mc_run_sharded` accepts (include/mc_b200.h): no fan chart, an explicit histogram range."""
        if spec.fan_chart or world > 16:
            return False
        return not spec.hist_bins or (spec.hist_range is not None and spec.hist_bins <= 16384)

    def attach_peers(self, group) -> bool:
        """Collective over `group`, once per engine: every rank allocates its exchange buffer, the 64-byte CUDA IPC
        handles are all-gathered (the one NCCL call of this engine's sharded runs) and every rank maps its peers'
        buffers.  False -- on every rank alike -- when any rank could not, and the staged protocol is used instead."""
        import torch.distributed as dist
        key = id(group)
        state = getattr(self, "_peer_state", None)
        if state is not None:
            return state[0] == key and state[1]
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        ok = 1
        with torch.cuda.device(self.device):
            raw = (C.c_ubyte * 64)()
            local, nbytes = C.c_void_p(), C.c_uint64()
            if self.lib.mc_peer_export(self.handle, world, raw, C.byref(local), C.byref(nbytes)) != 0:
                ok = 0
            mine = torch.tensor(list(raw), dtype=torch.uint8, device=self.device)
            everyone = torch.empty(64 * world, dtype=torch.uint8, device=self.device)
            dist.all_gather_into_tensor(everyone, mine, group=group)
            handles = bytes(everyone.cpu().numpy().tobytes())
            if ok and self.lib.mc_peer_attach(self.handle, rank, world, handles, None) != 0:
                ok = 0
            flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            ok = int(flag.item())
        self._peer_state = (key, bool(ok))
        return bool(ok)

    def debug_set_tuning(self, side_cap: int = -1, no_side_lists: bool = False, waves: int = 0, nblock: int = 0,
                         no_overlap: bool = False):
        """Test knobs (`mc_debug_set_tuning`), applied to this engine and the chunk engines it spawned."""
        self._tuning = (int(side_cap), int(bool(no_side_lists)) | (2 if no_overlap else 0), int(waves), int(nblock))
        for e in [self] + [x for x in getattr(self, "_chunk_engines", []) if x is not self]:
            e.check(e.lib.mc_debug_set_tuning(e.handle, *self._tuning))

    def debug_counter(self, which: int) -> int:
        """0 fused runs, 1 fused runs redone by the staged selection, 2 status of the last one, 3 re-simulations."""
        return int(self.lib.mc_debug_counter(self.handle, int(which)))

    def _simulate_regenerating(self, spec, pool_t, group, world, rank, offset, n_local, dt) -> RunOutput:
        """Fan chart (app.py:643-645) for a path range whose [(T+1), N] matrix does not fit: chunks of
        `spec.fan_chunk_paths` paths take turns in one buffer.  Every chunk is a virtual rank of the staged
        protocol (own handle; n_ranks = ranks x chunks) and is simulated AGAIN -- same seed, same global path
        ids, hence the same values -- before each stage that streams its rows."""
        import torch.distributed as dist
        chunk = int(spec.fan_chunk_paths)
        n_chunks = (n_local + chunk - 1) // chunk
        if world > 1:       # every rank must run the same number of virtual ranks
            t = torch.tensor([n_chunks], device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            n_chunks = int(t.item())
        vworld = world * n_chunks
        engines = getattr(self, "_chunk_engines", [])
        while len(engines) < n_chunks:
            engines.append(self if not engines else Engine(self.device_index))
            if engines[-1] is not self and getattr(self, "_tuning", None):
                engines[-1].check(engines[-1].lib.mc_debug_set_tuning(engines[-1].handle, *self._tuning))
        self._chunk_engines = engines
        with torch.cuda.device(self.device):
            terminal = torch.empty(n_local, dtype=dt, device=self.device)
            buf = torch.empty((spec.n_steps + 1, min(chunk, n_local)), dtype=dt, device=self.device)
            pf = make_portfolio(spec)
            backs = []
            for c in range(n_chunks):
                lo = min(c * chunk, n_local)
                cnt = min(chunk, n_local - lo)
                params = make_params(spec, path_offset=offset + lo, n_local=cnt, paths_ld=buf.shape[1], n_ranks=vworld,
                                     paths_volatile=True, pass_group=n_chunks)
                backs.append(_CudaStageBackend(engines[c], params, pf, pool_t, None, buf, terminal[lo:lo + cnt], 1))
            launches = 0
            for b in backs:                      # simulate + moments + pilot sample, chunk by chunk
                b.begin()
            done = False
            lead = backs[0].e
            while not done:
                words = backs[0].stage_words()
                # a narrowing pass of all chunks has gone into ONE buffer, the leader's (mc_params.pass_group): this
                # GPU is one rank of that stage; every other stage has one buffer per chunk
                shared = n_chunks > 1 and lead.lib.mc_stage_shared(lead.handle) == 1
                local = backs[0].stage_view(words) if shared else torch.cat([b.stage_view(words) for b in backs])
                n_gathered = world if shared else vworld
                if n_chunks == 1:
                    n_gathered = world
                if world > 1:
                    gathered = torch.empty(n_gathered * words, dtype=torch.int64, device=self.device)
                    dist.all_gather_into_tensor(gathered, local, group=group)
                else:
                    gathered = local
                flags = []
                for b in backs:
                    e = b.e
                    reads = e.lib.mc_stage_reads_paths(e.handle)
                    if reads < 0:
                        e.check(reads)
                    if reads:
                        e.check(e.lib.mc_stage_resimulate(e.handle, e.stream_ptr()))
                    flags.append(b.next(gathered, n_gathered, leader=backs[0]))   # one scan per GPU, shared by its chunks
                done = all(flags)
                if any(flags) and not done:
                    raise RuntimeError("virtual ranks disagree on the stage count")
            res = L.McResult()
            hist = np.zeros(spec.hist_bins, dtype=np.uint64) if spec.hist_bins else None
            fan = np.zeros((len(spec.path_pcts), spec.n_steps + 1))
            hist_p = C.c_void_p(hist.ctypes.data) if hist is not None else C.c_void_p(0)
            sim_ms = 0.0
            for i, b in enumerate(backs):        # every handle holds the same merged result; sum the sim times
                r_i = L.McResult()
                b.e.check(b.e.lib.mc_fetch_result(b.e.handle, C.byref(r_i if i else res), hist_p if i == 0 else C.c_void_p(0),
                                                  C.c_void_p(fan.ctypes.data) if i == 0 else C.c_void_p(0), b.e.stream_ptr()))
                r = r_i if i else res
                sim_ms += float(r.sim_ms)
                launches += int(r.launches)
            res.sim_ms = sim_ms
            res.launches = launches
        return self._pack(spec, res, hist, fan, terminal, None, offset)

    def _pack(self, spec, res, hist, fan, terminal, paths, offset) -> RunOutput:
        names = _pct_names(tuple(spec.pcts))
        k = len(names)
        pct = dict(zip(names, res.pct[:k]))                    # a slice of a ctypes array is one call, not one per element
        bounds = dict(zip(names, zip(res.pct_lo[:k], res.pct_hi[:k])))
        fan_d = dict(zip(_pct_names(tuple(spec.path_pcts)), fan)) if fan is not None else None
        return RunOutput(spec=spec, n=int(res.n), n_above=int(res.n_above), mean=float(res.mean), std=float(res.std),
                         min=float(res.min), max=float(res.max), pct=pct, pct_bounds=bounds,
                         tail_count=int(res.tail_count), tail_sum=float(res.tail_sum),
                         hist=(hist if hist.dtype == np.int64 else hist.astype(np.int64)) if hist is not None else None,
                         hist_range=(float(res.hist_lo), float(res.hist_hi)) if hist is not None else None,
                         fan=fan_d, terminal=terminal, paths=paths, sim_ms=float(res.sim_ms),
                         total_ms=float(res.total_ms), launches=int(res.launches), path_offset=offset)

    # ------------------------------------------------------------------ building blocks
    def percentiles(self, x: torch.Tensor, pcts) -> np.ndarray:
        """np.percentile(x, pcts, axis=-1) for a device vector or [rows, n] matrix -> [rows, len(pcts)]."""
        x2 = x if x.dim() == 2 else x.view(1, -1)
        if x2.stride(1) != 1:
            x2 = x2.contiguous()
        rows, n = x2.shape
        pc = (C.c_double * len(pcts))(*[float(p) for p in pcts])
        out = np.zeros((rows, len(pcts)))
        dtype = L.F64 if x2.dtype == torch.float64 else L.F32
        if x2.dtype not in (torch.float32, torch.float64):
            raise TypeError("percentiles needs float32/float64")
        self.check(self.lib.mc_percentiles(self.handle, C.c_void_p(x2.data_ptr()), dtype, rows, n, x2.stride(0), pc,
                                           len(pcts), C.c_void_p(out.ctypes.data), self.stream_ptr()))
        return out

    def histogram(self, x: torch.Tensor, bins: int, lo: float, hi: float) -> np.ndarray:
        x = x.contiguous().view(-1)
        out = np.zeros(bins, dtype=np.uint64)
        dtype = L.F64 if x.dtype == torch.float64 else L.F32
        self.check(self.lib.mc_histogram(self.handle, C.c_void_p(x.data_ptr()), dtype, x.numel(), bins, lo, hi,
                                         C.c_void_p(out.ctypes.data), self.stream_ptr()))
        return out.astype(np.int64)

    def prepare_inputs(self, close, lengths=None, first_row: int = 199, dtype: str = "float64"):
        """What the routine reads, from close prices [n_series, n_obs] (host array or device tensor), on the
        device (producers app.py:224, 233, 392-393, 592, 606, 609).  Returns (pool, info): `pool`
        a device tensor [n_series, pool_ld] in `dtype`, `info` a host structured array with the fields of
        `mc_series_inputs` (one small device->host copy, the only synchronisation)."""
        close_t = self.to_device(close, torch.float64)
        if close_t.dim() == 1:
            close_t = close_t.view(1, -1)
        if close_t.stride(1) != 1:
            close_t = close_t.contiguous()
        n_series, n_obs = close_t.shape
        len_t = None
        if lengths is not None:
            len_t = self.to_device(np.ascontiguousarray(lengths, dtype=np.int32) if not isinstance(lengths, torch.Tensor)
                                   else lengths, torch.int32)
            if len_t.numel() != n_series:
                raise ValueError("lengths must have one entry per series")
        pool_ld = max(1, n_obs - max(int(first_row), 20))
        pool_ld = (pool_ld + 3) // 4 * 4
        dt = torch.float64 if dtype == "float64" else torch.float32
        with torch.cuda.device(self.device):
            pool = torch.empty((n_series, pool_ld), dtype=dt, device=self.device)
            info_t = torch.empty((n_series, C.sizeof(L.McSeriesInputs) // 8), dtype=torch.int64, device=self.device)
            self.check(self.lib.mc_prepare_inputs(self.handle, C.c_void_p(close_t.data_ptr()), _CudaStageBackend._ptr(len_t),
                                                  n_series, n_obs, close_t.stride(0), int(first_row),
                                                  L.F64 if dtype == "float64" else L.F32, C.c_void_p(pool.data_ptr()),
                                                  pool_ld, C.c_void_p(info_t.data_ptr()), self.stream_ptr()))
            info = info_t.cpu().numpy().view(np.dtype(L.SERIES_INPUTS_DTYPE)).reshape(n_series)
        return pool, info

    def calibrate_garch(self, pool: torch.Tensor, info: np.ndarray, levels: int = 8, min_history: int = 100):
        """`mc_calibrate_garch`: per-history (omega, alpha, beta) by Gaussian quasi-maximum
        likelihood under the reference's recursion (app.py:632-633).  Returns (fit_device [n_series, 5] int64
        view of mc_garch_fit, fit_host structured array)."""
        if pool.dtype != torch.float64:
            raise ValueError("calibration reads the fp64 pool")
        n_series = int(pool.shape[0])
        with torch.cuda.device(self.device):
            info_t = torch.from_numpy(info.view(np.int64).reshape(n_series, -1)).to(self.device)
            fit_t = torch.empty((n_series, C.sizeof(L.McGarchFit) // 8), dtype=torch.int64, device=self.device)
            self.check(self.lib.mc_calibrate_garch(self.handle, C.c_void_p(pool.data_ptr()), pool.stride(0),
                                                   C.c_void_p(info_t.data_ptr()), n_series, int(min_history), int(levels),
                                                   C.c_void_p(fit_t.data_ptr()), self.stream_ptr()))
            fit = fit_t.cpu().numpy().view(np.dtype(L.GARCH_FIT_DTYPE)).reshape(n_series)
        return fit_t, fit

    def simulate_batch(self, pool: torch.Tensor, info: np.ndarray, n_paths: int, n_steps: int, seed: int, *,
                       omega: float, alpha: float, beta: float, pcts, path_pcts, tail_pct, store_paths: bool,
                       fan_chart: bool, min_history: int = 100, fit=None, raw: bool = False):
        """`mc_run_batch`: the reference's model (fp64) for every price history of `prepare_inputs` in one
        simulation launch.  Returns a list with one RunOutput per history (None where it was skipped)."""
        if pool.dtype != torch.float64:
            raise ValueError("the batched run is fp64 (the reference's precision)")
        n_series = int(pool.shape[0])
        p = L.McBatchParams()
        p.struct_size, p.version = C.sizeof(L.McBatchParams), L.MC_ABI_VERSION
        p.n_paths, p.seed, p.n_steps, p.n_series = int(n_paths), int(seed) & 0xFFFFFFFFFFFFFFFF, int(n_steps), n_series
        p.min_history = int(min_history)
        p.flags = (L.FLAG_STORE_PATHS if store_paths else 0) | (L.FLAG_FAN_CHART if fan_chart else 0)
        p.omega, p.alpha, p.beta = float(omega), float(alpha), float(beta)
        p.n_pcts, p.n_path_pcts = len(pcts), len(path_pcts) if fan_chart else 0
        for i, v in enumerate(pcts):
            p.pcts[i] = float(v)
        for i, v in enumerate(path_pcts if fan_chart else ()):
            p.path_pcts[i] = float(v)
        p.tail_pct = -1.0 if tail_pct is None else float(tail_pct)
        with torch.cuda.device(self.device):
            info_t = torch.from_numpy(info.view(np.int64).reshape(n_series, -1)).to(self.device)
            terminal = torch.empty((n_series, n_paths), dtype=torch.float64, device=self.device)
            paths = torch.empty((n_series, n_steps + 1, n_paths), dtype=torch.float64, device=self.device) if store_paths else None
            results = (L.McResult * n_series)()
            fan = np.zeros((n_series, n_steps + 1, len(path_pcts))) if fan_chart else None
            self.check(self.lib.mc_run_batch(self.handle, C.byref(p), C.c_void_p(pool.data_ptr()), pool.stride(0),
                                             C.c_void_p(info_t.data_ptr()),
                                             C.c_void_p(fit[0].data_ptr()) if fit is not None else C.c_void_p(0),
                                             _CudaStageBackend._ptr(paths),
                                             C.c_void_p(terminal.data_ptr()), C.cast(results, C.c_void_p),
                                             C.c_void_p(fan.ctypes.data) if fan is not None else C.c_void_p(0),
                                             self.stream_ptr()))
        if raw:      # (mc_result records as a structured array, fan [n_series, T+1, n_path_pcts], terminal, paths)
            return np.frombuffer(results, dtype=np.dtype(L.RESULT_DTYPE)).copy(), fan, terminal, paths
        outs = []
        for b in range(n_series):
            if results[b].n == 0:
                outs.append(None)
                continue
            if fit is not None:
                omega, alpha, beta = (float(fit[1][k][b]) for k in ("omega", "alpha", "beta"))
            spec = SimSpec(n_paths=n_paths, n_steps=n_steps, s0=float(info["s0"][b]), var0=float(info["var0"][b]),
                           omega=omega, alpha=alpha, beta=beta, dtype="float64", seed=(int(seed) + b) & 0xFFFFFFFFFFFFFFFF,
                           store_paths=store_paths, fan_chart=fan_chart, pcts=tuple(pcts), path_pcts=tuple(path_pcts),
                           tail_pct=tail_pct, pool_len=int(info["pool_len"][b]))
            outs.append(self._pack(spec, results[b], None, np.ascontiguousarray(fan[b].T) if fan is not None else None,
                                   terminal[b], paths[b] if paths is not None else None, 0))
        return outs

    def debug_shocks(self, spec: SimSpec, pool=None, path_offset=0, n_local=None) -> torch.Tensor:
        dt = self._torch_dtype(spec)
        pool_t = self.to_device(pool, dt) if pool is not None else None
        if pool_t is not None:
            spec.pool_len = int(pool_t.numel())
        n_local = spec.n_paths if n_local is None else n_local
        params = make_params(spec, path_offset=path_offset, n_local=n_local)
        out = torch.empty((spec.n_steps, n_local), dtype=dt, device=self.device)
        self.check(self.lib.mc_debug_shocks(self.handle, C.byref(params), _CudaStageBackend._ptr(pool_t),
                                            C.c_void_p(out.data_ptr()), self.stream_ptr()))
        return out

    def debug_philox(self, counters: np.ndarray, key0: int, key1: int) -> np.ndarray:
        ctr = torch.from_numpy(np.ascontiguousarray(counters, dtype=np.uint32).view(np.int32)).to(self.device)
        n = ctr.numel() // 4
        out = torch.empty(4 * n, dtype=torch.int32, device=self.device)
        self.check(self.lib.mc_debug_philox(self.handle, C.c_void_p(ctr.data_ptr()), n, key0, key1,
                                            C.c_void_p(out.data_ptr()), self.stream_ptr()))
        return out.cpu().numpy().view(np.uint32).reshape(n, 4)


_engines = threading.local()


class Pipeline:
    """`depth` engines, each on a stream of its own, taking jobs in turn: while one job's result is fetched and unpacked
    by the host and its closing kernels run, the next job's simulation already fills the GPU.

        pipe = Pipeline(2)
        for out in pipe.map(specs): ...          # results in submission order
    """

    def __init__(self, depth: int = 2, device: Optional[int] = None):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.engines = [Engine(device) for _ in range(depth)]
        self.streams = [torch.cuda.Stream(device=e.device) for e in self.engines]
        self._queue = []
        self._next = 0

    def submit(self, spec: SimSpec, pool=None, shocks=None) -> None:
        if len(self._queue) >= len(self.engines):
            raise RuntimeError("every engine of the pipeline has a job in flight: collect() first")
        i = self._next
        self._next = (i + 1) % len(self.engines)
        # the job starts after whatever the caller's stream holds now (inputs it has just produced)
        self.streams[i].wait_stream(torch.cuda.current_stream(self.engines[i].device))
        with torch.cuda.stream(self.streams[i]):
            ticket = self.engines[i].submit(spec, pool=pool, shocks=shocks)
        self._queue.append((i, ticket))

    def collect(self) -> RunOutput:
        i, ticket = self._queue.pop(0)
        return self.engines[i].collect(ticket)

    def map(self, specs, pool=None):
        """RunOutputs of `specs` in order, `depth` jobs in flight."""
        for spec in specs:
            if len(self._queue) >= len(self.engines):
                yield self.collect()
            self.submit(spec, pool=pool)
        while self._queue:
            yield self.collect()

    def close(self):
        while self._queue:
            self.collect()
        for e in self.engines:
            e.close()


def get_engine(device: Optional[int] = None) -> Engine:
    """Per-thread, per-device cached Engine (the app runs sessions in threads, app.py:1392)."""
    if not torch.cuda.is_available():
        raise RuntimeError("montecarlo_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    idx = torch.cuda.current_device() if device is None else int(device)
    cache = getattr(_engines, "cache", None)
    if cache is None:
        cache = _engines.cache = {}
    if idx not in cache:
        cache[idx] = Engine(idx)
    return cache[idx]
