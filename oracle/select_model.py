"""Host model of the multi-rank reduction protocol (TEST INFRASTRUCTURE, see oracle/__init__).

A NumPy stage backend so that the product's orchestration loop
`montecarlo_b200.engine.run_stages` (begin -> [stage_words, all-gather of one small stage
buffer, next] until done) can be exercised on CPU with the gloo backend, world_size > 1.
The stage CONTRACT is the device's (montecarlo_b200/csrc/mc_api.cu: every rank computes the
same decisions from the gathered buffers; only counters and candidate keys cross ranks); the
selection inside is the simpler exact method the device used first -- a multi-target radix
select on order-preserving keys (digits 12/10/10 for fp32 keys) -- where the device now
narrows pilot brackets (mc_select.cu).  Both are exact, so both must reproduce NumPy: moments
about s0, np.percentile's `_lerp`, tail sums, fixed-bin histogram (app.py:636-656).
"""
from __future__ import annotations

import numpy as np
import torch

DIGIT0_BITS, DIGIT_BITS = 12, 10


def keys_of(x: np.ndarray) -> np.ndarray:
    """Order-preserving unsigned keys (mc_common.cuh: key_of)."""
    if x.dtype == np.float32:
        b = x.view(np.uint32)
        return np.where(b >> np.uint32(31), ~b, b | np.uint32(0x80000000)).astype(np.uint64)
    b = np.ascontiguousarray(x, dtype=np.float64).view(np.uint64)
    return np.where(b >> np.uint64(63), ~b, b | np.uint64(1 << 63))


def values_of(keys: np.ndarray, dtype) -> np.ndarray:
    if np.dtype(dtype) == np.float32:
        k = keys.astype(np.uint32)
        b = np.where(k >> np.uint32(31), k & np.uint32(0x7FFFFFFF), ~k)
        return b.astype(np.uint32).view(np.float32).astype(np.float64)
    k = keys.astype(np.uint64)
    b = np.where(k >> np.uint64(63), k & np.uint64((1 << 63) - 1), ~k)
    return b.astype(np.uint64).view(np.float64)


def plan_passes(key_bits: int):
    left, out = key_bits, []
    while left > 0:
        want = DIGIT0_BITS if not out else DIGIT_BITS
        b = min(left, want)
        left -= b
        out.append((b, left))
    return out


def plan_targets(pcts, n_global):
    """np.percentile(method='linear') indices: (sorted distinct ranks, lo idx, hi idx, gamma)."""
    lo, hi, gamma = [], [], []
    for p in pcts:
        virt = (n_global - 1) * (p / 100.0)
        prev = np.floor(virt)
        gamma.append(virt - prev)
        if virt >= n_global - 1:
            lo.append(n_global - 1); hi.append(n_global - 1)
        else:
            lo.append(int(prev)); hi.append(int(prev) + 1)
    ranks = sorted(set(lo) | set(hi))
    return ranks, [ranks.index(v) for v in lo], [ranks.index(v) for v in hi], gamma


def numpy_lerp(a, b, t):
    diff = b - a
    return b - diff * (1 - t) if t >= 0.5 else a + diff * t


class NumpyStageBackend:
    """Same stage protocol as _CudaStageBackend, on CPU tensors."""

    def __init__(self, values, n_global, s0, pcts=(), tail_pct=None, hist_bins=0, hist_range=None):
        self.x = np.ascontiguousarray(values)
        self.keys = keys_of(self.x)
        self.key_bits = 32 if self.x.dtype == np.float32 else 64
        self.n_global, self.s0 = int(n_global), float(s0)
        self.pcts, self.tail_pct = list(pcts), tail_pct
        self.hist_bins, self.hist_range = hist_bins, hist_range
        self.passes = plan_passes(self.key_bits) if self.pcts else []
        if self.pcts:
            self.ranks, self.lo_idx, self.hi_idx, self.gamma = plan_targets(self.pcts, self.n_global)
            self.prefix = [0] * len(self.ranks)
            self.remaining = list(self.ranks)
        cap = max(6 + (1 << DIGIT0_BITS), 2 * len(self.pcts) * (1 << DIGIT_BITS), 2 + hist_bins)
        self.stage = torch.zeros(cap, dtype=torch.int64)
        self.gathered = None
        self.call = 0
        self.words = 0
        self.result = {}

    # -- buffer protocol ---------------------------------------------------------------
    def stage_words(self):
        return self.words

    def stage_view(self, words):
        return self.stage[:words]

    def gather_view(self, total):
        if self.gathered is None or self.gathered.numel() < total:
            self.gathered = torch.zeros(total, dtype=torch.int64)
        return self.gathered[:total]

    def _groups(self):
        gp = sorted(set(self.prefix))
        return gp, [gp.index(p) for p in self.prefix]

    def _emit_hist(self, k):
        bits, shift = self.passes[k]
        nb = 1 << bits
        if k == 0:
            h = np.bincount((self.keys >> np.uint64(shift)).astype(np.int64), minlength=nb)
            self.stage[6:6 + nb] = torch.from_numpy(h.astype(np.int64))
            self.words = 6 + nb
            return
        gp, _ = self._groups()
        pfx = self.keys >> np.uint64(shift + bits)
        out = np.zeros(len(self.ranks) * nb, dtype=np.int64)
        for g, p in enumerate(gp):
            sel = self.keys[pfx == np.uint64(p)]
            digit = ((sel >> np.uint64(shift)) & np.uint64(nb - 1)).astype(np.int64)
            out[g * nb:(g + 1) * nb] = np.bincount(digit, minlength=nb)
        self.stage[:out.size] = torch.from_numpy(out)
        self.words = out.size

    def begin(self):
        d = self.x.astype(np.float64) - self.s0
        hdr = np.zeros(6, dtype=np.int64)
        hdr[0] = self.x.size
        hdr[1] = int((self.x > self.x.dtype.type(self.s0)).sum())
        f = np.array([d.sum(), (d * d).sum(),
                      self.x.min() if self.x.size else np.inf, self.x.max() if self.x.size else -np.inf])
        hdr[2:6] = f.view(np.int64)
        self.stage[:6] = torch.from_numpy(hdr)
        if self.passes:
            self._emit_hist(0)
        else:
            self.words = 6 + (1 << DIGIT0_BITS)
        self.call = 0

    def _scan(self, k, g, n_ranks, stride):
        bits, _ = self.passes[k]
        nb = 1 << bits
        off = 6 if k == 0 else 0
        gp, grp = self._groups() if k else ([0], [0] * len(self.ranks))
        hist = np.zeros((len(gp), nb), dtype=np.int64)
        for r in range(n_ranks):
            part = g[r * stride + off: r * stride + off + len(self.ranks if k else [0]) * nb]
            hist += part.reshape(-1, nb)[:len(gp)]
        cum = np.cumsum(hist, axis=1)
        for j in range(len(self.ranks)):
            c = cum[grp[j]]
            digit = int(np.searchsorted(c, self.remaining[j], side="right"))
            digit = min(digit, nb - 1)
            self.remaining[j] -= int(c[digit - 1]) if digit else 0
            self.prefix[j] = (self.prefix[j] << bits) | digit

    def next(self, gathered, n_ranks):
        g = gathered.numpy()
        stride = self.words
        c, self.call = self.call, self.call + 1
        n_sel = len(self.passes)
        if c == (n_sel if n_sel else 1):
            tc = sum(int(g[r * stride]) for r in range(n_ranks))
            ts = sum(float(g[r * stride + 1:r * stride + 2].view(np.float64)[0]) for r in range(n_ranks))
            self.result["tail_count"], self.result["tail_sum"] = tc, ts
            if self.hist_bins:
                self.result["hist"] = sum(g[r * stride + 2:r * stride + 2 + self.hist_bins] for r in range(n_ranks))
            return True
        if c == 0:
            n = above = 0
            sd = sd2 = 0.0
            mn, mx = np.inf, -np.inf
            for r in range(n_ranks):
                w = g[r * stride:r * stride + 6]
                f = w[2:6].view(np.float64)
                n += int(w[0]); above += int(w[1]); sd += f[0]; sd2 += f[1]
                mn, mx = min(mn, f[2]), max(mx, f[3])
            md = sd / n
            self.result.update(n=n, n_above=above, mean=self.s0 + md, std=np.sqrt(max(sd2 / n - md * md, 0.0)),
                               min=mn, max=mx)
        if c < n_sel:
            self._scan(c, g, n_ranks, stride)
            if c + 1 < n_sel:
                self._emit_hist(c + 1)
                return False
            vals = values_of(np.array(self.prefix, dtype=np.uint64), self.x.dtype)
            self.result["pct"] = {p: numpy_lerp(vals[self.lo_idx[i]], vals[self.hi_idx[i]], self.gamma[i])
                                  for i, p in enumerate(self.pcts)}
        # final stage
        out = np.zeros(2 + self.hist_bins, dtype=np.int64)
        if self.tail_pct is not None and self.pcts:
            thr = self.result["pct"][self.tail_pct]
            xd = self.x.astype(np.float64)
            sel = xd[xd <= thr]
            out[0] = sel.size
            out[1:2] = np.array([sel.sum()]).view(np.int64)
        if self.hist_bins:
            lo, hi = self.hist_range if self.hist_range else (self.result["min"], self.result["max"])
            out[2:] = np.histogram(self.x.astype(np.float64), bins=self.hist_bins, range=(lo, hi))[0]
        self.stage[:out.size] = torch.from_numpy(out)
        self.words = out.size
        return False
