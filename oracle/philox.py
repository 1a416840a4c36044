"""Bit model of the device random-number path (TEST INFRASTRUCTURE, see oracle/__init__).

Philox-4x32-10 follows the published Random123 algorithm (Salmon et al., SC'11);
constants match /usr/local/cuda/include/curand_philox4x32_x.h:88-91.  The
counter/key layout and the uniform / normal transforms are THIS repo's
conventions (montecarlo_b200/csrc/mc_rng.cuh) -- the reference has no counter
based generator (it bootstraps with np.random.choice, app.py:620).

Counter layout:  ctr = (path_lo, path_hi, call_index, stream), key = (seed_lo, seed_hi)
  fp32 normals:  EIGHT per call -- one Box-Muller pair per 32-bit word:
                 radius uniform v1 = 1 - word[31:12] * 2^-20  in (0, 1]
                 direction d = word[11:0] (12 bits), theta = 2*pi*(d/4096 - 1/2); the device reads
                 (cos, sin) of it from a 4096-entry table rounded once from fp64
                 r = sqrt(-2 ln v1);  z[2k] = r*cos(theta), z[2k+1] = r*sin(theta)   -> steps 8j..8j+7
  fp64 normals:  FOUR per call: v1 = (u+1) * 2^-32 in (0,1],  theta = 2*pi*(u2 * 2^-32) - pi
  fp64 bootstrap: FOUR per call: idx = (word * H) >> 32   (mulhi)
  fp32 bootstrap (H <= 65536): FIVE per call from 23-bit fields (word_k[31:9], k = 0..3, and 23 bits of
                 the low bytes of words 0,1,2): idx = bits(fl32(fma(2^23 + field, H*2^-23, 2^23 - H + 1/2)))
                 - 0x4B000000 - 1 = floor(field*H/2^23), a tie (exact integer) rounding to even
"""
from __future__ import annotations

import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)

STREAM_SHOCK = 0


def philox4x32(ctr, key, rounds: int = 10):
    """ctr: 4 arrays (uint32-valued), key: 2 scalars/arrays.  Returns 4 uint32 arrays."""
    c = [np.asarray(x, dtype=np.uint64) & MASK32 for x in ctr]
    c = list(np.broadcast_arrays(*c))
    k0 = np.uint64(int(key[0]) & 0xFFFFFFFF)
    k1 = np.uint64(int(key[1]) & 0xFFFFFFFF)
    for _ in range(rounds):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = np.uint64((int(k0) + W0) & 0xFFFFFFFF)
        k1 = np.uint64((int(k1) + W1) & 0xFFFFFFFF)
    return [x.astype(np.uint32) for x in c]


def path_words(seed: int, path_ids, call_index, stream: int = STREAM_SHOCK):
    """Four words of the call `call_index` of each path (arrays broadcast)."""
    pid = np.asarray(path_ids, dtype=np.uint64)
    ctr = (pid & MASK32, pid >> np.uint64(32),
           np.asarray(call_index, dtype=np.uint64), np.uint64(stream))
    key = (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return philox4x32(ctr, key)


def mantissa_float(u):
    """bits(0x3f800000 | (u >> 9)) as float32 in [1, 2)."""
    u = np.asarray(u, dtype=np.uint32)
    return ((u >> np.uint32(9)) | np.uint32(0x3F800000)).view(np.float32)


def normals_per_call(dtype, full_res: bool = False) -> int:
    return 8 if (np.dtype(dtype) == np.float32 and not full_res) else 4


DIR_BITS = 12
BOOT32_MAX_POOL = 65536


def direction_table():
    """The device's (cos, sin) table: float32(cospi/sinpi(2*(d/4096 - 1/2))), d = 0..4095."""
    d = np.arange(1 << DIR_BITS, dtype=np.float64)
    theta = 2.0 * np.pi * (d / (1 << DIR_BITS) - 0.5)
    c, s = np.cos(theta), np.sin(theta)
    # exact zeros / ones where sincospi is exact (multiples of a quarter turn)
    q = (1 << DIR_BITS) // 4
    c[[q, 3 * q]] = 0.0
    s[[0, 2 * q]] = 0.0
    return c.astype(np.float32).astype(np.float64), s.astype(np.float32).astype(np.float64)


_DIR = None


def _pair_f32(word):
    global _DIR
    if _DIR is None:
        _DIR = direction_table()
    w = np.asarray(word, dtype=np.uint32)
    v1 = 1.0 - (w >> np.uint32(12)).astype(np.float64) * 2.0 ** -20
    d = (w & np.uint32(0xFFF)).astype(np.int64)
    r = np.sqrt(-2.0 * np.log(v1))
    return r * _DIR[0][d], r * _DIR[1][d]


TWO_PI_F = float(np.float32(6.28318530717958647692))
THREE_PI_F = float(np.float32(9.42477796076937971538))


def _pair_f32_full(wr, wd):
    """MC_FLAG_FULL_RES_NORMALS (mc_rng.cuh: pair_full): radius from all 32 bits of one word -- v1 = fl32(fl32(u) 2^-32 +
    2^-32), one fused rounding --, direction theta = fl32(m 2 pi - 3 pi) from the top 23 bits of the next word."""
    f = np.asarray(wr, dtype=np.uint32).astype(np.float32).astype(np.float64)          # cvt.rn.f32.u32
    v1 = (f * 2.0 ** -32 + 2.0 ** -32).astype(np.float32).astype(np.float64)           # exact in fp64, then one rounding
    m = mantissa_float(wd).astype(np.float64)
    th = (m * TWO_PI_F - THREE_PI_F).astype(np.float32).astype(np.float64)
    r = np.sqrt(-2.0 * np.log(v1))
    return r * np.cos(th), r * np.sin(th)


def normals_from_words(w, dtype, full_res: bool = False):
    """All normals of one call (list of arrays): 8 for float32 (4 at full resolution), 4 for float64."""
    if np.dtype(dtype) == np.float32 and full_res:
        return list(_pair_f32_full(w[0], w[1])) + list(_pair_f32_full(w[2], w[3]))
    if np.dtype(dtype) == np.float32:
        out = []
        for k in range(4):
            out += list(_pair_f32(w[k]))
        return out
    out = []
    for ua, ub in ((w[0], w[1]), (w[2], w[3])):
        v1 = (np.asarray(ua, dtype=np.float64) + 1.0) * 2.0 ** -32
        theta = 2.0 * np.pi * (np.asarray(ub, dtype=np.float64) * 2.0 ** -32) - np.pi
        r = np.sqrt(-2.0 * np.log(v1))
        out += [r * np.cos(theta), r * np.sin(theta)]
    return out


def normal_shocks(seed: int, path_offset: int, n_paths: int, n_steps: int,
                  dtype=np.float32, n_assets: int = 1, full_res: bool = False):
    """Z[t, i] (single asset) or Z[t, a, i] (n_assets>1), float64 values, exactly the
    draws the device makes for global paths path_offset..path_offset+n_paths-1."""
    pid = np.arange(path_offset, path_offset + n_paths, dtype=np.uint64)
    k = normals_per_call(dtype, full_res and n_assets == 1)
    if n_assets == 1:
        n_calls = (n_steps + k - 1) // k
        z = np.empty((n_calls * k, n_paths), dtype=np.float64)
        for j in range(n_calls):
            z[k * j:k * (j + 1)] = normals_from_words(path_words(seed, pid, j), dtype, full_res)
        return z[:n_steps]
    q_per_step = (n_assets + k - 1) // k
    z = np.empty((n_steps, q_per_step * k, n_paths), dtype=np.float64)
    for t in range(n_steps):
        for q in range(q_per_step):
            z[t, k * q:k * (q + 1)] = normals_from_words(path_words(seed, pid, t * q_per_step + q), dtype)
    return z[:, :n_assets]


def indices_per_call(dtype, pool_len: int) -> int:
    return 5 if (np.dtype(dtype) == np.float32 and pool_len <= BOOT32_MAX_POOL) else 4


def _boot_fields(w):
    """The five 23-bit fields of one call (fp32 kernels)."""
    low = lambda x: (np.asarray(x, dtype=np.uint32) & np.uint32(0xFF)).astype(np.uint32)
    f4 = low(w[0]) | (low(w[1]) << np.uint32(8)) | (low(w[2]) << np.uint32(16))
    return [np.asarray(w[k], dtype=np.uint32) >> np.uint32(9) for k in range(4)] + [f4 & np.uint32(0x7FFFFF)]


def _boot_index32(field, pool_len: int):
    """floor(field*H/2^23).  The device takes it with ONE fp32 FMA that rounds down:
    rd((2^23 + field) * (H 2^-23) + (2^23 - H + 1)) = 2^23 + floor(field*H/2^23) + 1 -- the product and the sum are
    exact inside the FMA and the result lies in [2^23, 2^24), where the fp32 grid is the integers.  Restated here
    with that same single rounding (exact in fp64, then floor), and checked against the integer floor in
    tests/test_oracle_golden.py."""
    g = 8388608.0 + field.astype(np.float64)                       # exact
    scale = float(np.float32(pool_len) * np.float32(2.0 ** -23))     # exact in fp32
    bias = float(np.float32(8388608.0 - pool_len + 1.0))             # exact in fp32
    t = np.floor(g * scale + bias)                                   # exact in fp64; rounding down to the fp32 grid
    return (t.astype(np.float32).view(np.uint32).astype(np.int64) - 0x4B000000) - 1


def bootstrap_indices(seed: int, path_offset: int, n_paths: int, n_steps: int, pool_len: int,
                      dtype=np.float64):
    """idx[t, i]: the device's stand-in for np.random.choice(std_returns, size=n_sims) (app.py:620).
    fp64 kernels: 4 per call, mulhi(word, H).  fp32 kernels: 5 per call from 23-bit fields."""
    pid = np.arange(path_offset, path_offset + n_paths, dtype=np.uint64)
    k = indices_per_call(dtype, pool_len)
    n_calls = (n_steps + k - 1) // k
    idx = np.empty((n_calls * k, n_paths), dtype=np.int64)
    for j in range(n_calls):
        w = path_words(seed, pid, j)
        if k == 4:
            for q in range(4):
                idx[4 * j + q] = (w[q].astype(np.uint64) * np.uint64(pool_len)) >> np.uint64(32)
        else:
            for q, f in enumerate(_boot_fields(w)):
                idx[5 * j + q] = _boot_index32(f, pool_len)
    return idx[:n_steps]


KAT_VECTORS = [
    # (counter, key, expected)  -- Random123 kat_vectors, philox4x32-10
    ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
     (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
]
