"""Bit model of the device random-number path (TEST INFRASTRUCTURE, see oracle/__init__).

Philox-4x32-10 follows the published Random123 algorithm (Salmon et al., SC'11);
constants match /usr/local/cuda/include/curand_philox4x32_x.h:88-91.  The
counter/key layout and the uniform / normal transforms are THIS repo's
conventions (montecarlo_b200/csrc/mc_rng.cuh) -- the reference has no counter
based generator (it bootstraps with np.random.choice, app.py:620).

Counter layout:  ctr = (path_lo, path_hi, call_index, stream), key = (seed_lo, seed_hi)
  fp32 normals:  SIX per call (three Box-Muller pairs cut from the 128 bits):
                 radius uniform k = 2 - bits(0x3f800000 | word_k[22:0])      k = 0,1,2
                 angle A field = word_3 (top 23 bits used), B = byte_perm(w0,w1,0x3733),
                 C = byte_perm(w2,w3,0x3433);  m = bits(0x3f800000 | (field >> 9))
                 r = sqrt(-2 ln(radius uniform)), theta = 2*pi*m - 3*pi in [-pi, pi)
                 z[2p] = r*cos(theta), z[2p+1] = r*sin(theta)   -> steps 6j..6j+5
  fp64 normals:  FOUR per call: v1 = (u+1) * 2^-32 in (0,1],  theta = 2*pi*(u2 * 2^-32) - pi
  bootstrap   :  FOUR per call: idx = (u * H) >> 32   (mulhi)
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


def normals_per_call(dtype) -> int:
    return 6 if np.dtype(dtype) == np.float32 else 4


def _byte_perm(x, y, selector):
    """CUDA __byte_perm: result byte i = byte (selector nibble i) of the 8 bytes {x0..x3, y0..y3}."""
    x = np.asarray(x, dtype=np.uint32).astype(np.uint64)
    y = np.asarray(y, dtype=np.uint32).astype(np.uint64)
    both = x | (y << np.uint64(32))
    out = np.zeros_like(x)
    for i in range(4):
        sel = (selector >> (4 * i)) & 0x7
        out |= ((both >> np.uint64(8 * sel)) & np.uint64(0xFF)) << np.uint64(8 * i)
    return out.astype(np.uint32)


def _pair_f32(radius_word, angle_field):
    m1 = ((np.asarray(radius_word, dtype=np.uint32) & np.uint32(0x007FFFFF)) | np.uint32(0x3F800000)).view(np.float32)
    v1 = 2.0 - m1.astype(np.float64)
    m2 = mantissa_float(angle_field).astype(np.float64)
    # the device forms theta = fma(m2, 2pi_f32, -3pi_f32) in fp32; model it in fp64
    theta = m2 * float(np.float32(2 * np.pi)) - float(np.float32(3 * np.pi))
    r = np.sqrt(-2.0 * np.log(v1))
    return r * np.cos(theta), r * np.sin(theta)


def normals_from_words(w, dtype):
    """All normals of one call (list of arrays): 6 for float32, 4 for float64."""
    if np.dtype(dtype) == np.float32:
        za = _pair_f32(w[0], w[3])
        zb = _pair_f32(w[1], _byte_perm(w[0], w[1], 0x3733))
        zc = _pair_f32(w[2], _byte_perm(w[2], w[3], 0x3433))
        return [za[0], za[1], zb[0], zb[1], zc[0], zc[1]]
    out = []
    for ua, ub in ((w[0], w[1]), (w[2], w[3])):
        v1 = (np.asarray(ua, dtype=np.float64) + 1.0) * 2.0 ** -32
        theta = 2.0 * np.pi * (np.asarray(ub, dtype=np.float64) * 2.0 ** -32) - np.pi
        r = np.sqrt(-2.0 * np.log(v1))
        out += [r * np.cos(theta), r * np.sin(theta)]
    return out


def normal_shocks(seed: int, path_offset: int, n_paths: int, n_steps: int,
                  dtype=np.float32, n_assets: int = 1):
    """Z[t, i] (single asset) or Z[t, a, i] (n_assets>1), float64 values, exactly the
    draws the device makes for global paths path_offset..path_offset+n_paths-1."""
    pid = np.arange(path_offset, path_offset + n_paths, dtype=np.uint64)
    k = normals_per_call(dtype)
    if n_assets == 1:
        n_calls = (n_steps + k - 1) // k
        z = np.empty((n_calls * k, n_paths), dtype=np.float64)
        for j in range(n_calls):
            z[k * j:k * (j + 1)] = normals_from_words(path_words(seed, pid, j), dtype)
        return z[:n_steps]
    q_per_step = (n_assets + k - 1) // k
    z = np.empty((n_steps, q_per_step * k, n_paths), dtype=np.float64)
    for t in range(n_steps):
        for q in range(q_per_step):
            z[t, k * q:k * (q + 1)] = normals_from_words(path_words(seed, pid, t * q_per_step + q), dtype)
    return z[:, :n_assets]


def bootstrap_indices(seed: int, path_offset: int, n_paths: int, n_steps: int, pool_len: int):
    """idx[t, i] = mulhi(word, pool_len): the device's stand-in for
    np.random.choice(std_returns, size=n_sims) (app.py:620)."""
    pid = np.arange(path_offset, path_offset + n_paths, dtype=np.uint64)
    n_calls = (n_steps + 3) // 4
    idx = np.empty((n_calls * 4, n_paths), dtype=np.int64)
    for j in range(n_calls):
        w = path_words(seed, pid, j)
        for k in range(4):
            idx[4 * j + k] = (w[k].astype(np.uint64) * np.uint64(pool_len)) >> np.uint64(32)
    return idx[:n_steps]


KAT_VECTORS = [
    # (counter, key, expected)  -- Random123 kat_vectors, philox4x32-10
    ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
     (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
]
