"""TEST INFRASTRUCTURE (oracle): NumPy restatement of the device RNG (mopo_b200/csrc/philox.cuh).

The reference draws from NumPy's global MT19937 (np.random.normal at mopo/models/fake_env.py:72, np.random.choice at
mopo/models/bnn.py:343); the CUDA path replaces it with counter-based Philox4x32-10 (Salmon et al., SC'11; known-answer
vectors from Random123 are pinned in tests/test_oracle_golden.py) keyed by the seed, with counter
(row id, step id, draw block, stream).  Elite picks are integer arithmetic and must match the device bit-exactly;
Box-Muller normals match to ~1e-5 (the device uses approximate lg2 / sin / cos).
"""
import numpy as np

STREAM_ELITE, STREAM_NOISE, STREAM_POOL, STREAM_POLICY, STREAM_SAC = 0, 1, 2, 3, 4     # SAC: policy noise inside the SAC update
_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10; all arguments broadcastable uint32-valued arrays. Returns 4 uint32 arrays."""
    c0, c1, c2, c3 = [np.asarray(c, dtype=np.uint64) & _MASK for c in np.broadcast_arrays(c0, c1, c2, c3)]
    k0 = np.uint64(int(k0) & 0xFFFFFFFF)
    k1 = np.uint64(int(k1) & 0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(_M0) * c0
        p1 = np.uint64(_M1) * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0 = (k0 + np.uint64(_W0)) & _MASK
        k1 = (k1 + np.uint64(_W1)) & _MASK
    return [c.astype(np.uint32) for c in (c0, c1, c2, c3)]


def u01(x):
    """uniform in (0,1): 24 high bits + half an ulp (mopo_u01)."""
    return ((np.asarray(x, np.uint32) >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)


def elite_index(rows, step_id, seed, n_elites):
    """philox_elite: index into the elite list for each global row id."""
    x = philox4x32_10(np.asarray(rows, np.uint32), np.uint32(step_id & 0xFFFFFFFF), 0, STREAM_ELITE,
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)[0]
    return ((x.astype(np.uint64) * np.uint64(n_elites)) >> np.uint64(32)).astype(np.int64)


def normals(rows, step_id, seed, d1, stream=STREAM_NOISE):
    """philox_normal4 over draw blocks 0..ceil(d1/4): standard normals [len(rows), d1] (float64 arithmetic)."""
    rows = np.asarray(rows, np.uint32)
    out = np.empty((len(rows), 4 * ((d1 + 3) // 4)))
    for blk in range((d1 + 3) // 4):
        r = philox4x32_10(rows, np.uint32(step_id & 0xFFFFFFFF), blk, stream, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        u = [u01(x).astype(np.float64) for x in r]
        r0, r1 = np.sqrt(-2.0 * np.log(u[0])), np.sqrt(-2.0 * np.log(u[2]))
        a0, a1 = 2 * np.pi * u[1], 2 * np.pi * u[3]
        out[:, 4 * blk + 0], out[:, 4 * blk + 1] = r0 * np.cos(a0), r0 * np.sin(a0)
        out[:, 4 * blk + 2], out[:, 4 * blk + 3] = r1 * np.cos(a1), r1 * np.sin(a1)
    return out[:, :d1]


def uniforms_pm1(rows, step_id, seed, d):
    """rollout_random actions (mopo.py:738): 2 * u01 - 1 over draw blocks of 4, stream STREAM_POLICY (float32 arithmetic)."""
    rows = np.asarray(rows, np.uint32)
    out = np.empty((len(rows), 4 * ((d + 3) // 4)), np.float32)
    for blk in range((d + 3) // 4):
        r = philox4x32_10(rows, np.uint32(step_id & 0xFFFFFFFF), blk, STREAM_POLICY, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        for u in range(4):
            out[:, 4 * blk + u] = np.float32(2.0) * u01(r[u]).astype(np.float32) - np.float32(1.0)
    return out[:, :d]


def pool_indices(n, size, seed, offset):
    """mopo_pool_sample / the SAC sampler: n indices in [0, size) for draw id ``offset`` (philox_pool_draw): counter =
    (row, offset lo, offset hi, STREAM_POOL), index = floor(((x << 32) | y) * size / 2^64).  Draws with different offsets
    share no counter, as np.random.randint per batch (flexible_replay_pool.py:85-87) shares no state."""
    r = philox4x32_10(np.arange(n, dtype=np.uint32), np.uint32(offset & 0xFFFFFFFF), np.uint32((offset >> 32) & 0xFFFFFFFF),
                      STREAM_POOL, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    u = [(int(x) << 32) | int(y) for x, y in zip(r[0], r[1])]
    return np.asarray([(v * int(size)) >> 64 for v in u], np.int64)
