"""NumPy restatement of the counter-based dropout stream of the kernels (csrc/common.cuh: philox4x32-10,
philox_seed, philox_keep8): mask[i] = scale if field16(i) >= round(p * 65536) else 0, where the 16-bit fields of
element group i // 8 come from Philox(counter = (group, site, 0), key = seed * golden + rank). The GPU tests use
it to rebuild, on the host, exactly the masks a kernel applied."""
import numpy as np

_MASK = np.uint64(0xFFFFFFFF)
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint64(0x9E3779B9), np.uint64(0xBB67AE85)


def philox4x32(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    for _ in range(10):
        p0, p1 = _M0 * c0, _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + _W0) & _MASK, (k1 + _W1) & _MASK
    return c0, c1, c2, c3


def keep_mask(seed, rank, site, numel, p):
    """float32 [numel]: 1/(1-p) where the element is kept, 0 where it is dropped (numel % 8 == 0)."""
    assert numel % 8 == 0
    key = (int(seed) * 0x9E3779B97F4A7C15 + int(rank)) & 0xFFFFFFFFFFFFFFFF
    groups = np.arange(numel // 8, dtype=np.uint64)
    r = philox4x32(groups & _MASK, groups >> np.uint64(32), np.full_like(groups, site), np.zeros_like(groups),
                   key & 0xFFFFFFFF, key >> 32)
    fields = np.empty((numel // 8, 8), dtype=np.uint64)
    for w in range(4):
        fields[:, 2 * w] = r[w] & np.uint64(0xFFFF)
        fields[:, 2 * w + 1] = r[w] >> np.uint64(16)
    thr = np.uint64(int(p * 65536.0 + 0.5))
    return np.where(fields >= thr, np.float32(1.0 / (1.0 - p)), np.float32(0.0)).reshape(-1)
