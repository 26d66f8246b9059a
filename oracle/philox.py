"""Philox4x32-10 in numpy — oracle for the dropout mask of the fused linear-forward epilogue.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference has no dropout (ResNet-18
has none; BASELINE.json's north_star asks for it), so there is no reference file:line to follow; this
restates the published Philox4x32-10 algorithm (Salmon et al., SC'11: multipliers 0xD2511F53 / 0xCD9E8D57,
Weyl constants 0x9E3779B9 / 0xBB67AE85, 10 rounds) with the keying the kernel documents:
    counter = (e >> 2 low 32, e >> 34, layer_id, 0)   e = row * N + col  (linear element index)
    key     = (seed low 32, seed high 32 ^ step)
    element e uses word (e & 3) of the 4-word output; keep iff word >= uint32(p * 2**32).
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = np.uint32(0x9E3779B9)
W1 = np.uint32(0xBB67AE85)
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint32) for c in (c0, c1, c2, c3))
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & MASK).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & MASK).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32((int(k0) + int(W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def dropout_threshold(p: float) -> int:
    """uint32 threshold the kernel uses: min(float32(p) * 2**32, 4294967040) truncated."""
    t = np.float32(p) * np.float32(4294967296.0)
    t = min(t, np.float32(4294967040.0))
    return int(t)


def dropout_keep_mask(rows: int, cols: int, p: float, seed: int, step: int, layer_id: int) -> np.ndarray:
    """Boolean [rows, cols] keep mask identical to the GPU epilogue's (cols % 4 == 0)."""
    assert cols % 4 == 0
    e = np.arange(rows * cols // 4, dtype=np.uint64)  # counter index = element index / 4
    c0 = (e & MASK).astype(np.uint32)
    c1 = (e >> np.uint64(32)).astype(np.uint32)
    c2 = np.full_like(c0, np.uint32(layer_id))
    c3 = np.zeros_like(c0)
    k0 = seed & 0xFFFFFFFF
    k1 = ((seed >> 32) & 0xFFFFFFFF) ^ (step & 0xFFFFFFFF)
    r = philox4x32_10(c0, c1, c2, c3, k0, k1)
    words = np.stack(r, axis=1).reshape(rows, cols)
    return words >= np.uint32(dropout_threshold(p))
