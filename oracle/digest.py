"""Order-sensitive 64-bit digest of integer state arrays (numpy, vectorised).  Test infrastructure."""
import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix(x):
    x = x.astype(np.uint64)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30)
        x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27)
        x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
    return x


def digest(*arrays):
    """sum_i mix(x_i + (i+1) * C) mod 2^64 over the concatenation of the arrays (as int64)."""
    h = np.uint64(0)
    off = 0
    with np.errstate(over="ignore"):
        for a in arrays:
            a = np.asarray(a).astype(np.int64).ravel().view(np.uint64)
            idx = (np.arange(off + 1, off + 1 + len(a), dtype=np.uint64)) * np.uint64(0x9E3779B97F4A7C15)
            h = h + _mix(a + idx).sum(dtype=np.uint64)
            off += len(a)
    return int(h)
