"""The CUDA kernels replace the reference's per-point f32 divisions by per-level constants (hex.rs:68-69, :79-80, :83;
metadata.rs:101) with  (float)((double)a * (1.0 / (double)b))  (common.cuh fdiv_inv).  This must be the IEEE binary32
quotient bit for bit; checked here on random bit patterns, on the divisors the default grid produces at every level,
and on the constant 3.0 (an exhaustive run over all 2^32 dividends for those divisors was done once)."""
import numpy as np


def _quotients(a, b):
    with np.errstate(all="ignore"):
        q_ref = (a / b).astype(np.float32)
        q_fast = (a.astype(np.float64) * (1.0 / b.astype(np.float64))).astype(np.float32)
    both_nan = np.isnan(q_ref) & np.isnan(q_fast)
    return q_ref.view(np.uint32), q_fast.view(np.uint32), both_nan


def test_random_bit_patterns():
    rng = np.random.default_rng(5)
    for _ in range(4):
        a = rng.integers(0, 2 ** 32, 5_000_000, dtype=np.uint64).astype(np.uint32).view(np.float32)
        b = rng.integers(0, 2 ** 32, 5_000_000, dtype=np.uint64).astype(np.uint32).view(np.float32)
        ok = np.isfinite(a) & np.isfinite(b) & (b != 0)
        r, f, nan = _quotients(a[ok], b[ok])
        assert int(((r != f) & ~nan).sum()) == 0


def test_level_divisors_and_three():
    rng = np.random.default_rng(6)
    sqrt3 = np.float32(1.73205080757)
    divisors = [np.float32(3.0)]
    for max_cell, dim in ((1000.0, 96), (333.3, 7), (10.0, 8), (4.0, 96), (16.0, 20)):
        for h in range(0, 24, 3):
            S = np.float32(max_cell) / np.float32(2 ** h)
            R = (S / np.float32(dim)) / np.float32(2.0)
            divisors += [S, R, R * sqrt3, (-R) * sqrt3]
    for b in divisors:
        a = np.concatenate([(rng.random(400_000) * 2e4 - 1e4).astype(np.float32),
                            rng.integers(0, 2 ** 32, 400_000, dtype=np.uint64).astype(np.uint32).view(np.float32)])
        a = a[np.isfinite(a)]
        r, f, nan = _quotients(a, np.full(a.shape, b, dtype=np.float32))
        assert int(((r != f) & ~nan).sum()) == 0, float(b)
