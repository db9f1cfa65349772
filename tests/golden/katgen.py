"""Deterministic, platform-independent input generator for known-answer tests.

Values are exact in fp32 (integers / 2^k) and include near-cancelling pairs so that a fused
multiply-add or a different summation order changes low-order bits of the FedAvg result.
"""
import numpy as np


def kat_rows(n_rows: int, P: int, salt: int = 0) -> np.ndarray:
    i = np.arange(n_rows, dtype=np.uint64)[:, None]
    j = np.arange(P, dtype=np.uint64)[None, :]
    h = (i * np.uint64(0x9E3779B97F4A7C15) + j * np.uint64(0xC2B2AE3D27D4EB4F) + np.uint64(salt * 7919 + 1))
    h ^= h >> np.uint64(29)
    h *= np.uint64(0xBF58476D1CE4E5B9)
    h ^= h >> np.uint64(32)
    m = (h & np.uint64(0xFFFFFF)).astype(np.int64) - (1 << 23)          # 24-bit signed mantissa
    e = ((h >> np.uint64(24)) & np.uint64(7)).astype(np.int64)            # exponent spread 2^-23..2^-16
    x = (m.astype(np.float64) * np.exp2(e - 23.0)).astype(np.float32)
    # near-cancelling structure: every odd row mostly negates the previous row
    x[1::2] = -x[0:(n_rows // 2) * 2:2] + x[1::2] * np.float32(2.0 ** -12)
    return x
