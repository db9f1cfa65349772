#!/usr/bin/env python
"""Measures SM cycles per tcgen05.mma for the operand forms used by the convolution kernels
(dls_tc_probe with its `repeat` counter).  Diagnostic only."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import dls_b200  # noqa: F401
from dls_b200 import _lib


def desc(addr, lbo, sbo):
    return (addr >> 4) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1 << 46)


def idesc(M, N, a_mn, b_mn):
    return (1 << 4) | (1 << 7) | (1 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def main():
    dev = torch.device("cuda:0")
    img = torch.zeros(64 * 1024, dtype=torch.uint8, device=dev)          # zeros: values are irrelevant
    out = torch.zeros((128, 32), device=dev)
    cyc = torch.zeros(1, dtype=torch.int64, device=dev)
    lib = _lib.lib()
    rep = 400
    for name, M, N, amn, bmn, la, sa, lb, sb, *rest in [
            ("K-major  M128 N32 ", 128, 32, 0, 0, 4096, 128, 512, 128),
            ("K-major  M128 N32 A start +16 B (shifted window)", 128, 32, 0, 0, 4096, 128, 512, 128, 16),
            ("K-major  M128 N32 A start +64 B", 128, 32, 0, 0, 4096, 128, 512, 128, 64),
            ("K-major  M128 N64 A start +16 B", 128, 64, 0, 0, 4096, 128, 1024, 128, 16),
            ("K-major  M128 N128 A start +16 B", 128, 128, 0, 0, 4096, 128, 2048, 128, 16),
            ("MN-major M64  N32 B start +16 B", 64, 32, 1, 1, 128, 4096, 128, 4096, 0, 16),
            ("K-major  M128 N64 ", 128, 64, 0, 0, 4096, 128, 1024, 128),
            ("K-major  M128 N128", 128, 128, 0, 0, 4096, 128, 2048, 128),
            ("K-major  M64  N32 ", 64, 32, 0, 0, 4096, 128, 512, 128),
            ("MN-major M64  N32 ", 64, 32, 1, 1, 128, 4096, 128, 4096),
            ("MN-major M64  N64 ", 64, 64, 1, 1, 128, 4096, 128, 4096),
            ("MN-major M128 N32 ", 128, 32, 1, 1, 128, 2048, 128, 4096),
            ("K-major  M128 N96 A start +16 B", 128, 96, 0, 0, 4096, 128, 1536, 128, 16),
            ("K-major  M128 N160 A start +16 B", 128, 160, 0, 0, 4096, 128, 2560, 128, 16),
            ("K-major  M128 N256 A start +16 B", 128, 256, 0, 0, 4096, 128, 4096, 128, 16),
            ("MN-major M128 N64 B start +16 B", 128, 64, 1, 1, 128, 2048, 128, 4096, 0, 16),
            ("MN-major M128 N96 B start +16 B", 128, 96, 1, 1, 128, 2048, 128, 4096, 0, 16),
            ("MN-major M128 N128 B start +16 B", 128, 128, 1, 1, 128, 2048, 128, 4096, 0, 16),
            ("MN-major M128 N160 B start +16 B", 128, 160, 1, 1, 128, 2048, 128, 4096, 0, 16)]:
        a_off = rest[0] if len(rest) > 0 else 0
        b_off = rest[1] if len(rest) > 1 else 0
        ad = torch.tensor(np.array([desc(i * 256 + a_off, la, sa) for i in range(8)], dtype=np.uint64).view(np.int64), device=dev)
        bd = torch.tensor(np.array([desc(i * 256 + b_off, lb, sb) for i in range(8)], dtype=np.uint64).view(np.int64), device=dev)
        rc = lib.dls_tc_probe(img.data_ptr(), img.numel(), img.data_ptr(), img.numel(), ad.data_ptr(), bd.data_ptr(), 8,
                              idesc(M, N, amn, bmn), 32, out.data_ptr(), rep, cyc.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        print("%s rc=%d  %.1f cycles per MMA" % (name, rc, cyc.item() / (rep * 8)), flush=True)


if __name__ == "__main__":
    main()
