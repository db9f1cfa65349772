#!/usr/bin/env python
"""Micro-benchmark of the aggregate wave alone: k-regular topology, GNLeNet rows, L2 flushed
between timed launches (a 512 MB write), CUDA events, median over iterations.

    python scripts/bench_agg.py [--nodes 256 1024] [--variants 1 2] [--iters 20]
"""
import argparse
import json
import math
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import dls_b200  # noqa: F401
from dls_b200 import _lib, kernels as K, synthetic
from dls_b200.layout import GNLENET as L


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, nargs="+", default=[256, 1024])
    ap.add_argument("--variants", type=int, nargs="+", default=[1, 3, 4])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--sweep", action="store_true", help="sweep staged tile width / stage count (tuning)")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
    for n in args.nodes:
        k = int(math.floor(math.log2(n)))
        if (n * k) % 2:
            k += 1
        nb = synthetic.kreg_topology(n, k, 42)
        rowptr = np.arange(n + 1, dtype=np.int32) * (k + 1)
        col = np.array([list(nb[i]) + [i] for i in range(n)], dtype=np.int32).reshape(-1)
        w = np.full(col.shape, np.float32(1.0 / (k + 1)), dtype=np.float32)
        X = torch.randn((2 * n, L.ld), device=dev)
        rp, ci, wv = (torch.from_numpy(a).to(dev) for a in (rowptr, col, w))
        out_rows = torch.arange(n, 2 * n, dtype=torch.int32, device=dev)
        ref = None
        configs = [(v, None, None) for v in args.variants]
        if args.sweep:
            configs = [(1, None, None)] + [(v, tw, st) for v in (2, 3) for tw in (4, 8, 16) for st in (2, 3, 4)]
        for v, tw, stg in configs:
            _lib.lib().dls_tune(0, int(tw or 0))            # aggregate tile width (0 = auto)
            _lib.lib().dls_tune(1, int(stg or 0))           # aggregate pipeline stages (0 = auto)
            try:
                times = []
                for it in range(args.iters + 3):
                    if not args.no_flush:
                        flush.fill_(it & 0xff)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    K.agg_spmm(X, X, rp, ci, wv, L.P, out_rows=out_rows, n_src_rows=n, variant=v)
                    e1.record()
                    torch.cuda.synchronize()
                    if it >= 3:
                        times.append(e0.elapsed_time(e1))
                y = X[n:, :L.P].clone()
                same = True if ref is None else bool(torch.equal(y, ref))
                ref = y if ref is None else ref
                ms = statistics.median(times)
                gbs = 2 * n * L.P * 4 / 1e9 / (ms / 1e3)
                print(json.dumps({"nodes": n, "k": k, "variant": v, "tw4": tw, "stages": stg, "ms": round(ms, 4), "min_ms": round(min(times), 4),
                                  "GBps": round(gbs, 1), "frac_of_peak": round(gbs / peak, 3), "bit_equal": same}))
            except Exception as exc:
                print(json.dumps({"nodes": n, "variant": v, "tw4": tw, "stages": stg, "error": str(exc)[:60]}))


if __name__ == "__main__":
    main()
