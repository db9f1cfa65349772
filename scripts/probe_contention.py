#!/usr/bin/env python
"""How tcgen05.mma, tcgen05.ld and warp shuffles contend on one SM (dls_tc_contention).  Diagnostic only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dls_b200
from dls_b200 import _lib
dev = torch.device("cuda:0")
cyc = torch.zeros(2, dtype=torch.int64, device=dev); sink = torch.zeros(256, device=dev)
lib = _lib.lib()
for n_mma, loops, shfl in [(2000, 0, 0), (0, 200, 0), (2000, 200, 0), (2000, 100, 0), (0, 200, 64), (2000, 200, 64), (2000, 100, 64), (0, 200, 256), (2000, 100, 256), (0, 200, -1), (0, 200, -2)]:
    cyc.zero_()
    rc = lib.dls_tc_contention(n_mma, loops, shfl, cyc.data_ptr(), sink.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    c = cyc.tolist()
    print("mma %5d  ld rounds %4d  shfl/round %3d  ->  MMA stream %8d cyc (%.1f / MMA)   load loop %8d cyc (%.0f / round)"
          % (n_mma, loops, shfl, c[0], c[0] / max(n_mma, 1), c[1], c[1] / max(loops, 1)), "rc", rc, flush=True)
