"""B200-native batched executor for the decentralized-learning simulator's compute-graph solve.

Importable as `dls_b200` (see dls_b200.py at the repo root).  Layout:
  csrc/         hand-written sm_100a CUDA kernels + the C-ABI (include/dls_b200.h)
  _lib.py       ctypes binding (fails loudly when the library is missing; no CPU fallback)
  kernels.py    torch-tensor front end of the C-ABI
  executor.py   stacked [N, P] parameter tensor, momentum, batched waves
  solver.py     DAG -> wavefronts -> waves (drop-in for the broker/worker loop)
  functions.py  API-compatible train / aggregate / test entry points
"""
__version__ = "0.1.0"
