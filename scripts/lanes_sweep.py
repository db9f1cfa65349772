"""Train-wave lane count sweep: python scripts/lanes_sweep.py NODES"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dls_b200
from dls_b200 import _lib, synthetic
from dls_b200.dataset import DeviceDataset
from dls_b200.executor import WaveExecutor
from dls_b200.layout import GNLENET as L
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device("cuda:0")
st = SessionSettings(participants=n, seed=42, learning=LearningSettings(0.002, 0.9, 0.0, 32, 5), precision=1)
ds = DeviceDataset.synthetic(L, n, 320, 16, dev, seed=42, per_node_seed=True)
nb = synthetic.kreg_topology(n, 8, 42)
for lanes in (0, 1, 2, 3, 4, 5, 7, 0):
    _lib.lib().dls_tune(2, lanes)
    ex = WaveExecutor(st, ds, device=dev)
    s = BatchedSolver(ex, synthetic.dpsgd_dag(n, 4, nb)).plan()
    for w in s.waves[:2]: s.run_wave(w)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for w in s.waves[2:]: s.run_wave(w)
    e1.record(); torch.cuda.synchronize()
    print("lanes %d: %.3f ms per round" % (lanes, e0.elapsed_time(e1) / 3), flush=True)
    del ex, s
    torch.cuda.empty_cache()
