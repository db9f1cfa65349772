"""A/B timing of the train wave with a dls_tune switch: python scripts/ab_train.py KEY [nodes]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import dls_b200
from dls_b200 import _lib, synthetic
from dls_b200.dataset import DeviceDataset
from dls_b200.executor import WaveExecutor
from dls_b200.layout import GNLENET as L
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver
key = int(sys.argv[1]); n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dev = torch.device("cuda:0")
st = SessionSettings(participants=n, seed=42, learning=LearningSettings(0.002, 0.9, 0.0, 32, 5), precision=1)
ds = DeviceDataset.synthetic(L, n, 320, 16, dev, seed=42, per_node_seed=True)
nb = synthetic.kreg_topology(n, 8, 42)
res = {}
for val in (1, 0, 1, 0):
    _lib.lib().dls_tune(key, val)
    ex = WaveExecutor(st, ds, device=dev)
    s = BatchedSolver(ex, synthetic.dpsgd_dag(n, 6, nb)).plan()
    tw = [w for w in s.waves]
    for w in tw[:4]: s.run_wave(w)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for w in tw[4:]: s.run_wave(w)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 4
    res.setdefault(val, []).append(ms)
    chk = float(ex.X[s.waves[-1].out_slots[0], :L.P].double().sum())
    print("tune[%d]=%d: %.3f ms per round, checksum %.9f" % (key, val, ms, chk), flush=True)
