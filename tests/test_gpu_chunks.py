"""chunk / reconstruct_from_chunks on the GPU (csrc/chunk.cu) against fixtures written by the reference's
functions.chunk / functions.reconstruct_from_chunks (tests/golden/make_golden.py gen_chunks) and against the
oracle's restatement of torch.mean(torch.stack(...), 0).  Bit-exact."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
import dls_b200
from dls_b200 import functions as F, kernels as K, synthetic
from dls_b200.dataset import DeviceDataset
from dls_b200.executor import RowHandle, WaveExecutor
from dls_b200.layout import GNLENET as L
from dls_b200.settings import SessionSettings, LearningSettings
from dls_b200.solver import BatchedSolver
from katgen import kat_rows

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_reference_known_answer_through_functions_api(cuda, golden_dir, tmp_path):
    g = np.load(os.path.join(golden_dir, "chunk_kat.npz"))
    meta = json.loads(str(g["meta"]))
    sets = json.loads(str(g["sets"]))
    X = kat_rows(meta["n_models"], L.P, salt=meta["salt"])
    st = SessionSettings(participants=meta["n_models"], seed=42, data_dir=str(tmp_path), learning=LearningSettings())
    ds = DeviceDataset.synthetic(L, meta["n_models"], 32, 8, cuda, seed=42)
    F.set_dataset(ds)
    ex = F.get_executor(st)
    slots = ex.alloc(meta["n_models"])
    for i, s in enumerate(slots):
        ex.retain(s)
        ex.X[s, :L.P] = torch.from_numpy(X[i]).to(cuda)
    pieces = [F.chunk(st, {"model": RowHandle(ex, s), "n": meta["n_chunks"]}) for s in slots]
    assert [p.c1 - p.c0 for p in pieces[0]] == list(g["sizes"])
    assert np.array_equal(pieces[1][0].numpy(), g["piece0_of_model1"])
    res = F.reconstruct_from_chunks(st, {"chunks": [[pieces[mi][c] for mi in sets[c]]
                                                    for c in range(meta["n_chunks"])]})
    assert isinstance(res, list) and len(res) == 1
    assert np.array_equal(bits(res[0].numpy()), bits(g["y"]))
    with pytest.raises(AssertionError, match="No chunks received at index 1"):
        F.reconstruct_from_chunks(st, {"chunks": [[pieces[0][0]], []]})
    F.set_dataset(None)


def test_random_segments_bit_exact(cuda):
    """Ragged segments: 1..70 contributors (cascade levels at 16, 32, ...), lengths around the 8 / 32-column
    boundaries, unaligned first columns."""
    rs = np.random.RandomState(5)
    n_rows, ld = 96, 4096
    Xh = (rs.standard_normal((n_rows, ld)) * np.exp2(rs.randint(-8, 8, (n_rows, ld)))).astype(np.float32)
    segs = []
    col = 3
    for ln, k in [(1, 2), (3, 5), (7, 9), (8, 6), (31, 17), (32, 33), (33, 4), (100, 18), (257, 70), (1025, 7),
                  (2049 - 1497 + 3, 1)]:
        src = rs.choice(64, k, replace=(k > 64)).tolist()
        segs.append((64 + len(segs), col, ln, src))
        col += ln
    assert col <= ld
    X = torch.from_numpy(Xh).to(cuda)
    Y = X.clone()
    ptr = np.zeros(len(segs) + 1, np.int32)
    src_all = []
    for i, s in enumerate(segs):
        src_all += s[3]
        ptr[i + 1] = len(src_all)
    t = lambda a, dt: torch.from_numpy(np.asarray(a, dtype=dt)).to(cuda)
    K.chunk_mean(X, Y, t(ptr, np.int32), t(src_all, np.int32), t([s[0] for s in segs], np.int32),
                 t([s[1] for s in segs], np.int64), t([s[2] for s in segs], np.int32), max(s[2] for s in segs),
                 int(np.diff(ptr).max()))
    torch.cuda.synchronize()
    Yh = Y.cpu().numpy()
    for o, c0, ln, src in segs:
        if ln == 1:
            continue          # a [k, 1] stack is an inner reduction in torch (not a chunk shape: P // n >= 2)
        want = O.mean_stack([Xh[s, c0:c0 + ln] for s in src])
        assert np.array_equal(bits(Yh[o, c0:c0 + ln]), bits(want)), (ln, len(src))
    untouched = np.ones_like(Yh, dtype=bool)
    for o, c0, ln, _ in segs:
        untouched[o, c0:c0 + ln] = False
    assert np.array_equal(Yh[untouched], Xh[untouched])


def test_too_many_contributors_is_an_error(cuda):
    X = torch.zeros((2, 64), device=cuda)
    z = torch.zeros(2, dtype=torch.int32, device=cuda)
    with pytest.raises(dls_b200._lib.DlsError, match="TOO_BIG"):
        K.chunk_mean(X, X, z, z, z, torch.zeros(1, dtype=torch.int64, device=cuda), z[:1], 8, 5000)


def test_conflux_shaped_graph(cuda, tmp_path):
    """Train -> chunk -> reconstruct_from_chunks over three rounds with varying contributor sets: every sink
    within 3e-4 of the oracle's sequential solve, and each reconstructed model bit-exact given its inputs."""
    n, n_i = 8, 64
    dag = synthetic.conflux_dag(n, rounds=3, sample_size=4, n_chunks=7, seed=2)
    st = SessionSettings(participants=n, seed=42, data_dir=str(tmp_path), learning=LearningSettings(0.002, 0.9, 0.0, 32, 2))
    ds = DeviceDataset.synthetic(L, n, n_i, 16, cuda, seed=42)
    ex = WaveExecutor(st, ds)
    solver = BatchedSolver(ex, dag).plan()
    # bit-exactness of every reconstruct wave on the rows the GPU itself produced
    checked = 0
    for w in solver.waves:
        before = ex.X.clone() if w.func == "reconstruct_from_chunks" else None
        solver.run_wave(w)
        if before is not None:
            torch.cuda.synchronize()
            bh, ah = before.cpu().numpy(), ex.X.cpu().numpy()
            for o, c0, ln, src in w.segments:
                want = O.mean_stack([bh[s, c0:c0 + ln] for s in src])
                assert np.array_equal(bits(ah[o, c0:c0 + ln]), bits(want))
                checked += 1
    assert checked >= 3 * 4 * 7
    imgs, labs = synthetic.make_partitions(n, n_i, 32, 3, 10)
    timg, tlab = synthetic.make_test_split(16, 32, 3, 10)
    torch.set_num_threads(2)
    sinks, _ = O.solve_dag(L, dag, imgs, labs, timg, tlab, seed=42, batch=32, local_steps=2, lr=0.002, mu=0.9, wd=0.0,
                           n_i=n_i, minibatch_permutation=synthetic.minibatch_permutation)
    assert sorted(solver.sink_slots) == sorted(sinks)
    for name, row in sinks.items():
        got = ex.X[solver.sink_slots[name], :L.P].cpu().numpy().astype(np.float64)
        r = np.linalg.norm(got - row[:L.P]) / np.linalg.norm(row[:L.P])
        assert r < 3e-4, (name, r)
