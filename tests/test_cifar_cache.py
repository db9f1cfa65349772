"""CIFAR-10 partition cache (dls_b200/cifar_cache.py) on a fabricated `cifar-10-batches-py` directory: the pickle
format, the HWC conversion torchvision's ToTensor expects (datasets/transforms.py:16-18), the IID and Dirichlet
splits, and ragged minibatches (a node's epoch ends where its data ends, model_trainer.py:99-102)."""
import os
import pickle

import numpy as np

import dls_b200
from dls_b200 import cifar_cache as CC


def fake_cifar(root, n_per_batch=200):
    rs = np.random.RandomState(0)
    os.makedirs(root, exist_ok=True)
    allx, ally = [], []
    for name in ["data_batch_%d" % i for i in range(1, 6)] + ["test_batch"]:
        x = rs.randint(0, 256, size=(n_per_batch, 3072)).astype(np.uint8)
        y = rs.randint(0, 10, size=n_per_batch).tolist()
        with open(os.path.join(root, name), "wb") as f:
            pickle.dump({b"data": x, b"labels": y}, f)
        if name != "test_batch":
            allx.append(x); ally += y
    return np.concatenate(allx), np.array(ally)


def test_iid_and_dirichlet_cache(tmp_path):
    x, y = fake_cifar(str(tmp_path / "c10"))
    img, lab, timg, tlab = CC.load_cifar10_batches(str(tmp_path / "c10"))
    assert img.shape == (1000, 32, 32, 3) and timg.shape == (200, 32, 32, 3)
    # planar CHW rows -> HWC: pixel (y, x) channel c is byte c*1024 + y*32 + x of the row
    assert img[7, 5, 9, 2] == x[7, 2 * 1024 + 5 * 32 + 9] and np.array_equal(lab, y)
    p = CC.build_cache(str(tmp_path / "c10"), str(tmp_path / "iid.npz"), n_nodes=8)
    z = np.load(p)
    assert z["part_images"].shape == (8, 125, 32, 32, 3) and (z["sizes"] == 125).all()
    assert np.array_equal(z["part_images"][3, 10], img[3 * 125 + 10]) and z["part_labels"][3, 10] == lab[3 * 125 + 10]
    p = CC.build_cache(str(tmp_path / "c10"), str(tmp_path / "dir.npz"), n_nodes=8, partitioner="dirichlet", alpha=0.3, seed=1)
    z = np.load(p)
    sizes = z["sizes"]
    assert sizes.sum() == 1000 and sizes.min() >= 10 and sizes.max() > sizes.min()        # skewed, complete, no empty node
    # every training sample appears exactly once
    seen = np.concatenate([z["part_labels"][i, :sizes[i]] for i in range(8)])
    assert np.array_equal(np.bincount(seen, minlength=10), np.bincount(lab, minlength=10))
    # deterministic
    p2 = CC.build_cache(str(tmp_path / "c10"), str(tmp_path / "dir2.npz"), n_nodes=8, partitioner="dirichlet", alpha=0.3, seed=1)
    assert np.array_equal(np.load(p2)["sizes"], sizes)


def test_ragged_minibatches(tmp_path):
    import torch
    fake_cifar(str(tmp_path / "c10"))
    p = CC.build_cache(str(tmp_path / "c10"), str(tmp_path / "dir.npz"), n_nodes=4, partitioner="dirichlet", alpha=0.5, seed=3)
    ds = CC.open_cache(p, torch.device("cpu"), seed=42)
    sizes = ds.sizes
    idx, cnt = ds.minibatches([2, 0], [1, 1], 32, [5, 5])
    assert idx.shape == (5, 2, 32) and cnt.shape == (5, 2)
    for t, q in enumerate([2, 0]):
        n_q = int(sizes[q])
        want_steps = min(5, -(-n_q // 32))
        assert int((cnt[:, t] > 0).sum()) == want_steps
        taken = idx[:, t][idx[:, t] >= 0]
        assert len(set(taken.tolist())) == len(taken) == min(n_q, 5 * 32)
        assert taken.min() >= q * ds.n_i and taken.max() < q * ds.n_i + n_q                 # only the node's own real samples
        if n_q < 160:
            assert cnt[want_steps - 1, t] == n_q - 32 * (want_steps - 1)                    # partial last batch
