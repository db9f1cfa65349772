"""Real-data path: a uint8 partition cache of CIFAR-10 for the device-resident executor.

The reference draws every node's partition through `flwr_datasets.FederatedDataset("uoft-cs/cifar10")` with an
`IidPartitioner` or `DirichletPartitioner(partition_by="label", alpha)` (dasklearn/datasets/__init__.py:7-20) and
converts images with torchvision `ToTensor` inside every train call (datasets/transforms.py:16-18).  Here the
partitions are cut ONCE from a local copy of CIFAR-10 (the python pickle batches `data_batch_1..5`, `test_batch`:
rows of 3072 bytes, channel-planar R, G, B) into the executor's layout — uint8 HWC `[n_nodes, n_i, 32, 32, 3]`,
the `/255` of ToTensor happens inside the kernels — and stored as one `.npz`.

Partitioners (restatements of the flwr_datasets 0.3 algorithms, which are not installed here — the split is
documented, deterministic and NOT bit-pinned to flwr):
  iid        sample i of the training set goes to node i // (n / N): contiguous shards of equal size, the
             remainder (n % N samples) dropped — `IidPartitioner` shards the dataset contiguously;
  dirichlet  per class, the class's samples (in dataset order) are divided among the nodes in proportions drawn
             from Dir(alpha) (numpy Generator seeded with `seed`); as in `DirichletPartitioner`, a draw is
             repeated (up to 10 times) until every node holds at least `min_partition_size` samples.
Partitions of unequal size are padded to the largest with `-1` labels; `DeviceDataset.from_cache` passes the true
sizes so that a node's epoch ends where its data ends (model_trainer.py:99-102).
"""
import os
import pickle
from typing import Optional, Sequence, Tuple

import numpy as np


def load_cifar10_batches(root: str) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """(train images uint8 [50000,32,32,3], train labels, test images [10000,32,32,3], test labels) from the
    extracted `cifar-10-batches-py` directory."""
    def read(name):
        with open(os.path.join(root, name), "rb") as f:
            d = pickle.load(f, encoding="bytes")
        data = np.asarray(d[b"data"], dtype=np.uint8).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)   # planar CHW -> HWC
        labels = np.asarray(d.get(b"labels", d.get(b"fine_labels")), dtype=np.int64)
        return np.ascontiguousarray(data), labels
    parts = [read("data_batch_%d" % i) for i in range(1, 6)]
    timg, tlab = read("test_batch")
    return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts]), timg, tlab


def partition_iid(n_samples: int, n_nodes: int) -> Sequence[np.ndarray]:
    per = n_samples // n_nodes
    return [np.arange(i * per, (i + 1) * per, dtype=np.int64) for i in range(n_nodes)]


def partition_dirichlet(labels: np.ndarray, n_nodes: int, alpha: float, seed: int = 42,
                        min_partition_size: int = 10) -> Sequence[np.ndarray]:
    rng = np.random.default_rng(seed)
    classes = np.unique(labels)
    for _ in range(10):
        parts = [[] for _ in range(n_nodes)]
        for c in classes:
            idx = np.nonzero(labels == c)[0]
            prop = rng.dirichlet(np.full(n_nodes, alpha))
            cuts = (np.cumsum(prop) * len(idx)).astype(np.int64)[:-1]
            for node, chunk in enumerate(np.split(idx, cuts)):
                parts[node].append(chunk)
        out = [np.sort(np.concatenate(p)) for p in parts]
        if min(len(p) for p in out) >= min_partition_size:
            return out
    raise ValueError("could not draw a Dirichlet(%g) split with every partition >= %d samples" % (alpha, min_partition_size))


def build_cache(cifar_root: str, out_path: str, n_nodes: int, partitioner: str = "iid", alpha: float = 1.0,
                seed: int = 42) -> str:
    """Cut CIFAR-10 into `n_nodes` partitions and write the cache `out_path` (.npz)."""
    img, lab, timg, tlab = load_cifar10_batches(cifar_root)
    parts = partition_dirichlet(lab, n_nodes, alpha, seed) if partitioner == "dirichlet" else partition_iid(len(lab), n_nodes)
    sizes = np.array([len(p) for p in parts], dtype=np.int64)
    n_i = int(sizes.max())
    pimg = np.zeros((n_nodes, n_i, 32, 32, 3), dtype=np.uint8)
    plab = -np.ones((n_nodes, n_i), dtype=np.int64)
    for i, p in enumerate(parts):
        pimg[i, :len(p)] = img[p]
        plab[i, :len(p)] = lab[p]
    np.savez(out_path, part_images=pimg, part_labels=plab, sizes=sizes, test_images=timg, test_labels=tlab,
             meta=np.array("cifar10 %s alpha=%g seed=%d nodes=%d" % (partitioner, alpha, seed, n_nodes)))
    return out_path


def open_cache(path: str, device, seed: int = 42, node_ids: Optional[Sequence[int]] = None):
    """DeviceDataset over a cache written by `build_cache` (all nodes, or the rank's `node_ids`)."""
    from .dataset import DeviceDataset
    z = np.load(path)
    pimg, plab, sizes = z["part_images"], z["part_labels"], z["sizes"]
    if node_ids is not None:
        sel = np.asarray(list(node_ids), dtype=np.int64)
        pimg, plab, sizes = pimg[sel], plab[sel], sizes[sel]
    ds = DeviceDataset(pimg, np.maximum(plab, 0), z["test_images"], z["test_labels"], device, seed, node_ids=node_ids)
    ds.sizes = sizes
    return ds
