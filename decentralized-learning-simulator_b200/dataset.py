"""Device-resident datasets for the batched executor.

The reference loads CIFAR-10 through flwr_datasets + torchvision ToTensor inside every train
call (dasklearn/model_trainer.py:67-87, dasklearn/datasets/transforms.py:16-18).  Here the
nodes' partitions live in HBM as uint8 HWC tensors (ToTensor's /255 happens inside the
kernels) and minibatches are drawn through an index: `minibatches(peers, rounds, ...)`."""
from typing import Callable, Optional, Sequence, Tuple

import numpy as np
import torch

from . import synthetic
from .layout import ModelLayout


class DeviceDataset:
    """uint8 partitions `[n_nodes, n_i, H, W, C]`, labels `[n_nodes, n_i]`, and a test split.

    `permutation(seed, node, round, n_i)` defines the sample order a node sees in a round
    (stand-in for the DataLoader's shuffle); the default is `synthetic.minibatch_permutation`.
    `first_node` is the global id of partition 0 (multi-GPU: a rank holds only its nodes)."""

    def __init__(self, part_images: np.ndarray, part_labels: np.ndarray, test_images: np.ndarray,
                 test_labels: np.ndarray, device, seed: int,
                 permutation: Optional[Callable] = None, first_node: int = 0, node_ids: Optional[Sequence[int]] = None):
        self.n_nodes, self.n_i = part_labels.shape
        self.device = device
        self.seed = seed
        self.permutation = permutation or synthetic.minibatch_permutation
        self.img_bytes = int(np.prod(part_images.shape[2:]))
        self.images = torch.from_numpy(np.ascontiguousarray(part_images).reshape(-1, self.img_bytes)).to(device)
        self.labels = torch.from_numpy(np.ascontiguousarray(part_labels).reshape(-1).astype(np.int64)).to(device)
        self.test_images = torch.from_numpy(np.ascontiguousarray(test_images).reshape(len(test_labels), -1)).to(device)
        self.test_labels = torch.from_numpy(np.ascontiguousarray(test_labels).astype(np.int64)).to(device)
        # local partition index of a global node id
        ids = list(node_ids) if node_ids is not None else list(range(first_node, first_node + self.n_nodes))
        self.local_of = {int(n): i for i, n in enumerate(ids)}
        self.sizes = None           # per-partition sample counts when partitions are ragged (cifar_cache.open_cache)

    @classmethod
    def synthetic(cls, layout: ModelLayout, n_nodes: int, n_i: int, n_test: int, device, seed: int = 42,
                  data_seed: int = 1234, node_ids: Optional[Sequence[int]] = None,
                  per_node_seed: bool = False) -> "DeviceDataset":
        """Synthetic partitions for `node_ids` (default: all `n_nodes`), identical on every rank.
        `per_node_seed`: node p's partition is `make_images(n_i, seed=data_seed + p)` — generated node by node
        (host memory stays O(one partition)), for large runs that are not compared with a whole-population fixture."""
        if per_node_seed:
            ids = list(node_ids) if node_ids is not None else list(range(n_nodes))
            imgs = np.empty((len(ids), n_i, layout.img_hw, layout.img_hw, layout.img_c), dtype=np.uint8)
            labs = np.empty((len(ids), n_i), dtype=np.int64)
            for i, p in enumerate(ids):
                imgs[i], labs[i] = synthetic.make_images(n_i, layout.img_hw, layout.img_c, layout.num_classes,
                                                         seed=data_seed + int(p))
            timg, tlab = synthetic.make_test_split(n_test, layout.img_hw, layout.img_c, layout.num_classes,
                                                   seed=data_seed)
            return cls(imgs, labs, timg, tlab, device, seed, node_ids=node_ids)
        imgs, labs = synthetic.make_partitions(n_nodes, n_i, layout.img_hw, layout.img_c, layout.num_classes,
                                               seed=data_seed)
        timg, tlab = synthetic.make_test_split(n_test, layout.img_hw, layout.img_c, layout.num_classes,
                                               seed=data_seed)
        if node_ids is not None:
            sel = np.array(list(node_ids), dtype=np.int64)
            imgs, labs = imgs[sel], labs[sel]
        return cls(imgs, labs, timg, tlab, device, seed, node_ids=node_ids)

    def minibatches(self, peers: Sequence[int], rounds: Sequence[int], batch: int, steps: Sequence[int]
                    ) -> Tuple[np.ndarray, np.ndarray]:
        """Host arrays `idx[step, task, slot]` (global image id in `self.images`, -1 = empty) and
        `count[step, task]`.  A task runs min(steps[t], ceil(n_i / batch)) steps; the last batch
        may be partial (dasklearn/model_trainer.py:99-102)."""
        T, S = len(peers), (max(steps) if len(steps) else 0)
        base = np.array([self.local_of[int(q)] * self.n_i for q in peers], dtype=np.int64).reshape(T, 1)
        if self.sizes is not None:
            return self._minibatches_ragged(peers, rounds, batch, steps, base)
        if self.permutation is synthetic.minibatch_permutation:
            perm = synthetic.minibatch_permutations(self.seed, peers, rounds, self.n_i)     # one argsort per wave
        else:
            perm = np.stack([self.permutation(self.seed, int(q), int(r), self.n_i) for q, r in zip(peers, rounds)]) \
                if T else np.zeros((0, self.n_i), np.int32)
        steps_a = np.asarray(steps, dtype=np.int64).reshape(T, 1)
        m = min(self.n_i, S * batch)
        padded = -np.ones((T, S * batch), dtype=np.int64)
        padded[:, :m] = perm[:, :m] + base
        padded[np.arange(S * batch).reshape(1, -1) >= steps_a * batch] = -1          # per-task local_steps
        idx = np.ascontiguousarray(padded.reshape(T, S, batch).transpose(1, 0, 2)).astype(np.int32)
        cnt = (idx >= 0).sum(axis=2).astype(np.int32)
        return idx, cnt

    def _minibatches_ragged(self, peers, rounds, batch, steps, base):
        """Partitions of different sizes (real data, Dirichlet splits): node q draws a permutation of ITS n_q samples;
        it runs min(steps, ceil(n_q / batch)) steps and its last batch may be partial (model_trainer.py:99-102)."""
        T, S = len(peers), (max(steps) if len(steps) else 0)
        idx = -np.ones((S, T, batch), dtype=np.int32)
        for t, (q, r) in enumerate(zip(peers, rounds)):
            n_q = int(self.sizes[self.local_of[int(q)]])
            perm = self.permutation(self.seed, int(q), int(r), n_q)[:int(steps[t]) * batch]
            full = np.full(S * batch, -1, dtype=np.int64)
            full[:len(perm)] = perm + base[t, 0]
            idx[:, t, :] = full.reshape(S, batch)
        cnt = (idx >= 0).sum(axis=2).astype(np.int32)
        return idx, cnt
