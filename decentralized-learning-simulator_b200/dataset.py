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

    @classmethod
    def synthetic(cls, layout: ModelLayout, n_nodes: int, n_i: int, n_test: int, device, seed: int = 42,
                  data_seed: int = 1234, node_ids: Optional[Sequence[int]] = None) -> "DeviceDataset":
        """Synthetic partitions for `node_ids` (default: all `n_nodes`), identical on every rank."""
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
        idx = -np.ones((S, T, batch), dtype=np.int32)
        cnt = np.zeros((S, T), dtype=np.int32)
        for t, (peer, rnd) in enumerate(zip(peers, rounds)):
            perm = self.permutation(self.seed, int(peer), int(rnd), self.n_i)
            base = self.local_of[int(peer)] * self.n_i
            for s in range(steps[t]):
                c = perm[s * batch:(s + 1) * batch]
                if len(c) == 0:
                    break
                idx[s, t, :len(c)] = c + base
                cnt[s, t] = len(c)
        return idx, cnt
