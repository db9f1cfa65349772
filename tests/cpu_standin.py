"""Test-only stand-in for WaveExecutor that keeps rows in CPU tensors and does the wave
arithmetic with the oracle.  It exists so the multi-rank host logic (partitioning, halo plan,
all_to_all exchange, slot bookkeeping) can be exercised with world_size 2 on the gloo backend
without a GPU.  It is NOT a product path: the product executor refuses to run without CUDA."""
import numpy as np
import torch

from oracle import oracle as O
from dls_b200 import synthetic
from dls_b200.executor import WaveExecutor
from dls_b200.layout import layout_for


class OracleExecutor(WaveExecutor):
    def __init__(self, settings, part_images, part_labels, test_images, test_labels, node_ids, n_i):
        self.settings = settings
        self.layout = layout_for(settings.dataset, settings.model)
        self.device = torch.device("cpu")
        self.node_ids = list(node_ids)
        self.local_of = {int(n): i for i, n in enumerate(self.node_ids)}
        self.images, self.labels = part_images, part_labels          # indexed by GLOBAL node id
        self.test_images, self.test_labels = test_images, test_labels
        self.n_i = n_i
        self.n_slots = 2 * len(self.node_ids) + 2
        self.halo_rows, self.sendbuf = 0, None
        self.X = torch.zeros((self.n_slots, self.layout.ld), dtype=torch.float32)
        self.free = list(range(self.n_slots - 1, -1, -1))
        self.refs = np.zeros(self.n_slots, dtype=np.int64)
        self.moms = {}
        self.init_row = torch.from_numpy(O.initial_row(self.layout, settings.seed))
        self.stats = {"train_tasks": 0, "train_steps": 0, "agg_rows": 0, "test_tasks": 0}

    def _upload(self, *arrays):
        """CPU stand-in of WaveExecutor._upload (the product one stages through pinned memory to a CUDA device)."""
        return [torch.from_numpy(np.ascontiguousarray(a).copy()) for a in arrays]

    def _agg_variant(self, n_rows, n_src, nnz, uniform, one_weight=False):
        return 1

    def init_rows(self, slots):
        for s in slots:
            self.X[s] = self.init_row

    def copy_rows(self, src, dst):
        for s, d in zip(src, dst):
            self.X[d] = self.X[s]

    def gather_rows(self, slots_dev, n):
        self.sendbuf[:n] = self.X[slots_dev.long()]
        return self.sendbuf[:n]

    def prepare_train(self, peers, rounds, steps=None):
        ls = self.settings.learning
        steps = list(steps) if steps is not None else [ls.local_steps] * len(peers)
        return peers, rounds, steps

    def train_wave(self, slots, peers, rounds, steps=None, prepared=None, loss_out=None):
        ls = self.settings.learning
        for t, (slot, peer, rnd) in enumerate(zip(slots, peers, rounds)):
            n_steps = steps[t] if steps is not None else ls.local_steps
            idx, cnt = synthetic.minibatch_index(self.settings.seed, [peer], [rnd], self.n_i, ls.batch_size, n_steps)
            row, mom, _ = O.train_task(self.layout, self.X[slot].numpy(), self.moms.get(peer), self.images[peer],
                                       self.labels[peer], idx[0], cnt[0], ls.learning_rate, ls.momentum,
                                       ls.weight_decay)
            self.X[slot] = torch.from_numpy(row)
            self.moms[peer] = mom
            self.stats["train_steps"] += int((cnt[0] > 0).sum())

    def grad_wave(self, model_slots, grad_slots, peers, rounds):
        ls = self.settings.learning
        for slot, g, peer, rnd in zip(model_slots, grad_slots, peers, rounds):
            perm = synthetic.minibatch_permutation(self.settings.seed, peer, rnd, self.n_i)[:ls.batch_size]
            grad, _ = O.gradient_of(self.layout, self.X[slot].numpy(), self.images[peer][perm], self.labels[peer][perm])
            self.X[g] = 0
            self.X[g, :self.layout.P] = torch.from_numpy(grad)

    def gradient_update_wave(self, model_slots, grad_slots, out_slots):
        ls = self.settings.learning
        for s, g, o in zip(model_slots, grad_slots, out_slots):
            self.X[o] = torch.from_numpy(O.gradient_update_task(self.X[s].numpy(), self.X[g].numpy(), self.layout.P,
                                                                ls.learning_rate, ls.weight_decay))

    def aggregate_wave(self, model_slots, out_slots, weights=None, variant=0, prepared=None):
        rowptr, colidx, wval, src_rows, out_rows, n_src, _hint = prepared or self.prepare_aggregate(model_slots, out_slots,
                                                                                          weights)
        Xs = self.X[src_rows.long()].numpy()
        Y = O.fedavg_csr(Xs, rowptr.numpy(), colidx.numpy(), wval.numpy(), self.layout.P)
        for r, o in enumerate(out_rows.tolist()):
            self.X[o] = torch.from_numpy(Y[r])

    def reconstruct_wave(self, segments):
        for o, c0, ln, srcs in segments:
            self.X[o, c0:c0 + ln] = torch.from_numpy(O.mean_stack([self.X[s, c0:c0 + ln].numpy() for s in srcs]))

    def test_wave(self, slots):
        acc, loss = [], []
        for s in slots:
            a, l, _ = O.evaluate(self.layout, self.X[s].numpy(), self.test_images, self.test_labels)
            acc.append(a); loss.append(l)
        return np.array(acc), np.array(loss)


class PushOracleExecutor(OracleExecutor):
    """Stand-in for the symmetric-memory push transport, all ranks as THREADS of one process: every rank's X is a
    CPU tensor laid out like the symmetric allocation ([pool_cap pool rows | 2 x halo_cap halo rows]), the "peer
    pointers" are the tensors' real addresses, a push is a memmove to the address the planner computed, and the
    cross-GPU barrier is a threading.Barrier.  Exercises the planner's halo slot numbering, the two halo halves and
    the transfer lists without a GPU."""

    def setup_push(self, pool_cap, halo_cap, barrier):
        X2 = torch.zeros((pool_cap + 2 * halo_cap, self.layout.ld), dtype=torch.float32)
        X2[:self.n_slots] = self.X[:self.n_slots]
        self.X = X2
        self.symm = object()                       # symmetric mode: _grow stays inside the allocation, halo_base = pool_cap
        self.pool_cap, self.halo_cap = pool_cap, halo_cap
        self._barrier = barrier
        self.transport = "push"
        self.pushed_to = []                        # destination addresses in list order (what the test inspects)

    def push_rows(self, slots_dev, dst_ptrs_dev, cols=None):
        import ctypes
        assert cols is None
        for s, d in zip(slots_dev.tolist(), dst_ptrs_dev.tolist()):
            ctypes.memmove(d, self.X[s].data_ptr(), self.layout.P * 4)
            self.pushed_to.append(d)

    def symm_barrier(self):
        self._barrier.wait(timeout=120)
