"""Wave executor: the stacked parameter tensor in HBM plus batched train / aggregate / test.

Replaces the per-node learning core the reference worker calls one task at a time
(`ModelTrainer` dasklearn/model_trainer.py, `ModelManager`/`FedAvg`
dasklearn/model_manager.py + gradient_aggregation/fedavg.py, `ModelEvaluator`
dasklearn/model_evaluator.py).  State that the reference keeps in process globals
(`model_trainers[peer].optimizer`, dasklearn/functions.py:22-25,50-51) lives here as the momentum
matrix `M[node]` and the `mom_valid[node]` flags.
"""
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import kernels as K
from ._lib import DlsError
from .dataset import DeviceDataset
from .layout import ModelLayout, layout_for
from .models import initial_row
from .util import MICROSECONDS


class RowHandle:
    """Opaque result of a task: one row (slot) of the executor's stacked tensor.  This is what
    the API-compatible `train/aggregate/test` functions return in place of an `nn.Module`."""
    __slots__ = ("executor", "slot", "grad_slot")

    def __init__(self, executor: "WaveExecutor", slot: int, grad_slot: Optional[int] = None):
        self.executor = executor
        self.slot = slot
        self.grad_slot = grad_slot      # row holding `model.gradient` (results of compute_gradient tasks)

    def gradient(self) -> Optional[np.ndarray]:
        if self.grad_slot is None:
            return None
        return self.executor.X[self.grad_slot, :self.executor.layout.P].cpu().numpy()

    def numpy(self) -> np.ndarray:
        return self.executor.X[self.slot, :self.executor.layout.P].cpu().numpy()

    def state_dict(self) -> Dict[str, torch.Tensor]:
        from .layout import unflatten_row
        return unflatten_row(self.executor.layout, self.executor.X[self.slot].cpu())

    def __repr__(self):
        return "<RowHandle slot=%d>" % self.slot


class WaveExecutor:
    # multi-GPU state (class-level defaults so that test stand-ins that skip __init__ get them too)
    transport = None            # "a2a" | "push", settled by the first DistributedSolver
    symm = None                 # push transport: symmetric-memory handle of X
    pool_cap = 0
    halo_cap = 0

    def __init__(self, settings, dataset: DeviceDataset, device=None, n_slots: Optional[int] = None,
                 node_ids: Optional[Sequence[int]] = None):
        if not torch.cuda.is_available():
            raise DlsError("the B200 executor needs a CUDA device (there is no CPU fallback)")
        self.settings = settings
        self.layout: ModelLayout = layout_for(settings.dataset, settings.model)
        self.device = device or torch.device(settings.torch_device_name if "cuda" in str(settings.torch_device_name)
                                             else "cuda:0")
        self.dataset = dataset
        self.precision = int(getattr(settings, "precision", 0))
        # nodes this executor owns (multi-GPU: a contiguous block per rank)
        ids = list(node_ids) if node_ids is not None else list(range(settings.participants))
        self.node_ids = ids
        self.local_of = {int(n): i for i, n in enumerate(ids)}
        n_local = len(ids)
        self.n_slots = n_slots or (2 * n_local + 2)
        ld = self.layout.ld
        self.halo_rows = 0          # rows after the slot pool that receive other GPUs' models
        self.sendbuf = None         # staging rows for the halo exchange
        self.X = torch.zeros((self.n_slots, ld), dtype=torch.float32, device=self.device)
        self.free: List[int] = list(range(self.n_slots - 1, -1, -1))
        self.refs = np.zeros(self.n_slots, dtype=np.int64)
        mu = float(settings.learning.momentum)
        self.M = torch.zeros((n_local, ld), dtype=torch.float32, device=self.device) if mu != 0 else None
        self.mom_valid = np.zeros(n_local, dtype=bool)
        self.init_row = initial_row(self.layout, settings.seed).to(self.device)
        self.ws = K.Workspace(self.device)
        self.stats = {"train_tasks": 0, "train_steps": 0, "agg_rows": 0, "test_tasks": 0}

    # -- slot pool --------------------------------------------------------------------------
    def alloc(self, n: int = 1) -> List[int]:
        if len(self.free) < n:
            self._grow(max(n - len(self.free), self.n_slots // 2))
        return [self.free.pop() for _ in range(n)]

    def _grow(self, extra: int, halo_rows: int = None, send_rows: int = 0):
        """Add `extra` pool slots and keep `halo_rows` receive rows after the pool (X is one
        allocation so the aggregate kernel can index pool and halo rows alike)."""
        if self.symm is not None:
            # push transport: X is one fixed symmetric allocation [pool_cap + 2 * halo_cap rows]; the pool grows inside it
            if self.n_slots + extra > self.pool_cap:
                raise DlsError("slot pool of the symmetric allocation is full (%d rows); use transport='a2a'"
                               % self.pool_cap)
            if extra:
                self.refs = np.concatenate([self.refs, np.zeros(extra, dtype=np.int64)])
                self.free = list(range(self.n_slots + extra - 1, self.n_slots - 1, -1)) + self.free
                self.n_slots += extra
            return
        halo_rows = self.halo_rows if halo_rows is None else max(halo_rows, self.halo_rows)
        if extra or halo_rows != self.halo_rows:
            X2 = torch.zeros((self.n_slots + extra + halo_rows, self.layout.ld), dtype=torch.float32,
                             device=self.device)
            X2[:self.n_slots].copy_(self.X[:self.n_slots])
            self.X = X2
            self.refs = np.concatenate([self.refs, np.zeros(extra, dtype=np.int64)])
            self.free = list(range(self.n_slots + extra - 1, self.n_slots - 1, -1)) + self.free
            self.n_slots += extra
            self.halo_rows = halo_rows
        if send_rows and (self.sendbuf is None or self.sendbuf.shape[0] < send_rows):
            self.sendbuf = torch.empty((send_rows, self.layout.ld), dtype=torch.float32, device=self.device)

    @property
    def halo_base(self) -> int:
        return self.pool_cap if self.symm is not None else self.n_slots

    def enable_symmetric(self, group, pool_cap: int, halo_cap: int) -> bool:
        """Move X into symmetric memory (torch.distributed._symmetric_memory: one allocation per GPU, mapped
        into every peer of `group`) laid out as [pool_cap pool rows | 2 x halo_cap halo rows] — two halo halves
        so that rows pushed for exchange e+1 never overwrite rows a slower GPU still reads for exchange e.
        COLLECTIVE: every rank of the group calls it at the same point with the same sizes.  Returns False
        (X untouched, on every rank alike) if any rank fails to set it up."""
        import torch.distributed as dist
        ok, X2, hdl = 1, None, None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            pool_cap = max(int(pool_cap), self.n_slots)
            X2 = symm_mem.empty((pool_cap + 2 * int(halo_cap), self.layout.ld), dtype=torch.float32, device=self.device)
            hdl = symm_mem.rendezvous(X2, group if group is not None else dist.group.WORLD)
        except Exception as e:                                    # noqa: BLE001 — any failure means "use a2a"
            import logging
            logging.getLogger(__name__).warning("symmetric memory unavailable (%s): halo exchange uses all_to_all", e)
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            return False
        X2.zero_()
        X2[:self.n_slots].copy_(self.X[:self.n_slots])
        self.X, self.symm = X2, hdl
        self.pool_cap, self.halo_cap = pool_cap, int(halo_cap)
        self.peer_x_ptrs = [int(p) for p in hdl.buffer_ptrs]
        torch.cuda.synchronize(self.device)
        hdl.barrier(channel=0)                                    # nobody pushes before every X is in place
        return True

    def push_rows(self, slots_dev: torch.Tensor, dst_ptrs_dev: torch.Tensor, cols=None):
        """`cols` = (first column, count): push only that column range of the rows (pipelined exchange)."""
        if cols is None:
            K.push_rows(self.X, slots_dev, dst_ptrs_dev, self.layout.P)
        else:
            c0, n = cols
            K.push_rows(self.X[:, c0:c0 + n], slots_dev, dst_ptrs_dev + 4 * c0, n)

    def symm_barrier(self):
        self.symm.barrier(channel=0)

    def gather_rows(self, slots_dev: torch.Tensor, n: int) -> torch.Tensor:
        """Pack rows `slots` of X into the first n rows of the send buffer."""
        dst = torch.arange(n, dtype=torch.int32, device=self.device)
        K.copy_rows(self.X, self.sendbuf, slots_dev, dst, self.layout.P)
        return self.sendbuf[:n]

    def retain(self, slot: int, n: int = 1):
        self.refs[slot] += n

    def release(self, slot: int, n: int = 1):
        self.refs[slot] -= n
        assert self.refs[slot] >= 0
        if self.refs[slot] == 0:
            self.free.append(slot)

    def _upload(self, *arrays) -> List[torch.Tensor]:
        """One asynchronous host->device copy for several small descriptor arrays: they are packed
        into one pinned staging block (torch's caching host allocator keeps the block alive until the
        copy has run), so the host never waits for kernels queued on the stream — a pageable
        cudaMemcpyAsync would."""
        arrs = [np.ascontiguousarray(a) for a in arrays]
        offs, n = [], 0
        for a in arrs:
            offs.append(n)
            n += (a.nbytes + 15) & ~15
        host = torch.empty(max(n, 16), dtype=torch.uint8, pin_memory=True)
        hv = host.numpy()
        for a, o in zip(arrs, offs):
            hv[o:o + a.nbytes] = a.reshape(-1).view(np.uint8)
        dev = host.to(self.device, non_blocking=True)
        out = []
        for a, o in zip(arrs, offs):
            out.append(dev[o:o + a.nbytes].view(getattr(torch, a.dtype.name)).reshape(a.shape))
        return out

    def _dev_i32(self, a) -> torch.Tensor:
        return self._upload(np.asarray(a, dtype=np.int32))[0]

    # -- waves ------------------------------------------------------------------------------
    def init_rows(self, slots: Sequence[int]):
        """Identical initial model for every node (dasklearn/functions.py:46-48)."""
        if len(slots):
            K.broadcast_row(self.init_row, self.X, self._dev_i32(slots), self.layout.P)

    def copy_rows(self, src: Sequence[int], dst: Sequence[int]):
        if len(src):
            K.copy_rows(self.X, self.X, self._dev_i32(src), self._dev_i32(dst), self.layout.P)

    def train_wave(self, slots: Sequence[int], peers: Sequence[int], rounds: Sequence[int],
                   steps: Optional[Sequence[int]] = None, prepared=None, loss_out=None):
        """In-place local SGD on `slots` for nodes `peers` (functions.train for every task)."""
        ls = self.settings.learning
        T = len(slots)
        if T == 0:
            return
        if prepared is None:
            prepared = self.prepare_train(peers, rounds, steps)
        mb_index, mb_count, n_steps_done = prepared
        local = [self.local_of[int(p)] for p in peers]
        if len(set(local)) != len(local):
            raise DlsError("a train wave holds two tasks of the same node; they must run in different waves")
        first = np.array([0 if self.mom_valid[l] else 1 for l in local], dtype=np.uint8)
        d_slots, d_local, d_first = self._upload(np.asarray(slots, dtype=np.int32),
                                                 np.asarray(local, dtype=np.int32), first)
        K.train_wave(self.layout, self.X, self.M, d_slots, d_local, d_first, self.dataset.images,
                     self.dataset.labels, mb_index, mb_count, int(ls.batch_size), float(ls.learning_rate),
                     float(ls.momentum), float(ls.weight_decay), self.ws, loss_out=loss_out,
                     precision=self.precision)
        if self.M is not None:
            for l, n in zip(local, n_steps_done):
                if n > 0:
                    self.mom_valid[l] = True
        self.stats["train_tasks"] += T
        self.stats["train_steps"] += int(sum(n_steps_done))

    def prepare_train(self, peers, rounds, steps=None):
        """Host->device staging of a train wave's minibatch index (the analogue of the DataLoader)."""
        ls = self.settings.learning
        steps = list(steps) if steps is not None else [int(ls.local_steps)] * len(peers)
        idx, cnt = self.dataset.minibatches(peers, rounds, int(ls.batch_size), steps)
        done = (cnt > 0).sum(axis=0).tolist()
        ti, tc = self._upload(idx, cnt)
        return ti, tc, done

    def grad_wave(self, model_slots: Sequence[int], grad_slots: Sequence[int], peers: Sequence[int],
                  rounds: Sequence[int]):
        """`compute_gradient` for every task of the wave (dasklearn/functions.py:80-82 ->
        ModelTrainer.train(gradient_only=True), model_trainer.py:60-124): mean-loss gradient of the first
        minibatch of the (peer, round) loader order, written to row `grad_slots[t]`; models unchanged."""
        T = len(model_slots)
        if T == 0:
            return
        ls = self.settings.learning
        mb_index, mb_count, _ = self.prepare_train(peers, rounds, [1] * T)
        G = torch.empty((T, self.layout.ld), dtype=torch.float32, device=self.device)
        K.grad_wave(self.layout, self.X, G, self._dev_i32(model_slots), self.dataset.images, self.dataset.labels,
                    mb_index[0], mb_count[0], int(ls.batch_size), self.ws, precision=self.precision)
        K.copy_rows(G, self.X, self._dev_i32(np.arange(T)), self._dev_i32(grad_slots), self.layout.P)
        self.stats["grad_tasks"] = self.stats.get("grad_tasks", 0) + T

    def gradient_update_wave(self, model_slots: Sequence[int], grad_slots: Sequence[int], out_slots: Sequence[int]):
        """`gradient_update` for every task (functions.py:84-86 -> ModelTrainer.gradient_update,
        model_trainer.py:133-143): one step of a fresh SGD optimizer, p <- p - lr * (g + wd * p)."""
        if len(out_slots) == 0:
            return
        ls = self.settings.learning
        K.gradient_update(self.X, self._dev_i32(model_slots), self._dev_i32(grad_slots), self._dev_i32(out_slots),
                          self.layout.P, float(ls.learning_rate), float(ls.weight_decay))
        self.stats["gradient_updates"] = self.stats.get("gradient_updates", 0) + len(out_slots)

    def aggregate_wave(self, model_slots: Sequence[Sequence[int]], out_slots: Sequence[int],
                       weights: Optional[Sequence[Optional[Sequence[float]]]] = None, variant: int = 0,
                       prepared=None, cols=None):
        """FedAvg for every task of the wave: `model_slots[r]` is task r's `models` list as slots
        in list order (dasklearn/gradient_aggregation/fedavg.py:12-26)."""
        if len(out_slots) == 0:
            return
        if prepared is None:
            prepared = self.prepare_aggregate(model_slots, out_slots, weights)
        rowptr, colidx, wval, src_rows, out_rows, n_src, hint = prepared
        # auto: the variant the library says fits this wave shape (settled in prepare_aggregate)
        if cols is None:
            K.agg_spmm(self.X, self.X, rowptr, colidx, wval, self.layout.P, src_rows=src_rows,
                       out_rows=out_rows, n_src_rows=n_src, variant=variant or hint["variant"])
            self.stats["agg_rows"] += len(out_slots)
        else:                                                # a column range of the wave (pipelined exchange)
            c0, n = cols
            Xc = self.X[:, c0:c0 + n]
            K.agg_spmm(Xc, Xc, rowptr, colidx, wval, n, src_rows=src_rows, out_rows=out_rows, n_src_rows=n_src,
                       variant=variant or hint["variant"])
            if c0 == 0:
                self.stats["agg_rows"] += len(out_slots)

    def prepare_aggregate(self, model_slots, out_slots, weights=None):
        """CSR descriptors of an aggregate wave, built with array operations (a wave of 1024 tasks x fan-in 11
        has 11 k entries: a Python loop over them was a third of the host time of a round)."""
        import itertools
        n_rows = len(out_slots)
        lens = np.fromiter(map(len, model_slots), dtype=np.int64, count=n_rows)
        nnz = int(lens.sum())
        rowptr = np.zeros(n_rows + 1, dtype=np.int32)
        np.cumsum(lens, out=rowptr[1:])
        flat = np.fromiter(itertools.chain.from_iterable(model_slots), dtype=np.int64, count=nnz)
        src_rows64, col = np.unique(flat, return_inverse=True)
        wv = np.repeat(1.0 / np.maximum(lens, 1), lens)                  # fedavg.py:14-15, float(1/len) per task
        if weights is not None:
            for r, ws in enumerate(weights):
                if ws:
                    if len(ws) != lens[r]:
                        raise AssertionError("len(weights) != len(models)")   # fedavg.py:17
                    wv[rowptr[r]:rowptr[r + 1]] = ws
        outs = np.asarray(out_slots, dtype=np.int32)
        if np.intersect1d(src_rows64, outs).size:
            raise DlsError("aggregate wave would overwrite one of its own inputs")
        uniform = n_rows > 0 and int(lens.min()) == int(lens.max()) and 2 <= int(lens[0]) <= 16
        one_weight = uniform and (weights is None or not any(weights))
        # shared-memory staged kernels when the wave fits (uniform fan-in -> the specialised ones), else the
        # L2 gather kernel: asked of the library (dls_agg_variant_fits), no trial launches
        hint = {"variant": self._agg_variant(n_rows, int(src_rows64.size), nnz, uniform, one_weight)}
        up = self._upload(rowptr, col.astype(np.int32), wv.astype(np.float32), src_rows64.astype(np.int32), outs)
        return (*up, int(src_rows64.size), hint)

    def _agg_variant(self, n_rows: int, n_src: int, nnz: int, uniform: bool, one_weight: bool = False) -> int:
        for v in ((4, 3, 2) if one_weight else (3, 2) if uniform else (2,)):
            if K.agg_variant_fits(v, n_rows, n_src, nnz):
                return v
        return 1

    def reconstruct_wave(self, segments: Sequence[Tuple[int, int, int, Sequence[int]]]):
        """`reconstruct_from_chunks` for every task of the wave (dasklearn/functions.py:142-146 ->
        ChunkManager.reconstruct_model, simulation/conflux/chunk_manager.py:34-53): `segments` holds, per
        (task, chunk index), (output slot, first column, length, contributor slots in arrival order)."""
        if len(segments) == 0:
            return
        ptr = np.zeros(len(segments) + 1, dtype=np.int32)
        src: List[int] = []
        outs = set()
        for i, (o, c0, ln, ss) in enumerate(segments):
            src.extend(int(x) for x in ss)
            ptr[i + 1] = len(src)
            outs.add(int(o))
        if outs & set(src):
            raise DlsError("reconstruct wave would overwrite one of its own inputs")
        d_ptr, d_src, d_out, d_c0, d_len = self._upload(
            ptr, np.asarray(src, dtype=np.int32), np.asarray([s[0] for s in segments], dtype=np.int32),
            np.asarray([s[1] for s in segments], dtype=np.int64), np.asarray([s[2] for s in segments], dtype=np.int32))
        K.chunk_mean(self.X, self.X, d_ptr, d_src, d_out, d_c0, d_len, max(s[2] for s in segments),
                     int(np.diff(ptr).max()))
        self.stats["chunk_segments"] = self.stats.get("chunk_segments", 0) + len(segments)

    def test_wave(self, slots: Sequence[int]) -> Tuple[np.ndarray, np.ndarray]:
        """(accuracy %, mean loss) per slot on the whole test split
        (ModelEvaluator.evaluate_accuracy, dasklearn/model_evaluator.py:26-64)."""
        if len(slots) == 0:
            return np.zeros(0), np.zeros(0)
        correct, loss = K.eval_wave(self.layout, self.X, self._dev_i32(slots), self.dataset.test_images,
                                    self.dataset.test_labels, self.ws, precision=self.precision)
        n = float(self.dataset.test_labels.numel())
        acc = correct.cpu().numpy().astype(np.float64) / n * 100.0
        ls = loss.cpu().numpy() / n
        self.stats["test_tasks"] += len(slots)
        return acc, ls

    # -- accuracy CSV, same rows as dasklearn/functions.py:121-128 --------------------------------
    def write_accuracy(self, peer: int, round_nr: int, cur_time: int, accuracy: float, loss: float):
        s = self.settings
        data_dir = getattr(s, "data_dir", "") or "."
        os.makedirs(data_dir, exist_ok=True)
        path = os.path.join(data_dir, "accuracies_" + str(peer) + ".csv")
        if not os.path.exists(path):
            with open(path, "w") as f:
                f.write("algorithm,dataset,partitioner,alpha,peer,round,time,accuracy,loss\n")
        with open(path, "a") as f:
            f.write("%s,%s,%s,%g,%d,%d,%.2f,%f,%f\n" % (
                s.algorithm, s.dataset, getattr(s, "partitioner", "iid"), getattr(s, "alpha", 1), peer, round_nr,
                cur_time / MICROSECONDS, accuracy, loss))
