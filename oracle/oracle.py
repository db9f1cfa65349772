"""ORACLE — test infrastructure only.

CPU restatement (numpy / plain fp32 PyTorch) of the reference's train / aggregate / test task
arithmetic.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this module; it is the checker, never the product path.

Parity pin: `tests/test_oracle_golden.py` checks every function here against fixtures under
`tests/golden/` that were produced by running the UNMODIFIED reference package from
`/root/reference` (see `tests/golden/make_golden.py`).  The reference's own tests hold no
numerical vectors for this path, so reference-generated fixtures are the pin.
FEMNIST: the reference's trainer / evaluator refuse the dataset (`model_trainer.py:77-78`,
`model_evaluator.py:36-37`) and its model file imports a module that is not in the tree
(`dasklearn/models/femnist.py:5`).  With that module supplied as a two-constant shim the reference's own
`CNN` class runs, and `tests/golden/femnist_kat.npz` (init, logits, loss, autograd gradients, three
optimizer steps — all produced by the reference's model + optimizer classes) pins the FEMNIST functions
here; only the evaluation loop for FEMNIST (argmax accuracy + CE sum) remains a restatement without a pin.
"""
import ctypes
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

_HERE = os.path.dirname(os.path.abspath(__file__))
_clib = None


def c_lib():
    """The C restatement of FedAvg (oracle/agg_oracle.c), built by `make -C oracle`."""
    global _clib
    if _clib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE, "-s"])
        _clib = ctypes.CDLL(path)
        _clib.oracle_fedavg_csr.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_int, ctypes.c_int64]
        _clib.oracle_fedavg_csr.restype = None
    return _clib


# ---------------------------------------------------------------------------------------------
# aggregate  — FedAvg.aggregate, dasklearn/gradient_aggregation/fedavg.py:12-26
# ---------------------------------------------------------------------------------------------

def fedavg_numpy(rows: Sequence[np.ndarray], weights: Optional[Sequence[float]] = None) -> np.ndarray:
    """`center = models[0]*0; for m, w: center += fl32(w) * m` in list order; numpy never fuses the
    multiply and the add, so this has the reference's two roundings per term."""
    if not weights:
        weights = [float(1.0 / len(rows))] * len(rows)        # fedavg.py:14-15 (python double)
    assert len(weights) == len(rows)                           # fedavg.py:17
    acc = rows[0].astype(np.float32) * np.float32(0.0)        # fedavg.py:20-22
    for x, w in zip(rows, weights):                            # fedavg.py:23-25
        acc = acc + (np.float32(w) * x.astype(np.float32))
    return acc


def fedavg_csr(X: np.ndarray, rowptr: np.ndarray, colidx: np.ndarray, wval: np.ndarray,
               P: int, out_rows: Optional[np.ndarray] = None, n_out: Optional[int] = None) -> np.ndarray:
    """A whole aggregate wave through the C oracle. X is `[S, ld]` fp32 (C-contiguous)."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
    colidx = np.ascontiguousarray(colidx, dtype=np.int32)
    wval = np.ascontiguousarray(wval, dtype=np.float32)
    n_rows = len(rowptr) - 1
    if n_out is None:
        n_out = n_rows if out_rows is None else int(out_rows.max()) + 1
    Y = np.zeros((n_out, X.shape[1]), dtype=np.float32)
    orp = None
    if out_rows is not None:
        out_rows = np.ascontiguousarray(out_rows, dtype=np.int32)
        orp = out_rows.ctypes.data
    c_lib().oracle_fedavg_csr(X.ctypes.data, X.shape[1], Y.ctypes.data, Y.shape[1], rowptr.ctypes.data,
                              colidx.ctypes.data, wval.ctypes.data, orp, n_rows, P)
    return Y


# ---------------------------------------------------------------------------------------------
# models — GNLeNet.forward dasklearn/models/cifar10.py:119-142; CNN.forward femnist.py:80-85
# ---------------------------------------------------------------------------------------------

def _views(layout, row: torch.Tensor):
    out = {}
    for key, shp, off in layout.entries:
        n = int(np.prod(shp))
        out[key] = row[off:off + n].view(shp)
    return out


def gnlenet_forward(p, x):
    """x: [B,3,32,32] contiguous NCHW fp32.  Block 1 is conv->pool->GN->ReLU, blocks 2 and 3 are
    conv->GN->pool->ReLU (cifar10.py:119-131); classifier + log_softmax (:135,140-142)."""
    x = F.conv2d(x, p["features.0.weight"], p["features.0.bias"], padding=2)
    x = F.max_pool2d(x, 3, 2)
    x = F.relu(F.group_norm(x, 2, p["features.2.weight"], p["features.2.bias"], eps=1e-5))
    x = F.conv2d(x, p["features.4.weight"], p["features.4.bias"], padding=2)
    x = F.group_norm(x, 2, p["features.5.weight"], p["features.5.bias"], eps=1e-5)
    x = F.relu(F.max_pool2d(x, 3, 2))
    x = F.conv2d(x, p["features.8.weight"], p["features.8.bias"], padding=2)
    x = F.group_norm(x, 2, p["features.9.weight"], p["features.9.bias"], eps=1e-5)
    x = F.relu(F.max_pool2d(x, 3, 2))
    x = torch.flatten(x, 1)
    x = F.linear(x, p["classifier.0.weight"], p["classifier.0.bias"])
    return F.log_softmax(x, dim=1)


def bf16r(t: torch.Tensor) -> torch.Tensor:
    """Round fp32 values to bf16 (nearest even) and back: what a tensor-core operand sees."""
    return t.to(torch.bfloat16).to(torch.float32)


class _ConvBf16(torch.autograd.Function):
    """5x5/pad-2 convolution as the precision-1 (tcgen05) kernels compute it: operands rounded to
    bf16, products accumulated in fp32.  forward: conv(bf16(x), bf16(w)) + b;  data gradient:
    conv_transpose(bf16(dy), bf16(w));  weight gradient: from bf16(dy), bf16(x) when `wgrad_bf16`
    (tensor-core weight-gradient kernel) else from the fp32 tensors (CUDA-core kernel)."""

    @staticmethod
    def forward(ctx, x, w, b, wgrad_bf16):
        ctx.save_for_backward(x, w)
        ctx.wgrad_bf16 = wgrad_bf16
        return F.conv2d(bf16r(x), bf16r(w), b, padding=2)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dx = torch.nn.grad.conv2d_input(x.shape, bf16r(w), bf16r(dy), padding=2) if ctx.needs_input_grad[0] else None
        if ctx.wgrad_bf16:
            dw = torch.nn.grad.conv2d_weight(bf16r(x), w.shape, bf16r(dy), padding=2)
        else:
            dw = torch.nn.grad.conv2d_weight(x, w.shape, dy, padding=2)
        return dx, dw, dy.sum(dim=(0, 2, 3)), None


def gnlenet_forward_bf16(p, x, wgrad_bf16=False, conv1_wgrad_bf16=False):
    """GNLeNet with the three convolutions evaluated like the precision-1 kernels (GroupNorm,
    pooling, classifier stay fp32).  Used to separate implementation errors from bf16 rounding."""
    x = _ConvBf16.apply(x, p["features.0.weight"], p["features.0.bias"], conv1_wgrad_bf16)
    x = F.max_pool2d(x, 3, 2)
    x = F.relu(F.group_norm(x, 2, p["features.2.weight"], p["features.2.bias"], eps=1e-5))
    x = _ConvBf16.apply(x, p["features.4.weight"], p["features.4.bias"], wgrad_bf16)
    x = F.group_norm(x, 2, p["features.5.weight"], p["features.5.bias"], eps=1e-5)
    x = F.relu(F.max_pool2d(x, 3, 2))
    x = _ConvBf16.apply(x, p["features.8.weight"], p["features.8.bias"], wgrad_bf16)
    x = F.group_norm(x, 2, p["features.9.weight"], p["features.9.bias"], eps=1e-5)
    x = F.relu(F.max_pool2d(x, 3, 2))
    x = torch.flatten(x, 1)
    x = F.linear(x, p["classifier.0.weight"], p["classifier.0.bias"])
    return F.log_softmax(x, dim=1)


def femnist_forward(p, x):
    """x: [B,1,28,28]; logits out (femnist.py:80-85)."""
    x = F.max_pool2d(F.relu(F.conv2d(x, p["conv1.weight"], p["conv1.bias"], padding=2)), 2, 2)
    x = F.max_pool2d(F.relu(F.conv2d(x, p["conv2.weight"], p["conv2.bias"], padding=2)), 2, 2)
    x = torch.flatten(x, 1)
    x = F.relu(F.linear(x, p["fc1.weight"], p["fc1.bias"]))
    return F.linear(x, p["fc2.weight"], p["fc2.bias"])


class _LinearBf16(torch.autograd.Function):
    """Dense layer as the precision-1 kernels compute it: every GEMM operand rounded to bf16, fp32 accumulation;
    the bias gradient from the fp32 output gradient."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        return bf16r(x) @ bf16r(w).t() + b

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        return bf16r(dy) @ bf16r(w), bf16r(dy).t() @ bf16r(x), dy.sum(dim=0)


def femnist_forward_bf16(p, x):
    """FEMNIST CNN with conv2 and fc1 evaluated like the precision-1 kernels (bf16 operands, fp32 accumulate,
    forward / data gradient / weight gradient alike); conv1 and fc2 stay fp32."""
    x = F.max_pool2d(F.relu(F.conv2d(x, p["conv1.weight"], p["conv1.bias"], padding=2)), 2, 2)
    x = F.max_pool2d(F.relu(_ConvBf16.apply(x, p["conv2.weight"], p["conv2.bias"], True)), 2, 2)
    x = torch.flatten(x, 1)
    x = F.relu(_LinearBf16.apply(x, p["fc1.weight"], p["fc1.bias"]))
    return F.linear(x, p["fc2.weight"], p["fc2.bias"])


def to_tensor(images_u8: np.ndarray) -> torch.Tensor:
    """torchvision ToTensor + default_collate (dasklearn/datasets/transforms.py:9,16-18):
    uint8 HWC -> contiguous fp32 CHW / 255."""
    t = torch.from_numpy(np.ascontiguousarray(images_u8))
    return (t.permute(0, 3, 1, 2).contiguous().to(torch.float32) / 255.0).contiguous()


def forward(layout, row: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    p = _views(layout, row)
    return gnlenet_forward(p, x) if layout.name == "gnlenet" else femnist_forward(p, x)


def loss_fn(layout, out, target, reduction="mean"):
    # NLLLoss for cifar10 lenet/gnlenet, CrossEntropyLoss otherwise (model_trainer.py:112-118)
    if layout.log_softmax_output:
        return F.nll_loss(out, target, reduction=reduction)
    return F.cross_entropy(out, target, reduction=reduction)


# ---------------------------------------------------------------------------------------------
# train — ModelTrainer.train dasklearn/model_trainer.py:60-131 + SGDOptimizer optimizer/sgd.py:4-14
# ---------------------------------------------------------------------------------------------

def train_task(layout, row: np.ndarray, mom: Optional[np.ndarray], images_u8: np.ndarray,
               labels: np.ndarray, idx: np.ndarray, cnt: np.ndarray, lr: float, mu: float, wd: float
               ) -> Tuple[np.ndarray, Optional[np.ndarray], List[float]]:
    """One `train` task of one node.  `row`: flat params (len >= P); `mom`: the node's momentum
    buffers from its previous task or None (first ever task: torch.optim.SGD sets buf = grad);
    `idx[step, slot]`/`cnt[step]`: scripted minibatches into the node's partition.
    Update rule (torch.optim.SGD, dampening 0, no nesterov): g += wd*p; buf = g (first step) else
    buf = mu*buf + g; p -= lr*buf.  Momentum buffers persist across tasks of the same node
    (model_trainer.py:89-92).  Returns (new_row, new_mom, per-step losses)."""
    P = layout.P
    w = torch.from_numpy(np.array(row[:P], dtype=np.float32, copy=True))
    buf = None if mom is None else torch.from_numpy(np.array(mom[:P], dtype=np.float32, copy=True))
    losses = []
    for s in range(idx.shape[0]):
        b = int(cnt[s])
        if b <= 0:
            break                                                       # loader exhausted (:99-102)
        sel = idx[s, :b].astype(np.int64)
        x = to_tensor(images_u8[sel])
        y = torch.from_numpy(labels[sel].astype(np.int64))
        wl = w.clone().requires_grad_(True)
        out = forward(layout, wl, x)
        loss = loss_fn(layout, out, y)                                  # mean over delivered batch
        loss.backward()
        g = wl.grad
        with torch.no_grad():
            if wd != 0:
                g = g.add(w, alpha=wd)
            if mu != 0:
                if buf is None:
                    buf = g.clone()
                else:
                    buf.mul_(mu).add_(g)
                g = buf
            w = w.add(g, alpha=-lr)
        losses.append(float(loss.detach()))
    new_row = np.zeros_like(np.asarray(row, dtype=np.float32))
    new_row[:P] = w.numpy()
    new_mom = None
    if buf is not None:
        new_mom = np.zeros_like(new_row)
        new_mom[:P] = buf.numpy()
    return new_row, new_mom, losses


def gradient_of(layout, row: np.ndarray, images_u8: np.ndarray, labels: np.ndarray, emulate_bf16: bool = False,
                wgrad_bf16: bool = False, conv1_wgrad_bf16: bool = False) -> Tuple[np.ndarray, float]:
    """Mean-loss gradient of one minibatch (used to check the backward kernels in isolation).
    `emulate_bf16`: evaluate conv2/conv3 with bf16-rounded operands like the precision-1 kernels."""
    P = layout.P
    wl = torch.from_numpy(np.array(row[:P], dtype=np.float32, copy=True)).requires_grad_(True)
    if emulate_bf16 and layout.name != "gnlenet":
        out = femnist_forward_bf16(_views(layout, wl), to_tensor(images_u8))
    elif emulate_bf16:
        out = gnlenet_forward_bf16(_views(layout, wl), to_tensor(images_u8), wgrad_bf16, conv1_wgrad_bf16)
    else:
        out = forward(layout, wl, to_tensor(images_u8))
    loss = loss_fn(layout, out, torch.from_numpy(labels.astype(np.int64)))
    loss.backward()
    return wl.grad.numpy().copy(), float(loss.detach())


def gradient_update_task(row: np.ndarray, grad: np.ndarray, P: int, lr: float, wd: float) -> np.ndarray:
    """`gradient_update` (dasklearn/functions.py:84-86 -> ModelTrainer.gradient_update,
    model_trainer.py:133-143): ONE step of a FRESH torch.optim.SGD with the supplied gradient.  A fresh
    optimizer has no momentum buffer: buf = g (+ wd*p), p <- p - lr*buf whatever the momentum setting.
    torch's `p.add_(buf, alpha=-lr)` rounds once (fused); pinned by tests/golden/gradient_kat.npz."""
    p = torch.from_numpy(np.array(row[:P], dtype=np.float32, copy=True))
    g = torch.from_numpy(np.array(grad[:P], dtype=np.float32, copy=True))
    if wd != 0:
        g = g.add(p, alpha=wd)
    p = p.add(g, alpha=-lr)
    out = np.zeros_like(np.asarray(row, dtype=np.float32))
    out[:P] = p.numpy()
    return out


# ---------------------------------------------------------------------------------------------
# test — ModelEvaluator.evaluate_accuracy dasklearn/model_evaluator.py:26-64
# ---------------------------------------------------------------------------------------------

def evaluate(layout, row: np.ndarray, test_u8: np.ndarray, test_labels: np.ndarray,
             batch: int = 512) -> Tuple[float, float, int]:
    """(accuracy %, mean loss, n_correct): batches of 512, summed loss / n, argmax accuracy.
    (For FEMNIST the reference would raise — model_evaluator.py:36-37 — so CE-sum is a restatement.)"""
    P = layout.P
    w = torch.from_numpy(np.array(row[:P], dtype=np.float32, copy=True))
    correct = 0
    total = 0.0
    n = len(test_labels)
    with torch.no_grad():
        for s in range(0, n, batch):
            x = to_tensor(test_u8[s:s + batch])
            y = torch.from_numpy(test_labels[s:s + batch].astype(np.int64))
            out = forward(layout, w, x)
            total += float(loss_fn(layout, out, y, reduction="sum").item())
            pred = out.argmax(dim=1)
            correct += int(pred.eq(y).sum().item())
    return float(correct) / float(n) * 100.0, total / float(n), correct


def logits(layout, row: np.ndarray, images_u8: np.ndarray) -> np.ndarray:
    P = layout.P
    w = torch.from_numpy(np.array(row[:P], dtype=np.float32, copy=True))
    with torch.no_grad():
        return forward(layout, w, to_tensor(images_u8)).numpy()


def initial_row(layout, seed: int) -> np.ndarray:
    """The identical initial model every node starts from: default PyTorch init under
    `torch.manual_seed(settings.seed)` (dasklearn/functions.py:46-48).  Restated with plain
    torch.nn layers built in the reference's construction order so the RNG stream matches."""
    import torch.nn as nn
    torch.manual_seed(seed)
    if layout.name == "gnlenet":
        feats = nn.Sequential(
            nn.Conv2d(3, 32, 5, 1, 2), nn.MaxPool2d(3, 2), nn.GroupNorm(2, 32), nn.ReLU(True),
            nn.Conv2d(32, 32, 5, 1, 2), nn.GroupNorm(2, 32), nn.MaxPool2d(3, 2), nn.ReLU(True),
            nn.Conv2d(32, 64, 5, 1, 2), nn.GroupNorm(2, 64), nn.MaxPool2d(3, 2), nn.ReLU(True))
        cls = nn.Sequential(nn.Linear(576, 10))
        sd = {("features." + k): v for k, v in feats.state_dict().items()}
        sd.update({("classifier." + k): v for k, v in cls.state_dict().items()})
    else:
        conv1 = nn.Conv2d(1, 32, 5, padding=2)
        conv2 = nn.Conv2d(32, 64, 5, padding=2)
        fc1 = nn.Linear(7 * 7 * 64, 512)
        fc2 = nn.Linear(512, 62)
        sd = {}
        for name, m in (("conv1", conv1), ("conv2", conv2), ("fc1", fc1), ("fc2", fc2)):
            sd[name + ".weight"] = m.weight
            sd[name + ".bias"] = m.bias
    row = np.zeros(layout.ld, dtype=np.float32)
    for key, shp, off in layout.entries:
        t = sd[key].detach().reshape(-1).numpy()
        row[off:off + t.size] = t
    return row


# ---------------------------------------------------------------------------------------------
# chunk / reconstruct_from_chunks — dasklearn/functions.py:136-146 -> ChunkManager
# (dasklearn/simulation/conflux/chunk_manager.py:12-53)
# ---------------------------------------------------------------------------------------------

def chunk_bounds(P: int, n: int) -> List[Tuple[int, int]]:
    """Column ranges of ChunkManager.chunk_model (chunk_manager.py:13-25): n slices of P // n elements of the
    flat state_dict (== parameters() order for these buffer-free models); the last takes the remainder."""
    size = P // n
    b = [(i * size, (i + 1) * size) for i in range(n)]
    b[-1] = (b[-1][0], P)
    return b


def _cascade_sum(rows: np.ndarray) -> np.ndarray:
    """torch's CPU `multi_row_sum` (aten/src/ATen/native/cpu/SumKernel.cpp, torch 2.x) over axis 0: sequential
    accumulation from 0 with a 4-level cascade every 16 rows (level_step = max(16, ...))."""
    size = rows.shape[0]
    num_levels = 4
    level_power = max(4, (int(np.ceil(np.log2(size))) if size > 1 else 0) // num_levels)
    level_step = 1 << level_power
    level_mask = level_step - 1
    acc = [np.zeros(rows.shape[1], np.float32) for _ in range(num_levels)]
    i = 0
    while i + level_step <= size:
        for _ in range(level_step):
            acc[0] = acc[0] + rows[i]
            i += 1
        for j in range(1, num_levels):
            acc[j] = acc[j] + acc[j - 1]
            acc[j - 1] = np.zeros_like(acc[0])
            if (i & (level_mask << (j * level_power))) != 0:
                break
    while i < size:
        acc[0] = acc[0] + rows[i]
        i += 1
    for j in range(1, num_levels):
        acc[0] = acc[0] + acc[j]
    return acc[0]


def _ilp4_sum(rows: np.ndarray) -> np.ndarray:
    """torch's CPU `row_sum`: four interleaved partial sums (rows 4i+q), remainder rows into partial 0,
    then p0 + p1 + p2 + p3."""
    size = rows.shape[0]
    g = size // 4
    parts = [_cascade_sum(rows[q:4 * g:4]) if g else np.zeros(rows.shape[1], np.float32) for q in range(4)]
    for i in range(4 * g, size):
        parts[0] = parts[0] + rows[i]
    return ((parts[0] + parts[1]) + parts[2]) + parts[3]


def mean_stack(pieces: Sequence[np.ndarray]) -> np.ndarray:
    """`torch.mean(torch.stack(pieces), dim=0)` for fp32 pieces of equal length n, bit for bit as torch's CPU
    kernel (AVX2 dispatch, vector width 8) evaluates it — chunk_manager.py:40.  The reduction runs over the
    outer dimension of the contiguous [k, n] stack (`vectorized_outer_sum`): columns below the last multiple
    of 32 are summed sequentially over the contributors (with the 16-row cascade); the remaining columns go
    through `row_sum`, four interleaved partial sums.  The sum is then DIVIDED by k (sum(...).div_(k))."""
    rows = np.stack([np.asarray(p, dtype=np.float32) for p in pieces])
    k, n = rows.shape
    out = np.empty(n, np.float32)
    if n >= 8:
        nb = (n // 32) * 32
    else:
        nb = (n // 4) * 4                 # scalar_outer_sum: groups of four columns, then row_sum
    if nb:
        out[:nb] = _cascade_sum(rows[:, :nb])
    if nb < n:
        out[nb:] = _ilp4_sum(rows[:, nb:])
    return out / np.float32(k)


def reconstruct_from_chunks(P: int, chunks: Sequence[Sequence[np.ndarray]]) -> np.ndarray:
    """ChunkManager.reconstruct_model (chunk_manager.py:34-53): per chunk index the mean over the received
    pieces (in arrival order), concatenated back into the flat model."""
    for idx, c in enumerate(chunks):
        assert len(c), "No chunks received at index %d!" % idx            # chunk_manager.py:36
    flat = np.concatenate([mean_stack(c) for c in chunks])
    assert flat.size == P
    return flat


# ---------------------------------------------------------------------------------------------
# sequential compute-graph solve — Worker.start dasklearn/worker.py:21-38 and
# Broker.handle_task_result dasklearn/broker.py:277-290 ("1 broker / 1 worker")
# ---------------------------------------------------------------------------------------------

def solve_dag(layout, serialized_tasks, part_images, part_labels, test_images, test_labels, *, seed: int,
              batch: int, local_steps: int, lr: float, mu: float, wd: float, n_i: int,
              minibatch_permutation, init_row: Optional[np.ndarray] = None, timings: Optional[dict] = None):
    """Runs every task of a serialized WorkflowDAG one at a time, FIFO from the source tasks, the
    way one broker with one worker does.  Returns ({sink task name: row}, [accuracy rows])
    where an accuracy row is (peer, round, time_us, accuracy, loss)."""
    import time as _time
    tasks = {t["name"]: {"func": t["func"], "data": t["data"], "inputs": list(t["inputs"]),
                         "outputs": list(t["outputs"]), "resolved": 0} for t in serialized_tasks}
    values = {}
    grads = {}
    moms = {}
    acc_rows = []
    sinks = {}

    def _ref(r):
        if r is None:
            return None
        if isinstance(r, str):
            return (r, 0)
        r = tuple(r)
        return None if r[0] is None else r

    if init_row is None:
        init_row = initial_row(layout, seed)
    queue = [n for n, t in tasks.items() if not t["inputs"]]
    while queue:
        name = queue.pop(0)
        t = tasks[name]
        d = t["data"]
        t0 = _time.perf_counter()
        if t["func"] == "train":
            src = _ref(d["model"])
            row = init_row if src is None else values[src]
            peer, rnd = d["peer"], d["round"]
            steps = d.get("local_steps", local_steps)
            perm = minibatch_permutation(seed, peer, rnd, n_i)
            idx = -np.ones((steps, batch), dtype=np.int32)
            cnt = np.zeros((steps,), dtype=np.int32)
            for s in range(steps):
                c = perm[s * batch:(s + 1) * batch]
                idx[s, :len(c)] = c
                cnt[s] = len(c)
            row, mom, _ = train_task(layout, row, moms.get(peer), part_images[peer], part_labels[peer], idx, cnt,
                                     lr, mu, wd)
            moms[peer] = mom
            res = row
        elif t["func"] == "aggregate":
            rows = [values[tuple(m)] for m in d["models"]]
            res = fedavg_numpy(rows, d.get("weights"))
        elif t["func"] == "test":
            res = values[_ref(d["model"])]
            acc, loss, _ = evaluate(layout, res, test_images, test_labels)
            acc_rows.append((d["peer"], d["round"], d["time"], acc, loss))
        elif t["func"] == "compute_gradient":                       # functions.py:80-82
            src = _ref(d["model"])
            res = init_row if src is None else values[src]
            peer, rnd = d["peer"], d["round"]
            perm = minibatch_permutation(seed, peer, rnd, n_i)[:batch]
            g, _ = gradient_of(layout, res, part_images[peer][perm], part_labels[peer][perm])
            gfull = np.zeros_like(np.asarray(res, dtype=np.float32))
            gfull[:layout.P] = g
            grads[(name, 0)] = gfull
        elif t["func"] == "gradient_update":                        # functions.py:84-86
            res = gradient_update_task(values[_ref(d["model"])], grads[_ref(d["gradient_model"])], layout.P, lr, wd)
        elif t["func"] == "chunk":                                  # functions.py:136-140
            res = values[_ref(d["model"])]
            for ci, (c0, c1) in enumerate(chunk_bounds(layout.P, int(d["n"]))):
                values[(name, ci)] = res[c0:c1]
        elif t["func"] == "reconstruct_from_chunks":                # functions.py:142-146
            flat_ = reconstruct_from_chunks(layout.P, [[values[tuple(m)] for m in c] for c in d["chunks"]])
            res = np.zeros(layout.ld, dtype=np.float32)
            res[:layout.P] = flat_
        else:
            raise RuntimeError("Task function %s not found!" % t["func"])     # worker.py:26-27
        if timings is not None:
            timings.setdefault(t["func"], []).append(_time.perf_counter() - t0)
        if t["func"] != "chunk":
            values[(name, 0)] = res
        if not t["outputs"]:
            sinks[name] = res
        for nxt, idx_ in t["outputs"]:
            tasks[nxt]["resolved"] += 1
            if tasks[nxt]["resolved"] >= len(tasks[nxt]["inputs"]):
                queue.append(nxt)
    return sinks, acc_rows
