"""API-compatible task entry points: `train`, `compute_gradient`, `gradient_update`, `aggregate`, `test`,
`chunk`, `reconstruct_from_chunks`
with the signature `f(settings, params) -> [result]` of the reference (dasklearn/functions.py:28-77,
80-86, 89-106, 109-133, 136-146), looked up by name by the worker loop (`globals()[func_name]`, dasklearn/worker.py:27-31).

Differences that the DAG format hides: results are opaque `RowHandle`s (rows of the stacked
HBM tensor) instead of `nn.Module`s, and the dataset is a device-resident `DeviceDataset`
(set with `set_dataset`, the analogue of the module-global `dataset` of functions.py:23,40-41).
A single call is simply a wave of one task; the batched solver (`solver.BatchedSolver`) is the
fast path."""
import logging
import threading
from typing import Dict, List, Optional

from ._lib import DlsError
from .dataset import DeviceDataset
from .executor import RowHandle, WaveExecutor

logger = logging.getLogger(__name__)
executor: Optional[WaveExecutor] = None      # process-global like `model_trainers` (functions.py:22)
dataset: Optional[DeviceDataset] = None      # process-global like `dataset`        (functions.py:23)
lock = threading.Lock()


def set_dataset(ds: Optional[DeviceDataset]):
    global dataset, executor
    dataset = ds
    executor = None


def get_executor(settings) -> WaveExecutor:
    global executor
    with lock:
        if executor is None:
            if dataset is None:
                raise DlsError("no dataset: call dls_b200.functions.set_dataset(DeviceDataset...) first")
            executor = WaveExecutor(settings, dataset)
    return executor


def _slot_of(model, ex: WaveExecutor) -> int:
    if not isinstance(model, RowHandle) or model.executor is not ex:
        raise DlsError("model is not a RowHandle of this executor: %r" % (model,))
    return model.slot


def train(settings, params: Dict) -> List[RowHandle]:
    model = params["model"]
    if isinstance(model, tuple):        # first round: the tuple (None, 0)  (functions.py:30-32)
        model = model[0]
    peer_id = params["peer"]
    compute_grad = params["compute_gradient"] if "compute_gradient" in params else False
    gradient_model = params["gradient_model"] if "gradient_model" in params else None
    local_steps = params["local_steps"] if "local_steps" in params else settings.learning.local_steps
    ex = get_executor(settings)
    out = ex.alloc(1)[0]
    ex.retain(out)
    if not model:
        ex.init_rows([out])             # torch.manual_seed(seed); create_model(...)  (functions.py:46-48)
    else:
        ex.copy_rows([_slot_of(model, ex)], [out])   # the result is a detached copy (functions.py:64-70)
    if gradient_model:                  # functions.py:54-55 -> ModelTrainer.gradient_update
        if not isinstance(gradient_model, RowHandle) or gradient_model.grad_slot is None:
            raise DlsError("gradient_model carries no gradient: %r" % (gradient_model,))
        ex.gradient_update_wave([out], [gradient_model.grad_slot], [out])
        return [RowHandle(ex, out)]
    if compute_grad:                    # gradient_only=True: first minibatch, no optimizer step (model_trainer.py:122-124)
        g = ex.alloc(1)[0]
        ex.retain(g)
        ex.grad_wave([out], [g], [peer_id], [params["round"]])
        return [RowHandle(ex, out, g)]
    ex.train_wave([out], [peer_id], [params["round"]], [local_steps])
    return [RowHandle(ex, out)]


def compute_gradient(settings, params: Dict) -> List[RowHandle]:
    params["compute_gradient"] = True   # functions.py:80-82
    return train(settings, params)


def gradient_update(settings, params: Dict) -> List[RowHandle]:
    return train(settings, params)      # functions.py:84-86


def aggregate(settings, params: Dict) -> List[RowHandle]:
    models = params["models"]
    weights = params["weights"] if "weights" in params else None
    ex = get_executor(settings)
    out = ex.alloc(1)[0]
    ex.retain(out)
    ex.aggregate_wave([[_slot_of(m, ex) for m in models]], [out], [weights])
    return [RowHandle(ex, out)]


def test(settings, params: Dict) -> List[RowHandle]:
    model = params["model"]
    ex = get_executor(settings)
    slot = _slot_of(model, ex)
    acc, loss = ex.test_wave([slot])
    ex.write_accuracy(params["peer"], params["round"], params["time"], float(acc[0]), float(loss[0]))
    logger.info("[t=%.2f] Model accuracy (peer %d, round %d): %f, loss: %f", params["time"] / 1e6,
                params["peer"], params["round"], acc[0], loss[0])
    out = ex.alloc(1)[0]                # the reference returns a copy of the model (functions.py:132)
    ex.retain(out)
    ex.copy_rows([slot], [out])
    return [RowHandle(ex, out)]


class ChunkHandle:
    """One piece of a chunked model: columns [c0, c1) of a row (the analogue of the tensor slices returned by
    ChunkManager.chunk_model, simulation/conflux/chunk_manager.py:13-25)."""
    __slots__ = ("executor", "slot", "c0", "c1")

    def __init__(self, executor: WaveExecutor, slot: int, c0: int, c1: int):
        self.executor, self.slot, self.c0, self.c1 = executor, slot, c0, c1

    def numpy(self):
        return self.executor.X[self.slot, self.c0:self.c1].cpu().numpy()


def chunk(settings, params: Dict) -> List[ChunkHandle]:
    """functions.py:136-140: `n` pieces of P // n elements of the flat model, the last takes the remainder.
    The pieces alias the model's row (a copy of it, like the reference's slices of a fresh flat tensor)."""
    ex = get_executor(settings)
    n = int(params["n"])
    out = ex.alloc(1)[0]
    ex.retain(out)
    ex.copy_rows([_slot_of(params["model"], ex)], [out])
    P = ex.layout.P
    size = P // n
    return [ChunkHandle(ex, out, i * size, (i + 1) * size if i < n - 1 else P) for i in range(n)]


def reconstruct_from_chunks(settings, params: Dict) -> List[RowHandle]:
    """functions.py:142-146: params["chunks"][i] = the pieces received for chunk index i."""
    chunks = params["chunks"]
    ex = get_executor(settings)
    out = ex.alloc(1)[0]
    ex.retain(out)
    segs = []
    for idx, pieces in enumerate(chunks):
        assert pieces, "No chunks received at index %d!" % idx            # chunk_manager.py:36
        c0, c1 = pieces[0].c0, pieces[0].c1
        for p in pieces:
            if not isinstance(p, ChunkHandle) or p.executor is not ex or (p.c0, p.c1) != (c0, c1):
                raise DlsError("piece %r does not belong to chunk index %d" % (p, idx))
        segs.append((out, c0, c1 - c0, [p.slot for p in pieces]))
    if sum(s[2] for s in segs) != ex.layout.P:
        raise DlsError("the pieces do not cover the model")
    ex.reconstruct_wave(segs)
    return [RowHandle(ex, out)]


def release(handle: RowHandle):
    """Give a row back to the pool (the reference relies on Python GC of nn.Modules)."""
    handle.executor.release(handle.slot)
    if handle.grad_slot is not None:
        handle.executor.release(handle.grad_slot)
