"""Thin torch-tensor front end of the C-ABI wave kernels.  PyTorch only provides device memory
and the stream; all arithmetic happens in csrc/ (hand-written sm_100a CUDA)."""
from typing import Optional

import torch

from . import _lib
from .layout import ModelLayout


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _lib.DlsError("dls_b200 kernels need CUDA tensors (no CPU fallback)")


def _i32(t):
    assert t.dtype == torch.int32 and t.is_contiguous()
    return t


def agg_spmm(X: torch.Tensor, Y: torch.Tensor, rowptr, colidx, wval, P: int, src_rows=None, out_rows=None,
             n_src_rows: Optional[int] = None, variant: int = 0):
    """Aggregate wave (FedAvg.aggregate for every row of the wave, fedavg.py:12-26)."""
    _need_cuda(X, Y, rowptr, colidx, wval, src_rows, out_rows)
    assert X.dtype == torch.float32 and Y.dtype == torch.float32 and wval.dtype == torch.float32
    n_rows = rowptr.numel() - 1
    if n_src_rows is None:
        n_src_rows = src_rows.numel() if src_rows is not None else X.shape[0]
    rc = _lib.lib().dls_agg_spmm_f32(X.data_ptr(), X.stride(0), Y.data_ptr(), Y.stride(0),
                                     _i32(rowptr).data_ptr(), _i32(colidx).data_ptr(), wval.data_ptr(),
                                     _lib.ptr(src_rows), _lib.ptr(out_rows), n_rows, n_src_rows, int(colidx.numel()), P, variant,
                                     _stream())
    _lib.check(rc, "dls_agg_spmm_f32")


def agg_variant_fits(variant: int, n_rows: int, n_src_rows: int, nnz: int) -> bool:
    """Does aggregate `variant` (2 staged, 3 staged + uniform fan-in) accept this wave shape on the current device?"""
    rc = _lib.lib().dls_agg_variant_fits(variant, n_rows, n_src_rows, nnz)
    if rc < 0:
        _lib.check(rc, "dls_agg_variant_fits")
    return rc == 1


def chunk_mean(X, Y, seg_ptr, seg_src_rows, seg_out_row, seg_col0, seg_len, max_len: int, max_contributors: int):
    """reconstruct_from_chunks wave: per segment the torch-order mean of a column range over its contributor rows."""
    _need_cuda(X, Y, seg_ptr, seg_src_rows, seg_out_row, seg_col0, seg_len)
    assert seg_col0.dtype == torch.int64
    rc = _lib.lib().dls_chunk_mean_f32(X.data_ptr(), X.stride(0), Y.data_ptr(), Y.stride(0), _i32(seg_ptr).data_ptr(),
                                       _i32(seg_src_rows).data_ptr(), _i32(seg_out_row).data_ptr(),
                                       seg_col0.data_ptr(), _i32(seg_len).data_ptr(), seg_len.numel(), int(max_len),
                                       int(max_contributors), _stream())
    _lib.check(rc, "dls_chunk_mean_f32")


def copy_rows(X, Y, src_rows, dst_rows, P: int):
    _need_cuda(X, Y, src_rows, dst_rows)
    rc = _lib.lib().dls_copy_rows(X.data_ptr(), X.stride(0), Y.data_ptr(), Y.stride(0), _lib.ptr(src_rows),
                                  _lib.ptr(dst_rows), dst_rows.numel(), P, _stream())
    _lib.check(rc, "dls_copy_rows")


def push_rows(X, src_rows, dst_ptrs, P: int):
    """Halo push: X[src_rows[e]] -> the (possibly peer-GPU) row at device address dst_ptrs[e] (int64 tensor)."""
    _need_cuda(X, src_rows, dst_ptrs)
    assert dst_ptrs.dtype == torch.int64 and dst_ptrs.numel() == src_rows.numel()
    rc = _lib.lib().dls_push_rows(X.data_ptr(), X.stride(0), _i32(src_rows).data_ptr(), dst_ptrs.data_ptr(),
                                  src_rows.numel(), P, _stream())
    _lib.check(rc, "dls_push_rows")


def broadcast_row(src, Y, dst_rows, P: int):
    _need_cuda(src, Y, dst_rows)
    rc = _lib.lib().dls_broadcast_row(src.data_ptr(), Y.data_ptr(), Y.stride(0), dst_rows.data_ptr(),
                                      dst_rows.numel(), P, _stream())
    _lib.check(rc, "dls_broadcast_row")


def gradient_update(X, p_rows, g_rows, out_rows, P: int, lr: float, wd: float):
    """X[out] = X[p] - lr * (X[g] + wd * X[p]) row-wise (functions.gradient_update for every task of a wave)."""
    _need_cuda(X, p_rows, g_rows, out_rows)
    rc = _lib.lib().dls_gradient_update(X.data_ptr(), X.stride(0), _i32(p_rows).data_ptr(), _i32(g_rows).data_ptr(),
                                        _i32(out_rows).data_ptr(), out_rows.numel(), P, float(lr), float(wd), _stream())
    _lib.check(rc, "dls_gradient_update")


class Workspace:
    """Caller-owned scratch buffer, grown on demand (torch allocates, the library never does)."""

    def __init__(self, device):
        self.device = device
        self.buf = None

    def get(self, nbytes: int) -> torch.Tensor:
        if nbytes < 0:
            _lib.check(int(nbytes), "workspace size query")
        if self.buf is None or self.buf.numel() < nbytes:
            self.buf = None
            self.buf = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self.buf


def train_wave(layout: ModelLayout, X, M, task_row, task_peer, mom_first, images, labels, mb_index, mb_count,
               batch: int, lr: float, mu: float, wd: float, ws: Workspace, loss_out=None, precision: int = 0):
    """Train wave: `mb_index` is `[steps, n_tasks, batch]` int32 global image ids, `mb_count`
    `[steps, n_tasks]` (functions.train -> ModelTrainer.train for every task of the wave)."""
    _need_cuda(X, M, task_row, task_peer, mom_first, images, labels, mb_index, mb_count, loss_out)
    n_steps, n_tasks = mb_count.shape
    assert mb_index.shape == (n_steps, n_tasks, batch) and mb_index.is_contiguous() and mb_count.is_contiguous()
    assert labels.dtype == torch.int64 and images.dtype == torch.uint8
    l = _lib.lib()
    nbytes = l.dls_train_workspace_bytes(layout.arch_id, n_tasks, batch)
    buf = ws.get(nbytes)
    rc = l.dls_train_wave(layout.arch_id, X.data_ptr(), X.stride(0), _lib.ptr(M), 0 if M is None else M.stride(0),
                          _i32(task_row).data_ptr(), _i32(task_peer).data_ptr(), _lib.ptr(mom_first),
                          images.data_ptr(), labels.data_ptr(), _i32(mb_index).data_ptr(),
                          _i32(mb_count).data_ptr(), n_tasks, n_steps, batch, lr, mu, wd, _lib.ptr(loss_out),
                          precision, buf.data_ptr(), buf.numel(), _stream())
    _lib.check(rc, "dls_train_wave")


def grad_wave(layout: ModelLayout, X, G, task_row, images, labels, mb_index, mb_count, batch: int,
              ws: Workspace, loss_out=None, precision: int = 0):
    _need_cuda(X, G, task_row, images, labels, mb_index, mb_count, loss_out)
    n_tasks = mb_count.numel()
    l = _lib.lib()
    nbytes = l.dls_train_workspace_bytes(layout.arch_id, n_tasks, batch)
    buf = ws.get(nbytes)
    rc = l.dls_grad_wave(layout.arch_id, X.data_ptr(), X.stride(0), G.data_ptr(), G.stride(0),
                         _i32(task_row).data_ptr(), images.data_ptr(), labels.data_ptr(),
                         _i32(mb_index).data_ptr(), _i32(mb_count).data_ptr(), n_tasks, batch,
                         _lib.ptr(loss_out), precision, buf.data_ptr(), buf.numel(), _stream())
    _lib.check(rc, "dls_grad_wave")


def eval_wave(layout: ModelLayout, X, model_rows, test_images, test_labels, ws: Workspace, out_logits=None,
              precision: int = 0):
    """Test wave: returns (correct int32 [n_models], loss_sum float64 [n_models]) device tensors."""
    _need_cuda(X, model_rows, test_images, test_labels, out_logits)
    n_models = model_rows.numel()
    n_test = test_labels.numel()
    l = _lib.lib()
    nbytes = l.dls_eval_workspace_bytes(layout.arch_id, n_models, n_test)
    buf = ws.get(nbytes)
    correct = torch.zeros(n_models, dtype=torch.int32, device=X.device)
    loss = torch.zeros(n_models, dtype=torch.float64, device=X.device)
    rc = l.dls_eval_wave(layout.arch_id, X.data_ptr(), X.stride(0), _i32(model_rows).data_ptr(), n_models,
                         test_images.data_ptr(), test_labels.data_ptr(), n_test, correct.data_ptr(),
                         loss.data_ptr(), _lib.ptr(out_logits), precision, buf.data_ptr(), buf.numel(), _stream())
    _lib.check(rc, "dls_eval_wave")
    return correct, loss
