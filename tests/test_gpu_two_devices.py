"""One process, two GPUs: the library caches per-device state (opt-in shared-memory attributes, SM count, side
streams, scratch) by device ordinal, so a C-ABI caller that moves from cuda:0 to cuda:1 must get the same bits on
both.  Skipped on a one-GPU box (the multi-GPU product path runs one process per GPU and never needs this)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
import dls_b200
from dls_b200 import kernels as K, synthetic
from dls_b200.layout import FEMNIST, GNLENET

pytestmark = pytest.mark.gpu


def _grad(layout, dev, precision, hw, c, classes):
    imgs, labs = synthetic.make_partitions(2, 40, hw, c, classes)
    base = O.initial_row(layout, 42)
    X = np.tile(base, (2, 1)).astype(np.float32)
    rs = np.random.RandomState(5)
    X[:, :layout.P] += 0.01 * rs.standard_normal((2, layout.P)).astype(np.float32)
    idx = np.stack([np.arange(32), 40 + np.arange(32)]).astype(np.int32)
    with torch.cuda.device(dev):
        Xd = torch.from_numpy(X).to(dev)
        G = torch.zeros((2, layout.ld), device=dev)
        loss = torch.zeros(2, device=dev)
        K.grad_wave(layout, Xd, G, torch.arange(2, dtype=torch.int32, device=dev),
                    torch.from_numpy(imgs.reshape(-1, hw * hw * c)).to(dev), torch.from_numpy(labs.reshape(-1)).to(dev),
                    torch.from_numpy(idx).to(dev), torch.tensor([32, 32], dtype=torch.int32, device=dev), 32,
                    K.Workspace(dev), loss_out=loss, precision=precision)
        # aggregate on the same device: the SpMM caches its own per-device limits
        Y = torch.zeros_like(Xd)
        K.agg_spmm(Xd, Y, torch.tensor([0, 2, 4], dtype=torch.int32, device=dev),
                   torch.tensor([0, 1, 1, 0], dtype=torch.int32, device=dev),
                   torch.tensor([0.5, 0.5, 0.5, 0.5], dtype=torch.float32, device=dev), layout.P)
        torch.cuda.synchronize(dev)
        return G.cpu().numpy(), loss.cpu().numpy(), Y.cpu().numpy()


@pytest.mark.parametrize("layout,hw,c,classes", [(GNLENET, 32, 3, 10), (FEMNIST, 28, 1, 62)])
@pytest.mark.parametrize("precision", [0, 1])
def test_second_device_in_one_process(cuda, layout, hw, c, classes, precision):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    a = _grad(layout, torch.device("cuda:0"), precision, hw, c, classes)
    b = _grad(layout, torch.device("cuda:1"), precision, hw, c, classes)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert np.isfinite(a[0]).all() and np.abs(a[0]).sum() > 0
