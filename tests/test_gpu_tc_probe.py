"""Pins the tcgen05 shared-memory / instruction descriptor encodings used by the tensor-core
convolution kernels (no-swizzle core-matrix layouts, K-major and MN-major operands, shifted
windows) by running single MMAs through dls_tc_probe and comparing with a float64 matmul of the
bf16-rounded operands."""
import numpy as np
import pytest
import torch

import dls_b200
from dls_b200 import _lib

pytestmark = pytest.mark.gpu


def bf16_bits(x):
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    return r


def bf16_round(x):
    return (bf16_bits(x).astype(np.uint32) << 16).view(np.float32)


def desc(addr, lbo, sbo):
    return (addr >> 4) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1 << 46)


def idesc(M, N, a_mn, b_mn):
    return (1 << 4) | (1 << 7) | (1 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def run(cuda, a_img, b_img, a_desc, b_desc, idsc, ncols, repeat=0):
    out = torch.full((128, ncols), float("nan"), device=cuda)
    cyc = torch.zeros(1, dtype=torch.int64, device=cuda)
    ta = torch.from_numpy(a_img.view(np.uint8)).to(cuda)
    tb = torch.from_numpy(b_img.view(np.uint8)).to(cuda)
    da = torch.from_numpy(np.array(a_desc, dtype=np.uint64).view(np.int64)).to(cuda)
    db = torch.from_numpy(np.array(b_desc, dtype=np.uint64).view(np.int64)).to(cuda)
    rc = _lib.lib().dls_tc_probe(ta.data_ptr(), ta.numel(), tb.data_ptr(), tb.numel(), da.data_ptr(), db.data_ptr(),
                                 len(a_desc), idsc, ncols, out.data_ptr(), repeat, cyc.data_ptr(),
                                 torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    if repeat:
        return out.cpu().numpy(), int(cyc.item())
    return out.cpu().numpy()


def test_kmajor_planar_shifted_window(cuda):
    """A: channel-chunk planar activations [chunk][pixel][8 ch] (pixel stride 16 B), rows of the MMA
    = 128 consecutive pixels starting at an arbitrary pixel shift; B: weights [k/8][n][8]."""
    rs = np.random.RandomState(0)
    npix, nchunk, N = 300, 4, 32
    act = rs.standard_normal((npix, nchunk * 8)).astype(np.float32)          # [pixel][channel]
    wgt = rs.standard_normal((N, nchunk * 8)).astype(np.float32)             # [n][k]
    plane = npix * 16
    a_img = np.zeros(nchunk * npix * 8, np.uint16)
    ab = bf16_bits(act)
    for c in range(nchunk):
        a_img[c * npix * 8:(c + 1) * npix * 8] = ab[:, c * 8:(c + 1) * 8].reshape(-1)
    b_img = np.zeros(nchunk * N * 8, np.uint16)
    wb = bf16_bits(wgt)
    for c in range(nchunk):
        b_img[c * N * 8:(c + 1) * N * 8] = wb[:, c * 8:(c + 1) * 8].reshape(-1)
    for shift in (0, 3, 41):
        # K = 32 channels = two MMAs of K = 16 (two 8-channel chunks each)
        a_desc = [desc(shift * 16 + h * 2 * plane, plane, 128) for h in range(2)]
        b_desc = [desc(h * 2 * N * 16, N * 16, 128) for h in range(2)]
        D = run(cuda, a_img, b_img, a_desc, b_desc, idesc(128, N, 0, 0), N)
        want = bf16_round(act[shift:shift + 128]).astype(np.float64) @ bf16_round(wgt).astype(np.float64).T
        assert np.abs(D - want).max() < 1e-3, (shift, np.abs(D - want).max())


def test_mn_major_both_operands(cuda):
    """Weight-gradient form: reduction over pixels (K), A = dY^T and B = input, both stored
    [chunk][pixel][8 ch] so that the M/N index is the contiguous one (MN-major)."""
    rs = np.random.RandomState(1)
    npix, M, N = 64, 64, 32
    dy = rs.standard_normal((npix, M)).astype(np.float32)                     # [pixel][co]
    xin = rs.standard_normal((npix + 16, N)).astype(np.float32)               # [pixel][ci]
    pa, pb = npix * 16, (npix + 16) * 16
    a_img = np.concatenate([bf16_bits(dy)[:, c * 8:(c + 1) * 8].reshape(-1) for c in range(M // 8)])
    b_img = np.concatenate([bf16_bits(xin)[:, c * 8:(c + 1) * 8].reshape(-1) for c in range(N // 8)])
    for shift in (0, 5):
        nk = 2                                                               # K = 32 pixels: two MMAs
        a_desc = [desc(k * 16 * 16, 128, pa) for k in range(nk)]
        b_desc = [desc((shift + k * 16) * 16, 128, pb) for k in range(nk)]
        D = run(cuda, a_img, b_img, a_desc, b_desc, idesc(M, N, 1, 1), N)
        want = bf16_round(dy[:32]).astype(np.float64).T @ bf16_round(xin[shift:shift + 32]).astype(np.float64)
        # an M = 64 accumulator keeps 16 rows per 32-lane TMEM quadrant: row m -> lane (m//16)*32 + m%16
        got = D[[(m // 16) * 32 + (m % 16) for m in range(M)]]
        if np.abs(got - want).max() >= 1e-3:
            # report where the 64 rows of an M = 64 tile live in TMEM, to help fix the epilogue
            lanes = [int(np.argmin(np.abs(D - want[m]).sum(1))) for m in range(0, M, 8)]
            raise AssertionError("M=64 tile mismatch, max err %g; lanes of rows 0,8,..: %s"
                                 % (np.abs(got - want).max(), lanes))

