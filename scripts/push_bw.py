#!/usr/bin/env python
"""Halo-exchange bandwidth on this box (run under torchrun, one rank per GPU): every rank sends ROWS rows of LD floats
to every peer — the traffic pattern of one C3 aggregate level — through (1) NCCL all_to_all_single from a packed
buffer, (2) dls_push_rows into symmetric memory in three transfer-list orders (ascending destination rank, shifted =
all of rank + 1 first, interleaved = round-robin over the destinations) x the kernel forms / CTA counts of dls_tune
keys 6 / 7, and (3) one pair alone (rank 0 -> rank 1), by stores and by the copy engine.  Prints one JSON line on
rank 0: GB/s per GPU, outbound, max over ranks.  Results: profiles/r02_halo_bandwidth_8gpu_*.json.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/push_bw.py [ROWS] [LD]
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dls_b200  # noqa: E402
from dls_b200 import _lib, kernels as K  # noqa: E402


def main():
    rows_per_peer = int(sys.argv[1]) if len(sys.argv) > 1 else 85
    ld = int(sys.argv[2]) if len(sys.argv) > 2 else 1690048
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm_mem
    lib = _lib.lib()
    n_send = rows_per_peer * (world - 1)
    X = symm_mem.empty((2 * n_send, ld), dtype=torch.float32, device=dev)          # [own rows | halo rows]
    hdl = symm_mem.rendezvous(X, dist.group.WORLD)
    X.normal_()
    ptrs = [int(p) for p in hdl.buffer_ptrs]
    peers = [p for p in range(world) if p != rank]
    bytes_out = n_send * ld * 4

    def timed(fn, iters=4):
        fn(); fn()
        torch.cuda.synchronize(); dist.barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {"world": world, "rows_per_peer": rows_per_peer, "row_bytes": ld * 4, "bytes_out_per_gpu": bytes_out}

    def transfers(order):
        # transfer: own row (j-th peer, r) -> that peer's halo row (n_send + slot of this rank at the peer)
        items = []
        for j, p in enumerate(peers):
            my_index_at_peer = rank if rank < p else rank - 1
            for r in range(rows_per_peer):
                items.append((p, r, j * rows_per_peer + r, ptrs[p] + (n_send + my_index_at_peer * rows_per_peer + r) * ld * 4))
        if order == "shifted":                         # phase k: everyone writes to (rank + 1 + k) mod N
            items.sort(key=lambda it: ((it[0] - rank) % world, it[1]))
        elif order == "interleaved":                   # consecutive transfers go to consecutive peers
            items.sort(key=lambda it: (it[1], (it[0] - rank) % world))
        return (torch.tensor([it[2] for it in items], dtype=torch.int32, device=dev),
                torch.tensor([it[3] for it in items], dtype=torch.int64, device=dev))

    for order in ("ascending", "shifted", "interleaved"):
        s_d, d_d = transfers(order)

        def push():
            K.push_rows(X, s_d, d_d, ld)
            hdl.barrier(channel=0)

        for mode, ctas in ((0, 8), (1, 4), (2, 1), (2, 2), (2, 4), (2, 8)):
            lib.dls_tune(6, ctas); lib.dls_tune(7, mode)
            ms = timed(push)
            out["%s_mode%d_ctas%d" % (order, mode, ctas)] = {"ms": round(ms, 3), "GBps": round(bytes_out / ms / 1e6, 1)}
    lib.dls_tune(6, 0); lib.dls_tune(7, 0)

    # NCCL all_to_all from a packed buffer (rows already contiguous per peer)
    rbuf = X[n_send:].reshape(-1)
    sbuf = X[:n_send].reshape(-1)
    splits = [rows_per_peer * ld if p != rank else 0 for p in range(world)]

    def a2a():
        dist.all_to_all_single(rbuf, sbuf, output_split_sizes=splits, input_split_sizes=splits)

    ms = timed(a2a)
    out["nccl_all_to_all"] = {"ms": ms, "GBps": bytes_out / ms / 1e6}

    # one pair alone: rank 0 -> rank 1, the whole n_send rows
    if world > 1:
        dst1 = torch.tensor([ptrs[1 if rank == 0 else 0] + (n_send + r) * ld * 4 for r in range(n_send)], dtype=torch.int64,
                            device=dev)
        src1 = torch.arange(n_send, dtype=torch.int32, device=dev)

        def pair():
            if rank == 0:
                K.push_rows(X, src1, dst1, ld)
            hdl.barrier(channel=0)

        for mode in (0, 1):
            lib.dls_tune(7, mode)
            ms = timed(pair)
            out["pair_0_to_1_mode%d" % mode] = {"ms": ms, "GBps": bytes_out / ms / 1e6}
        lib.dls_tune(7, 0)

        # peer-to-peer DMA (copy engines): contiguous block to one peer
        peer_buf = hdl.get_buffer(1 if rank == 0 else 0, (2 * n_send, ld), torch.float32)

        def dma():
            if rank == 0:
                peer_buf[n_send:].copy_(X[:n_send])
            hdl.barrier(channel=0)

        ms = timed(dma)
        out["pair_0_to_1_copy"] = {"ms": ms, "GBps": bytes_out / ms / 1e6}
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
