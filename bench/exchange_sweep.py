#!/usr/bin/env python
"""BASELINE config 5: fused reduce-scatter + sgd update + all-gather kernel vs NCCL all-reduce +
separate update, swept over the AlexNet per-layer message sizes.

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/exchange_sweep.py [--grid G]

Device-timed with CUDA events, max over ranks.  Reported bandwidth is the per-GPU, per-direction NVLink traffic of
the unicast algorithm, (N-1)/N * bytes (reduce-scatter in, all-gather out), divided by time, and its fraction of the
measured peer-copy bandwidth (770 GB/s per direction, B200_PROFILING.md; 900 nominal).  With the NVSwitch multicast
data plane (CXN_SYMM=auto on a fabric that supports it) the switch reduces and replicates, so fewer bytes cross the
GPU's own links than the unicast figure assumes: the fraction then measures the speed-up over a perfect unicast
exchange, and may pass 1.  CXN_SYMM=ipc forces the unicast peer loads / stores.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

SIZES_MB = [0.139, 1.229, 3.539, 2.654, 1.769, 150.995, 67.109, 16.384]
NAMES = ["conv1", "conv2", "conv3", "conv4", "conv5", "fc6", "fc7", "fc8"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=0)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    os.environ["CXN_XCHG_MIN_BUCKET_MB"] = "0"       # one bucket per layer: this is a sweep over message sizes
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from cxxnet_b200 import ops
    from cxxnet_b200.parallel.arena import ParamArena
    from cxxnet_b200.parallel.exchange import GradExchange
    from cxxnet_b200.parallel.symm import SymmetricBlock
    from cxxnet_b200.updaters import UpdaterParam
    dev = torch.device("cuda", torch.cuda.current_device())

    class FakeNet:
        pass
    net = FakeNet()
    net.rank, net.world_size, net.device = rank, world, dev
    net.tp = None
    arena = ParamArena(dev, "sgd")
    for i, mb in enumerate(SIZES_MB):
        n = int(mb * 1e6 / 4)
        up = UpdaterParam("wmat")
        up.base_lr = 0.01
        arena.add(None, i, "wmat", 1, n, 0, up)
    nn = max(arena.total, arena.chunk)
    flag_bytes = ops.lib().cxn_xchg_flag_bytes()
    block = SymmetricBlock(3 * (nn * 4 + 4096) + 4096 + flag_bytes + 65536 + 1024, dev)
    offs = {}
    offs["flags"], _ = block.carve(flag_bytes)
    names = iter(["w", "g", "m", "shadow"])

    def alloc(count, dtype):
        o, t = block.alloc(count, dtype)
        offs[next(names)] = o
        return t
    arena._alloc = alloc
    arena.finalize()
    net.arena, net.symm_block, net.symm_offsets = arena, block, offs
    ex = GradExchange(net, grid=args.grid)
    arena.g.normal_()
    torch.cuda.synchronize()
    dist.barrier()

    def time_it(fn, iters, stream=None):
        stream = stream or torch.cuda.current_stream()
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(iters):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    rows = []
    for bi, b in enumerate(ex.buckets):
        n = (b["ce"] - b["cb"]) * arena.chunk
        nbytes = n * 4

        def fused(b=b):
            # the kernel alone, back to back on the exchange stream (in training it is one node of the step's graph; the
            # main-stream handshake of GradExchange.layer_done is not part of what is measured here)
            ex.seq += 1
            ops.check(ops.lib().cxn_fused_rs_update_ag(
                ex.block.peers_c, world, rank, offs["w"], offs["g"], offs["m"], 0, offs["shadow"], offs["flags"],
                arena.seg_dev.data_ptr(), arena.chunk_seg.data_ptr(), b["cb"], b["ce"], arena.kind, ex.seq, b["slot"],
                ex.counter.data_ptr(), max(1, min(ex.grid, (b["ce"] - b["cb"] + 3) // 4)), ex.stream.cuda_stream),
                "cxn_fused_rs_update_ag")
        gview = arena.g[b["cb"] * arena.chunk:b["ce"] * arena.chunk]
        wview = arena.w[b["cb"] * arena.chunk:b["ce"] * arena.chunk]
        mview = arena.m[b["cb"] * arena.chunk:b["ce"] * arena.chunk]

        def nccl():
            dist.all_reduce(gview)
            mview.mul_(0.9).add_(gview, alpha=-0.01)
            wview.add_(mview)
            gview.zero_()
        tf = time_it(fused, args.iters, ex.stream)
        tn = time_it(nccl, args.iters)
        traffic = 2.0 * (world - 1) / world * nbytes
        per_dir = traffic / 2 / (tf * 1e-3) / 1e9
        rows.append({"layer": NAMES[b["trigger"]], "MB": nbytes / 1e6, "fused_us": tf * 1e3, "nccl_allreduce_update_us": tn * 1e3,
                     "fused_GBps_per_dir": per_dir, "frac_of_770GBps": per_dir / 770.0, "frac_of_900GBps": per_dir / 900.0,
                     "nccl_GBps_per_dir": traffic / 2 / (tn * 1e-3) / 1e9, "speedup_vs_nccl": tn / tf})
    if rank == 0:
        tot_f = sum(r["fused_us"] for r in rows)
        tot_n = sum(r["nccl_allreduce_update_us"] for r in rows)
        print(json.dumps({"n_gpus": world, "grid": ex.grid, "symmetric_memory": block.kind, "rows": rows,
                          "total_fused_us": tot_f, "total_nccl_us": tot_n,
                          "note": "fused = reduce-scatter + sgd update + all-gather in one kernel; nccl = all_reduce + "
                                  "torch update ops; device-timed, max over ranks"}, indent=1))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
