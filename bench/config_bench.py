#!/usr/bin/env python
"""Training throughput of the reference's smaller example nets (BASELINE.json configs 2 and 4), measured like
bench.py measures AlexNet: NetTrainer.update through the public trainer API, synthetic data of the example's shape,
CUDA events around K steps after W warm-up steps, barrier + synchronize on both sides, max over ranks, clocks sampled
during the timed region.  `value` has the batch already on the device; `e2e` feeds every step from pinned host memory
and reads the train-error counter back.

  python bench/config_bench.py --model lenet --gpus 1                       # example/MNIST/MNIST_CONV.conf
  torchrun --nproc-per-node 2 ... bench/config_bench.py --model lenet --gpus 2
  torchrun --nproc-per-node 4 ... bench/config_bench.py --model bowl --gpus 4   # example/kaggle_bowl net, nag updater
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench as B   # noqa: E402  (ClockSampler, timed, dist_setup of the headline benchmark)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="lenet", choices=["lenet", "bowl"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--batch", type=int, default=0, help="global batch (0 = the example's: 100 / 64)")
    ap.add_argument("--extra", default="")
    args = ap.parse_args()
    import torch
    from cxxnet_b200 import ops
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import DataBatch
    from cxxnet_b200.models import bowl_conf, lenet_conf
    from cxxnet_b200.trainer import NetTrainer
    rank, world, local = B.dist_setup(args.gpus)
    dev = "gpu:%d" % local if world == 1 else "gpu:" + ",".join(str(i) for i in range(world))
    extra = args.extra.replace(";", "\n") + "\ncuda_graph = 1\n"
    if args.model == "lenet":
        batch = args.batch or 100
        shape, nclass = (1, 28, 28), 10
        conf = lenet_conf(batch_size=batch, dev=dev, extra=extra)
        what = "MNIST LeNet (example/MNIST/MNIST_CONV.conf), sgd momentum 0.9"
    else:
        batch = args.batch or 64
        shape, nclass = (1, 48, 48), 121
        conf = bowl_conf(batch_size=batch, dev=dev, updater="nag", input_shape="1,48,48", extra=extra)
        what = "kaggle_bowl net (example/kaggle_bowl/bowl.conf layers, 1x48x48 input), nag updater"
    tr = NetTrainer()
    for k, v in parse_config_string(conf):
        tr.set_param(k, v)
    tr.set_param("silent", "1")
    tr.init_model()
    g = torch.Generator().manual_seed(7)
    host = []
    for _ in range(3):
        b = DataBatch()
        b.data = torch.randn((batch,) + shape, generator=g).pin_memory()
        b.label = torch.randint(0, nclass, (batch, 1), generator=g).float().pin_memory()
        b.batch_size = batch
        host.append(b)
    devb = DataBatch()
    devb.data, devb.label, devb.batch_size = host[0].data.cuda(), host[0].label.cuda(), batch
    result = torch.zeros(1, dtype=torch.float64).pin_memory()

    def step_dev(i):
        tr.update(devb)

    def step_e2e(i):
        tr.update(host[i % 3])
        result.copy_(tr.train_metric.evals[0].sum_metric.reshape(1), non_blocking=True)
    for i in range(max(args.warmup, 6)):
        step_dev(i)
    torch.cuda.synchronize()
    c0 = ops.launch_count()
    sampler = B.ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = B.timed(step_dev, args.steps, world)
    clocks = sampler.stop() if rank == 0 else None
    launches = ops.launch_count() - c0
    for i in range(3):
        step_e2e(i)
    ms_e2e = B.timed(step_e2e, args.steps, world)
    lo, hi = tr._slice(batch)
    out = {"metric": "training images/sec (device-timed, max over ranks)", "value": batch * args.steps / (ms / 1e3),
           "unit": "images/sec", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
           "dtype": "bf16", "data": "synthetic", "impl": "ours", "scaling": "strong",
           "config": {"model": what, "global_batch": batch, "per_gpu_batch": hi - lo,
                      "l2": "no flush: the whole working set of these nets fits the 126 MB L2 - as it does in real training"},
           "e2e": {"value": batch * args.steps / (ms_e2e / 1e3), "unit": "images/sec", "ms_per_step": ms_e2e / args.steps,
                   "h2d_bytes_per_step": (hi - lo) * (shape[0] * shape[1] * shape[2] * 4 + 4), "d2h_bytes_per_step": 8},
           "gpu_launches": int(launches), "clocks": clocks, "train_error_running": tr.train_metric.evals[0].get()}
    if world > 1:
        out["replica_max_diff"] = tr.net.exchange.check_replicas()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
