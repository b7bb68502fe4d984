#!/usr/bin/env python
"""Headline benchmark: AlexNet (3x227x227, global batch 256) training images/sec.

  python bench.py --gpus N --steps K --warmup W [--impl ours|reference|library]

* ``ours``      cxxnet_b200 NetTrainer.update (forward + backward + sgd update) through the
                public trainer API.  ``value`` is device-timed with the batch already staged
                on the device; ``e2e`` feeds every step from pinned host memory (H2D of the
                step's images + labels) and reads the step's train-error counter back (D2H).
* ``reference`` the unmodified reference from baseline/_ref.  It cannot be installed offline
                (no setup.py/pyproject; all kernels live in the un-vendored mshadow), so this
                arm prints ``{"impl": "reference", "unavailable": ...}`` and exits 0.
* ``library``   stock PyTorch (cuDNN/cuBLAS, bf16 autocast, channels_last) + NCCL all-reduce
                + torch SGD: the "library baseline" (BASELINE.md B3), for context only.

Batch semantics follow the reference: ``batch_size`` is the GLOBAL batch, split
ceil(256/N) per GPU (strong scaling).  Multi-GPU runs are launched by torchrun, timed on the
device with CUDA events, max over ranks.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "AlexNet images/sec at 1/2/4/8 B200 (device-timed, max over ranks) vs reference build"


class ClockSampler:
    """samples nvidia-smi SM clocks / throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_setup(n):
    import torch
    import torch.distributed as dist
    rank, world, local = 0, 1, 0
    if n > 1:
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        local = int(os.environ.get("LOCAL_RANK", "0"))
        if world != n:
            raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (n, n))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
    return rank, world, local


def max_over_ranks(ms, world):
    import torch
    import torch.distributed as dist
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, steps, world):
    import torch
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for i in range(steps):
        fn(i)
    timed.host_ms = (time.perf_counter() - t0) * 1e3      # host time to enqueue the K steps (this rank)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier(world)
    return max_over_ranks(ms, world)


def run_ours(args, rank, world, local):
    import torch
    from cxxnet_b200 import ops
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import DataBatch
    from cxxnet_b200.models import alexnet_conf
    from cxxnet_b200.trainer import NetTrainer
    B = args.batch
    tr = NetTrainer()
    dev = "gpu:%d" % local if world == 1 else "gpu:" + ",".join(str(i) for i in range(world))
    extra = args.extra
    if "cuda_graph" not in extra:
        extra += "\ncuda_graph = 1\n"        # one captured graph per step instead of ~80 launches
    if world > 1 and "fullc_gather" not in extra and "fullc_shard" not in extra:
        # Fully connected layers on several GPUs.  Four or more: sharded by output columns (`fullc_shard = 1`, weights
        # never cross a link, activations move on the critical path).  Two: the reference's own option,
        # `fullc_gather = 1` (row-sharded weight gradient / update, new bf16 rows published on a side stream) - with
        # two ranks that publish hides under the convolution backward pass and beats the sharded layers' nine
        # barriers per step (profiles/r2_scaling.md).  Either can be forced through --extra.
        extra += "\nfullc_shard = 1\n" if world >= 4 else "\nfullc_gather = 1\n"
    for k, v in parse_config_string(alexnet_conf(batch_size=B, dev=dev, extra=extra)):
        tr.set_param(k, v)
    tr.set_param("silent", "1")
    tr.set_param("eval_train", "1")
    tr.init_model()
    g = torch.Generator().manual_seed(1234)
    nbuf = 3
    host = []
    for i in range(nbuf):
        b = DataBatch()
        # the input pipeline delivers bf16 batches (`input_dtype = bf16` of the batch iterator): the model
        # computes in bf16, so rounding on the host instead of on the device halves the PCIe bytes
        if args.input == "u8":
            b.data = torch.randint(0, 256, (B, 3, 227, 227), generator=g, dtype=torch.uint8).pin_memory()
            b.norm = ([104.0, 117.0, 123.0], 1.0 / 64)
        else:
            b.data = torch.randn((B, 3, 227, 227), generator=g).to(torch.bfloat16).pin_memory()
        b.label = torch.randint(0, 1000, (B, 1), generator=g).float().pin_memory()
        b.batch_size = B
        host.append(b)
    lo, hi = tr._slice(B)
    devb = DataBatch()
    devb.data = host[0].data.cuda(non_blocking=True)
    devb.label = host[0].label.cuda(non_blocking=True)
    devb.norm = host[0].norm
    devb.batch_size = B

    def step_dev(i):
        tr.update(devb)

    result = torch.zeros(2, dtype=torch.float64).pin_memory()

    def step_e2e(i):
        tr.update(host[i % nbuf])
        m = tr.train_metric.evals[0]
        # D2H of the step's result: running train-error numerator
        result[0:1].copy_(m.sum_metric.reshape(1), non_blocking=True)

    # cuda_graph captures one graph per input staging slot after three eager steps: keep the captures
    # out of the timed region whatever W is
    for i in range(max(args.warmup, 6)):
        step_dev(i)
    torch.cuda.synchronize()
    c0 = ops.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(step_dev, args.steps, world)
    clocks = sampler.stop() if rank == 0 else None
    host_ms = timed.host_ms
    launches = ops.launch_count() - c0
    for i in range(max(args.warmup, 3)):
        step_e2e(i)
    ms_e2e = timed(step_e2e, args.steps, world)
    err = tr.train_metric.evals[0].get()
    h2d = (hi - lo) * (3 * 227 * 227 * host[0].data.element_size() + 4)
    out = {
        "metric": METRIC, "value": B * args.steps / (ms / 1e3), "unit": "images/sec", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ours",
        "config": {"model": "AlexNet (example/ImageNet/ImageNet.conf)", "global_batch": B,
                   "input": "3x227x227 (%s NCHW batches from the iterator, fp32 labels)" % args.input,
                   "per_gpu_batch": hi - lo,
                   "parallelism": ("dp%d" % world) if world == 1 else
                                  ("dp%d (conv) + tp%d (fc6-8 column-sharded)" % (world, world)) if "fullc_shard" in extra
                                  else ("dp%d (fc6-8: fullc_gather, row-sharded weight gradient and update)" % world),
                   "updater": "sgd momentum 0.9 wd 5e-4",
                   "l2": "no flush: per-step working set (activations + gradients + 244 MB of fp32 weights, 122 MB of "
                         "bf16 weights, optimiser state) far exceeds the 126 MB L2; e2e rotates 3 distinct %d MB "
                         "(%s) input batches" % (B * 3 * 227 * 227 * (1 if args.input == "u8" else 2) // 1000000,
                                                args.input),
                   "timing": "CUDA events around K steps, barrier+synchronize both sides, max over ranks",
                   "host_enqueue_note": "with cuda_graph the host runs exactly one step ahead (it waits for step N-1 "
                                        "before rewriting that step's pinned hyper-parameter table), so its enqueue "
                                        "time per step tracks the device time; the launch itself is one graph"},
        "host_enqueue_ms_per_step": host_ms / args.steps,
        "e2e": {"value": B * args.steps / (ms_e2e / 1e3), "unit": "images/sec", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8},
        "gpu_launches": int(launches), "clocks": clocks, "train_error_running": err,
    }
    if tr.net.trace_buf is not None:
        sys.stderr.write("rank %d\n%s\n" % (rank, tr.net.trace_report()))
    if world > 1:
        # outside the timed regions: are the replicas of the timed trainer identical, and does a multi-rank run
        # reproduce a single-GPU run of the same batches?
        out["replica_max_diff"] = tr.net.exchange.check_replicas()
        out["symmetric_memory"] = tr.net.symm_block.kind       # ipc | vmm | nvls (NVSwitch multicast: multimem.*)
        if not args.no_verify:
            rep, rel_w, rel_dw = verify_multi(args, rank, world, local, extra)
            out["verify"] = {"replica_max_diff_after_3_steps": rep, "max_rel_err_vs_1gpu": rel_w,
                             "max_rel_err_of_updates_vs_1gpu": rel_dw,
                             "note": "fresh nets, dropout off, 3 global batches; relative L1 per weight tensor of the "
                                     "weights and of (w - w0); bf16 rounding differs with the batch split"}
    return out


def verify_multi(args, rank, world, local, extra):
    """Correctness of the multi-GPU path, outside every timed region (the reference's `test_on_server` idea,
    src/updater/async_updater-inl.hpp:144-154, plus a single-GPU twin): a fresh multi-rank trainer and, on every
    rank's own device, a single-process twin start from identical weights (dropout off: its masks depend on the
    batch split) and take the same 3 global batches.  Returns (max |w_rank - w_rank0| over all weights after the
    steps, max relative L1 difference of any weight tensor against the twin)."""
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import DataBatch
    from cxxnet_b200.models import alexnet_conf
    from cxxnet_b200.trainer import NetTrainer
    B = args.batch

    def make(dev, **over):
        t = NetTrainer()
        conf = alexnet_conf(batch_size=B, dev=dev, extra=extra).replace("threshold = 0.5", "threshold = 0.0")
        for k, v in parse_config_string(conf):
            t.set_param(k, v)
        for k, v in over.items():
            t.set_param(k, v)
        t.set_param("silent", "1")
        t.init_model()
        return t
    multi = make("gpu:" + ",".join(str(i) for i in range(world)))
    single = make("gpu:%d" % local, param_server="NONE")
    names = [(i, tag) for i, c in enumerate(multi.net.connections) for tag, _, _ in c.layer.weights()]
    w0 = {}
    for i, tag in names:
        w, _ = multi.net.get_weight(i, tag)
        single.net.set_weight(i, w, tag)
        w0[(i, tag)] = w.double()
    g = torch.Generator().manual_seed(4321)
    for s_ in range(3):
        b = DataBatch()
        b.data = torch.randn((B, 3, 227, 227), generator=g).to(torch.bfloat16)
        b.label = torch.randint(0, 1000, (B, 1), generator=g).float()
        b.batch_size = B
        multi.update(b)
        single.update(b)
    torch.cuda.synchronize()
    replica = multi.net.exchange.check_replicas()
    worst_w, worst_dw = 0.0, 0.0
    for i, tag in names:
        a, _ = multi.net.get_weight(i, tag)
        c, _ = single.net.get_weight(i, tag)
        a, c = a.double(), c.double()
        denom = float(c.abs().sum())
        worst_w = max(worst_w, float((a - c).abs().sum()) / denom if denom > 0 else float((a - c).abs().sum()))
        # what moved: w - w0 is small against w, so the weights alone would hide a wrong update
        da, dc = a - w0[(i, tag)], c - w0[(i, tag)]
        denom = float(dc.abs().sum())
        e_dw = float((da - dc).abs().sum()) / denom if denom > 0 else float((da - dc).abs().sum())
        worst_dw = max(worst_dw, e_dw)
        if rank == 0:
            sys.stderr.write("verify %s.%s: |w| %.4g  rel err w %.3g  rel err (w - w0) %.3g\n" % (
                multi.net_cfg.layers[i].name or i, tag, float(c.abs().sum()),
                float((a - c).abs().sum()) / max(float(c.abs().sum()), 1e-30), e_dw))
    t = torch.tensor([worst_w, worst_dw], dtype=torch.float64, device="cuda")
    import torch.distributed as dist
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return replica, float(t[0].item()), float(t[1].item())


def run_library(args, rank, world, local):
    """stock PyTorch AlexNet (cuDNN/cuBLAS bf16) + DDP-style all-reduce + SGD: BASELINE.md B3."""
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    B = args.batch
    per = (B + world - 1) // world

    class Net(nn.Module):
        def __init__(s):
            super().__init__()
            s.f = nn.Sequential(
                nn.Conv2d(3, 96, 11, 4), nn.ReLU(True), nn.MaxPool2d(3, 2, ceil_mode=True), nn.LocalResponseNorm(5, 0.001, 0.75, 1.0),
                nn.Conv2d(96, 256, 5, 1, 2, groups=2), nn.ReLU(True), nn.MaxPool2d(3, 2, ceil_mode=True), nn.LocalResponseNorm(5, 0.001, 0.75, 1.0),
                nn.Conv2d(256, 384, 3, 1, 1), nn.ReLU(True), nn.Conv2d(384, 384, 3, 1, 1, groups=2), nn.ReLU(True),
                nn.Conv2d(384, 256, 3, 1, 1, groups=2), nn.ReLU(True), nn.MaxPool2d(3, 2, ceil_mode=True))
            s.c = nn.Sequential(nn.Flatten(), nn.Linear(9216, 4096), nn.ReLU(True), nn.Dropout(0.5),
                                nn.Linear(4096, 4096), nn.ReLU(True), nn.Dropout(0.5), nn.Linear(4096, 1000))

        def forward(s, x):
            return s.c(s.f(x))
    torch.backends.cudnn.benchmark = True
    net = Net().cuda().to(memory_format=torch.channels_last)
    opt = torch.optim.SGD(net.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
    x = torch.randn(per, 3, 227, 227, device="cuda").to(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (per,), device="cuda")
    lossf = nn.CrossEntropyLoss(reduction="sum")
    params = [p for p in net.parameters()]

    def step(i):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = net(x)
        loss = lossf(out.float(), y) / B
        loss.backward()
        if world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            o = 0
            for p in params:
                n = p.numel()
                p.grad.copy_(flat[o:o + n].view_as(p))
                o += n
        opt.step()
    for i in range(args.warmup):
        step(i)
    ms = timed(step, args.steps, world)
    return {"metric": METRIC, "value": B * args.steps / (ms / 1e3), "unit": "images/sec", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "library",
            "config": {"model": "AlexNet", "global_batch": B, "parallelism": "dp%d" % world,
                       "note": "stock PyTorch cuDNN/cuBLAS + NCCL all-reduce + torch SGD (library baseline B3)"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "library"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default, the reference's semantics): --batch is the global batch, split over the "
                         "GPUs; weak: --batch images per GPU, global batch = batch x GPUs (secondary series)")
    ap.add_argument("--extra", default="", help="extra config lines for the ours arm, ';' separated")
    ap.add_argument("--no-verify", action="store_true",
                    help="skip the multi-GPU correctness check against a single-GPU twin (after the timed regions)")
    ap.add_argument("--input", default="u8", choices=["bf16", "u8"],
                    help="host batches of the end-to-end measurement: u8 (default: raw pixels as the image iterators "
                         "decode them, mean / scale applied on the device - a quarter of the fp32 bytes over PCIe) or "
                         "bf16 (normalised on the host)")
    args = ap.parse_args()
    args.extra = args.extra.replace(";", "\n")
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:
            print(json.dumps({"impl": "reference", "unavailable":
                              "reference has no setup.py/pyproject (pip install fails) and its whole tensor/GPU/"
                              "parameter-server backend (mshadow, mshadow-ps) is fetched by git clone at build time "
                              "and is absent offline; see DESIGN.md section 'Reference arm'"}))
        return 0
    rank, world, local = dist_setup(args.gpus)
    if args.scaling == "weak":
        args.batch *= world
    out = run_ours(args, rank, world, local) if args.impl == "ours" else run_library(args, rank, world, local)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
