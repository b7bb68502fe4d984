"""Multi-GPU fused exchange (needs >= 2 GPUs; run with `gpurun --gpus 2`)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

def _run(nproc, script_args, timeout=300):
    env = dict(os.environ)
    env.update({"MASTER_ADDR": "127.0.0.1"})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", "29613"] + list(script_args)
    return subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)


@pytest.mark.skipif(torch.cuda.is_available() and torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("updater,mode,graph", [("sgd", "lenet", ""), ("nag", "lenet", ""), ("sgd", "gather", ""),
                                                ("sgd", "lenet", "graph"), ("sgd", "gather", "graph"),
                                                ("sgd", "shard", ""), ("nag", "shard", "graph")])
def test_fused_exchange_matches_single_gpu(updater, mode, graph):
    script = os.path.join(ROOT, "tests", "mp_exchange_worker.py")
    r = _run(2, [script, updater, mode] + ([graph] if graph else []))
    assert r.returncode == 0 and r.stdout.count("OK") == 2, r.stdout[-4000:]


@pytest.mark.skipif(torch.cuda.is_available() and torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.skipif(os.environ.get("CXN_TEST_COLLECTIVE", "0") != "1",
                    reason="collective (param_server = dist) exchange on GPUs: opt-in with CXN_TEST_COLLECTIVE=1")
def test_collective_exchange_matches_single_gpu():
    script = os.path.join(ROOT, "tests", "mp_exchange_worker.py")
    r = _run(2, [script, "sgd", "collective"])
    assert r.returncode == 0 and r.stdout.count("OK") == 2, r.stdout[-4000:]


PS_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
dist.init_process_group("gloo")
rank = dist.get_rank()
from util import make_trainer, make_batch
from cxxnet_b200.models import lenet_conf
conf = lenet_conf(batch_size=32, dev="gpu:0", extra="updater = %(updater)s\nwmat:wd = 0.001\n")
tr = make_trainer(conf, param_server="dist", update_on_server=1)
if tr.is_server:
    stats = tr.serve()
    assert stats["push"] == 3 * 6 and stats["init"] == 6          # 3 steps x 6 keys
    print("OK server", stats)
    sys.exit(0)
assert tr.net.backend == "b200" and type(tr.net.exchange).__name__ == "ServerExchange"
single = make_trainer(conf, param_server="NONE")
names = [(n, t) for n in ("cv1", "fc1", "fc2") for t in ("wmat", "bias")]
for n, t in names:
    single.set_weight(tr.get_weight(n, t)[0], n, t)
assert tr.net.exchange.check_replicas() == 0.0
for s in range(3):
    b = make_batch((32, 1, 28, 28), seed=s, pin=True)
    tr.update(b); single.update(b)
torch.cuda.synchronize()
for n, t in names:
    a = tr.get_weight(n, t)[0]; c = single.get_weight(n, t)[0]
    err = float(np.abs(a - c).sum() / max(np.abs(c).sum(), 1e-9))
    assert err < 2e-3, (n, t, err)        # same gradients (same kernels), the server's fp32 CPU updater against the arena kernel
assert tr.net.exchange.check_replicas() == 0.0
tr.finish()
print("OK worker")
'''


@pytest.mark.parametrize("updater", ["sgd", "nag"])
def test_parameter_server_with_a_gpu_worker(tmp_path, updater):
    """`param_server = dist`, `update_on_server = 1` with a b200-backend worker on the GPU and a server rank on the
    host: gradients leave through pinned memory, the server's CPU updater answers with the weights - same weights as
    the local arena updater after 3 steps (one GPU is enough)"""
    script = tmp_path / "ps.py"
    script.write_text(PS_WORKER % {"root": ROOT, "updater": updater})
    r = _run(2, [str(script)])
    assert r.returncode == 0 and r.stdout.count("OK") == 2, r.stdout[-4000:]
