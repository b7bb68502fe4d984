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
