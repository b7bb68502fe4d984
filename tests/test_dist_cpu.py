"""Data-parallel trainer logic on CPU: two gloo ranks, each with half of every global batch, must
reproduce single-process training (batch slicing, global-batch loss scaling, gradient SUM)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
dist.init_process_group("gloo")
rank = dist.get_rank()
from util import make_trainer, make_batch
from cxxnet_b200.models import mlp_conf
conf = mlp_conf(dev="cpu", batch_size=60, extra="wmat:wd = 0.01\n")
tr = make_trainer(conf)
assert tr.world_size == 2 and tr.step == 30
single = make_trainer(conf, param_server="NONE")
assert single.world_size == 1 and single.step == 60
for name in ("fc1", "fc2"):
    for tag in ("wmat", "bias"):
        single.set_weight(tr.get_weight(name, tag)[0], name, tag)
for s in range(3):
    b = make_batch((60, 1, 1, 784), seed=s)
    tr.update(b); single.update(b)
for name in ("fc1", "fc2"):
    for tag in ("wmat", "bias"):
        a = tr.get_weight(name, tag)[0]; c = single.get_weight(name, tag)[0]
        assert np.allclose(a, c, atol=2e-6), (name, tag, float(np.abs(a - c).max()))
b = make_batch((60, 1, 1, 784), seed=9)
assert np.array_equal(tr.predict(b), single.predict(b))
line = tr.evaluate(None, "train")
assert line.startswith("\ttrain-error:")
dist.barrier()
print("OK", rank)
'''


def test_two_gloo_ranks_match_single_process(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % {"root": ROOT})
    env = dict(os.environ)
    env["MASTER_ADDR"] = "127.0.0.1"
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29633", str(script)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.count("OK") == 2, r.stdout[-3000:]
