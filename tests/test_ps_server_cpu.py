"""Asynchronous parameter server (`param_server = dist`, `update_on_server = 1`; parallel/ps_server.py) over gloo on
CPUs: servers own the weights and the updaters, workers push gradients and pull weights (reference
src/nnet/nnet_ps_server.cpp, src/updater/async_updater-inl.hpp:109-113)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
from util import make_trainer, make_batch
from cxxnet_b200.models import mlp_conf
NS = %(servers)d
conf = mlp_conf(dev="cpu", batch_size=60, extra=%(extra)r)
tr = make_trainer(conf, param_server="dist", update_on_server=1, num_servers=NS)
W = world - NS
assert tr.world_size == W and tr.is_server == (rank >= W)
if tr.is_server:
    stats = tr.serve()
    # every key of the two layers lives on exactly one server; 4 keys in all
    print("SERVER", rank, stats)
    assert stats["push"] > 0 and stats["init"] > 0
    print("OK", rank)
    sys.exit(0)
assert tr.step == 60 // W
single = make_trainer(conf, param_server="NONE")
for name in ("fc1", "fc2"):
    for tag in ("wmat", "bias"):
        single.set_weight(tr.get_weight(name, tag)[0], name, tag)
assert tr.net.exchange.check_replicas() == 0.0        # what this worker holds is what the servers hold
if W > 1:
    dist.barrier(group=tr.ps_wgroup)                  # (... as long as nobody has pushed yet)
STEPS = 4
for s in range(STEPS):
    b = make_batch((60, 1, 1, 784), seed=s)
    tr.update(b); single.update(b)
    if W > 1:
        # asynchronous by design: a worker's pull may miss the other worker's push of the same step.  With a
        # rendezvous and a full pull after every step the run IS synchronous sgd (no momentum, no decay: two half
        # steps in either order add up to the full step)
        dist.barrier(group=tr.ps_wgroup)
        tr.net.exchange.pull_all()
        dist.barrier(group=tr.ps_wgroup)
for name in ("fc1", "fc2"):
    for tag in ("wmat", "bias"):
        a = tr.get_weight(name, tag)[0]; c = single.get_weight(name, tag)[0]
        assert np.allclose(a, c, atol=%(atol)g), (name, tag, float(np.abs(a - c).max()))
b = make_batch((60, 1, 1, 784), seed=9)
assert np.array_equal(tr.predict(b), single.predict(b))
line = tr.evaluate(None, "train")
assert line.startswith("\ttrain-error:")
# a checkpoint written by a worker loads into a plain trainer
import io
buf = io.BytesIO(); tr.save_model(buf); buf.seek(0)
plain = make_trainer(conf, param_server="NONE"); plain.load_model(buf)
assert np.array_equal(plain.get_weight("fc1", "wmat")[0], tr.get_weight("fc1", "wmat")[0])
tr.finish()
print("OK", rank)
'''


@pytest.mark.parametrize("workers,servers,extra,atol", [
    (1, 1, "wmat:wd = 0.01\nmomentum = 0.9\n", 1e-6),       # one worker: the server's updater IS the local one
    (1, 1, "updater = nag\nwmat:wd = 0.01\nlr:schedule = expdecay\nlr:gamma = 0.5\nlr:step = 2\n", 1e-6),
    (1, 1, "updater = adam\n", 1e-6),
    (2, 1, "momentum = 0\nwd = 0\n", 2e-6),                   # two workers, one server
    (2, 2, "momentum = 0\nwd = 0\n", 2e-6),                   # keys spread over two servers
])
def test_parameter_server_matches_local_training(tmp_path, workers, servers, extra, atol):
    script = tmp_path / "w.py"
    script.write_text(WORKER % {"root": ROOT, "servers": servers, "extra": extra, "atol": atol})
    env = dict(os.environ)
    env["MASTER_ADDR"] = "127.0.0.1"
    port = 29640 + workers * 3 + servers + (5 if 'nag' in extra else 9 if 'adam' in extra else 0)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(workers + servers),
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.count("OK") == workers + servers, r.stdout[-4000:]


def test_roles_and_keys():
    from cxxnet_b200.parallel.ps_server import Roles, encode_key
    r = Roles(5, 3, 2)
    assert r.num_workers == 3 and r.is_server and r.servers == [3, 4]
    assert Roles(5, 2, 2).is_server is False
    assert r.server_of(encode_key(0, "wmat")) == 3 and r.server_of(encode_key(0, "bias")) == 4
    assert encode_key(7, "bias") == 29            # layer * 4 + tag index, as the reference's EncodeDataKey
    with pytest.raises(ValueError):
        Roles(2, 0, 2)


def test_task_driver_with_two_workers_and_a_server(tmp_path):
    """`python -m cxxnet_b200 conf` under torchrun with 3 ranks: ranks 0-1 train on their halves of every batch, rank 2
    serves; the error falls, rank 0 writes the checkpoints, and `continue = 1` starts the servers from the checkpoint"""
    import re
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_cli_tasks import CONF
    mdir = tmp_path / "models"
    conf = tmp_path / "mnist.conf"
    conf.write_text(CONF % {"rounds": 2, "model_dir": mdir, "pred": tmp_path / "pred.txt"} +
                    "param_server = dist\nupdate_on_server = 1\nnum_servers = 1\nsilent = 0\n")
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    env["MASTER_ADDR"] = "127.0.0.1"

    def run(extra, port):
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "3",
                            "--master-addr", "127.0.0.1", "--master-port", str(port), "-m", "cxxnet_b200", str(conf)] + extra,
                           cwd=tmp_path, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
        return r
    r = run([], 29661)
    assert sorted(os.listdir(mdir)) == ["0000.model", "0001.model", "0002.model"]
    lines = [l for l in r.stderr.splitlines() if l.startswith("[")]
    errs = [float(re.match(r"\[\d+\]\ttrain-error:([0-9.eE+-]+)", l).group(1)) for l in lines]
    assert len(errs) == 2 and errs[1] < min(errs[0], 0.3), lines
    m = re.search(r"parameter server 2: \{'init': (\d+), 'push': (\d+), 'pull': (\d+)\}", r.stdout)
    assert m and int(m.group(1)) == 4 and int(m.group(2)) == 2 * 2 * 30 * 4, r.stdout[-2000:]    # 2 workers x 2 rounds x 30 batches x 4 keys
    r2 = run(["continue=1", "num_round=3", "max_round=3"], 29662)
    assert "0003.model" in os.listdir(mdir)
    l3 = [l for l in r2.stderr.splitlines() if l.startswith("[3]")]
    assert l3 and float(re.match(r"\[3\]\ttrain-error:([0-9.eE+-]+)", l3[0]).group(1)) < errs[0]
