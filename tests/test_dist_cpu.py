"""Data-parallel trainer logic on CPU: two gloo ranks, each with half of every global batch, must
reproduce single-process training (batch slicing, global-batch loss scaling, gradient SUM)."""
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
rank = dist.get_rank()
from util import make_trainer, make_batch
from cxxnet_b200.models import mlp_conf
conf = mlp_conf(dev="cpu", batch_size=60, extra="wmat:wd = 0.01\n")
tr = make_trainer(conf%(over)s)
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


@pytest.mark.parametrize("pserver", ["local", "dist"])
def test_two_gloo_ranks_match_single_process(tmp_path, pserver):
    """`param_server = dist` (the reference's multi-node key) selects the collective exchange; on the CPU backend
    both settings sum the gradients with an all-reduce."""
    script = tmp_path / "w.py"
    script.write_text(WORKER % {"root": ROOT, "over": "" if pserver == "local" else ', param_server="dist"'})
    env = dict(os.environ)
    env["MASTER_ADDR"] = "127.0.0.1"
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29633" if pserver == "local" else "29634", str(script)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.count("OK") == 2, r.stdout[-3000:]


def test_param_server_dist_needs_a_process_group():
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.models import mlp_conf
    from cxxnet_b200.trainer import NetTrainer
    tr = NetTrainer()
    for k, v in parse_config_string(mlp_conf(dev="cpu", batch_size=10)):
        tr.set_param(k, v)
    tr.set_param("param_server", "dist")
    with pytest.raises(ValueError, match="torchrun"):
        tr.init_model()


def test_host_detection_reads_the_launcher_environment(monkeypatch):
    from cxxnet_b200.parallel.dist_exchange import ranks_share_one_host
    monkeypatch.setenv("WORLD_SIZE", "16")
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    assert not ranks_share_one_host()
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "16")
    assert ranks_share_one_host()
    monkeypatch.delenv("LOCAL_WORLD_SIZE")
    assert ranks_share_one_host()


def test_two_ranks_over_packed_jpegs_decode_their_own_rows_and_match_one_process(tmp_path):
    """task driver under torchrun (gloo) on imgbin data: every rank decodes only its rows of each batch
    (slice_world / slice_rank), random crops and mirrors included, and the trained weights equal a single process's"""
    PIL = pytest.importorskip("PIL")
    import numpy as np
    from PIL import Image
    im2bin = os.path.join(ROOT, "bin", "im2bin")
    if not os.path.exists(im2bin):
        pytest.skip("bin/im2bin not built")
    rs = np.random.RandomState(0)
    os.makedirs(tmp_path / "img")
    lines = []
    for i in range(60):
        cls = i % 2
        a = np.clip(rs.randn(20, 22, 3) * 20 + (70 if cls == 0 else 180), 0, 255).astype(np.uint8)
        Image.fromarray(a).save(tmp_path / "img" / ("%03d.jpg" % i), quality=90)
        lines.append("%d\t%d\timg/%03d.jpg\n" % (i, cls, i))
    (tmp_path / "all.lst").write_text("".join(lines))
    subprocess.run([im2bin, "all.lst", "./", "all.bin"], cwd=tmp_path, check=True, stdout=subprocess.PIPE, timeout=120)
    (tmp_path / "net.conf").write_text("""
data = train
iter = imgbin
  image_list = "all.lst"
  image_bin = "all.bin"
  rand_crop = 1
  rand_mirror = 1
  mean_value = 128,128,128
  divideby = 64
iter = threadbuffer
iter = end
eval = val
iter = imgbin
  image_list = "all.lst"
  image_bin = "all.bin"
  mean_value = 128,128,128
  divideby = 64
iter = end
netconfig = start
layer[+1] = flatten
layer[+1] = fullc:fc
  nhidden = 2
  init_sigma = 0.01
layer[+0] = softmax
netconfig = end
input_shape = 3,16,16
batch_size = 20
round_batch = 1
dev = cpu
eta = 0.05
momentum = 0.9
metric = error
num_round = 2
max_round = 2
silent = 1
""")
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    env["MASTER_ADDR"] = "127.0.0.1"
    one = subprocess.run([sys.executable, "-m", "cxxnet_b200", "net.conf", "model_dir=m1"], cwd=tmp_path, env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert one.returncode == 0, one.stderr[-2000:]
    two = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29635", "-m", "cxxnet_b200", "net.conf",
                          "model_dir=m2"], cwd=tmp_path, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                         text=True, timeout=300)
    assert two.returncode == 0, two.stderr[-2000:]
    lines1 = [ln for ln in one.stderr.splitlines() if ln.startswith("[")]
    lines2 = [ln for ln in two.stderr.splitlines() if ln.startswith("[")]
    assert lines1 == lines2 and len(lines1) == 2
    a = np.frombuffer((tmp_path / "m1" / "0002.model").read_bytes()[-1500 * 4:], dtype=np.float32)
    b = np.frombuffer((tmp_path / "m2" / "0002.model").read_bytes()[-1500 * 4:], dtype=np.float32)
    assert np.abs(a - b).max() < 1e-5


def test_launcher_stops_the_other_ranks_when_one_fails(tmp_path):
    """parallel/launch.py wait_all: a failing rank ends the job promptly with its exit code instead of leaving the
    others waiting"""
    import time
    from cxxnet_b200.parallel.launch import wait_all
    sleeper = subprocess.Popen([sys.executable, "-c", "import time; time.sleep(120)"])
    failing = subprocess.Popen([sys.executable, "-c", "import sys, time; time.sleep(0.5); sys.exit(7)"])
    t0 = time.time()
    rc = wait_all([sleeper, failing], grace_seconds=3.0)
    assert rc == 7 and time.time() - t0 < 30
    assert sleeper.poll() is not None                       # stopped by the launcher
    ok = [subprocess.Popen([sys.executable, "-c", "pass"]) for _ in range(3)]
    assert wait_all(ok) == 0
