"""The task driver on a B200 with the fast options switched on: captured training steps
(`cuda_graph = 1`), bf16 batches from the iterator (`input_dtype = bf16`), a threaded buffer, a
convolutional net.  Training must converge as on the reference path, evaluation and prediction
(eager forward passes between replayed steps) must see the updated weights, checkpoints must load."""
import os
import re

import pytest

from test_cli_tasks import _run

pytestmark = pytest.mark.gpu

CONF = """
data = train
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 3000
    shuffle = 1
    input_flat = 0
    input_dtype = bf16
iter = threadbuffer
iter = end
eval = test
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 200
    index_offset = 7
    input_flat = 0
iter = end
pred = %(pred)s
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 200
    index_offset = 7
    input_flat = 0
iter = end
netconfig=start
layer[+1] = conv:cv1
  kernel_size = 5
  stride = 2
  nchannel = 16
layer[+1] = relu
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = flatten
layer[+1:fc1] = fullc:fc1
  nhidden = 64
layer[+1] = relu
layer[+0] = dropout
  threshold = 0.2
layer[+1] = fullc:fc2
  nhidden = 10
layer[+0] = softmax
netconfig=end
input_shape = 1,28,28
batch_size = 100
dev = gpu
cuda_graph = 1
random_type = xavier
save_model = 1
num_round = 2
max_round = 2
eta = 0.05
momentum = 0.9
metric[label] = error
model_dir = %(model_dir)s
silent = 1
"""


def test_cli_trains_with_captured_steps_and_bf16_batches(tmp_path):
    mdir = tmp_path / "models"
    conf = tmp_path / "conv.conf"
    conf.write_text(CONF % {"model_dir": mdir, "pred": tmp_path / "pred.txt"})
    r = _run([str(conf)], tmp_path)
    assert sorted(os.listdir(mdir)) == ["0000.model", "0001.model", "0002.model"]
    lines = [l for l in r.stderr.splitlines() if l.startswith("[")]
    assert len(lines) == 2, r.stderr
    err = [(float(re.search(r"train-error:([0-9.eE+-]+)", l).group(1)),
            float(re.search(r"test-error:([0-9.eE+-]+)", l).group(1))) for l in lines]
    assert err[1][0] < 0.1 and err[1][1] < 0.1, err          # the synthetic digits are easy
    # prediction from the checkpoint, on the reference-math backend of the same GPU: same answers
    _run([str(conf), "task=pred", "model_in=%s" % (mdir / "0002.model")], tmp_path)
    fast = (tmp_path / "pred.txt").read_text().split()
    _run([str(conf), "task=pred", "model_in=%s" % (mdir / "0002.model"), "backend=ref"], tmp_path)
    ref = (tmp_path / "pred.txt").read_text().split()
    assert len(fast) == len(ref) == 200
    assert sum(a != b for a, b in zip(fast, ref)) <= 4      # bf16 vs fp32 may flip a borderline instance


def test_cli_layer_timing_on_the_gpu(tmp_path):
    """`layer_timing = 1` brackets every layer with CUDA events (and runs eagerly: a captured step
    cannot be bracketed)"""
    conf = tmp_path / "conv.conf"
    conf.write_text(CONF % {"model_dir": tmp_path / "models", "pred": tmp_path / "pred.txt"})
    r = _run([str(conf), "layer_timing=1", "num_round=1", "max_round=1"], tmp_path)
    assert "layer timing (mean us per step" in r.stdout
    rows = [l for l in r.stdout.splitlines() if re.match(r"\s+\d+ (conv|relu|max_pooling|flatten|fullc|dropout|softmax):", l)]
    assert len(rows) == 9, r.stdout
    conv = [l for l in rows if " conv:" in l][0].split()
    assert float(conv[2]) > 0 and float(conv[3]) > 0
