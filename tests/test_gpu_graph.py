"""`cuda_graph = 1`: a training step captured into a CUDA graph must train exactly like the eager
step (same kernels, same order), keep the train metric's device accumulator and host count going,
and fall back to the eager path for a short last batch."""
import os

import pytest
import torch

from util import copy_weights, make_batch, make_trainer, rel_err

pytestmark = pytest.mark.gpu

CONF = """
netconfig=start
layer[+1] = conv:cv1
  kernel_size = 5
  stride = 2
  nchannel = 16
layer[+1] = relu
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = conv:cv2
  kernel_size = 3
  pad = 1
  nchannel = 64
layer[+1] = relu
layer[+1] = conv:cv3
  kernel_size = 3
  pad = 1
  nchannel = 32
layer[+1] = flatten
layer[+0] = dropout
  threshold = 0.5
layer[+1] = fullc:fc1
  nhidden = 10
layer[+0] = softmax
netconfig=end
batch_size = 16
input_shape = 3,33,33
eta = 0.05
momentum = 0.9
wd = 0.0005
random_type = xavier
metric = error
eval_train = 1
dev = gpu:0
"""


def test_captured_step_trains_like_eager_step():
    eager = make_trainer(CONF)
    graph = make_trainer(CONF, cuda_graph=1)
    copy_weights(eager, graph)
    batches = [make_batch((16, 3, 33, 33), seed=s, pin=True) for s in range(3)]
    steps = 12
    for i in range(steps):
        b = batches[i % 3]
        eager.update(b)
        graph.update(b)
    torch.cuda.synchronize()
    assert len(graph._graphs) == 2, "one captured graph per input staging slot"
    for i, c in enumerate(eager.net.connections):
        for tag, _, _ in c.layer.weights():
            we, _ = eager.net.get_weight(i, tag)
            wg, _ = graph.net.get_weight(i, tag)
            # the weight-gradient kernels reduce with atomics, so the two runs agree to rounding only
            assert rel_err(we, wg) < 2e-2, "layer %d %s diverged" % (i, tag)
    me, mg = eager.train_metric.evals[0], graph.train_metric.evals[0]
    assert me.cnt_inst == mg.cnt_inst == steps * 16
    assert abs(me.get() - mg.get()) <= 0.1
    # a short last batch takes the eager path and the graphs stay valid afterwards
    short = make_batch((16, 3, 33, 33), seed=9)
    short.batch_size = 11
    graph.update(short)
    graph.update(batches[0])
    torch.cuda.synchronize()
    assert mg.cnt_inst == steps * 16 + 11 + 16
    assert graph.epoch_counter == steps + 2


def test_bf16_batches_train_exactly_like_fp32_batches():
    """a bf16 batch is the fp32 batch rounded on the host instead of on the device: same bits reach the
    first layer, so the two runs may differ only through the atomics of the weight-gradient kernels"""
    a = make_trainer(CONF, cuda_graph=1)
    b = make_trainer(CONF, cuda_graph=1)
    copy_weights(a, b)
    batches = [make_batch((16, 3, 33, 33), seed=s, pin=True) for s in range(3)]
    halves = []
    for x in batches:
        h = make_batch((16, 3, 33, 33), seed=0, pin=True)
        h.data = x.data.to(torch.bfloat16).pin_memory()
        h.label = x.label
        halves.append(h)
    for i in range(8):
        a.update(batches[i % 3])
        b.update(halves[i % 3])
    torch.cuda.synchronize()
    assert b.net.in_stage_bf16 is not None and len(b._graphs) == 2
    for i, c in enumerate(a.net.connections):
        for tag, _, _ in c.layer.weights():
            wa, _ = a.net.get_weight(i, tag)
            wb, _ = b.net.get_weight(i, tag)
            assert rel_err(wa, wb) < 2e-2, "layer %d %s diverged" % (i, tag)
    assert a.train_metric.evals[0].cnt_inst == b.train_metric.evals[0].cnt_inst == 8 * 16


def test_u8_batches_train_like_host_normalised_bf16_batches():
    """`input_dtype = u8`: raw pixels cross PCIe and cxn_nchw_u8_to_nhwc4_pitched applies (pixel - mean) * scale before
    the bf16 rounding, so the first layer sees the bits of a bf16 batch normalised on the host"""
    a = make_trainer(CONF, cuda_graph=1)
    b = make_trainer(CONF, cuda_graph=1)
    copy_weights(a, b)
    mean, scale = [120.0, 125.0, 130.0], 1.0 / 128          # inputs in about [-1, 1], like the other tests' batches
    g = torch.Generator().manual_seed(5)
    raws, halves = [], []
    for s in range(3):
        lab = make_batch((16, 3, 33, 33), seed=s, pin=True)
        px = torch.randint(0, 256, (16, 3, 33, 33), generator=g, dtype=torch.uint8)
        r = make_batch((16, 3, 33, 33), seed=s, pin=True)
        r.data, r.label, r.norm = px.pin_memory(), lab.label, (mean, scale)
        h = make_batch((16, 3, 33, 33), seed=s, pin=True)
        h.data = ((px.float() - torch.tensor(mean).view(1, 3, 1, 1)) * scale).to(torch.bfloat16).pin_memory()
        h.label = lab.label
        raws.append(r)
        halves.append(h)
    for i in range(8):
        a.update(halves[i % 3])
        b.update(raws[i % 3])
    torch.cuda.synchronize()
    assert b.net.in_stage_u8 is not None and len(b._graphs) == 2
    for i, c in enumerate(a.net.connections):
        for tag, _, _ in c.layer.weights():
            wa, _ = a.net.get_weight(i, tag)
            wb, _ = b.net.get_weight(i, tag)
            assert rel_err(wa, wb) < 2e-2, "layer %d %s diverged" % (i, tag)
    # prediction takes the same route
    assert (a.predict(halves[0]) == b.predict(raws[0])).all()
