"""Planning of the column-sharded fully connected layers (cxxnet_b200/parallel/tensor_parallel.py) without a GPU:
which layers form a chain, how the output columns are split over the ranks, which nodes change their storage."""
import pytest
import torch

from cxxnet_b200 import ops
from cxxnet_b200.config import parse_config_string
from cxxnet_b200.executor import NeuralNet
from cxxnet_b200.models import alexnet_conf
from cxxnet_b200.netconfig import NetConfig

pytestmark = pytest.mark.skipif(not ops.available(), reason="kernel library not built")


def _planned(conf, world, rank, batch):
    cfg = NetConfig()
    cfg.configure(parse_config_string(conf))
    net = NeuralNet(cfg, batch, 0, torch.device("cpu"), "ref", silent=1, world_size=world, rank=rank)
    net._init_net()
    net._config_connections()
    net._init_connections()
    consumers = {}
    for c in net.connections:
        c.native = True
        for n in c.nodes_in:
            if not (c.layer.self_loop and n in c.nodes_out):
                consumers[n.index] = consumers.get(n.index, 0) + 1
    from cxxnet_b200.parallel.tensor_parallel import TensorParallel
    tp = TensorParallel(net)
    tp.plan(consumers)
    return net, tp


def test_alexnet_fc_stack_is_one_chain_of_column_shards():
    net, tp = _planned(alexnet_conf(batch_size=256, dev="cpu", extra="fullc_shard = 1\n"), world=8, rank=7, batch=32)
    assert len(tp.chains) == 1
    fc6, fc7, fc8 = tp.chains[0]
    names = [net.cfg.layers[t.layer_index].name for t in (fc6, fc7, fc8)]
    assert names == ["fc6", "fc7", "fc8"]
    assert (fc6.first, fc6.last, fc7.first, fc7.last, fc8.first, fc8.last) == (True, False, False, False, False, True)
    assert (fc6.ns, fc6.c0, fc6.my_cols) == (512, 3584, 512) and (fc7.ns, fc7.my_cols) == (512, 512)
    # 1000 classes over 8 ranks: 128 columns each, the last rank owns the remaining 104
    assert (fc8.ns, fc8.c0, fc8.my_cols) == (128, 896, 104)
    assert fc7.prev is fc6 and fc6.next is fc7 and fc8.next is None
    # the nodes between the layers (fc output, dropout's self-loop node) are stored [global batch][own columns]
    assert tp.shard_nodes == {17, 18, 19, 20} and tp.total_rows() == 256
    assert tp.shard_width(17) == 512 and tp.shard_width(20) == 512
    assert sum(1 for c in net.connections if getattr(c.layer, "tp", None) is not None) == 3


def test_layers_that_cannot_be_sharded_stay_data_parallel():
    conf = """
netconfig=start
layer[+1] = fullc:a
  nhidden = 96
layer[+1] = relu
layer[+1] = fullc:b
  nhidden = 40
layer[+1] = sigmoid
layer[+1] = fullc:c
  nhidden = 10
layer[+0] = softmax
netconfig=end
input_shape = 1,1,1024
batch_size = 64
fullc_shard = 1
"""
    net, tp = _planned(conf, world=2, rank=1, batch=32)
    # a: 48 columns x 1024 inputs per rank sits on update chunks; b: 24 x 96 does not; c: 10 outputs are no multiple of 8
    assert [net.cfg.layers[t.layer_index].name for ch in tp.chains for t in ch] == ["a"]
    a = tp.chains[0][0]
    assert a.first and a.last and (a.ns, a.c0, a.my_cols) == (48, 48, 48) and not tp.shard_nodes
    # without the key nothing is planned
    net, tp = _planned(conf.replace("fullc_shard = 1", ""), world=2, rank=0, batch=32)
    assert not tp.active()
