"""config tokenizer (reference src/utils/config.h) and netconfig grammar (src/nnet/nnet_config.h)."""
import io

import pytest

from cxxnet_b200 import netconfig as nc
from cxxnet_b200.config import ConfigError, parse_cli_overrides, parse_config_string
from cxxnet_b200.netconfig import NetConfig


def test_tokenizer_basic_and_comments():
    cfg = parse_config_string("""
# comment line
a = 1   # trailing comment
b=2
 c   =   hello
path = "./data/x y.gz"
multi = 'line1
line2'
esc = "a\\"b"
""")
    assert cfg == [("a", "1"), ("b", "2"), ("c", "hello"), ("path", "./data/x y.gz"),
                   ("multi", "line1\nline2"), ("esc", 'a"b')]


def test_tokenizer_duplicates_and_order_preserved():
    cfg = parse_config_string("eta = 0.1\neta = 0.2\nlayer[+1] = fullc:fc1\n  nhidden = 3\n")
    assert [k for k, _ in cfg] == ["eta", "eta", "layer[+1]", "nhidden"]
    assert cfg[1][1] == "0.2"


def test_tokenizer_unterminated_string():
    with pytest.raises(ConfigError):
        parse_config_string('a = "abc\n')


def test_cli_overrides():
    assert parse_cli_overrides(["a=1", "junk", "model_dir=/tmp/x"]) == [("a", "1"), ("model_dir", "/tmp/x")]


MLP = """
netconfig=start
layer[+1:fc1] = fullc:fc1
  nhidden = 100
layer[+1:sg1] = sigmoid:se1
layer[sg1->fc2] = fullc:fc2
  nhidden = 10
layer[+0] = softmax
netconfig=end
input_shape = 1,1,784
eta = 0.1
"""


def test_netconfig_sugar_and_scoping():
    n = NetConfig()
    n.configure(parse_config_string(MLP))
    assert n.node_names == ["in", "fc1", "sg1", "fc2"]
    assert [(l.type, l.name, l.nindex_in, l.nindex_out) for l in n.layers] == [
        (nc.kFullConnect, "fc1", [0], [1]), (nc.kSigmoid, "se1", [1], [2]),
        (nc.kFullConnect, "fc2", [2], [3]), (nc.kSoftmax, "", [3], [3])]
    assert n.input_shape == (1, 1, 784)
    assert ("nhidden", "100") in n.layercfg[0] and ("nhidden", "10") in n.layercfg[2]
    assert ("eta", "0.1") in n.defcfg and all(k != "nhidden" for k, _ in n.defcfg)
    assert n.layer_name_map == {"fc1": 0, "se1": 1, "fc2": 2}


def test_netconfig_anonymous_plus_one_and_multi_io():
    txt = """
netconfig=start
layer[0->1] = conv:c1
  nchannel = 8
  kernel_size = 3
layer[1->2,3] = split
layer[2->4] = relu
layer[3->5] = tanh
layer[4,5->6] = ch_concat
layer[+1] = flatten
layer[+1] = fullc:fc
  nhidden = 4
layer[+0] = softmax
netconfig=end
input_shape = 3,8,8
"""
    n = NetConfig()
    n.configure(parse_config_string(txt))
    assert n.layers[1].nindex_out == [2, 3]
    assert n.layers[4].nindex_in == [4, 5] and n.layers[4].type == nc.kChConcat
    assert n.node_names[7].startswith("!node-after-6")
    assert n.num_nodes == 9 and n.num_layers == 8


def test_netconfig_shared_layer_and_label_vec():
    txt = """
label_vec[0,1) = cls
label_vec[1,4) = reg
netconfig=start
layer[0->1] = fullc:enc
  nhidden = 5
layer[1->2] = share[enc]
layer[2->2] = softmax
netconfig=end
input_shape = 1,1,5
"""
    n = NetConfig()
    n.configure(parse_config_string(txt))
    assert n.layers[1].type == nc.kSharedLayer and n.layers[1].primary_layer_index == 0
    assert n.label_name_map == {"label": 0, "cls": 1, "reg": 2}
    assert n.label_range == [(0, 1), (0, 1), (1, 4)]
    with pytest.raises(ConfigError):
        NetConfig().configure(parse_config_string("netconfig=start\nlayer[0->1] = share[nope]\nnetconfig=end\n"))


def test_netconfig_binary_roundtrip_and_structure_check():
    n = NetConfig()
    n.configure(parse_config_string(MLP))
    buf = io.BytesIO()
    n.save_net(buf)
    raw = buf.getvalue()
    assert len(raw) >= 152
    m = NetConfig()
    m.load_net(io.BytesIO(raw))
    assert m.node_names == n.node_names and m.num_layers == 4 and m.init_end == 1
    m.configure(parse_config_string(MLP))          # same structure: accepted, hyper-parameters re-read
    assert ("nhidden", "100") in m.layercfg[0]
    bad = MLP.replace("nhidden = 10", "nhidden = 10\nlayer[+1] = relu")
    with pytest.raises(ConfigError):
        k = NetConfig()
        k.load_net(io.BytesIO(raw))
        k.configure(parse_config_string(bad))


def test_layer_type_ids_and_pairtest_encoding():
    assert nc.get_layer_type("conv") == 10 and nc.get_layer_type("batch_norm") == 30
    assert nc.get_layer_type("share[x]") == 0
    assert nc.get_layer_type("pairtest-conv-conv") == 1024 * 10 + 10
    with pytest.raises(ConfigError):
        nc.get_layer_type("nonsense")


def test_parameter_server_tuning_keys_are_accepted():
    """keys of the reference's async updater that have nothing to act on here must not be errors"""
    from cxxnet_b200.trainer import NetTrainer
    tr = NetTrainer()
    for k, v in (("pull_at_backprop", "0"), ("bigarray_bound", "1000000"), ("update_on_server", "0"),
                 ("init_on_worker", "1"), ("force_contiguous", "1"), ("test_on_server", "1")):
        tr.set_param(k, v)
    assert tr.test_on_server == 1
