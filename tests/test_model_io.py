""".model checkpoint layout: byte-level checks and a golden file produced by
an independent C++ writer built from the documented structs."""
import io
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest
import torch

from cxxnet_b200.layers.base import LayerParam
from cxxnet_b200.models import lenet_conf, mlp_conf
from util import make_batch, make_trainer

HERE = os.path.dirname(os.path.abspath(__file__))


def _save(tr):
    buf = io.BytesIO()
    buf.write(struct.pack("<i", 0))
    tr.save_model(buf)
    return buf.getvalue()


def test_layerparam_is_328_bytes_and_roundtrips():
    p = LayerParam()
    p.set_param("nhidden", "100"); p.set_param("kernel_size", "5"); p.set_param("pad", "2"); p.set_param("ngroup", "2")
    raw = p.pack()
    assert len(raw) == 328
    v = struct.unpack("<ififf", raw[:20])
    assert v[0] == 100 and v[1] == pytest.approx(0.01) and v[2] == 10 and v[3] == -1.0
    q = LayerParam()
    q.unpack(raw)
    assert (q.num_hidden, q.kernel_height, q.kernel_width, q.pad_y, q.pad_x, q.num_group) == (100, 5, 5, 2, 2, 2)
    assert q.temp_col_max == 64 << 18


def test_model_bytes_match_documented_layout():
    tr = make_trainer(mlp_conf(dev="cpu"))
    raw = _save(tr)
    off = 0
    assert struct.unpack_from("<i", raw, off)[0] == 0; off += 4                  # net_type
    num_nodes, num_layers, c, h, w, init_end, extra = struct.unpack_from("<ii3Iii", raw, off)
    assert (num_nodes, num_layers, c, h, w, init_end, extra) == (4, 4, 1, 1, 784, 1, 0)
    off += 152
    names = []
    for _ in range(num_nodes):
        (n,) = struct.unpack_from("<Q", raw, off); off += 8
        names.append(raw[off:off + n].decode()); off += n
    assert names == ["in", "fc1", "sg1", "fc2"]
    types = []
    for _ in range(num_layers):
        t, prim = struct.unpack_from("<ii", raw, off); off += 8
        (n,) = struct.unpack_from("<Q", raw, off); off += 8 + n
        for _ in range(2):
            (m,) = struct.unpack_from("<Q", raw, off); off += 8 + 4 * m
        types.append((t, prim))
    assert types == [(1, -1), (4, -1), (1, -1), (2, -1)]
    (epoch,) = struct.unpack_from("<q", raw, off); off += 8
    assert epoch == 0
    (blob_len,) = struct.unpack_from("<Q", raw, off); off += 8
    assert off + blob_len == len(raw)
    # fullc payload: LayerParam(328) + wmat (uint32 shape[2], data) + bias (uint32 shape[1], data)
    assert blob_len == (328 + 8 + 100 * 784 * 4 + 4 + 100 * 4) + (328 + 8 + 10 * 100 * 4 + 4 + 10 * 4)
    assert struct.unpack_from("<i", raw, off)[0] == 100                          # LayerParam.num_hidden
    assert struct.unpack_from("<II", raw, off + 328) == (100, 784)
    w_file = np.frombuffer(raw, dtype="<f4", count=100 * 784, offset=off + 336).reshape(100, 784)
    assert np.array_equal(w_file, tr.get_weight("fc1", "wmat")[0])


def test_save_load_roundtrip_and_conv_layout():
    tr = make_trainer(lenet_conf(dev="cpu", batch_size=8))
    b = make_batch((8, 1, 28, 28))
    tr.update(b)
    raw = _save(tr)
    from cxxnet_b200.trainer import NetTrainer
    from cxxnet_b200.config import parse_config_string
    tr2 = NetTrainer()
    for k, v in parse_config_string(lenet_conf(dev="cpu", batch_size=8)):
        tr2.set_param(k, v)
    tr2.set_param("silent", "1")
    f = io.BytesIO(raw)
    f.read(4)
    tr2.load_model(f)
    assert tr2.epoch_counter == 1
    for name in ("cv1", "fc1", "fc2"):
        for tag in ("wmat", "bias"):
            a, sa = tr.get_weight(name, tag)
            c, sc = tr2.get_weight(name, tag)
            assert sa == sc and np.array_equal(a, c)
    assert tr.get_weight("cv1", "wmat")[1] == [1, 32, 9]                         # (g, oc/g, ic/g*kh*kw)
    assert np.array_equal(tr.predict(b), tr2.predict(b))
    assert _save(tr2) == raw                                                     # byte-identical re-save


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_golden_file_from_independent_cpp_writer(tmp_path):
    exe = tmp_path / "write_golden"
    subprocess.check_call(["g++", "-O0", "-o", str(exe), os.path.join(HERE, "golden", "write_golden.cpp")])
    model = tmp_path / "golden.model"
    subprocess.check_call([str(exe), str(model)])
    from cxxnet_b200.trainer import NetTrainer
    tr = NetTrainer()
    for k, v in [("dev", "cpu"), ("batch_size", "2"), ("silent", "1"), ("netconfig", "start"),
                 ("layer[+1:fc1]", "fullc:fc1"), ("nhidden", "2"), ("layer[+0]", "softmax"), ("netconfig", "end"),
                 ("input_shape", "1,1,3")]:
        tr.set_param(k, v)
    with open(model, "rb") as f:
        assert struct.unpack("<i", f.read(4))[0] == 0
        tr.load_model(f)
    assert tr.epoch_counter == 42
    w, shape = tr.get_weight("fc1", "wmat")
    assert shape == [2, 3] and w.tolist() == [[0.5, 1.5, 2.5], [10.5, 11.5, 12.5]]
    assert tr.get_weight("fc1", "bias")[0].reshape(-1).tolist() == [-1.0, -2.0]
    b = make_batch((2, 1, 1, 3))
    x = b.data.view(2, 3).numpy()
    ref = (x @ w.T + np.array([-1.0, -2.0])).argmax(1)
    assert tr.predict(b).tolist() == ref.tolist()
    # and our writer reproduces the golden bytes exactly
    buf = io.BytesIO()
    buf.write(struct.pack("<i", 0))
    tr.save_model(buf)
    assert buf.getvalue() == open(model, "rb").read()
