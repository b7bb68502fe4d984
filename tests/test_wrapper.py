"""Python wrapper (reference wrapper/cxxnet.py surface, usage of example/MNIST/mnist.py) and the
22-function C ABI driven from a plain C program."""
import os
import shutil
import subprocess
import sys
import sysconfig

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ITER = """iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = %d
    batch_size = 100
    silent = 1
iter = end
"""
NET = """
netconfig=start
layer[+1:fc1] = fullc:fc1
  nhidden = 24
  init_sigma = 0.01
layer[+1:sg1] = sigmoid:se1
layer[sg1->fc2] = fullc:fc2
  nhidden = 10
  init_sigma = 0.01
layer[+0] = softmax
netconfig=end
input_shape = 1,1,784
batch_size = 100
random_type = gaussian
"""


def test_python_api_like_mnist_py(tmp_path):
    from cxxnet_b200.wrapper import cxxnet
    data = cxxnet.DataIter(ITER % 2000)
    deval = cxxnet.DataIter(ITER % 200)
    with pytest.raises(Exception):
        data.get_data()                                     # head state
    param = {"eta": 0.1, "dev": "cpu", "momentum": 0.9, "metric[label]": "error", "silent": 1}
    net = cxxnet.train(NET, data, 5, param, eval_data=deval)
    # weight get / set round trip
    w = net.get_weight("fc1", "wmat")
    b = net.get_weight("fc1", "bias")
    assert w.shape == (24, 784) and b.shape == (24,)
    assert net.get_weight("se1", "wmat") is None
    data.before_first()
    assert data.next()
    pred = net.predict(data)
    dbatch, lbatch = data.get_data(), data.get_label()
    assert dbatch.shape == (100, 1, 1, 784) and lbatch.shape == (100, 1)
    pred2 = net.predict(dbatch)
    assert np.abs(pred - pred2).sum() == 0.0                # iterator path == ndarray path
    assert (pred == lbatch[:, 0]).mean() > 0.8
    f1, f2 = net.extract(data, "sg1"), net.extract(dbatch, "sg1")
    assert f1.shape == (100, 1, 1, 24) and np.abs(f1 - f2).sum() == 0.0
    assert net.extract(dbatch, "top[-1]").shape == (100, 1, 1, 10)
    # zero a weight -> predictions change; set it back -> identical predictions
    net.set_weight(np.zeros_like(w), "fc1", "wmat")
    assert np.abs(net.predict(dbatch) - pred).sum() > 0
    net.set_weight(w, "fc1", "wmat")
    assert np.abs(net.predict(dbatch) - pred).sum() == 0.0
    # manual update loop with ndarrays (1-D label accepted) and evaluate string
    net.update(dbatch, lbatch[:, 0])
    s = net.evaluate(deval, "eval")
    assert "\teval-error:" in s
    # save / load
    net.save_model(str(tmp_path / "m.model"))
    net2 = cxxnet.Net(dev="cpu", cfg=NET)
    net2.set_param("silent", 1)
    net2.load_model(str(tmp_path / "m.model"))
    assert np.array_equal(net2.predict(dbatch), net.predict(dbatch))
    # second train signature: (cfg, data, label, num_round, param)
    net3 = cxxnet.train(NET, dbatch, lbatch, 2, dict(param))
    assert net3.predict(dbatch).shape == (100,)


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_c_abi_from_plain_c(tmp_path):
    lib = os.path.join(ROOT, "cxxnet_b200", "lib", "libcxxnetwrapper.so")
    if not os.path.exists(lib):
        import build as _b
        _b.build_host()
    exe = tmp_path / "capi_smoke"
    libdir = os.path.dirname(lib)
    subprocess.check_call(["gcc", "-O1", "-o", str(exe), os.path.join(ROOT, "tests", "capi", "capi_smoke.c"),
                           "-I", os.path.join(ROOT, "include"), "-L", libdir, "-lcxxnetwrapper", "-lm",
                           "-Wl,-rpath," + libdir])
    env = dict(os.environ)
    env["CXXNET_B200_ROOT"] = ROOT
    env["PYTHONPATH"] = os.pathsep.join([ROOT] + [p for p in sys.path if p])
    env["PYTHONHOME"] = sys.base_prefix
    r = subprocess.run([str(exe), str(tmp_path / "capi.model")], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and "C ABI OK" in r.stdout, r.stdout[-3000:]


def test_predict_takes_more_rows_than_the_batch_size_and_update_refuses():
    from cxxnet_b200.wrapper import cxxnet
    net = cxxnet.Net(dev="cpu", cfg=NET)
    net.set_param("silent", 1)
    net.init_model()
    x = np.random.RandomState(0).rand(250, 1, 1, 784).astype(np.float32)
    whole = net.predict(x)
    assert whole.shape == (250,)
    parts = np.concatenate([net.predict(x[:100]), net.predict(x[100:200]), net.predict(x[200:])])
    assert np.array_equal(whole, parts)
    assert net.extract(x, "sg1").shape == (250, 1, 1, 24)
    with pytest.raises(ValueError, match="batch_size"):
        net.update(x, np.zeros(250, dtype=np.float32))
