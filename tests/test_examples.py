"""Every shipped example configuration parses and builds its net (reference-math backend, tiny batch)."""
import glob
import os

import pytest

from cxxnet_b200.config import parse_config_file
from cxxnet_b200.trainer import NetTrainer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFS = sorted(glob.glob(os.path.join(ROOT, "examples", "*", "*.conf")))


@pytest.mark.parametrize("path", CONFS, ids=[os.path.relpath(p, ROOT) for p in CONFS])
def test_example_conf_builds(path):
    tr = NetTrainer()
    in_iter = False
    for k, v in parse_config_file(path):
        # the trainer takes the global keys; iterator blocks belong to the task driver
        if k in ("data", "eval", "pred"):
            in_iter = True
            continue
        if k == "iter":
            in_iter = v != "end"
            continue
        if not in_iter:
            tr.set_param(k, v)
    tr.set_param("dev", "cpu")
    tr.set_param("batch_size", "2")
    tr.set_param("silent", "1")
    tr.init_model()
    assert len(tr.net.connections) >= 3
    nparam = sum(w.numel() for c in tr.net.connections for _, w, _ in c.layer.weights())
    assert nparam > 0
    if path.endswith("ImageNet.conf"):
        assert nparam == 60965224          # AlexNet
