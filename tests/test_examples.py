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
    tr.set_param("param_server", "local")      # dist.conf asks for a multi-node job
    tr.set_param("batch_size", "2")
    tr.set_param("silent", "1")
    tr.init_model()
    assert len(tr.net.connections) >= 3
    nparam = sum(w.numel() for c in tr.net.connections for _, w, _ in c.layer.weights())
    assert nparam > 0
    if path.endswith("ImageNet.conf"):
        assert nparam == 60965224          # AlexNet


def test_mnist_python_api_example_runs(tmp_path):
    """examples/MNIST/mnist.py: the tour of the Python API (surrogate data, CPU)."""
    import subprocess
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "MNIST", "mnist.py"), "cpu"], cwd=tmp_path,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "predict: sum|iter - ndarray| = 0.0" in r.stdout
    assert "reloaded model agrees: True" in r.stdout
    assert os.path.exists(tmp_path / "models" / "python_api.model")


def test_kaggle_resize_and_submission_scripts(tmp_path):
    """gen_train.py / gen_img_list.py / make_submission.py on a two-class toy tree"""
    import subprocess
    import sys
    from PIL import Image
    kb = os.path.join(ROOT, "examples", "kaggle_bowl")
    for cls in ("acantharia", "copepod"):
        os.makedirs(tmp_path / "raw" / cls)
        for i in range(3):
            Image.new("L", (30 + i, 20)).save(tmp_path / "raw" / cls / ("%s_%d.jpg" % (cls, i)))
    (tmp_path / "sample.csv").write_text("image,acantharia,copepod\nx.jpg,0.5,0.5\n")
    run = lambda *a: subprocess.run([sys.executable] + [str(x) for x in a], cwd=tmp_path, check=True, timeout=120)
    run(os.path.join(kb, "gen_train.py"), str(tmp_path / "raw") + "/", str(tmp_path / "small") + "/", 48)
    assert Image.open(tmp_path / "small" / "copepod" / "copepod_2.jpg").size == (48, 48)
    run(os.path.join(kb, "gen_img_list.py"), "train", tmp_path / "sample.csv", str(tmp_path / "small") + "/",
        tmp_path / "train.lst")
    rows = [ln.split("\t") for ln in (tmp_path / "train.lst").read_text().strip().split("\n")]
    assert len(rows) == 6 and sorted(r[1] for r in rows) == ["0"] * 3 + ["1"] * 3
    (tmp_path / "probs.txt").write_text("".join("0.25 0.75 \n" for _ in rows))
    run(os.path.join(kb, "make_submission.py"), tmp_path / "sample.csv", tmp_path / "train.lst",
        tmp_path / "probs.txt", tmp_path / "out.csv")
    out = (tmp_path / "out.csv").read_text().strip().split("\n")
    assert out[0] == "image,acantharia,copepod" and len(out) == 7 and out[1].endswith(",0.25,0.75")


def test_make_image_list_resizes_splits_and_packs(tmp_path):
    """tools/make_image_list.py on a class-per-directory tree, then bin/im2bin raw=1 and an imgbin read-back"""
    import subprocess
    import sys
    import numpy as np
    from PIL import Image
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    for ci, cls in enumerate(("n01", "n02", "n03")):
        os.makedirs(tmp_path / "raw" / cls)
        for i in range(8):
            a = np.full((40 + 3 * i, 50, 3), 60 * ci + 20, dtype=np.uint8)
            Image.fromarray(a).save(tmp_path / "raw" / cls / ("im%d.%s" % (i, "png" if i % 2 else "jpg")))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_image_list.py"), str(tmp_path / "raw"),
                        str(tmp_path / "train.lst"), "--val-fraction", "0.25", "--resize", "32", "--resized-root",
                        str(tmp_path / "small"), "--synsets", str(tmp_path / "synsets.txt")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0, r.stdout
    assert (tmp_path / "synsets.txt").read_text().split() == ["n01", "n02", "n03"]
    tr = (tmp_path / "train.lst").read_text().strip().split("\n")
    va = (tmp_path / "train.val.lst").read_text().strip().split("\n")
    assert len(tr) + len(va) == 24 and len(va) >= 1 and [ln.split("\t")[0] for ln in tr] == [str(i) for i in range(len(tr))]
    assert Image.open(tmp_path / "small" / tr[0].split("\t")[2]).size == (32, 32)
    im2bin = os.path.join(ROOT, "bin", "im2bin")
    if not os.path.exists(im2bin):
        pytest.skip("bin/im2bin not built")
    subprocess.run([im2bin, str(tmp_path / "train.lst"), str(tmp_path / "small") + "/", str(tmp_path / "train.bin"), "raw=1"],
                   check=True, stdout=subprocess.PIPE, timeout=120)
    it = create_iterator(parse_config_string(
        'iter = imgbin\n image_list = "%s"\n image_bin = "%s"\n batch_size = 4\n input_shape = 3,32,32\n silent = 1\niter = end\n'
        % (tmp_path / "train.lst", tmp_path / "train.bin")))
    it.init()
    it.before_first()
    assert it.next()
    b = it.value()
    for row in range(4):                                       # pixel level = 60 * label + 20 (flat images survive JPEG)
        assert abs(float(b.data[row].mean()) - (60 * float(b.label[row, 0]) + 20)) < 3


def test_check_conf_knows_every_documented_key_and_flags_typos(tmp_path):
    """python -m cxxnet_b200.tools.check_conf: the shipped examples are clean, every key of the user guide's tables is
    known to it, misspelt keys are reported with a suggestion"""
    import re
    from cxxnet_b200.config import parse_config_file
    from cxxnet_b200.tools import check_conf
    for path in CONFS:
        assert check_conf.check(list(parse_config_file(path))) == [], path
    documented = set()
    for md in glob.glob(os.path.join(ROOT, "doc", "*.md")):
        in_key_table = False
        for line in open(md):
            if not line.startswith("|"):
                in_key_table = False
            elif line.startswith("| Key"):
                in_key_table = True
            elif in_key_table:
                documented.update(re.findall(r"`([a-z_:0-9@]+)`", line.split("|")[1]))
    assert len(documented) > 50
    unknown = sorted(k for k in documented if not check_conf.known(k) and k not in ("fp32", "bf16", "u8"))
    assert unknown == [], unknown
    bad = tmp_path / "bad.conf"
    bad.write_text("slient = 1\ntrain_eval = 3\nlayer[+1] = fullc:fc\n  nhiden = 3\nwmat:lr = 0.1\n")
    found = check_conf.check(list(parse_config_file(str(bad))))
    assert [(f[0], f[3]) for f in found] == [("slient", "silent"), ("train_eval", "eval_train"), ("nhiden", "nhidden")]
    assert found[2][2].startswith("layer[+1]")
    assert check_conf.main([str(bad)]) == 1
