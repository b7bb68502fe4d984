#!/usr/bin/env python
"""Tour of the Python API on the MNIST perceptron (what the reference's example/MNIST/mnist.py
exercises: iterator-driven training, weight get/set, iterator-vs-ndarray prediction and feature
extraction, a hand-written update loop).  Python 3; run from this directory:

    python mnist.py [dev]          # dev defaults to cpu; use gpu:0 on a B200

Without ./data/*.gz the mnist iterator falls back to a deterministic surrogate set of the same
shape, so the script also works offline.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from cxxnet_b200.wrapper import cxxnet  # noqa: E402

ITER = """
iter = mnist
    path_img = "./data/%s-images-idx3-ubyte.gz"
    path_label = "./data/%s-labels-idx1-ubyte.gz"
    %s
iter = end
input_shape = 1,1,784
batch_size = 100
"""

NET = """
netconfig=start
layer[+1:fc1] = fullc:fc1
  nhidden = 100
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


def error_rate(net, it):
    """fraction of instances of `it` whose predicted class differs from the label."""
    wrong = total = 0
    it.before_first()
    while it.next():
        label = it.get_label()[:, 0]
        wrong += int(np.sum(net.predict(it) != label))
        total += label.shape[0]
    return wrong / max(total, 1)


def main(dev="cpu"):
    train = cxxnet.DataIter(ITER % ("train", "train", "shuffle = 1"))
    test = cxxnet.DataIter(ITER % ("t10k", "t10k", "index_offset = 100000"))

    # 1. one round driven by the iterator, evaluation on the test set after the round
    net = cxxnet.train(NET, train, 1, {"eta": 0.1, "dev": dev, "momentum": 0.9, "metric[label]": "error"},
                       eval_data=test)

    # 2. snapshot of the weights
    snapshot = [(layer, tag, net.get_weight(layer, tag)) for layer in ("fc1", "fc2") for tag in ("wmat", "bias")]
    print("fc1.wmat", snapshot[0][2].shape, "fc2.bias", snapshot[3][2].shape)

    # 3. the iterator path and the ndarray path give the same answers
    train.before_first()
    train.next()
    batch = train.get_data()
    print("predict: sum|iter - ndarray| =", float(np.abs(net.predict(train) - net.predict(batch)).sum()))
    print("extract sg1: sum|iter - ndarray| =",
          float(np.abs(net.extract(train, "sg1") - net.extract(batch, "sg1")).sum()))
    print("test error after round 1: %f" % error_rate(net, test))

    # 4. a second round written by hand: ndarray batches into net.update
    net.start_round(2)
    train.before_first()
    while train.next():
        net.update(train.get_data(), train.get_label())
    print("test error after round 2: %f" % error_rate(net, test))

    # 5. putting the snapshot back returns the net to its state after round 1
    for layer, tag, w in snapshot:
        net.set_weight(w, layer, tag)
    print("test error after restoring the snapshot: %f" % error_rate(net, test))

    # 6. checkpoint round trip in the .model format of the command line tool
    os.makedirs("models", exist_ok=True)
    net.save_model("models/python_api.model")
    twin = cxxnet.Net(dev=dev, cfg=NET)
    twin.load_model("models/python_api.model")
    print("reloaded model agrees:", bool(np.array_equal(twin.predict(batch), net.predict(batch))))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1] if len(sys.argv) > 1 else "cpu"))
