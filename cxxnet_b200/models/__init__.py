"""Model zoo: the reference's example networks as ``.conf`` text builders.

``alexnet_conf``   reference ``example/ImageNet/ImageNet.conf`` (3x227x227, batch 256, 60 965 224 params)
``lenet_conf``     reference ``example/MNIST/MNIST_CONV.conf`` (1x28x28, batch 100)
``mlp_conf``       reference ``example/MNIST/MNIST.conf`` (784 -> 100 -> 10)
``bowl_conf``      reference ``example/kaggle_bowl/bowl.conf`` (3x40x40, 121 classes)

Each returns the text of a complete configuration (net + hyper-parameters, no iterators) that
``config.parse_config_string`` / ``NetTrainer.set_param`` accept.
"""
from __future__ import annotations


def alexnet_conf(batch_size: int = 256, dev: str = "gpu", nclass: int = 1000, extra: str = "") -> str:
    return """
netconfig=start
layer[0->1] = conv:conv1
  kernel_size = 11
  stride = 4
  nchannel = 96
layer[1->2] = relu:relu1
layer[2->3] = max_pooling:pool1
  kernel_size = 3
  stride = 2
layer[3->4] = lrn:lrn1
  local_size = 5
  alpha = 0.001
  beta = 0.75
  knorm = 1
layer[4->5] = conv:conv2
  ngroup = 2
  nchannel = 256
  kernel_size = 5
  pad = 2
layer[5->6] = relu:relu2
layer[6->7] = max_pooling:pool2
  kernel_size = 3
  stride = 2
layer[7->8] = lrn:lrn2
  local_size = 5
  alpha = 0.001
  beta = 0.75
  knorm = 1
layer[8->9] = conv:conv3
  nchannel = 384
  kernel_size = 3
  pad = 1
layer[9->10]= relu:relu3
layer[10->11] = conv:conv4
  nchannel = 384
  ngroup = 2
  kernel_size = 3
  pad = 1
layer[11->12] = relu:relu4
layer[12->13] = conv:conv5
  nchannel = 256
  ngroup = 2
  kernel_size = 3
  pad = 1
  init_bias = 1.0
layer[13->14] = relu:relu5
layer[14->15] = max_pooling:pool5
  kernel_size = 3
  stride = 2
layer[15->16] = flatten:flatten
layer[16->17] = fullc:fc6
  nhidden = 4096
  init_sigma = 0.005
  init_bias = 1.0
layer[17->18] = relu:relu6
layer[18->18] = dropout:drop6
  threshold = 0.5
layer[18->19] = fullc:fc7
  nhidden = 4096
  init_sigma = 0.005
  init_bias = 1.0
layer[19->20] = relu:relu7
layer[20->20] = dropout:drop7
  threshold = 0.5
layer[20->21] = fullc:fc8
  nhidden = %(nclass)d
layer[21->21] = softmax:softmax
netconfig=end
metric = error
input_shape = 3,227,227
batch_size = %(batch)d
dev = %(dev)s
momentum = 0.9
wmat:lr  = 0.01
wmat:wd  = 0.0005
bias:wd  = 0.000
bias:lr  = 0.02
lr:schedule = expdecay
lr:gamma = 0.1
lr:step = 100000
random_type = xavier
%(extra)s
""" % {"batch": batch_size, "dev": dev, "nclass": nclass, "extra": extra}


def lenet_conf(batch_size: int = 100, dev: str = "gpu", extra: str = "") -> str:
    return """
netconfig=start
layer[0->1] = conv:cv1
  kernel_size = 3
  pad = 1
  stride = 2
  nchannel = 32
  random_type = xavier
  no_bias=0
layer[1->2] = max_pooling
  kernel_size = 3
  stride = 2
layer[2->3] = flatten
layer[3->3] = dropout
  threshold = 0.5
layer[3->4] = fullc:fc1
  nhidden = 100
  init_sigma = 0.01
layer[4->5] = sigmoid:se1
layer[5->6] = fullc:fc2
  nhidden = 10
  init_sigma = 0.01
layer[6->6] = softmax
netconfig=end
input_shape = 1,28,28
batch_size = %(batch)d
dev = %(dev)s
random_type = gaussian
eta = 0.1
momentum = 0.9
wd  = 0.0
metric = error
%(extra)s
""" % {"batch": batch_size, "dev": dev, "extra": extra}


def mlp_conf(batch_size: int = 100, dev: str = "cpu", extra: str = "") -> str:
    return """
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
batch_size = %(batch)d
dev = %(dev)s
random_type = gaussian
eta = 0.1
momentum = 0.9
wd  = 0.0
metric[label] = error
%(extra)s
""" % {"batch": batch_size, "dev": dev, "extra": extra}


def bowl_conf(batch_size: int = 64, dev: str = "gpu", updater: str = "sgd", input_shape: str = "3,40,40",
              extra: str = "") -> str:
    """kaggle national data science bowl net (reference ``example/kaggle_bowl/bowl.conf:32-111``)."""
    return """
netconfig=start
layer[+1] = conv:cv1
  kernel_size = 4
  pad = 2
  nchannel = 48
layer[+1] = relu
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = conv:cv2
  nchannel = 96
  kernel_size = 3
  pad = 1
layer[+1] = relu
layer[+1] = conv:cv3
  nchannel = 96
  kernel_size = 3
  pad = 1
layer[+1] = relu
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = conv:cv4
  nchannel = 128
  kernel_size = 2
layer[+1] = relu
layer[+1] = conv:cv5
  nchannel = 128
  kernel_size = 3
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 256
layer[+0] = dropout
  threshold = 0.5
layer[+1] = fullc:fc2
  nhidden = 121
layer[+0] = softmax
netconfig=end
input_shape = %(shape)s
batch_size = %(batch)d
dev = %(dev)s
updater = %(updater)s
random_type = xavier
init_sigma = 0.01
momentum = 0.9
wmat:lr  = 0.001
wmat:wd  = 0.0005
bias:wd  = 0.000
bias:lr  = 0.002
lr:schedule = expdecay
lr:gamma = 0.1
lr:step = 20000
metric = error
metric = logloss
%(extra)s
""" % {"batch": batch_size, "dev": dev, "updater": updater, "shape": input_shape, "extra": extra}
