#!/usr/bin/env python
"""cProfile of the host side of training steps (where does the enqueue time go?)."""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200.config import parse_config_string
from cxxnet_b200.io.data import DataBatch
from cxxnet_b200.models import alexnet_conf
from cxxnet_b200.trainer import NetTrainer

B = int(os.environ.get("BATCH", "32"))
tr = NetTrainer()
for k, v in parse_config_string(alexnet_conf(batch_size=B, dev="gpu:0", extra="")):
    tr.set_param(k, v)
tr.set_param("silent", "1")
tr.set_param("eval_train", "1")
tr.init_model()
b = DataBatch()
b.data = torch.randn((B, 3, 227, 227)).cuda()
b.label = torch.randint(0, 1000, (B, 1)).float().cuda()
b.batch_size = B
for _ in range(5):
    tr.update(b)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    tr.update(b)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
