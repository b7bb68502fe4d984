"""Inspect a ``.model`` checkpoint without a configuration file:

    python -m cxxnet_b200.tools.model_info models/0010.model [--weights]

prints the net type, the update counter, the input shape, every node and layer of the stored structure
(reference layout: SURVEY.md section 2.11 / ``src/nnet/nnet_config.h:126-191``) and, for each layer with
parameters, their shapes - with ``--weights`` also mean / std / min / max.  The weights are read by rebuilding the
net on the CPU backend from the stored structure alone (layer keys such as kernel sizes live in each layer's
stored parameter block).
"""
from __future__ import annotations

import argparse
import io
import struct
import sys

import torch

from .. import model_io as mio
from .. import netconfig as nc
from ..executor import NeuralNet


def describe(path: str, weights: bool = False, out=sys.stdout) -> int:
    with open(path, "rb") as f:
        net_type = struct.unpack("<i", f.read(4))[0]
        cfg = nc.NetConfig()
        cfg.load_net(f)
        epoch = mio.read_i64(f)
        blob = mio.read_bytes(f)
    names = {v: k for k, v in vars(nc).items() if k.startswith("k") and isinstance(v, int)}
    print("file: %s" % path, file=out)
    print("net_type %d, update counter %d, input %s, %d nodes, %d layers, %d bytes of weights"
          % (net_type, epoch, "x".join(str(v) for v in cfg.input_shape), cfg.num_nodes, cfg.num_layers, len(blob)),
          file=out)
    print("nodes: " + ", ".join("%d:%s" % (i, n) for i, n in enumerate(cfg.node_names)), file=out)
    net = NeuralNet(cfg, 1, 0, torch.device("cpu"), "ref", silent=1)
    net.load_model(io.BytesIO(blob), infer_shapes=False)
    total = 0
    for i, (info, c) in enumerate(zip(cfg.layers, net.connections)):
        tname = names.get(info.type, str(info.type))
        if info.type == nc.kSharedLayer:
            tname = "share[layer %d]" % info.primary_layer_index
        line = "layer %2d  %-22s %-12s %s -> %s" % (
            i, tname[1:] if tname.startswith("k") else tname, info.name or "-",
            ",".join(cfg.node_names[j] for j in info.nindex_in), ",".join(cfg.node_names[j] for j in info.nindex_out))
        print(line, file=out)
        if info.type == nc.kSharedLayer:
            continue
        for tag, w, _ in c.layer.weights():
            total += w.numel()
            desc = "    %-5s %s" % (tag, tuple(w.shape))
            if weights:
                wf = w.float()
                desc += "  mean %.4g std %.4g min %.4g max %.4g" % (float(wf.mean()), float(wf.std()) if wf.numel() > 1 else 0.0,
                                                                   float(wf.min()), float(wf.max()))
            print(desc, file=out)
    print("parameters: %d" % total, file=out)
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="describe a cxxnet .model checkpoint")
    ap.add_argument("model")
    ap.add_argument("--weights", action="store_true", help="statistics of every weight tensor")
    args = ap.parse_args(argv)
    return describe(args.model, args.weights)


if __name__ == "__main__":
    sys.exit(main())
