"""Network structure configuration: the ``netconfig=start ... netconfig=end`` graph grammar.

Behavioural parity with reference ``src/nnet/nnet_config.h:26-411``:
``layer[a->b]=type:name``, ``layer[+1]`` / ``layer[+0]`` / ``layer[+1:tag]`` sugar, multi-node
``layer[a,b->c]``, ``share[tag]`` layers, ``label_vec[a,b)=name``, ``extra_data_num`` /
``extra_data_shape[i]``; pairs are split into global ``defcfg`` and per-layer ``layercfg``;
``save_net`` / ``load_net`` write the binary structure block of the ``.model`` file.
"""
from __future__ import annotations

import re
import struct
from typing import BinaryIO, Dict, List, Tuple

from . import model_io as mio
from .config import ConfigError, ConfigPairs

# layer type ids (reference src/layer/layer.h:284-315)
kSharedLayer = 0
kFullConnect = 1
kSoftmax = 2
kRectifiedLinear = 3
kSigmoid = 4
kTanh = 5
kSoftplus = 6
kFlatten = 7
kDropout = 8
kConv = 10
kMaxPooling = 11
kSumPooling = 12
kAvgPooling = 13
kLRN = 15
kBias = 17
kConcat = 18
kXelu = 19
kCaffe = 20
kReluMaxPooling = 21
kMaxout = 22
kSplit = 23
kInsanity = 24
kInsanityPooling = 25
kL2Loss = 26
kMultiLogistic = 27
kChConcat = 28
kPRelu = 29
kBatchNorm = 30
kFixConnect = 31
kPairTestGap = 1024

_TYPE_BY_NAME = {
    "fullc": kFullConnect, "fixconn": kFixConnect, "bias": kBias, "softmax": kSoftmax,
    "relu": kRectifiedLinear, "sigmoid": kSigmoid, "tanh": kTanh, "softplus": kSoftplus,
    "flatten": kFlatten, "dropout": kDropout, "conv": kConv, "relu_max_pooling": kReluMaxPooling,
    "max_pooling": kMaxPooling, "sum_pooling": kSumPooling, "avg_pooling": kAvgPooling, "lrn": kLRN,
    "concat": kConcat, "xelu": kXelu, "maxout": kMaxout, "split": kSplit, "insanity": kInsanity,
    "insanity_max_pooling": kInsanityPooling, "l2_loss": kL2Loss, "multi_logistic": kMultiLogistic,
    "ch_concat": kChConcat, "prelu": kPRelu, "batch_norm": kBatchNorm,
}
_NAME_BY_TYPE = {v: k for k, v in _TYPE_BY_NAME.items()}


def get_layer_type(name: str) -> int:
    """String -> layer type id (reference ``layer.h:322-361``)."""
    if name.startswith("share"):
        return kSharedLayer
    if name in _TYPE_BY_NAME:
        return _TYPE_BY_NAME[name]
    if name.startswith("pairtest-"):
        m = re.match(r"([^-]+)-([^:]+)", name[9:])
        if not m:
            raise ConfigError('unknown layer type: "%s"' % name)
        return kPairTestGap * get_layer_type(m.group(1)) + get_layer_type(m.group(2))
    raise ConfigError('unknown layer type: "%s"' % name)


def layer_type_name(t: int) -> str:
    if t == kSharedLayer:
        return "share"
    if t >= kPairTestGap:
        return "pairtest-%s-%s" % (layer_type_name(t // kPairTestGap), layer_type_name(t % kPairTestGap))
    return _NAME_BY_TYPE.get(t, "unknown(%d)" % t)


class LayerInfo:
    __slots__ = ("type", "primary_layer_index", "name", "nindex_in", "nindex_out")

    def __init__(self):
        self.type = 0
        self.primary_layer_index = -1
        self.name = ""
        self.nindex_in: List[int] = []
        self.nindex_out: List[int] = []

    def same(self, o: "LayerInfo") -> bool:
        return (self.type == o.type and self.primary_layer_index == o.primary_layer_index
                and self.name == o.name and self.nindex_in == o.nindex_in and self.nindex_out == o.nindex_out)


class NetConfig:
    NETPARAM_BYTES = 152  # sizeof(NetParam): 2 int + 3 uint + 2 int + int reserved[31]

    def __init__(self):
        # NetParam
        self.num_nodes = 0
        self.num_layers = 0
        self.input_shape = (0, 0, 0)  # (c, h, w)
        self.init_end = 0
        self.extra_data_num = 0
        # structure
        self.layers: List[LayerInfo] = []
        self.node_names: List[str] = []
        self.extra_shape: List[int] = []
        # training-time state (not saved)
        self.node_name_map: Dict[str, int] = {}
        self.layer_name_map: Dict[str, int] = {}
        self.updater_type = "sgd"
        self.sync_type = "simple"
        self.label_name_map: Dict[str, int] = {"label": 0}
        self.label_range: List[Tuple[int, int]] = [(0, 1)]
        self.defcfg: ConfigPairs = []
        self.layercfg: List[ConfigPairs] = []

    # ------------------------------------------------------------------ binary structure
    def save_net(self, fo: BinaryIO) -> None:
        assert self.num_layers == len(self.layers), "model inconsistent"
        assert self.num_nodes == len(self.node_names), "num_nodes is inconsistent with node_names"
        c, h, w = self.input_shape
        fo.write(struct.pack("<ii3Iii31i", self.num_nodes, self.num_layers, c, h, w, self.init_end,
                             self.extra_data_num, *([0] * 31)))
        if self.extra_data_num != 0:
            mio.write_i32_vec(fo, self.extra_shape)
        for n in self.node_names:
            mio.write_str(fo, n)
        for l in self.layers:
            mio.write_i32(fo, l.type)
            mio.write_i32(fo, l.primary_layer_index)
            mio.write_str(fo, l.name)
            mio.write_i32_vec(fo, l.nindex_in)
            mio.write_i32_vec(fo, l.nindex_out)

    def load_net(self, fi: BinaryIO) -> None:
        raw = fi.read(self.NETPARAM_BYTES)
        if len(raw) != self.NETPARAM_BYTES:
            raise mio.ModelFormatError("NetConfig: invalid model file")
        v = struct.unpack("<ii3Iii31i", raw)
        self.num_nodes, self.num_layers = v[0], v[1]
        self.input_shape = (v[2], v[3], v[4])
        self.init_end, self.extra_data_num = v[5], v[6]
        if self.extra_data_num != 0:
            self.extra_shape = mio.read_i32_vec(fi)
        self.node_names = [mio.read_str(fi) for _ in range(self.num_nodes)]
        self.node_name_map = {n: i for i, n in enumerate(self.node_names)}
        self.layers = []
        self.layer_name_map = {}
        for i in range(self.num_layers):
            l = LayerInfo()
            l.type = mio.read_i32(fi)
            l.primary_layer_index = mio.read_i32(fi)
            l.name = mio.read_str(fi)
            l.nindex_in = mio.read_i32_vec(fi)
            l.nindex_out = mio.read_i32_vec(fi)
            if l.type == kSharedLayer:
                if l.name:
                    raise mio.ModelFormatError("SharedLayer must not have name")
            elif l.name:
                if l.name in self.layer_name_map:
                    raise mio.ModelFormatError("NetConfig: invalid model file, duplicated layer name: %s" % l.name)
                self.layer_name_map[l.name] = i
            self.layers.append(l)
        self.layercfg = [[] for _ in range(self.num_layers)]
        self._clear_config()

    # ------------------------------------------------------------------ configuration
    def _set_global_param(self, name: str, val: str) -> None:
        if name == "updater":
            self.updater_type = val
        if name == "sync":
            self.sync_type = val
        m = re.match(r"label_vec\[(\d+),(\d+)\)", name)
        if m:
            self.label_range.append((int(m.group(1)), int(m.group(2))))
            self.label_name_map[val] = len(self.label_range) - 1

    def configure(self, cfg: ConfigPairs) -> None:
        self._clear_config()
        if not self.node_names and not self.node_name_map:
            self.node_names.append("in")
            self.node_name_map["in"] = 0
        self.node_name_map["0"] = 0
        netcfg_mode = 0
        cfg_top_node = 0
        cfg_layer_index = 0
        for name, val in cfg:
            if name == "extra_data_num":
                num = int(val)
                for i in range(num):
                    nm = "in_%d" % (i + 1)
                    if nm not in self.node_name_map:
                        self.node_names.append(nm)
                        self.node_name_map[nm] = i + 1
                self.extra_data_num = num
            if name.startswith("extra_data_shape["):
                m = re.match(r"extra_data_shape\[(\d+)", name)
                parts = val.split(",")
                if not m or len(parts) != 3:
                    raise ConfigError("extra data shape config incorrect")
                self.extra_shape.extend(int(p) for p in parts)
            if self.init_end == 0 and name == "input_shape":
                parts = val.split(",")
                if len(parts) != 3:
                    raise ConfigError("input_shape must be three consecutive integers without space example: 1,1,200 ")
                self.input_shape = (int(parts[0]), int(parts[1]), int(parts[2]))
            if netcfg_mode != 2:
                self._set_global_param(name, val)
            if name == "netconfig" and val == "start":
                netcfg_mode = 1
            if name == "netconfig" and val == "end":
                netcfg_mode = 0
            if name.startswith("layer["):
                info = self._get_layer_info(name, val, cfg_top_node, cfg_layer_index)
                netcfg_mode = 2
                if self.init_end == 0:
                    assert len(self.layers) == cfg_layer_index, "NetConfig inconsistent"
                    self.layers.append(info)
                    self.layercfg.append([])
                else:
                    if cfg_layer_index >= len(self.layers):
                        raise ConfigError("config layer index exceed bound")
                    if not info.same(self.layers[cfg_layer_index]):
                        raise ConfigError("config setting does not match existing network structure")
                cfg_top_node = info.nindex_out[0] if len(info.nindex_out) == 1 else -1
                cfg_layer_index += 1
                continue
            if netcfg_mode == 2:
                if self.layers[cfg_layer_index - 1].type == kSharedLayer:
                    raise ConfigError("please do not set parameters in shared layer, set them in primary layer")
                self.layercfg[cfg_layer_index - 1].append((name, val))
            else:
                self.defcfg.append((name, val))
        if self.init_end == 0:
            self._init_net()

    def get_layer_index(self, name: str) -> int:
        if name not in self.layer_name_map:
            raise KeyError("unknown layer name %s" % name)
        return self.layer_name_map[name]

    # ------------------------------------------------------------------ helpers
    def _get_layer_info(self, name: str, val: str, top_node: int, cfg_layer_index: int) -> LayerInfo:
        inf = LayerInfo()
        m_inc = re.match(r"layer\[\+(\d+)", name)
        if m_inc:
            inc = int(m_inc.group(1))
            if top_node < 0:
                raise ConfigError("ConfigError: layer[+1] is used, but last layer have more than one output"
                                  "use layer[input-name->output-name] instead")
            inf.nindex_in.append(top_node)
            m_tag = re.match(r"layer\[\+1:([^\]]+)\]", name)
            if m_tag:
                inf.nindex_out.append(self._get_node_index(m_tag.group(1), True))
            elif inc == 0:
                inf.nindex_out.append(top_node)
            else:
                inf.nindex_out.append(self._get_node_index("!node-after-%d" % top_node, True))
        else:
            m = re.match(r"layer\[([^-]+)->([^\]]+)\]", name)
            if not m:
                raise ConfigError("ConfigError: invalid layer format %s" % name)
            for s in m.group(1).split(","):
                if s:
                    inf.nindex_in.append(self._get_node_index(s, False))
            for s in m.group(2).split(","):
                if s:
                    inf.nindex_out.append(self._get_node_index(s, True))
        layer_name = ""
        if ":" in val and not val.startswith(":"):
            ltype, rest = val.split(":", 1)
            rest = rest.split()[0] if rest.split() else ""
            if rest:
                layer_name = rest
                inf.type = get_layer_type(ltype)
            else:
                inf.type = get_layer_type(val)
                ltype = val
        else:
            ltype = val
            inf.type = get_layer_type(val)
        if inf.type == kSharedLayer:
            lb = ltype.find("[")
            if lb < 0:
                raise ConfigError("ConfigError: shared layer must specify tag of layer to share with")
            s_tag = ltype[lb + 1:]
            s_tag = s_tag[:-1]
            if s_tag not in self.layer_name_map:
                raise ConfigError("ConfigError: shared layer tag %s is not defined before" % s_tag)
            inf.primary_layer_index = self.layer_name_map[s_tag]
        elif layer_name:
            if layer_name in self.layer_name_map:
                if self.layer_name_map[layer_name] != cfg_layer_index:
                    raise ConfigError("ConfigError: layer name in the configuration file do not "
                                      "match the name stored in model")
            else:
                self.layer_name_map[layer_name] = cfg_layer_index
            inf.name = layer_name
        return inf

    def _get_node_index(self, name: str, alloc_unknown: bool) -> int:
        if name in self.node_name_map:
            return self.node_name_map[name]
        if not alloc_unknown:
            raise ConfigError("ConfigError: undefined node name %s, input node of a layer must be specified as "
                              "output of another layer presented before the layer declaration" % name)
        idx = len(self.node_names)
        self.node_name_map[name] = idx
        self.node_names.append(name)
        return idx

    def _init_net(self) -> None:
        self.num_nodes = 0
        self.num_layers = len(self.layers)
        for l in self.layers:
            for j in l.nindex_in + l.nindex_out:
                self.num_nodes = max(self.num_nodes, j + 1)
        assert self.num_nodes == len(self.node_names), "num_nodes is inconsistent with node_names"
        self.init_end = 1

    def _clear_config(self) -> None:
        self.defcfg = []
        self.layercfg = [[] for _ in self.layercfg]
