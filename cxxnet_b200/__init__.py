"""cxxnet_b200: a Blackwell-native CNN training framework with cxxnet's capabilities.

Front end (``.conf`` grammar, task modes, trainer, layer zoo, updaters, ``.model`` format,
``wrapper.cxxnet`` Python API) over hand-written sm_100a kernels (``csrc/kernels``) and a fused
NVLink gradient exchange (``parallel``).  See DESIGN.md.
"""
__version__ = "0.1.0"

from .config import parse_config_file, parse_config_string  # noqa: F401
from .netconfig import NetConfig  # noqa: F401
from .trainer import NetTrainer, create_net  # noqa: F401
