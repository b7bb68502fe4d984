from .base import LabelInfo, Layer, LayerParam, Node  # noqa: F401
from .factory import create_layer  # noqa: F401
