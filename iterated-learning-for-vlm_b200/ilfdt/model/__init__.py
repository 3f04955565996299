"""Mirror of prototype/model/__init__.py:1-6 for the classes on the hot path."""
from .clip import CLIP  # noqa: F401
from .clip_fdt import AllGather, Clip_FDT, Query_model, clip_fdt_vitb16, clip_fdt_vitb32  # noqa: F401
from .declip_fdt import DECLIP_FDT, declip_fdt_vitb32  # noqa: F401
from .sparsemax import Sparsemax  # noqa: F401


def model_entry(config):
    """model_entry(config) = globals()[config['type']](**config['kwargs'])   (prototype/model/__init__.py:5-6)"""
    return globals()[config['type']](**config['kwargs'])
