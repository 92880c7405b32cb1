"""dcnn_b200 -- a Blackwell-native (sm_100a) pipeline-parallel CNN training framework with the
capabilities, public API and checkpoint format of tungphambasement/DCNN ("TNN")."""
__version__ = "0.1.0"

from . import ops  # noqa: F401
from .nn import *  # noqa: F401,F403
from .nn import Sequential, SequentialBuilder, Partition  # noqa: F401


_LAZY = ("models", "train", "io", "data", "partitioner", "parallel", "utils", "serve")


def __getattr__(name):
    """Subpackages on first use (``dcnn_b200.models``, ``dcnn_b200.train`` ...): importing the package stays cheap."""
    if name in _LAZY:
        import importlib
        mod = importlib.import_module(f"{__name__}.{name}")
        globals()[name] = mod
        return mod
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
