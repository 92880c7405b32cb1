"""dcnn_b200 -- a Blackwell-native (sm_100a) pipeline-parallel CNN training framework with the
capabilities, public API and checkpoint format of tungphambasement/DCNN ("TNN")."""
__version__ = "0.1.0"

from . import ops  # noqa: F401
from .nn import *  # noqa: F401,F403
from .nn import Sequential, SequentialBuilder, Partition  # noqa: F401
