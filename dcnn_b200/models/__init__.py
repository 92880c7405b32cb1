from .zoo import *  # noqa: F401,F403
from .zoo import MODEL_ZOO  # noqa: F401
