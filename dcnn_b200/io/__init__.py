from .tensor_codec import read_tensor, write_tensor, tensor_record_bytes  # noqa: F401
from .training_state import load_training_state, save_training_state  # noqa: F401,E402
from .torch_export import load_from_torch_module, to_torch_module  # noqa: F401,E402
