from .tensor_codec import read_tensor, write_tensor, tensor_record_bytes  # noqa: F401
