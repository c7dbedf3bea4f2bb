# Same module path as the reference (flash_attention/load_cpp_extention.py).
from flash_attention_cute_b200.load_cpp_extention import load_extension  # noqa: F401
