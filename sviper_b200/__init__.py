"""sviper_b200 -- B200 (sm_100a) implementation of SViper's per-variant polishing hot path.
The product is the C-ABI library sviper_b200/csrc/libsviper_b200.so (include/sviper_b200.h);
sviper_b200.api is the ctypes host side used by tests and bench.py."""
from . import api, synth  # noqa: F401
