"""lara_b200 -- B200-native implementation of LaRA 2's all-pairs Lagrangian subgradient loop.

The package holds only what the hot path needs: csrc/ (CUDA kernels + C ABI + thin host C++) and a
ctypes mirror of the reference's interface (api.py).  See DESIGN.md.
"""
from .api import (Context, InputStorage, LaraError, OutputLibrary, Parameters, results_to_list,  # noqa: F401
                  solve)
