"""B200 (sm_100a) lattice-Boltzmann hot path behind the C ABI of include/lbm_b200.h.

The package is a thin ctypes binding over ``liblbm_b200.so`` (hand-written CUDA,
built in-tree by ``latticeboltzmannmethod_b200.build``).  There is no CPU path:
importing :mod:`latticeboltzmannmethod_b200.api` raises if the library is
missing, and every compute call fails with the library's own error text when no
sm_100a GPU is present.
"""
from .api import (Context, Config, LbmError, MODELS, FIELDS, HALO, load_library, library_path,  # noqa: F401
                  shanchen_config, binary2d_config, mrt_config, microchannel_config)
