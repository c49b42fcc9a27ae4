"""steviapp_b200 -- B200-native (sm_100a, FP64 CUDA) bundle-adjustment engine behind the solver /
projector-module interface of french-paragon/steviapp's libs/sparsesolver.

The product is libsba_b200.so (C ABI: include/sba_b200.h); this package is its host-side mirror.
"""
from .solver import SBASolver, SBAError, default_options, comm_unique_id, partition_points  # noqa: F401
from . import synthetic  # noqa: F401
from . import bal  # noqa: F401
from . import project  # noqa: F401

__all__ = ["SBASolver", "SBAError", "default_options", "comm_unique_id", "partition_points", "synthetic", "bal", "project"]
