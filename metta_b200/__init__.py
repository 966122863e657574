"""metta_b200: a batched MettaGrid simulator for NVIDIA B200 (sm_100a).

Only the step hot path of the reference is implemented: ordered action
pass, assembler/chest/inventory/reward update, 11x11 token observation encode, auto-reset from a
device-resident map pool. Host code lowers the reference's MettaGridConfig to a flat blob and
drives hand-written CUDA kernels through the C-ABI in include/mettagrid_b200.h.
"""

from .config import MettaGridConfig  # noqa: F401

__all__ = ["MettaGridConfig"]
