"""B200-native batched physics backend behind dm_control's Physics / suite API.

Hot path: `BatchedPhysics.step()` (= mj_step2 + mj_step1 of the reference,
/root/reference/dm_control/mujoco/engine.py:149-166) over a leading env axis,
executed by hand-written sm_100a CUDA kernels behind a C-ABI (`include/bphys.h`).
"""

__version__ = "0.1.0"
