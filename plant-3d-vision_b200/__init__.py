"""B200-native voxel back-projection: drop-in for ``plant3dvision.cl.Backprojection`` and the
``Voxels`` task of romi/plant-3d-vision (reference: plant3dvision/cl.py:47-311,
plant3dvision/tasks/cl.py:18-186).

    from plant3dvision_b200.cl import Backprojection      # was: from plant3dvision.cl import ...
    from plant3dvision_b200.tasks.cl import Voxels, run_voxels

Everything numeric runs in hand-written sm_100a CUDA behind the C ABI of ``include/bp_b200.h``
(``libbp_b200.so``).  There is no CPU fallback: importing ``cl`` without the built library, or
constructing a ``Backprojection`` without a CUDA device, raises.
"""
__version__ = "0.1"
