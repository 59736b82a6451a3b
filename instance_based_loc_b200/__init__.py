"""instance_based_loc_b200 -- B200 (sm_100a) implementation of the per-frame object-localisation
hot path of instance-based-loc: hand-written CUDA kernels behind the C ABI in
include/ibl.h (libibl.so), plus the Python mirror of the reference's own module signatures
(utils/depth_utils.py, utils/similarity_volume.py, utils/fpfh_register.py,
object_memory.ObjectMemory.localise).  There is no CPU fallback anywhere in this package."""
__version__ = "0.1.0"
