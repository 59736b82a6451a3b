"""instance_based_loc_b200 -- B200 (sm_100a) implementation of the per-frame object-localisation
hot path of instance-based-loc: hand-written CUDA kernels behind the C ABI in
include/ibl.h (libibl.so), plus the Python mirror of the reference's own module signatures
(utils/depth_utils.py, utils/similarity_volume.py, utils/fpfh_register.py,
object_memory.ObjectMemory.localise).  There is no CPU fallback anywhere in this package."""
__version__ = "0.1.0"

import os as _os
import sys as _sys

# Serving several frames at once (ObjectMemory.localise_batch) allocates per-frame tensors on several CUDA streams.  torch's
# default caching allocator then keeps calling cudaMalloc (a block size it has not cached for that stream), which stalls
# the whole device for the length of a step every few steps; the stream-ordered backend shares the driver pool libibl.so
# uses and stops that.  It can only be chosen before torch initialises CUDA, so it is set here when torch has not been
# imported yet and never overrides the user's own setting.
if "torch" not in _sys.modules:
    _os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "backend:cudaMallocAsync")
