"""parsip_b200: B200-native (sm_100a) polygonization path of etrigger/Parsip.

The product is the C-ABI library `libpsgpu.so` (include/psgpu.h) built from parsip_b200/csrc; this
package holds its ctypes binding (`api`), a host-side BlobTree description (`blobtree`) and the seeded
scene generators of SURVEY.md section 8(d) (`scenes`). There is no CPU fallback: every compute entry point
raises if the CUDA extension is missing or no GPU is present.
"""
from . import blobtree, scenes, sharding  # noqa: F401
from .api import (  # noqa: F401
    PsGpuError, Context, Program, Params, linearize, prepare_bboxes, binarize_tree, count_mpus, lattice_dims, lib_path, load_library, comm_unique_id, slab_bounds, flatten_transforms, tree_octrees, scene_load, scene_save,
    SEM_SIMD, SEM_SCALAR, ROOT_FOURSAMPLE, ROOT_NEWTON_REF, ROOT_BISECT, COORD_MPU_LOCAL, COORD_GLOBAL,
    OUT_INDEXED, OUT_MPU, CULL_NONE, CULL_SUPPORT, CULL_INTERVAL, LIN_REASSOCIATE, LIN_FULL_AFFINE, LIN_AS_SHIPPED, NOALLOC_MAX_POINTS,
)
