/* Node-type numbering and flat BlobTree description shared by the product C ABI (psgpu.h) and the
 * test oracles. The reference has two incompatible BlobNodeType numberings
 * (Parsip100/PS_SimdPoly/include/PS_Polygonizer.h:86-90 standalone vs
 * Parsip100/PS_BlobTree/include/_constSettings.h:26-38 with the scene API), so the
 * C ABI carries its own; the C++ facade translates from whichever reference enum is in force. */
#ifndef PSGPU_TYPES_H
#define PSGPU_TYPES_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum psgpu_node_type {
    /* skeletal primitives (Wyvill field of the squared distance to the skeleton) */
    PSGPU_PRIM_POINT = 0,
    PSGPU_PRIM_LINE = 1,      /* pos = start, dir = END point (not a direction) */
    PSGPU_PRIM_CYLINDER = 2,  /* pos, dir (unit), res[0] = radius, res[1] = height */
    PSGPU_PRIM_DISC = 3,      /* pos, dir = normal (unit), res[0] = radius */
    PSGPU_PRIM_RING = 4,      /* pos, dir = normal (unit), res[0] = radius */
    PSGPU_PRIM_CUBE = 5,      /* pos, res[0] = half side */
    PSGPU_PRIM_NULL = 6,      /* field == 0 */
    /* non-skeletal primitives of the scene API; PSGPU_SEM_SCALAR only, the SIMD evaluator has neither */
    PSGPU_PRIM_QUADRICPOINT = 7, /* pos, res[0] = field radius R, res[1] = field scale s: f = s (1 - d2/R2)^2 inside R
                                    (PS_BlobTree/include/CQuadricPoint.h:43-52, CompactBlobTree.cpp:1051-1064) */
    PSGPU_PRIM_INSTANCE = 8,     /* res[0] = flat index of the instanced node; its own xform is applied first
                                    (CompactBlobTree.cpp:1069-1078) */
    /* n-ary operators */
    PSGPU_OP_UNION = 16,
    PSGPU_OP_INTERSECT = 17,
    PSGPU_OP_DIF = 18,
    PSGPU_OP_SMOOTHDIF = 19,
    PSGPU_OP_BLEND = 20,
    PSGPU_OP_RICCI = 21,      /* res[0] = n (res[1] = 1/n is derived) */
    /* unary space warps: res as in PS_HighPerformanceRender.cpp:128-159 */
    PSGPU_OP_WARP_TWIST = 32, /* res[0] = factor, res[1] = major axis (0 x, 1 y, 2 z) */
    PSGPU_OP_WARP_TAPER = 33, /* res[0] = factor, res[1] = axis along, res[2] = axis taper */
    PSGPU_OP_WARP_BEND = 34,  /* res[0] = rate, res[1] = center, res[2] = left, res[3] = right */
    PSGPU_OP_WARP_SHEAR = 35  /* res[0] = factor, res[1] = axis along, res[2] = axis dependent */
};
#define PSGPU_IS_PRIM(t) ((t) < 16)
#define PSGPU_IS_WARP(t) ((t) >= 32)

/* Status codes: the reference's own (PS_Polygonizer.h:47-50) plus the linearizer's
 * (PS_HighPerformanceRender.cpp:10-12) and two for conditions the reference cannot express. */
#define PSGPU_SUCCESS 1
#define PSGPU_PARAM_ERROR -1
#define PSGPU_NOT_ENOUGH_MEM -2
#define PSGPU_INVALID_BVH -3
#define PSGPU_ERROR_UNSUPPORTED_NODE -4
#define PSGPU_ERROR_CUDA -5
#define PSGPU_ERROR_NO_DEVICE -6

/* Flat n-ary BlobTree: one record per node, children as index ranges into child[].
 * This is what a CBlobNode* tree (Parsip100/PS_BlobTree/include/CBlobTree.h:29-394) flattens to. */
typedef struct psgpu_tree {
    int32_t n_nodes;
    int32_t root;
    const int32_t* type;        /* [n_nodes] psgpu_node_type */
    const int32_t* first_child; /* [n_nodes] index into child[] */
    const int32_t* n_children;  /* [n_nodes] */
    const int32_t* child;       /* concatenated child lists */
    const float* pos;           /* [n_nodes*3] primitives */
    const float* dir;           /* [n_nodes*3] primitives */
    const float* res;           /* [n_nodes*4] primitive res.xyz / operator params res.xyzw */
    const float* color;         /* [n_nodes*3] primitive diffuse colour */
    const float* xform;         /* [n_nodes*12] backward (world->local) 3x4 row-major; NULL = all identity */
    const uint8_t* has_xform;   /* [n_nodes] 0 = identity; NULL = all identity */
    /* optional (NULL = absent): the scene API's own transform parametrisation, CAffineTransformation = scale xyz, rotation
     * quaternion xyzw, translate xyz (PS_FrameWork/include/PS_AffineTransformation.h). Needed by psgpu_tree_flatten_transforms
     * (the reference composes these triples, not matrices) and written / read by the .scene I/O. */
    const float* srt;           /* [n_nodes*10] */
    const uint8_t* has_srt;     /* [n_nodes] */
    const int32_t* node_id;     /* [n_nodes] CBlobNode::getID() (the BLOBNODE section numbers of a .scene file); NULL = index */
    /* optional (NULL = absent), read by PSGPU_LIN_AS_SHIPPED only: operator boxes as the reference's PrepareBBoxes leaves them in
     * SOABlobOps.vBoxLo / vBoxHi (PS_SimdPoly/include/PS_Polygonizer.cpp:238-307); entries of primitives are ignored */
    const float* op_box_lo;     /* [n_nodes*3] */
    const float* op_box_hi;     /* [n_nodes*3] */
} psgpu_tree;

#ifdef __cplusplus
}
#endif
#endif
