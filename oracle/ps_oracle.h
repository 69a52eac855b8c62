/* TEST INFRASTRUCTURE ONLY - CPU restatement ("port") of the reference's polygonization path in plain C.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the checker. The product (parsip_b200/csrc) never links it.
 *
 * Two semantics are restated, because the reference has two CPU polygonizers:
 *   pso_simd_*   : PS_SimdPoly (Parsip100/PS_SimdPoly/include/PS_Polygonizer.cpp) - binary SOA tree,
 *                  infinite lines, warps are pass-through, 4-sample root + lerp, inside = f >= iso.
 *                  PINNED: bit-exact against the reference's own code compiled unmodified into
 *                  oracle/_ref/libpsref.so (tests/test_oracle_pinning.py) and against the golden
 *                  vectors under tests/golden/ generated from that build. Documented deviations:
 *                  _mm_rsqrt_ps / _mm_rcp_ps (12-bit, CPU-vendor specific; disc, ring, Ricci,
 *                  normal normalisation) are evaluated exactly.
 *   pso_scalar_* : the scalar twin (Parsip100/ParsipHaptics/include/CompactBlobTree.cpp +
 *                  CPolyParsipOptimized.cpp) - n-ary tree, clamped segments, warps, Newton root,
 *                  inside = f > iso. That code needs Qt + Loki and cannot be compiled here, and the
 *                  reference has no tests: PARITY UNPINNED for this half (pinned only where it must
 *                  agree with the SIMD half: point/cylinder/cube primitives and blend/CSG operators).
 */
#ifndef PS_ORACLE_H
#define PS_ORACLE_H
#include <stdint.h>
#include "../include/psgpu_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Binary SOA tree, the content of SOABlobPrims/SOABlobOps/SOABlobPrimMatrices
 * (PS_Polygonizer.h:96-172) as SimdPoly::linearizeBlobTree fills them. Node types use psgpu_node_type. */
typedef struct pso_soa_desc {
    int32_t n_prims, n_ops, n_mats;
    const int32_t* prim_type;
    const float* prim_pos;   /* n*3 */
    const float* prim_dir;   /* n*3 */
    const float* prim_res;   /* n*3 */
    const float* prim_color; /* n*3 */
    const int32_t* prim_mat; /* matrix slot, 0 = identity */
    const int32_t* op_type;
    const int32_t* op_left;
    const int32_t* op_right;
    const int32_t* op_kind;  /* bit1: left is op, bit0: right is op */
    const float* op_res;     /* n*4 */
    const float* mats_bwd;   /* n_mats*12, slot 0 identity */
    const float* mats_fwd;   /* n_mats*16, slot 0 identity */
} pso_soa_desc;

typedef struct pso_simd pso_simd;
typedef struct pso_scalar pso_scalar;

#define PSO_GRID_DIM 8
#define PSO_ISO_VALUE 0.5f
#define PSO_NORMAL_DELTA 0.001f
#define PSO_FIELDVALUE_EPSILON 0.001f

/* ---- SIMD-path semantics ---- */
pso_simd* pso_simd_create(const pso_soa_desc* d);
void pso_simd_destroy(pso_simd* s);
int pso_simd_prepare_bboxes(pso_simd* s, float cellsize);
void pso_simd_get_bbox(const pso_simd* s, float* lo, float* hi);
void pso_simd_set_bbox(pso_simd* s, const float* lo, const float* hi);
void pso_simd_get_prim_boxes(const pso_simd* s, float* out6);
void pso_simd_get_op_boxes(const pso_simd* s, float* out6);
void pso_simd_pad_op_boxes(pso_simd* s, float pad);
uint32_t pso_count_mpus(float cellsize, const float* lo, const float* hi);
void pso_simd_field(const pso_simd* s, int64_t n, const float* xyz, float* out);
void pso_simd_newton(const pso_simd* s, int64_t n, const float* p1, const float* p2, const float* fp1, const float* fp2,
                     float target, int iterations, float* root, float* root_field, int32_t* evals);
void pso_simd_field_color_normal(const pso_simd* s, int64_t n, const float* xyz, float* f, float* col, float* nrm);
/* field cache (z contiguous: fv[i*8 + j*64 + k]) and cube cases (cell i*49 + j*7 + k) of one MPU */
void pso_simd_mpu_cache(const pso_simd* s, float cellsize, const int32_t* ijk, float* fv512, uint8_t* case343);
/* Polygonize over the lattice sub-block [b0,e0) (NULL = whole lattice).
 * stats: [0] MPUs processed [1] with triangles [2] vertices [3] triangles [4] SIMD evals [5..7] lattice dims
 *        [8] MPUs whose vertex or triangle count exceeds the reference's 512 capacity */
int pso_simd_polygonize(pso_simd* s, float cellsize, int nthreads, const int32_t* b0, const int32_t* e0,
                        int keep_mesh, int64_t* stats);
void pso_simd_result_counts(const pso_simd* s, int64_t* out3);
void pso_simd_result_copy(const pso_simd* s, float* pos, float* nrm, float* col, uint32_t* tri,
                          int32_t* mpuIdx, uint32_t* mpuCounts);

/* ---- scalar-twin semantics ---- */
pso_scalar* pso_scalar_create(const psgpu_tree* t);
void pso_scalar_destroy(pso_scalar* s);
void pso_scalar_set_prim_boxes(pso_scalar* s, const float* lo3n, const float* hi3n);
void pso_scalar_field(const pso_scalar* s, int64_t n, const float* xyz, float* out);
void pso_scalar_field_color_normal(const pso_scalar* s, int64_t n, const float* xyz, float* f, float* col, float* nrm);
/* ComputeRootNewtonRaphsonVEC4 for n edges: p1,p2 (n*3), f1,f2 (n) -> root (n*3), field (n), evals (n) */
void pso_scalar_newton(const pso_scalar* s, int64_t n, const float* p1, const float* p2, const float* f1,
                       const float* f2, float iso, int iterations, float* root, float* froot, int32_t* evals);
void pso_scalar_mpu_cache(const pso_scalar* s, const float* lo, float cellsize, const int32_t* ijk,
                          float* fv512, uint8_t* case343);
int pso_scalar_polygonize(pso_scalar* s, const float* lo, const float* hi, float cellsize, int nthreads,
                          const int32_t* b0, const int32_t* e0, int keep_mesh, int64_t* stats);
void pso_scalar_result_counts(const pso_scalar* s, int64_t* out3);
void pso_scalar_result_copy(const pso_scalar* s, float* pos, float* nrm, float* col, uint32_t* tri,
                            int32_t* mpuIdx, uint32_t* mpuCounts);

#ifdef __cplusplus
}
#endif
#endif
