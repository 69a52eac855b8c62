/* psgpu.h - C ABI of the B200-native polygonizer (libpsgpu.so).
 *
 * Drop-in boundary for Parsip's parallel polygonization path. The C++ facade in
 * include/PS_SimdPolyCompat.h (same names/structs as the reference's PS::SIMDPOLY, see
 * INTEGRATION.md) forwards to these entry points; nothing in the signatures is a C++ or torch type.
 *
 * Reference interfaces replaced (all under /root/reference/Parsip100):
 *   psgpu_linearize          <- SimdPoly::linearizeBlobTree        ParsipHaptics/include/PS_HighPerformanceRender.cpp:42-371
 *                               COMPACTBLOBTREE::convert           ParsipHaptics/include/CompactBlobTree.cpp:25-408
 *   psgpu_prepare_bboxes     <- PrepareBBoxes                      PS_SimdPoly/include/PS_Polygonizer.cpp:55-309
 *   psgpu_count_mpus         <- CountMPUNeeded                     PS_SimdPoly/include/PS_Polygonizer.cpp:388-412
 *   psgpu_polygonize         <- Polygonize / CParsipOptimized::run PS_SimdPoly/include/PS_Polygonizer.cpp:315-385,
 *                                                                  ParsipHaptics/include/CPolyParsipOptimized.cpp:392-410
 *   psgpu_field_eval         <- FieldComputer::fieldValue          PS_SimdPoly/include/PS_Polygonizer.cpp:1184-1376
 *                               COMPACTBLOBTREE::fieldvalue        ParsipHaptics/include/CompactBlobTree.cpp:476-487
 *   psgpu_field_color_normal <- FieldComputer::fieldValueAndColor + normal   PS_Polygonizer.cpp:1378-1622
 *   psgpu_newton_roots       <- FieldComputer::computeRootNewtonRaphson   PS_SimdPoly/include/PS_Polygonizer.cpp:1625-1681
 *                               ComputeRootNewtonRaphsonVEC4       ParsipHaptics/include/CompactBlobTree.cpp:1581-1622
 *   psgpu_get_stats          <- PrintThreadResults / MPUSTATS      PS_SimdPoly/include/PS_Polygonizer.cpp:414-428
 *
 * Conventions: every function returns an int status (PSGPU_SUCCESS == 1, errors negative, the
 * reference's RET_* values where one exists, PS_Polygonizer.h:47-50). No exceptions cross the ABI.
 * One context per host thread; calls on a context are serialised by the caller. There is no CPU
 * fallback: without a CUDA device psgpu_create fails with PSGPU_ERROR_NO_DEVICE.
 */
#ifndef PSGPU_H
#define PSGPU_H
#include <stddef.h>
#include <stdint.h>
#include "psgpu_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct psgpu_ctx psgpu_ctx;
typedef struct psgpu_program psgpu_program; /* host-side linearized BlobTree */

/* which reference polygonizer's semantics to reproduce (SURVEY.md "two CPU polygonizers") */
#define PSGPU_SEM_SIMD 0   /* PS_SimdPoly: infinite lines, fast_pow Ricci, warps pass-through, f >= iso, SIMD colour rule */
#define PSGPU_SEM_SCALAR 1 /* CompactBlobTree: clamped segments, true Ricci, warps, f > iso, normalised colours (canonical) */

#define PSGPU_ROOT_FOURSAMPLE 0 /* PS_Polygonizer.cpp:703-762: t = 0,1/3,2/3,1, bracket, lerp */
#define PSGPU_ROOT_NEWTON_REF 1 /* CompactBlobTree.cpp:1581-1622: shrink-wrap Newton, <= 8 iterations */
#define PSGPU_ROOT_BISECT 2     /* bracketed section search (x5 per round) + secant on the edge, re-evaluating the field (native default) */

#define PSGPU_COORD_MPU_LOCAL 0 /* corner = (lo + I*side) + i*cellsize, the reference's expressions */
#define PSGPU_COORD_GLOBAL 1    /* corner = lo + (7I+i)*cellsize: identical across MPUs, slabs and GPUs */

#define PSGPU_OUT_INDEXED 0 /* one indexed mesh, vertices deduplicated by global edge id */
#define PSGPU_OUT_MPU 1     /* per-MPU meshes in the reference's vertex/triangle order (PolyMPUs content) */

#define PSGPU_CULL_NONE 0     /* every MPU evaluates the whole program (brute-force cross-check) */
#define PSGPU_CULL_SUPPORT 1  /* per-MPU pruned lists: primitives whose support misses the MPU are dropped */
#define PSGPU_CULL_INTERVAL 2 /* + MPUs whose field interval excludes the iso value are skipped (default) */

#define PSGPU_LIN_REASSOCIATE 1 /* flatten blend-in-blend chains (changes summation order by a few ulp) */
#define PSGPU_LIN_FULL_AFFINE 2 /* PSGPU_SEM_SCALAR: apply an operator's whole backward matrix. Default (0) reproduces the compact
                                   tree, whose vec4 w lane is 0 at the root and below warps, so an operator's translation column
                                   only acts directly below another operator matrix (CompactBlobTree.cpp:685-694) */
#define PSGPU_LIN_AS_SHIPPED 4  /* PSGPU_SEM_SIMD: reproduce the reference's prune of operators deeper than 3 levels
                                   (PS_Polygonizer.cpp:1228-1252): such an operator evaluates to 0 for a SIMD packet of which no
                                   point lies inside its box's x-, y- OR z-slab. Needs psgpu_tree.op_box_lo / op_box_hi = the operator
                                   boxes PrepareBBoxes computed (PS_Polygonizer.cpp:238-307; boxes padded by 0.504 although the field's
                                   support is 1, so contributions of up to 0.0135 vanish). Field values, positions and indices then equal
                                   the shipped reference's; its COLOURS where the prune fires come from uninitialised stack arrays
                                   (:1384-1385) and are not reproducible. Default (0): every contribution is kept. Binary trees without
                                   warps only; programs get one 3-quad instruction per deep operator. */

typedef struct psgpu_params {
    float cellsize;
    float bbox_lo[3], bbox_hi[3]; /* grid domain; the MPU lattice is laid out from bbox_lo like PS_Polygonizer.cpp:328-371 */
    float iso;                    /* 0 = default ISO_VALUE 0.5 */
    int32_t root_mode;
    int32_t root_iterations; /* 0 = default (8 Newton iterations / 4 section rounds) */
    int32_t coord_mode;
    int32_t output;
    int32_t cull;
    int32_t z_begin, z_end; /* MPU-layer slab [z_begin, z_end); z_end <= 0 = whole lattice (multi-GPU sharding) */
    int32_t want_colors, want_normals; /* 0 skips the colour walk / gradient samples; buffers stay unwritten */
    int32_t collect_stats;             /* 1: also run the op-count kernel that fills prim_evals / algorithmic_flops */
} psgpu_params;

/* Result mesh. In PSGPU_OUT_INDEXED tri[] indexes pos[] directly; in PSGPU_OUT_MPU indices are
 * MPU-local and mpu_* describe the slices (MPUs in lattice order, x outer, z inner). Pointers refer
 * to context-owned DEVICE memory valid until the next psgpu_polygonize / psgpu_destroy. */
typedef struct psgpu_mesh {
    int64_t n_vertices, n_triangles;
    const float* pos; /* 3*n_vertices */
    const float* nrm;
    const float* col;
    const uint32_t* tri; /* 3*n_triangles */
    int64_t n_mpus_out;         /* PSGPU_OUT_MPU: MPUs that produced triangles */
    const int32_t* mpu_ijk;     /* 3*n_mpus_out lattice index */
    const uint32_t* mpu_vbase;  /* n_mpus_out+1 */
    const uint32_t* mpu_tbase;  /* n_mpus_out+1 */
    int32_t lattice[3];
    int32_t over_capacity;      /* PSGPU_OUT_MPU: MPUs exceeding the reference's 512 vertex/triangle capacity */
} psgpu_mesh;

typedef struct psgpu_stats {
    int64_t mpus_total;       /* lattice size (slab) */
    int64_t mpus_evaluated;   /* MPUs whose 512 corners were evaluated */
    int64_t mpus_crossed;     /* MPUs that produced triangles */
    int64_t field_evals;      /* point evaluations of the (pruned) tree: corners + root/gradient samples */
    int64_t prim_evals;       /* primitive instructions executed, summed over points (grid phase) */
    int64_t algorithmic_flops;/* SURVEY 8(d) op counts x executions, grid phase */
    int64_t program_quads;    /* size of the full program */
    int64_t pruned_quads;     /* total size of all per-MPU pruned programs */
    float ms_cull, ms_field, ms_scan, ms_emit, ms_total; /* CUDA-event times of the last polygonize */
    int64_t gpu_launches;     /* kernels launched by the last polygonize */
    float ms_tri, ms_vertex, ms_color; /* the three kernels inside ms_emit */
    float ms_exchange;                 /* psgpu_polygonize_sharded: what the compute stream waited for the count all-gather + halo send/recv
                                          (they run on a stream of their own beside the vertex / colour kernels; inside ms_total) */
    int64_t vertex_eval_flops; /* collect_stats: sum over surface MPUs of vertices x SURVEY 8(d) op count of the MPU's pruned program,
                                  i.e. the algorithmic work of ONE field evaluation per vertex (a vertex costs 8: 4 root samples,
                                  4 gradient samples; its colour walk 1 more) */
    int64_t unresolved_edges;  /* PSGPU_OUT_INDEXED: shared edges missing from their owner's crossing mask (always 0; else the call fails) */
} psgpu_stats;

/* One rank's share of a mesh polygonized over several GPUs (psgpu_polygonize_sharded). The MPU lattice is cut into z-slabs of MPU
 * layers (the reference's parallel_for over independent MPUs, PS_Polygonizer.cpp:379-381, spread over devices); slab-boundary corners
 * come from global lattice indices, so both sides compute bit-identical values, the lower-z rank leaves the vertices of the shared
 * plane to the rank above and learns their ids from it. tri[] of every rank then holds GLOBAL vertex ids: rank r's vertices are
 * [vertex_offset, vertex_offset + n_vertices) of the concatenated array - the same mesh a single GPU produces. */
typedef struct psgpu_shard {
    int32_t nranks, rank;
    int32_t z_begin, z_end;          /* MPU layers of this rank */
    int64_t vertex_offset, triangle_offset;
    int64_t total_vertices, total_triangles;
    float ms_exchange;               /* what the compute stream waited for the count all-gather + halo exchange (CUDA events) */
    int32_t stitched;                /* 1: boundary vertices exist once and tri[] is globally indexed */
} psgpu_shard;

/* ---- context ---- */
int psgpu_create(int device, psgpu_ctx** out);
int psgpu_destroy(psgpu_ctx* ctx);
const char* psgpu_last_error(const psgpu_ctx* ctx);
int psgpu_device_count(void);

/* ---- host side: linearizer and bounding boxes (no GPU needed) ---- */
/* Flattens an n-ary BlobTree into the postfix instruction + parameter stream.
 * Errors: PSGPU_PARAM_ERROR (empty tree), PSGPU_ERROR_UNSUPPORTED_NODE, PSGPU_NOT_ENOUGH_MEM (stack depth). */
int psgpu_linearize(const psgpu_tree* tree, int semantics, int flags, psgpu_program** out);
int psgpu_program_free(psgpu_program* p);
/* info[0] quads, [1] primitives, [2] max value-stack depth, [3] max point-stack depth, [4] instructions */
int psgpu_program_info(const psgpu_program* p, int64_t* info5);
int psgpu_program_copy(const psgpu_program* p, uint32_t* quads_out /* 4 words per quad */);
/* diffuse colour table, 3 floats per primitive id (the id in each primitive's header word) */
int psgpu_program_colors(const psgpu_program* p, float* rgb_out /* 3 * primitives */);
/* Pre-linearization pass (host only): ConvertToBinaryTree::run, ParsipHaptics/include/CLayerManager.cpp:81-233. Operators with
 * more than two children become right-leaning chains of clones of themselves - op(c0, op(c1, ... op(c_{n-2}, c_{n-1}))) - unless
 * split_always == 0 and all the children are primitives; a non-unary operator with one child gets a PSGPU_PRIM_NULL sibling when
 * pad_with_null != 0. Output nodes are numbered in the reference's clone order (= the ids its converted tree carries).
 * n_converted (optional) = operators split + NULL pads. The result is what SimdPoly::linearizeBlobTree requires (exactly two
 * children per operator, PS_HighPerformanceRender.cpp:81-86); psgpu_linearize itself also accepts n-ary trees. */
typedef struct psgpu_tree_buf psgpu_tree_buf;
int psgpu_tree_binarize(const psgpu_tree* in, int split_always, int pad_with_null, psgpu_tree_buf** out, int32_t* n_converted);
const psgpu_tree* psgpu_tree_view(const psgpu_tree_buf* buf);
int psgpu_tree_free(psgpu_tree_buf* buf);
/* CLayer::flattenTransformations (CLayerManager.cpp:604-624): pushes the transforms of n-ary operators down to the primitives and
 * warps below them, composing scale / rotation / translation the way CAffineTransformation::add does (PS_AffineTransformation.h:
 * 46-52: scales multiply, quaternions multiply, translations ADD). The tree must carry `srt`. Output: same topology, operators
 * at identity, leaf srt + backward matrices rebuilt. The SIMD semantics require a flattened tree. */
int psgpu_tree_flatten_transforms(const psgpu_tree* in, psgpu_tree_buf** out);
/* Bounding boxes ("octrees") of all nodes as the scene API computes them: recompute = 0 CBlobNode::getOctree() on a fresh tree,
 * recompute = 1 after CLayer::setOctreeFromBlobTree (CLayerManager.cpp:421-427, 626-645). lo3n / hi3n: 3 floats per node; the root's
 * box is the grid domain the GUI polygonizes. */
int psgpu_tree_octrees(const psgpu_tree* tree, int recompute, float* lo3n, float* hi3n);
/* .scene files (INI, one [BLOBNODE id] section per node): reader = CLayer::recursive_ReadBlobNode (CLayerManager.cpp:684-800) incl.
 * the SkeletonType spelling of the shipped Distrib/train_corrected.scene; writer = CBlobNode::saveScript (CBlobTree.cpp:40-118).
 * root_id < 0: the first section of the file. */
int psgpu_scene_load(const char* path, int root_id, psgpu_tree_buf** out);
int psgpu_scene_save(const psgpu_tree* tree, const char* path);
/* Grid-domain box of the tree the way PrepareBBoxes computes it (iso distance 0.4542 + 0.05 padding). */
int psgpu_prepare_bboxes(const psgpu_tree* tree, float cellsize, float* lo3, float* hi3);
uint32_t psgpu_count_mpus(float cellsize, const float* lo3, const float* hi3);

/* ---- device side ---- */
int psgpu_upload_program(psgpu_ctx* ctx, const psgpu_program* p);
/* Runs cull -> field -> scan -> emit. The mesh stays on the device (see psgpu_mesh). */
int psgpu_polygonize(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out);
/* ---- multi-GPU: one process (or thread) per GPU, one context per rank, NCCL over NVLink. NCCL is bound at run time (libnccl.so.2);
 * single-GPU use never loads it. Rank 0 creates the id, the application hands the 128 bytes to the other ranks by any channel it has
 * (MPI, torch.distributed, a file), every rank calls psgpu_comm_init. */
#define PSGPU_UNIQUE_ID_BYTES 128
int psgpu_comm_get_unique_id(void* id128);
int psgpu_comm_init(psgpu_ctx* ctx, int nranks, int rank, const void* id128);
int psgpu_comm_destroy(psgpu_ctx* ctx);
/* [z_begin, z_end) of `rank`: equal slabs (layer_weights == NULL) or balanced by one weight per MPU layer. */
int psgpu_slab_bounds(int n_layers, int nranks, int rank, const float* layer_weights, int32_t* z_begin, int32_t* z_end);
/* MPUs evaluated / MPUs with triangles per MPU layer of the global lattice in the LAST polygonize of this context (zero outside its
 * slab). Summed over the ranks and weighted it feeds psgpu_slab_bounds for the next frame (re-polygonization, animation). */
int psgpu_layer_work(psgpu_ctx* ctx, int32_t* evaluated, int32_t* crossed, int32_t n_layers);
/* Collective over the communicator: polygonizes this rank's slab (params->z_begin/z_end, or an equal share if z_end <= 0), exchanges
 * counts and the boundary layer's vertex tables, and leaves a stitched share of ONE mesh on every rank (see psgpu_shard).
 * PSGPU_OUT_INDEXED + PSGPU_COORD_GLOBAL only. Without psgpu_comm_init it is psgpu_polygonize. */
int psgpu_polygonize_sharded(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, psgpu_shard* shard);
/* Same, with this rank's share copied into caller-provided HOST buffers while the emission kernels still run. */
int psgpu_polygonize_sharded_host(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, psgpu_shard* shard, float* pos, float* nrm,
                                  float* col, uint32_t* tri, int64_t cap_vertices, int64_t cap_triangles);
/* Collective: gathers the shares into DEVICE buffers on `root` (3 * total_vertices floats each, 3 * total_triangles indices);
 * every rank passes the same set of non-NULL arguments (any valid non-NULL value on the senders means "send this array"). */
int psgpu_mesh_gather(psgpu_ctx* ctx, int root, float* pos, float* nrm, float* col, uint32_t* tri);

/* Same, then copies the mesh into caller-provided HOST buffers (capacity in vertices / triangles;
 * PSGPU_NOT_ENOUGH_MEM if too small, with the needed sizes in out). NULL buffers are skipped. */
int psgpu_polygonize_host(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, float* pos, float* nrm,
                          float* col, uint32_t* tri, int64_t cap_vertices, int64_t cap_triangles);
/* Copies the last device mesh (and the PSGPU_OUT_MPU tables) to host buffers. */
int psgpu_mesh_download(psgpu_ctx* ctx, float* pos, float* nrm, float* col, uint32_t* tri, int32_t* mpu_ijk,
                        uint32_t* mpu_vbase, uint32_t* mpu_tbase);
/* Same content without a caller buffer: the arrays land in a context-owned PINNED host arena (full link rate, no pageable
 * staging) and the pointers stay valid until the next psgpu_polygonize* / psgpu_mesh_map_host on this context. NULL outputs are
 * skipped. This is what the PolyMPUs writer of the facade reads from (the scatter into the reference's MPU structs,
 * PS_SimdPoly/include/PS_Polygonizer.h:180-198, happens on the host). */
int psgpu_mesh_map_host(psgpu_ctx* ctx, const float** pos, const float** nrm, const float** col, const uint32_t** tri,
                        const int32_t** mpu_ijk, const uint32_t** mpu_vbase, const uint32_t** mpu_tbase);
/* Parity hooks: field (and colour / normal) at arbitrary HOST points with the full, unculled program. */
int psgpu_field_eval(psgpu_ctx* ctx, const float* xyz, int64_t n, float* f_out);
int psgpu_field_color_normal(psgpu_ctx* ctx, const float* xyz, int64_t n, float* f, float* col, float* nrm);
/* psgpu_field_eval without any allocation. NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving / haptics loops, harnesses that forbid it). Same inputs, outputs
 * and results as psgpu_field_eval (the same kernel); the device staging is a file-scope static device array of
 * 4 * PSGPU_NOALLOC_MAX_POINTS floats = 1 MiB that every GPU which loads the library holds until the process exits. It calls no
 * cudaMalloc / cudaFree or allocating library routine, on the first call or later (the program must already be uploaded).
 * n > PSGPU_NOALLOC_MAX_POINTS is rejected with PSGPU_NOT_ENOUGH_MEM and a message in psgpu_last_error: split the batch.
 * All calls on one device share the array and it keeps the last call's contents; nothing in it needs to start at zero (a call
 * overwrites every word it reads). Calls that OVERLAP on the GPU - two contexts on one device, two host threads - race on it and
 * must be serialised by the caller. */
#define PSGPU_NOALLOC_MAX_POINTS 65536
int psgpu_field_eval_noalloc(psgpu_ctx* ctx, const float* xyz, int64_t n, float* f_out);
/* Shrink-wrap Newton root search between n HOST point pairs (p1, p2: 3 floats each; f1, f2: their field values) with the full program, in
 * the uploaded program's semantics: PSGPU_SEM_SIMD = FieldComputer::computeRootNewtonRaphson (public in the reference, not called by its
 * Polygonize), PSGPU_SEM_SCALAR = ComputeRootNewtonRaphsonVEC4 (what CParsipOptimized places vertices with). Outputs per pair: root (3
 * floats), the field there, and the reference's return value (field evaluations spent, (rounds) * 4; -1 if iterations <= 0). */
int psgpu_newton_roots(psgpu_ctx* ctx, const float* p1, const float* p2, const float* f1, const float* f2, int64_t n, float target_field,
                       int iterations, float* root_out, float* field_out, int32_t* evals_out);
/* Field cache (fv[i*8 + j*64 + k]) and cube cases (cell i*49 + j*7 + k) of one MPU, for parity tests. */
int psgpu_mpu_cache(psgpu_ctx* ctx, const psgpu_params* params, const int32_t* ijk, float* fv512, uint8_t* case343);
/* Parity hook: runs cull + field for `params` and returns the lattice index and the 512-value field cache of EVERY evaluated
 * MPU, each computed with that MPU's own pruned program (compare with psgpu_mpu_cache = full program). n_out = MPUs evaluated;
 * nothing is copied if cap_mpus is too small. */
int psgpu_mpu_caches_culled(psgpu_ctx* ctx, const psgpu_params* params, int32_t* ijk_out, float* fv_out, int64_t cap_mpus, int64_t* n_out);
/* Parity hook for the culling pass: prunes the uploaded program against the box [lo, hi] on the device (one thread of the
 * same prune_walk the cull kernels run). len_out = pruned length in quads, 0 = region needs no evaluation. */
int psgpu_prune_region(psgpu_ctx* ctx, const float* lo3, const float* hi3, int cull, float iso, int color_exact, uint32_t* quads_out,
                       uint32_t cap_quads, uint32_t* len_out);
int psgpu_get_stats(const psgpu_ctx* ctx, psgpu_stats* out);
/* FP32 FMA-chain microbenchmark (dependent-free FFMA streams on every SM); returns measured TFLOP/s. */
int psgpu_measure_fp32_peak(psgpu_ctx* ctx, float* tflops_out, float* sm_clock_mhz_out);

#ifdef __cplusplus
}
#endif
#endif
