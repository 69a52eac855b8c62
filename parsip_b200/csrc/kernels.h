// Launch wrappers and POD parameter blocks shared by context.cu and the kernel translation units.
#ifndef PSGPU_KERNELS_H
#define PSGPU_KERNELS_H
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/psgpu.h"
#include "program.h"

namespace psgpu {

// A region of the MPU lattice (coordinates in units of the level's region size; for the last level = MPU
// lattice index) with its pruned program in the pool.
struct Region {
    int32_t x, y, z;
    uint32_t off, len;  // quads
};

struct CullLevel {
    int childSize;       // MPUs per child-region edge
    int lo[3], hi[3];    // slab bounds in MPUs (absolute lattice indices)
    float gridLo[3];
    float mpuSide;
    int cullMode;
    float iso;
    int colorExact;      // PS_SimdPoly semantics with colours wanted: blends keep OPC_CSCALE markers (program.h)
    int nonNeg;          // the program has no difference operators: no stack value can be negative (cull_walk.cuh: skippable)
};

struct GridParams {
    float gridLo[3];
    float cellsize;
    float mpuSide;       // cellsize * 7 as a float product (PS_Polygonizer.cpp:329)
    int lo[3], hi[3];    // slab bounds in MPUs
    int dims[3];         // full lattice
    int coordMode, sem, output, rootMode, rootIters;
    int insideGE;        // 1: inside = f >= iso (SIMD path), 0: f > iso (scalar twin)
    int wantColors, wantNormals;
    int forceFv;         // parity hook: write the field cache of every candidate
    int hasWarps;        // the uploaded program contains warp instructions (picks the kernels compiled with them)
    int hasMasks;        // the program contains OPC_BOXMASK (PSGPU_LIN_AS_SHIPPED): kernels compiled with the mask routine
    float iso;
    // z-slab stitching (psgpu_polygonize_sharded): the MPU layer whose closing z-plane this slab owns - its own last layer for a
    // stand-alone slab, the lattice's last layer when the slabs of several ranks form ONE mesh (then only the last rank owns a
    // closing plane and the others resolve those edges through the halo table of the rank above)
    int zOwnLast;
    uint32_t vertexOffset;  // added to every vertex id this rank writes into tri[] (global numbering across ranks)
    uint32_t haloVertexOffset;  // the same of the rank above: added to the (rank-local) vertex bases of its halo table
};

// per-candidate outputs of the field kernel
struct FieldOutputs {
    float4* fv;            // [cand][128] field cache, fv[i*8 + j*64 + k] layout
    uint32_t* vcount;      // [cand]
    uint32_t* tcount;      // [cand]
    uint32_t* edgeMask;    // [cand][48] crossing-edge bits, bit e = (i*8 + j*64 + k)*3 + axis
    uint16_t* edgePrefix;  // [cand][48] exclusive popcount prefix
    uint32_t* inside;      // [cand][16] inside bits of the 512 corners (bit i*8 + j*64 + k), written for candidates with triangles
    uint32_t* mixedList;   // candidates with triangles (arrival order)
    uint32_t* counters;    // [0] field work counter, [1] mixed count, [2] over-capacity MPUs, [3] tri work counter, [4] vertex work counter
};

struct MeshOutputs {
    float* pos; float* nrm; float* col; uint32_t* tri;
    const uint32_t* vbase;   // [cand] exclusive scan of vcount
    const uint32_t* tbase;   // [cand]
    const uint32_t* mixedIdx;  // [cand] rank among mixed candidates (PSGPU_OUT_MPU tables), may be null
    int32_t* mpuIjk; uint32_t* mpuVbase; uint32_t* mpuTbase;
    uint16_t* vperm;           // PSGPU_OUT_MPU: [vertex base + edge rank] -> reference-order local vertex id
    // halo of the rank above (first MPU layer of its slab), indexed x * ny + y; null when this slab owns its closing plane
    const uint32_t* haloVbase;   // vertex base of the MPU within ITS rank's arrays (+ GridParams.haloVertexOffset), 0xFFFFFFFF = no surface there
    const uint32_t* haloMask;    // [.][48]
    const uint16_t* haloPrefix;  // [.][48]
};

// cull.cu
void launch_all_mpus(const CullLevel& L, Region* out, uint32_t fullLen, int32_t* candOfMpu, int n, cudaStream_t st);
void launch_prune_count(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint8_t* qflag, uint32_t* childLen,
                        uint32_t* maxLen, cudaStream_t st);
void launch_flags(const uint32_t* len, uint32_t* flag, int n, cudaStream_t st);
void launch_prune_emit(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint8_t* qflag, const uint32_t* childLen,
                       const uint32_t* childOff, const uint32_t* childIdx, float4* outPool, uint8_t* outFlag, uint32_t outBase, Region* children,
                       int32_t* candOfMpu, uint32_t* maxLen, cudaStream_t st);
bool prune_level_is_fused(const CullLevel& L, int nParents);
void launch_prune_fused(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint8_t* qflag, float4* scratch,
                        uint8_t* scratchFlag, uint32_t stride, uint32_t* childLen, uint32_t* maxLen, bool ctaPerRegion, cudaStream_t st);
void launch_prune_compact(const CullLevel& L, const Region* parents, int nParents, const float4* scratch, const uint8_t* scratchFlag, uint32_t stride,
                          const uint32_t* childLen, const uint32_t* childOff, const uint32_t* childIdx, float4* outPool, uint8_t* outFlag,
                          uint32_t outBase, Region* children, int32_t* candOfMpu, cudaStream_t st);
void launch_prune_slab(const float4* prog, const uint8_t* progFlag, uint32_t progLen, const float* lo, const float* hi, float iso, int colorExact,
                       float4* out, uint8_t* outFlag, uint32_t cap, uint32_t* lenOut, cudaStream_t st);
void launch_prune_one(const float4* prog, const float* lo, const float* hi, int cullMode, float iso, int colorExact, float4* out, uint32_t cap,
                      uint32_t* lenOut, cudaStream_t st);
size_t scan_temp_bytes(int n);
void exclusive_scan_u32(void* temp, size_t tempBytes, const uint32_t* in, uint32_t* out, int n, cudaStream_t st);
// outA / outB[0..n] = exclusive sums of a[0..n) and of b[0..n) (b == nullptr: of the flags a[i] != 0) in ONE pass; entry n = the totals
void exclusive_scan_pair(void* temp, size_t tempBytes, const uint32_t* a, const uint32_t* b, uint32_t* outA, uint32_t* outB, int n, cudaStream_t st);

// field.cu
int field_kernel_smem_bytes(uint32_t maxProgQuads, uint32_t* progCapQuads);
void launch_field_kernel(const GridParams& G, const Region* cands, int nCand, const float4* pool, const FieldOutputs& out, uint32_t maxProgQuads,
                         bool hasWarps, int numSMs, cudaStream_t st);
void launch_points_field(const float4* prog, const float* xyz, int64_t n, float* f, int sem, cudaStream_t st);
void launch_points_newton(const float4* prog, const float* p1, const float* p2, const float* f1, const float* f2, int64_t n, float target, int iterations,
                          float* root, float* rootField, int32_t* evals, int sem, cudaStream_t st);
void launch_points_field_color_normal(const float4* prog, const float* colorTab, const float* xyz, int64_t n, float* f, float* col, float* nrm,
                                      int sem, cudaStream_t st);
void launch_program_stats(const Region* cands, int nCand, const float4* pool, const uint32_t* tcount, unsigned long long* acc, cudaStream_t st);
float run_fp32_peak(int numSMs, cudaStream_t st, float* ms_out);

// emit.cu
void launch_cube_cases(const GridParams& G, const float* fv, uint8_t* case343, cudaStream_t st);
void launch_halo_build(const GridParams& G, const int32_t* candOfMpu, const FieldOutputs& fo, const uint32_t* vbase, uint32_t* hVbase,
                       uint32_t* hMask, uint16_t* hPrefix, int nxy, cudaStream_t st);
bool normals_in_color_pass(const GridParams& G);
// phase: a stitched multi-rank run launches the vertex and colour kernels first (they need nothing from the other ranks) and
// tri_kernel once the exchange with the neighbours is through (vertex offsets, halo table of the rank above)
enum { PS_EMIT_ALL = 0, PS_EMIT_VERTEX_COLOR = 1, PS_EMIT_TRI = 2 };
void launch_emit_kernels(const GridParams& G, const Region* cands, const float4* pool, const float* colorTab, const FieldOutputs& fo,
                         const MeshOutputs& mo, const int32_t* candOfMpu, int nMixed, uint32_t maxProgQuads, int numSMs, cudaStream_t st,
                         cudaEvent_t evTriDone = nullptr, cudaEvent_t evVertexDone = nullptr, int phase = PS_EMIT_ALL);

}  // namespace psgpu
#endif
