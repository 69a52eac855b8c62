// Marching-cubes mesh emission (north-star subsystem 4), two kernels over the surface-crossing MPUs:
//
//   tri_kernel     one CTA per MPU: re-classifies the 343 cells from the field cache and writes the triangles with GLOBAL
//                  vertex indices. A vertex's index is (MPU's vertex base) + (rank of its edge bit in the owner MPU's
//                  crossing-edge mask), so neighbouring MPUs resolve shared edges with one popcount and no hashing - the
//                  deterministic replacement for the reference's per-MPU EDGETABLE (PS_Polygonizer.h:211-222, cpp:832-916).
//                  PSGPU_OUT_MPU reproduces the reference's per-MPU vertex/triangle ORDER (first use in the sequential
//                  i,j,k cell loop) with prefix sums and leaves the rank -> reference-order permutation for vertex_kernel.
//   vertex_kernel  one WARP per MPU, 32 crossing edges at a time (an MPU has ~50 of them, so CTA-wide loops would leave
//                  most lanes idle): rank -> edge bit -> root (4-sample + lerp / reference Newton / bracketed section search,
//                  all re-evaluating the field with the MPU's pruned program), forward-difference normal, colour walk.
//
// Reference: process_cells_simd phases 2-4 (PS_SimdPoly/include/PS_Polygonizer.cpp:612-829), doMarchingCubes +
// ComputeRootNewtonRaphsonVEC4 (ParsipHaptics/include/CPolyParsipOptimized.cpp:187-327, CompactBlobTree.cpp:1581-1622).
#include "device_util.cuh"
#include "interp_packed.cuh"
#include "kernels.h"
#include "mc_tables.h"

namespace psgpu {

namespace {

constexpr int kThreads = 128;

__device__ const int8_t g_triTable[256 * 16] = PS_MC_TRITABLE_INIT;
__constant__ uint8_t c_nTriE[256] = PS_MC_NTRI_INIT;

struct EdgePos { float e1[3]; float e2a; int axis; };

__device__ __forceinline__ EdgePos edge_endpoints(const GridParams& G, const Region& r, int ci, int cj, int ck, int axis) {
    EdgePos E;
    E.axis = axis;
    E.e1[0] = corner_coord(G, 0, r.x, ci); E.e1[1] = corner_coord(G, 1, r.y, cj); E.e1[2] = corner_coord(G, 2, r.z, ck);
    if (G.coordMode == PSGPU_COORD_MPU_LOCAL) E.e2a = __fadd_rn(E.e1[axis], G.cellsize);  // PS_Polygonizer.cpp:722-724
    else E.e2a = corner_coord(G, axis, axis == 0 ? r.x : (axis == 1 ? r.y : r.z), (axis == 0 ? ci : (axis == 1 ? cj : ck)) + 1);
    return E;
}

// In the STRICT (PS_SimdPoly) instantiations of this file the WARPS parameter - meaningless there, the SIMD path ignores warps - selects the
// variants that know OPC_BOXMASK (programs linearized with PSGPU_LIN_AS_SHIPPED; GridParams.hasMasks).
// ONE out-of-line copy of the plain 4-wide interpreter per kernel: the root search and the normal both call it (inlining it
// at every call site multiplied the hot code past the instruction caches)
template <bool STRICT, bool SMEM, bool WARPS>
__device__ __noinline__ void eval4(const float4* prog, const float (&X)[4], const float (&Y)[4], const float (&Z)[4], float (&F)[4], int sem,
                                   bool packet4 = false) {  // packet4: the 4 points are ONE SIMD call of the reference (OPC_BOXMASK, program.h)
    if constexpr (STRICT) {  // PS_SimdPoly semantics: packed-pair arithmetic (interp_packed.cuh), same bits
        PairPts<2, false> P;
        P.x[0] = pk(X[0], X[1]); P.x[1] = pk(X[2], X[3]);
        P.y[0] = pk(Y[0], Y[1]); P.y[1] = pk(Y[2], Y[3]);
        P.z[0] = pk(Z[0], Z[1]); P.z[1] = pk(Z[2], Z[3]);
        F2 FP[2];
        eval_program_packed<2, false, SMEM, false, WARPS>(prog, P, FP, nullptr, nullptr, packet4);  // STRICT: WARPS stands for MASKS
        up(FP[0], F[0], F[1]); up(FP[1], F[2], F[3]);
    } else
        eval_program<4, false, STRICT, SMEM, WARPS>(prog, X, Y, Z, F, nullptr, nullptr, sem);
}

// PS_Polygonizer.cpp:728-762: samples at t = 0, 1/3, 2/3, 1 in one 4-wide evaluation, first lane whose
// inside flag differs from lane 0 brackets the root, linear interpolation inside that sub-interval.
template <bool STRICT, bool SMEM, bool WARPS>
__device__ void root_foursample(const GridParams& G, const float4* prog, const EdgePos& E, float (&p)[3]) {
    float X[4], Y[4], Z[4], F[4];
    const float d = __fsub_rn(E.e2a, E.e1[E.axis]);
#pragma unroll
    for (int l = 0; l < 4; l++) {
        const float res = __fmul_rn((float)l, 1.0f / 3.0f);
        const float moved = __fadd_rn(E.e1[E.axis], __fmul_rn(d, res));
        X[l] = (E.axis == 0) ? moved : E.e1[0];
        Y[l] = (E.axis == 1) ? moved : E.e1[1];
        Z[l] = (E.axis == 2) ? moved : E.e1[2];
    }
    eval4<STRICT, SMEM, WARPS>(prog, X, Y, Z, F, G.sem, true);
    const bool state = is_inside(G, F[0]);
    // first lane whose flag differs from lane 0, else the last one (cpp:748-755)
    const int interval = (is_inside(G, F[1]) != state) ? 1 : ((is_inside(G, F[2]) != state) ? 2 : 3);
    float fa = F[0], fb = F[1], xa = X[0], xb = X[1], ya = Y[0], yb = Y[1], za = Z[0], zb = Z[1];
    if (interval == 2) { fa = F[1]; fb = F[2]; xa = X[1]; xb = X[2]; ya = Y[1]; yb = Y[2]; za = Z[1]; zb = Z[2]; }
    if (interval == 3) { fa = F[2]; fb = F[3]; xa = X[2]; xb = X[3]; ya = Y[2]; yb = Y[3]; za = Z[2]; zb = Z[3]; }
    const float scale = __fdiv_rn(__fsub_rn(G.iso, fa), __fsub_rn(fb, fa));
    p[0] = __fadd_rn(xa, __fmul_rn(scale, __fsub_rn(xb, xa)));
    p[1] = __fadd_rn(ya, __fmul_rn(scale, __fsub_rn(yb, ya)));
    p[2] = __fadd_rn(za, __fmul_rn(scale, __fsub_rn(zb, za)));
}

// Bracketed section search along the edge (native default, PSGPU_ROOT_BISECT): every round evaluates the field at the four
// interior fifth-points of the current bracket in ONE 4-wide interpreter pass and keeps the fifth that holds the sign
// change (x5 per round, a fixed trip count - a warp never waits for one slow lane the way per-lane false position does);
// the root is the secant point of the final bracket. The bracket is guaranteed: both edge ends are lattice corners whose
// cached values came from this same program. 4 rounds: bracket 1.6e-3 cell, secant error far below 1e-4 cell.
template <bool STRICT, bool SMEM, bool WARPS>
__device__ void root_section(const GridParams& G, const float4* prog, const EdgePos& E, float f1, float f2, int rounds, float (&p)[3]) {
    float ta = 0.f, tb = 1.f, fa = f1 - G.iso, fb = f2 - G.iso;
    const float d = E.e2a - E.e1[E.axis];
    const bool aIn = is_inside(G, f1);
    for (int it = 0; it < rounds; it++) {
        const float w = (tb - ta) * 0.2f;
        float X[4], Y[4], Z[4], F[4], T[4];
#pragma unroll
        for (int l = 0; l < 4; l++) {
            T[l] = fmaf((float)(l + 1), w, ta);
            const float moved = fmaf(T[l], d, E.e1[E.axis]);
            X[l] = (E.axis == 0) ? moved : E.e1[0];
            Y[l] = (E.axis == 1) ? moved : E.e1[1];
            Z[l] = (E.axis == 2) ? moved : E.e1[2];
        }
        eval4<STRICT, SMEM, WARPS>(prog, X, Y, Z, F, G.sem);
        // first sample on the far side of the surface closes the bracket; the one before it opens it
        float nta = ta, nfa = fa, ntb = tb, nfb = fb;
        bool found = false;
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const bool in = is_inside(G, F[l]);
            if (!found) {
                if (in == aIn) { nta = T[l]; nfa = F[l] - G.iso; }
                else { ntb = T[l]; nfb = F[l] - G.iso; found = true; }
            }
        }
        ta = nta; fa = nfa; tb = ntb; fb = nfb;
    }
    const float den = fa - fb;
    float t = (den != 0.f) ? ta + (tb - ta) * (fa / den) : 0.5f * (ta + tb);
    t = fminf(fmaxf(t, ta), tb);
    p[0] = E.e1[0]; p[1] = E.e1[1]; p[2] = E.e1[2];
    p[E.axis] = fmaf(t, d, E.e1[E.axis]);
}

// ComputeRootNewtonRaphsonVEC4 (CompactBlobTree.cpp:1581-1622) incl. the fp^2 damping of its vec4 dot
template <bool STRICT, bool SMEM, bool WARPS>
__device__ void root_newton_ref(const GridParams& G, const float4* prog, const EdgePos& E, float f1, float f2, int iters, float (&p)[3]) {
    float x[3] = {E.e1[0], E.e1[1], E.e1[2]};
    if (!(fabsf(f1 - G.iso) < fabsf(f2 - G.iso))) x[E.axis] = E.e2a;
    const float eps = PS_FIELDVALUE_EPSILON, inv = 1.0f / PS_FIELDVALUE_EPSILON;
    for (int it = 0; it < iters; it++) {
        float X[4] = {x[0], x[0] + eps, x[0], x[0]}, Y[4] = {x[1], x[1], x[1] + eps, x[1]}, Z[4] = {x[2], x[2], x[2], x[2] + eps}, F[4];
        eval4<STRICT, SMEM, WARPS>(prog, X, Y, Z, F, G.sem);
        const float fp = F[0];
        const float gx = (F[1] - fp) * inv, gy = (F[2] - fp) * inv, gz = (F[3] - fp) * inv;
        const float dd = G.iso - fp;
        const float g = 1.0f / (gx * gx + gy * gy + gz * gz + fp * fp);
        x[0] = x[0] + (dd * gx) * g; x[1] = x[1] + (dd * gy) * g; x[2] = x[2] + (dd * gz) * g;
        float X1[1] = {x[0]}, Y1[1] = {x[1]}, Z1[1] = {x[2]}, F1[1];
        eval_program<1, false, STRICT, SMEM, WARPS>(prog, X1, Y1, Z1, F1, nullptr, nullptr, G.sem);
        if (fabsf(F1[0] - G.iso) < eps) break;
    }
    p[0] = x[0]; p[1] = x[1]; p[2] = x[2];
}

// ---------------------------------------------------------------------------------------------------------------------
// tri_kernel: persistent warps, one surface MPU per warp at a time, no block barriers. A lane owns 11 consecutive cells of the
// 343 (cell order = the reference's i, j, k loop nest), cube cases come from the 512 inside bits read as bytes
// (byte (i, j) = the 8 corners along k), triangle slots from one warp scan.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kCellsPerLane = 11;  // 32 * 11 >= 343

struct TriWarpSmem {
    uint32_t cellList[344];  // surface cells in cell order: cell | case << 9 | exclusive triangle offset << 17
    uint32_t insideW[16];    // bit i*8 + j*64 + k
    uint32_t edgeW[48];
    uint16_t edgeP[48];
    uint16_t edgeVid[1536];  // PSGPU_OUT_MPU: reference-order local vertex id per edge bit
    int32_t nbCand[8];       // candidate index of the neighbour MPU at (+dI, +dJ, +dK), slot dI | dJ << 1 | dK << 2; -1 = culled / outside
    uint32_t nbVbase[8];     // ... and its vertex base: read once per MPU, not once per shared edge behind a dependent load
};

__device__ __forceinline__ uint32_t warp_excl_scan(uint32_t v, uint32_t& total) {
    const int lane = threadIdx.x & 31;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += a;
    }
    total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

// corner1[e] / edgeaxis[e] (_CellConfigTable.h:48-51, mc_tables.h) packed into immediates: lanes look up different edges, which a
// __constant__ table would serialise
__device__ __forceinline__ int edge_corner1(int e) { return (int)((0x321054641020ull >> (4 * e)) & 7ull); }
__device__ __forceinline__ int edge_axis(int e) { return (int)((0x5a5au >> (2 * e)) & 3u); }

// cube case of cell (x, y, z): corner b = 4*dx + 2*dy + dz (PS_Polygonizer.cpp:662-669)
__device__ __forceinline__ uint32_t cell_case(const uint8_t* B, int x, int y, int z) {
    const uint32_t b00 = B[y * 8 + x], b10 = B[y * 8 + x + 1], b01 = B[(y + 1) * 8 + x], b11 = B[(y + 1) * 8 + x + 1];
    return ((b00 >> z) & 3u) | (((b01 >> z) & 3u) << 2) | (((b10 >> z) & 3u) << 4) | (((b11 >> z) & 3u) << 6);
}

__global__ void __launch_bounds__(kThreads) tri_kernel(GridParams G, const Region* __restrict__ cands, FieldOutputs fo, MeshOutputs mo,
                                                       const int32_t* __restrict__ candOfMpu, int nMixed) {
    __shared__ int8_t sTri[256 * 16];
    __shared__ uint8_t sNTri[256];
    __shared__ TriWarpSmem sWarp[kThreads / 32];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 256 * 16; i += kThreads) sTri[i] = g_triTable[i];
    for (int i = threadIdx.x; i < 256; i += kThreads) sNTri[i] = c_nTriE[i];
    __syncthreads();
    TriWarpSmem& S = sWarp[wib];
    const uint8_t* B = reinterpret_cast<const uint8_t*>(S.insideW);
    const bool perMpu = (G.output == PSGPU_OUT_MPU);
    const int ny = G.hi[1] - G.lo[1], nz = G.hi[2] - G.lo[2];
    const int warpsTotal = gridDim.x * (kThreads / 32);
    const int c0 = lane * kCellsPerLane;
    const int x0 = c0 / 49, y0 = (c0 - x0 * 49) / 7, z0 = c0 - x0 * 49 - y0 * 7;

    int m = blockIdx.x * (kThreads / 32) + wib;
    while (m < nMixed) {
        const uint32_t cand = fo.mixedList[m];
        const Region r = cands[cand];
        const uint32_t vbase = mo.vbase[cand], tbase = mo.tbase[cand];
        __syncwarp();
        S.edgeW[lane] = fo.edgeMask[(size_t)cand * 48 + lane];
        S.edgeP[lane] = fo.edgePrefix[(size_t)cand * 48 + lane];
        if (lane < 16) {
            S.edgeW[32 + lane] = fo.edgeMask[(size_t)cand * 48 + 32 + lane];
            S.edgeP[32 + lane] = fo.edgePrefix[(size_t)cand * 48 + 32 + lane];
        }
        if (lane < 16) S.insideW[lane] = fo.inside[(size_t)cand * 16 + lane];  // 512 inside bits, as field_kernel classified them
        if (lane >= 16 && lane < 24 && !perMpu) {
            const int q = lane - 16, nx_ = r.x + (q & 1), ny_ = r.y + ((q >> 1) & 1), nz_ = r.z + (q >> 2);
            int32_t nc = -1;
            if (q != 0 && nx_ < G.hi[0] && ny_ < G.hi[1] && nz_ < G.hi[2])
                nc = candOfMpu[((size_t)(nx_ - G.lo[0]) * ny + (ny_ - G.lo[1])) * nz + (nz_ - G.lo[2])];
            S.nbCand[q] = nc;
            S.nbVbase[q] = (nc >= 0) ? mo.vbase[nc] : 0u;
        }
        __syncwarp();

        // ---- cube cases of this lane's cells, triangle count
        uint32_t cf[3] = {0u, 0u, 0u};
        uint32_t nT = 0;
        {
            int x = x0, y = y0, z = z0;
#pragma unroll
            for (int i = 0; i < kCellsPerLane; i++) {
                if (c0 + i < 343) {
                    const uint32_t cfg = cell_case(B, x, y, z);
                    cf[i >> 2] |= cfg << ((i & 3) * 8);
                    nT += sNTri[cfg];
                }
                if (++z == 7) { z = 0; if (++y == 7) { y = 0; x++; } }
            }
        }

        if (perMpu) {
            // reference vertex order: a vertex is created by the first cell (i outer, k inner) whose table row names its edge, in
            // row order (PS_Polygonizer.cpp:647-808). The first cell of edge (corner c, axis a) is c with the two non-axis
            // coordinates decremented (clipped at 0).
            uint32_t newMask[kCellsPerLane];
            uint32_t nNew = 0;
            {
                int x = x0, y = y0, z = z0;
#pragma unroll
                for (int i = 0; i < kCellsPerLane; i++) {
                    uint32_t nm = 0;
                    if (c0 + i < 343) {
                        const int8_t* row = &sTri[((cf[i >> 2] >> ((i & 3) * 8)) & 255u) * 16];
                        uint32_t seen = 0;
                        for (int k = 0; k < 16; k++) {
                            const int e = row[k];
                            if (e < 0) break;
                            if (seen & (1u << e)) continue;
                            seen |= 1u << e;
                            const int c1 = edge_corner1(e), ax = edge_axis(e);
                            const int sx = x + ((c1 >> 2) & 1), sy = y + ((c1 >> 1) & 1), sz = z + (c1 & 1);
                            const bool fx = (ax == 0) || (sx == x + 1) || (sx == 0), fy = (ax == 1) || (sy == y + 1) || (sy == 0),
                                       fz = (ax == 2) || (sz == z + 1) || (sz == 0);
                            if (fx && fy && fz) { nm |= 1u << k; nNew++; }
                        }
                    }
                    newMask[i] = nm;
                    if (++z == 7) { z = 0; if (++y == 7) { y = 0; x++; } }
                }
            }
            uint32_t total;
            uint32_t id = warp_excl_scan(nNew, total);
            {
                int x = x0, y = y0, z = z0;
#pragma unroll
                for (int i = 0; i < kCellsPerLane; i++) {
                    uint32_t nm = newMask[i];
                    if (nm) {
                        const int8_t* row = &sTri[((cf[i >> 2] >> ((i & 3) * 8)) & 255u) * 16];
                        while (nm) {
                            const int k = __ffs((int)nm) - 1;
                            nm &= nm - 1u;
                            const int e = row[k];
                            const int c1 = edge_corner1(e), ax = edge_axis(e);
                            const int sx = x + ((c1 >> 2) & 1), sy = y + ((c1 >> 1) & 1), sz = z + (c1 & 1);
                            const int eb = (sx * 8 + sy * 64 + sz) * 3 + ax;
                            S.edgeVid[eb] = (uint16_t)id;
                            // rank of this edge in the MPU's crossing-edge mask -> where vertex_kernel must put its vertex
                            const uint32_t rank = S.edgeP[eb >> 5] + __popc(S.edgeW[eb >> 5] & ((1u << (eb & 31)) - 1u));
                            mo.vperm[vbase + rank] = (uint16_t)id;
                            id++;
                        }
                    }
                    if (++z == 7) { z = 0; if (++y == 7) { y = 0; x++; } }
                }
            }
            __syncwarp();
        }

        // ---- surface cells compacted in cell order, then one TRIANGLE per lane (3 indices, 12 contiguous bytes per lane)
        uint32_t totalT, totalA;
        {
            uint32_t nAct = 0;
#pragma unroll
            for (int i = 0; i < kCellsPerLane; i++) nAct += (sNTri[(cf[i >> 2] >> ((i & 3) * 8)) & 255u] != 0) ? 1u : 0u;
            uint32_t ex = warp_excl_scan(nT, totalT);
            uint32_t slot = warp_excl_scan(nAct, totalA);
#pragma unroll
            for (int i = 0; i < kCellsPerLane; i++) {
                const uint32_t cfg = (cf[i >> 2] >> ((i & 3) * 8)) & 255u;
                const uint32_t nTri = sNTri[cfg];  // 0 for the padding cells (case 0)
                if (nTri) { S.cellList[slot++] = (uint32_t)(c0 + i) | (cfg << 9) | (ex << 17); ex += nTri; }
            }
        }
        __syncwarp();
        for (uint32_t t = lane; t < totalT; t += 32) {
            int lo = 0, hi = (int)totalA - 1;  // last surface cell whose triangle offset <= t
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if ((S.cellList[mid] >> 17) <= t) lo = mid; else hi = mid - 1;
            }
            const uint32_t ent = S.cellList[lo];
            const int cell = (int)(ent & 511u);
            const int x = cell / 49, y = (cell - x * 49) / 7, z = cell - x * 49 - y * 7;
            const int8_t* row = &sTri[((ent >> 9) & 255u) * 16 + (t - (ent >> 17)) * 3];
            uint32_t vid3[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const int e = row[k];
                const int c1 = edge_corner1(e), ax = edge_axis(e);
                int sx = x + ((c1 >> 2) & 1), sy = y + ((c1 >> 1) & 1), sz = z + (c1 & 1);
                uint32_t vid;
                if (perMpu) vid = S.edgeVid[(sx * 8 + sy * 64 + sz) * 3 + ax];
                else {
                    const int dI = (sx == 7 && ax != 0 && r.x != G.hi[0] - 1) ? 1 : 0;
                    const int dJ = (sy == 7 && ax != 1 && r.y != G.hi[1] - 1) ? 1 : 0;
                    const int dK = (sz == 7 && ax != 2 && r.z != G.zOwnLast) ? 1 : 0;
                    if ((dI | dJ | dK) == 0) {
                        const int eb = (sx * 8 + sy * 64 + sz) * 3 + ax;
                        vid = G.vertexOffset + vbase + S.edgeP[eb >> 5] + __popc(S.edgeW[eb >> 5] & ((1u << (eb & 31)) - 1u));
                    } else {
                        sx -= 7 * dI; sy -= 7 * dJ; sz -= 7 * dK;
                        const int eb = (sx * 8 + sy * 64 + sz) * 3 + ax;
                        const uint32_t bit = 1u << (eb & 31);
                        vid = 0u;
                        bool found = false;
                        if (r.z + dK >= G.hi[2]) {
                            // the owner lives in the slab of the rank above: its first-layer table came over NVLink (context.cu)
                            const size_t h = (size_t)(r.x + dI - G.lo[0]) * ny + (r.y + dJ - G.lo[1]);
                            const uint32_t hb = mo.haloVbase ? mo.haloVbase[h] : 0xFFFFFFFFu;
                            if (hb != 0xFFFFFFFFu) {
                                const uint32_t w = mo.haloMask[h * 48 + (eb >> 5)];
                                found = (w & bit) != 0u;
                                vid = G.haloVertexOffset + hb + mo.haloPrefix[h * 48 + (eb >> 5)] + __popc(w & (bit - 1u));
                            }
                        } else {
                            const int q = dI | (dJ << 1) | (dK << 2);
                            const int32_t nc = S.nbCand[q];
                            if (nc >= 0) {  // a culled neighbour has no masks at all (index -1: the out-of-bounds read ADVICE r1 flagged)
                                const uint32_t w = fo.edgeMask[(size_t)nc * 48 + (eb >> 5)];
                                found = (w & bit) != 0u;
                                vid = G.vertexOffset + S.nbVbase[q] + fo.edgePrefix[(size_t)nc * 48 + (eb >> 5)] + __popc(w & (bit - 1u));
                            }
                        }
                        // The owner must have seen the same crossing: shared corners are evaluated from the same lattice
                        // coordinates by programs that differ only in operands that are exactly zero there. If it ever did
                        // not, count it (psgpu_polygonize fails with the count) and keep the index in range.
                        if (!found) { atomicAdd(&fo.counters[6], 1u); vid = G.vertexOffset; }
                    }
                }
                vid3[k] = vid;
            }
            uint32_t* o = mo.tri + (size_t)(tbase + t) * 3;
            o[0] = vid3[0]; o[1] = vid3[1]; o[2] = vid3[2];
        }
        if (perMpu && lane == 0 && mo.mixedIdx) {
            const uint32_t mi = mo.mixedIdx[cand];
            mo.mpuIjk[mi * 3] = r.x; mo.mpuIjk[mi * 3 + 1] = r.y; mo.mpuIjk[mi * 3 + 2] = r.z;
            mo.mpuVbase[mi] = vbase; mo.mpuTbase[mi] = tbase;
        }
        int nxt = 0;
        if (lane == 0) nxt = (int)atomicAdd(&fo.counters[3], 1u) + warpsTotal;
        m = __shfl_sync(0xffffffffu, nxt, 0);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// vertex_kernel / color_kernel: persistent warps, one surface MPU per warp at a time
// ---------------------------------------------------------------------------------------------------------------------
// SMEM variants: every warp copies its MPU's pruned list into its own slice of dynamic shared memory (slice = the longest list of
// the run, chosen by the host) and interprets it with LDS; the other variants (very long lists) read from global memory.
__device__ __forceinline__ void stage_warp_program(float4* slice, const float4* __restrict__ src, uint32_t len, int lane) {
    __syncwarp();
    for (uint32_t i = lane; i < len; i += 32) slice[i] = src[i];
    __syncwarp();
}

// position (root on the edge) and forward-difference normal of one vertex; the plain 4-wide interpreter, the same code the
// field kernel runs (the colour walk lives in color_kernel so that neither hot loop outgrows the instruction caches: with
// both inlined here ncu r01b showed stall_no_instruction at 12.7 cycles per issue)
template <bool STRICT, int ROOT, bool SMEM, bool WARPS>
__device__ __forceinline__ void vertex_at(const GridParams& G, const float4* prog, const EdgePos& E, float f1, float f2, float (&p)[3], float (&nrm)[3]) {
    if (ROOT == PSGPU_ROOT_FOURSAMPLE) root_foursample<STRICT, SMEM, WARPS>(G, prog, E, p);
    else if (ROOT == PSGPU_ROOT_NEWTON_REF) root_newton_ref<STRICT, SMEM, WARPS>(G, prog, E, f1, f2, G.rootIters, p);
    else root_section<STRICT, SMEM, WARPS>(G, prog, E, f1, f2, G.rootIters, p);
    if (G.wantNormals) {
        const float dl = PS_NORMAL_DELTA;
        float X[4] = {p[0], p[0] + dl, p[0], p[0]}, Y[4] = {p[1], p[1], p[1] + dl, p[1]}, Z[4] = {p[2], p[2], p[2], p[2] + dl};
        float F[4];
        eval4<STRICT, SMEM, WARPS>(prog, X, Y, Z, F, G.sem);
        const float dinv = -1.0f / dl;
        float nx = (F[1] - F[0]) * dinv, nyv = (F[2] - F[0]) * dinv, nzv = (F[3] - F[0]) * dinv;
        const float l2 = nx * nx + nyv * nyv + nzv * nzv;
        if (l2 > 0.f) { const float il = rsqrtf(l2); nx *= il; nyv *= il; nzv *= il; }
        else if (G.sem == PSGPU_SEM_SCALAR) { nx = nyv = nzv = 1.0f; }  // normalizeXYZ, PS_Vector.h:1072-1086
        nrm[0] = nx; nrm[1] = nyv; nrm[2] = nzv;
    }
}

template <bool STRICT, int ROOT, bool SMEM, bool WARPS>
__global__ void __launch_bounds__(kThreads, 6) vertex_kernel(GridParams G, const Region* __restrict__ cands, const float4* __restrict__ pool,
                                                             FieldOutputs fo, MeshOutputs mo, int nMixed, uint32_t sliceQuads) {
    __shared__ uint32_t sW[kThreads / 32][48];
    __shared__ uint32_t sP[kThreads / 32][48];
    extern __shared__ __align__(16) float4 sProgDyn[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const bool perMpu = (G.output == PSGPU_OUT_MPU);
    const int warpsTotal = gridDim.x * (kThreads / 32);
    int m = blockIdx.x * (kThreads / 32) + wib;  // first item is static, the rest come from the work counter
    while (m < nMixed) {
        const uint32_t cand = fo.mixedList[m];
        const Region r = cands[cand];
        const uint32_t vbase = mo.vbase[cand];
        const uint32_t nV = fo.vcount[cand];
        __syncwarp();
        sW[wib][lane] = fo.edgeMask[(size_t)cand * 48 + lane];
        sP[wib][lane] = fo.edgePrefix[(size_t)cand * 48 + lane];
        if (lane < 16) {
            sW[wib][32 + lane] = fo.edgeMask[(size_t)cand * 48 + 32 + lane];
            sP[wib][32 + lane] = fo.edgePrefix[(size_t)cand * 48 + 32 + lane];
        }
        float4* slice = sProgDyn + (size_t)wib * sliceQuads;
        if (SMEM) stage_warp_program(slice, pool + r.off, r.len, lane);
        const float4* prog = SMEM ? slice : pool + r.off;
        const float* fv = reinterpret_cast<const float*>(fo.fv + (size_t)cand * 128);
        for (uint32_t rk = lane; rk < nV; rk += 32) {
            int lo = 0, hi = 47;  // last word whose exclusive prefix <= rk
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (sP[wib][mid] <= rk) lo = mid; else hi = mid - 1;
            }
            const int bit = __fns(sW[wib][lo], 0, (int)(rk - sP[wib][lo]) + 1);
            const int e = lo * 32 + bit;
            const int corner = e / 3, axis = e - corner * 3;
            const int ck = corner & 7, ci = (corner >> 3) & 7, cj = corner >> 6;
            const EdgePos E = edge_endpoints(G, r, ci, cj, ck, axis);
            float f1 = 0.f, f2 = 0.f;
            if (ROOT != PSGPU_ROOT_FOURSAMPLE) { f1 = fv[corner]; f2 = fv[corner + (axis == 0 ? 8 : (axis == 1 ? 64 : 1))]; }
            float p[3], nr[3] = {0.f, 0.f, 0.f};
            vertex_at<STRICT, ROOT, SMEM, WARPS>(G, prog, E, f1, f2, p, nr);
            const size_t vo = (size_t)(vbase + (perMpu ? (uint32_t)mo.vperm[vbase + rk] : rk)) * 3;
            mo.pos[vo] = p[0]; mo.pos[vo + 1] = p[1]; mo.pos[vo + 2] = p[2];
            if (G.wantNormals) { mo.nrm[vo] = nr[0]; mo.nrm[vo + 1] = nr[1]; mo.nrm[vo + 2] = nr[2]; }
        }
        int nxt = 0;
        if (lane == 0) nxt = (int)atomicAdd(&fo.counters[4], 1u) + warpsTotal;
        m = __shfl_sync(0xffffffffu, nxt, 0);
    }
}

// colour walk at the final vertex positions (fieldValueAndColor, PS_Polygonizer.cpp:1378-1551 / baseColorOp,
// CompactBlobTree.cpp:1124-1294): one vertex per lane, the 1-wide coloured interpreter
// NRM: the same walk 4 wide over (p, p + dx, p + dy, p + dz) - colour at p plus the forward-difference normal vertex_kernel would
// otherwise spend its second interpreter pass on (vertex_at): the primitive decode, the colour blends and the scale steps dominate
// this kernel, three more field lanes ride along for a fraction of a separate pass.
template <bool STRICT, bool SMEM, bool WARPS, bool NRM>
__global__ void __launch_bounds__(kThreads) color_kernel(GridParams G, const Region* __restrict__ cands, const float4* __restrict__ pool,
                                                         const float* __restrict__ colorTab, FieldOutputs fo, MeshOutputs mo, int nMixed,
                                                         uint32_t sliceQuads) {
    extern __shared__ __align__(16) float4 sProgDyn[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int warpsTotal = gridDim.x * (kThreads / 32);
    int m = blockIdx.x * (kThreads / 32) + wib;
    while (m < nMixed) {
        const uint32_t cand = fo.mixedList[m];
        const Region r = cands[cand];
        const uint32_t vbase = mo.vbase[cand];
        const uint32_t nV = fo.vcount[cand];
        float4* slice = sProgDyn + (size_t)wib * sliceQuads;
        if (SMEM) stage_warp_program(slice, pool + r.off, r.len, lane);
        const float4* prog = SMEM ? slice : pool + r.off;
        for (uint32_t rk = lane; rk < nV; rk += 32) {
            const size_t vo = (size_t)(vbase + rk) * 3;
            float rgb[3];
            if (NRM) {
                const float px = mo.pos[vo], py = mo.pos[vo + 1], pz = mo.pos[vo + 2], dl = PS_NORMAL_DELTA;
                float F[4];
                if constexpr (STRICT) {  // the packed-pair walk with the colour rules on point 0 (interp_packed.cuh), same bits
                    PairPts<2, false> P;
                    P.x[0] = pk(px, px + dl); P.x[1] = pk(px, px);
                    P.y[0] = pk(py, py); P.y[1] = pk(py + dl, py);
                    P.z[0] = pk(pz, pz); P.z[1] = pk(pz, pz + dl);
                    F2 FP[2];
                    eval_program_packed<2, false, SMEM, true, WARPS>(prog, P, FP, rgb, colorTab);  // STRICT: WARPS stands for MASKS
                    up(FP[0], F[0], F[1]); up(FP[1], F[2], F[3]);
                } else {
                    float X[4] = {px, px + dl, px, px}, Y[4] = {py, py, py + dl, py}, Z[4] = {pz, pz, pz, pz + dl};
                    eval_program<4, true, STRICT, SMEM, WARPS>(prog, X, Y, Z, F, rgb, colorTab, G.sem);
                }
                const float dinv = -1.0f / dl;  // as vertex_at
                float nx = (F[1] - F[0]) * dinv, nyv = (F[2] - F[0]) * dinv, nzv = (F[3] - F[0]) * dinv;
                const float l2 = nx * nx + nyv * nyv + nzv * nzv;
                if (l2 > 0.f) { const float il = rsqrtf(l2); nx *= il; nyv *= il; nzv *= il; }
                else if (G.sem == PSGPU_SEM_SCALAR) { nx = nyv = nzv = 1.0f; }
                mo.nrm[vo] = nx; mo.nrm[vo + 1] = nyv; mo.nrm[vo + 2] = nzv;
            } else {
                float X[1] = {mo.pos[vo]}, Y[1] = {mo.pos[vo + 1]}, Z[1] = {mo.pos[vo + 2]}, F[1];
                eval_program<1, true, STRICT, SMEM, WARPS>(prog, X, Y, Z, F, rgb, colorTab, G.sem);
            }
            mo.col[vo] = rgb[0]; mo.col[vo + 1] = rgb[1]; mo.col[vo + 2] = rgb[2];
        }
        int nxt = 0;
        if (lane == 0) nxt = (int)atomicAdd(&fo.counters[5], 1u) + warpsTotal;
        m = __shfl_sync(0xffffffffu, nxt, 0);
    }
}

}  // namespace

// first-layer table this rank sends to the rank below (see tri_kernel): per (x, y) the vertex base of the MPU at the slab's first
// z-layer WITHIN THIS RANK's arrays (the receiver adds this rank's global offset, GridParams.haloVertexOffset - so the table does not
// wait for the count exchange) and its crossing-edge mask / prefix, or 0xFFFFFFFF where that MPU has no surface
__global__ void halo_build_kernel(GridParams G, const int32_t* __restrict__ candOfMpu, FieldOutputs fo, const uint32_t* __restrict__ vbase,
                                  uint32_t* __restrict__ hVbase, uint32_t* __restrict__ hMask, uint16_t* __restrict__ hPrefix, int nxy) {
    const int h = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (h >= nxy) return;
    const int nz = G.hi[2] - G.lo[2];
    const int32_t c = candOfMpu[(size_t)h * nz];
    const bool ok = c >= 0 && fo.tcount[c] != 0u;
    if (lane == 0) hVbase[h] = ok ? vbase[c] : 0xFFFFFFFFu;
    for (int w = lane; w < 48; w += 32) {
        hMask[(size_t)h * 48 + w] = ok ? fo.edgeMask[(size_t)c * 48 + w] : 0u;
        hPrefix[(size_t)h * 48 + w] = ok ? fo.edgePrefix[(size_t)c * 48 + w] : (uint16_t)0;
    }
}

// parity hook (psgpu_mpu_cache): the cube cases of ONE MPU from its field cache, by the classifier the emission kernels use
// (inside bits as bytes, cell_case) - so the hook checks the device classification, not a host re-statement of it
__global__ void cube_cases_kernel(GridParams G, const float* __restrict__ fv, uint8_t* __restrict__ case343) {
    __shared__ uint8_t B[64];
    const int t = threadIdx.x;  // 64 threads: one z-column each
    uint32_t inb = 0;
    for (int n = 0; n < 8; n++) inb |= (is_inside(G, fv[t * 8 + n]) ? 1u : 0u) << n;
    B[t] = (uint8_t)inb;
    __syncthreads();
    for (int c = t; c < 343; c += 64) {
        const int x = c / 49, y = (c - x * 49) / 7, z = c - x * 49 - y * 7;
        case343[c] = (uint8_t)cell_case(B, x, y, z);
    }
}

void launch_cube_cases(const GridParams& G, const float* fv, uint8_t* case343, cudaStream_t st) { cube_cases_kernel<<<1, 64, 0, st>>>(G, fv, case343); }

void launch_halo_build(const GridParams& G, const int32_t* candOfMpu, const FieldOutputs& fo, const uint32_t* vbase, uint32_t* hVbase,
                       uint32_t* hMask, uint16_t* hPrefix, int nxy, cudaStream_t st) {
    if (nxy <= 0) return;
    halo_build_kernel<<<(nxy + 3) / 4, 128, 0, st>>>(G, candOfMpu, fo, vbase, hVbase, hMask, hPrefix, nxy);
}

// PS_SimdPoly semantics with colours and normals wanted: the normal comes out of the colour pass (color_kernel<.., NRM>)
bool normals_in_color_pass(const GridParams& G) {
    return G.sem == PSGPU_SEM_SIMD && G.wantColors && G.wantNormals && !getenv("PSGPU_SEPARATE_NORMAL_PASS");
}

void launch_emit_kernels(const GridParams& G, const Region* cands, const float4* pool, const float* colorTab, const FieldOutputs& fo,
                         const MeshOutputs& mo, const int32_t* candOfMpu, int nMixed, uint32_t maxProgQuads, int numSMs, cudaStream_t st,
                         cudaEvent_t evTriDone, cudaEvent_t evVertexDone, int phase) {
    if (nMixed <= 0) return;
    if (phase != PS_EMIT_VERTEX_COLOR) {
        int perSM = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, tri_kernel, kThreads, 0);
        if (perSM < 1) perSM = 1;
        int grid = numSMs * perSM;
        const int needT = (nMixed + kThreads / 32 - 1) / (kThreads / 32);
        if (grid > needT) grid = needT;
        tri_kernel<<<grid, kThreads, 0, st>>>(G, cands, fo, mo, candOfMpu, nMixed);
        if (evTriDone) cudaEventRecord(evTriDone, st);
    }
    if (phase == PS_EMIT_TRI) return;
    const bool strict = (G.sem == PSGPU_SEM_SIMD);
    // per-warp staging slice = the longest list (multiple of 8 quads); up to 32 KB of dynamic shared memory per CTA
    uint32_t slice = (maxProgQuads + 7u) & ~7u;
    const bool smemOk = (size_t)slice * 16u * (kThreads / 32) <= 32768u;
    if (!smemOk) slice = 0;
    const int smem = (int)((size_t)slice * 16u * (kThreads / 32));
    using KernelT = void (*)(GridParams, const Region*, const float4*, FieldOutputs, MeshOutputs, int, uint32_t);
    KernelT k;
    const bool warps = !strict && G.hasWarps != 0;
    const bool masks = strict && G.hasMasks != 0;
#define PS_PICK(ROOT) (strict ? (masks ? (smemOk ? vertex_kernel<true, ROOT, true, true> : vertex_kernel<true, ROOT, false, true>) \
                                       : (smemOk ? vertex_kernel<true, ROOT, true, false> : vertex_kernel<true, ROOT, false, false>)) \
                       : warps ? (smemOk ? vertex_kernel<false, ROOT, true, true> : vertex_kernel<false, ROOT, false, true>) \
                               : (smemOk ? vertex_kernel<false, ROOT, true, false> : vertex_kernel<false, ROOT, false, false>))
    if (G.rootMode == PSGPU_ROOT_FOURSAMPLE) k = PS_PICK(PSGPU_ROOT_FOURSAMPLE);
    else if (G.rootMode == PSGPU_ROOT_NEWTON_REF) k = PS_PICK(PSGPU_ROOT_NEWTON_REF);
    else k = PS_PICK(PSGPU_ROOT_BISECT);
#undef PS_PICK
    const int need = (nMixed + kThreads / 32 - 1) / (kThreads / 32);
    int perSM = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k, kThreads, smem);
    if (perSM < 1) perSM = 1;
    int grid = numSMs * perSM;
    if (grid > need) grid = need;
    const bool nrmInColor = normals_in_color_pass(G);
    GridParams Gv = G;
    if (nrmInColor) Gv.wantNormals = 0;  // vertex_kernel places the vertex only
    k<<<grid, kThreads, smem, st>>>(Gv, cands, pool, fo, mo, nMixed, slice);
    if (evVertexDone) cudaEventRecord(evVertexDone, st);
    if (G.wantColors) {
        using ColorT = void (*)(GridParams, const Region*, const float4*, const float*, FieldOutputs, MeshOutputs, int, uint32_t);
        const ColorT kc = nrmInColor ? (masks ? (smemOk ? color_kernel<true, true, true, true> : color_kernel<true, false, true, true>)
                                              : (smemOk ? color_kernel<true, true, false, true> : color_kernel<true, false, false, true>))
                          : strict ? (smemOk ? color_kernel<true, true, false, false> : color_kernel<true, false, false, false>)
                          : warps ? (smemOk ? color_kernel<false, true, true, false> : color_kernel<false, false, true, false>)
                                  : (smemOk ? color_kernel<false, true, false, false> : color_kernel<false, false, false, false>);
        perSM = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kc, kThreads, smem);
        if (perSM < 1) perSM = 1;
        grid = numSMs * perSM;
        if (grid > need) grid = need;
        kc<<<grid, kThreads, smem, st>>>(G, cands, pool, colorTab, fo, mo, nMixed, slice);
    }
}

}  // namespace psgpu
