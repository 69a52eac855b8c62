// Field-evaluation kernel (north-star subsystem 3) and the point-sampling parity hooks.
//
// One CTA of 128 threads evaluates the 8x8x8 corner lattice of an MPU: the MPU's pruned instruction list is
// staged into shared memory by a TMA bulk copy (cp.async.bulk + mbarrier, double-buffered so the next MPU's
// list lands while this one is evaluated), every thread interprets it for 4 consecutive-z corners, and the
// field cache leaves as one coalesced float4 per thread. The same pass classifies the MPU (mixed signs?),
// marks the crossing edges it owns and counts its triangles, so the mesh kernel knows its output offsets.
//
// Reference: CMPUProcessor::process_cells_simd phases 0-2 (PS_SimdPoly/include/PS_Polygonizer.cpp:483-701) and
// CSIMDMPURunBody::doMarchingCubes' cache + classify (ParsipHaptics/include/CPolyParsipOptimized.cpp:187-244).
#include <cstdlib>

#include "device_util.cuh"
#include "interp_packed.cuh"
#include "kernels.h"
#include "mc_tables.h"

// resident 64-thread CTAs per SM asked of ptxas for the STRICT 8-point kernel: 12 = 80 registers, 16 = 64 registers (A/B in
// profiles/r02_*: see DESIGN.md section 9)
#ifndef PSGPU_STRICT_CTAS
#define PSGPU_STRICT_CTAS 12
#endif

namespace psgpu {

__constant__ uint8_t c_nTri[256] = PS_MC_NTRI_INIT;

namespace {

// z-consecutive corners per thread: 8 = a whole z-column of the 8x8x8 lattice, 64 threads per MPU - the decode cost of an interpreted
// instruction is shared by 8 points (ncu r01c: issue-bound on decode). Programs with warps (sincos per point, point stack) spill at
// 8 points per thread and run the 4-point / 128-thread variant.
constexpr int kMaxThreads = 128;
constexpr uint32_t kMaxProgCapQuads = 1024;  // 16 KB per staging buffer; longer lists are read from global

struct __align__(16) FieldSmemHeader {
    float fv[512];
    uint64_t bar[2];
    uint32_t insideW[16];
    uint32_t edgeW[48];
    uint32_t warpSum[8];
    uint8_t nTri[256];  // triangles per cube case (lanes look up different cases: shared memory, not __constant__)
    uint32_t spread3[256];  // bit n of the index -> bit 3n (interleaves the x / y / z crossing bytes of a z-column into 24 edge bits)
    int next;
    int flags;
};

// STAGED: every candidate list fits the staging buffers (the host checks the longest one) and is interpreted from shared memory;
// the unstaged variant (brute-force mode, very long lists) reads the lists from global memory. One interpreter copy per kernel.
template <bool STRICT, bool STAGED, int kPts, bool MASKS = false>
__global__ void __launch_bounds__(512 / kPts, kPts == 4 ? 6 : (STRICT ? PSGPU_STRICT_CTAS : 16)) field_kernel(GridParams G, const Region* __restrict__ cands, int nCand,
                                                         const float4* __restrict__ pool, FieldOutputs out, uint32_t capQuads) {
    extern __shared__ __align__(16) unsigned char smemRaw[];
    FieldSmemHeader& S = *reinterpret_cast<FieldSmemHeader*>(smemRaw);
    float4* progBuf = reinterpret_cast<float4*>(smemRaw + sizeof(FieldSmemHeader));
    constexpr int kThreads = 512 / kPts;
    const int t = threadIdx.x;
    const int lane = t & 31;

    if (t == 0) {
        mbar_init(&S.bar[0], 1);
        mbar_init(&S.bar[1], 1);
        fence_mbar_init();
        if (blockIdx.x == 0) { out.vcount[nCand] = 0u; out.tcount[nCand] = 0u; }  // closing entries of the exclusive scans that follow
    }
    for (int i = t; i < 256; i += kThreads) {
        S.nTri[i] = c_nTri[i];
        uint32_t sp = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) sp |= ((uint32_t)(i >> b) & 1u) << (3 * b);
        S.spread3[i] = sp;
    }
    __syncthreads();

    int item = blockIdx.x;  // first item is static, the rest come from the work counter
    int stage = 0;
    uint32_t phase[2] = {0u, 0u};
    if (STAGED && t == 0 && item < nCand) {
        const Region r = cands[item];
        mbar_expect_tx(&S.bar[0], r.len * 16u);
        tma_load_bulk(progBuf, pool + r.off, r.len * 16u, &S.bar[0]);
    }
    while (item < nCand) {
        // fetch and prefetch the next work item
        if (t == 0) {
            const int nxt = (int)atomicAdd(&out.counters[0], 1u) + (int)gridDim.x;
            S.next = nxt;
            if (STAGED && nxt < nCand) {
                const Region rn = cands[nxt];
                fence_proxy_async();
                mbar_expect_tx(&S.bar[stage ^ 1], rn.len * 16u);
                tma_load_bulk(progBuf + (size_t)(stage ^ 1) * capQuads, pool + rn.off, rn.len * 16u, &S.bar[stage ^ 1]);
            }
        }
        if (kPts != 8) {  // the 8-point variant writes whole bytes of both masks
            if (t < 16) S.insideW[t] = 0u;
            if (t < 48) S.edgeW[t] = 0u;
        }
        const Region r = cands[item];
        if (STAGED) {
            mbar_wait(&S.bar[stage], phase[stage]);
            phase[stage] ^= 1u;
        }

        // ---- evaluate kPts corners: flat index kPts*t + n = i*8 + j*64 + k
        const int k0 = (kPts * t) & 7, ci = ((kPts * t) >> 3) & 7, cj = (kPts * t) >> 6;
        float X[kPts], Y[kPts], Z[kPts], F[kPts];
        {
            const float x = corner_coord(G, 0, r.x, ci), y = corner_coord(G, 1, r.y, cj);
#pragma unroll
            for (int n = 0; n < kPts; n++) { X[n] = x; Y[n] = y; Z[n] = corner_coord(G, 2, r.z, k0 + n); }
        }
        if constexpr (STRICT) {
            // PS_SimdPoly semantics: packed-pair arithmetic, the column shares x and y (interp_packed.cuh)
            PairPts<kPts / 2, true> P;
            P.x[0] = X[0]; P.y[0] = Y[0];
#pragma unroll
            for (int n = 0; n < kPts / 2; n++) P.z[n] = pk(Z[2 * n], Z[2 * n + 1]);
            F2 FP[kPts / 2];
            eval_program_packed<kPts / 2, true, STAGED, false, MASKS>(STAGED ? progBuf + (size_t)stage * capQuads : pool + r.off, P, FP, nullptr, nullptr, true);  // packets of 4 z-corners
#pragma unroll
            for (int n = 0; n < kPts / 2; n++) up(FP[n], F[2 * n], F[2 * n + 1]);
        } else
            eval_program<kPts, false, STRICT, STAGED, (kPts == 4)>(STAGED ? progBuf + (size_t)stage * capQuads : pool + r.off, X, Y, Z, F, nullptr, nullptr, G.sem);
        uint32_t inb = 0;
#pragma unroll
        for (int n = 0; n < kPts; n++) inb |= (is_inside(G, F[n]) ? 1u : 0u) << n;
#pragma unroll
        for (int q = 0; q < kPts / 4; q++)
            reinterpret_cast<float4*>(S.fv)[t * (kPts / 4) + q] = make_float4(F[4 * q], F[4 * q + 1], F[4 * q + 2], F[4 * q + 3]);
        uint32_t nV = 0, nT = 0;
        bool mixed;
        uint32_t myT = 0;
        if constexpr (kPts == 8) {
            // A thread owns one z-column = exactly one byte of the inside bits (byte i + 8j = the 8 corners along k), so the
            // classification works on bytes with plain stores: no shared-memory atomics, one barrier instead of four
            // (ncu r02a: this epilogue was ~30 % of the kernel's samples, barrier stalls 0.9 per issue at 2 warps per CTA).
            uint8_t* B = reinterpret_cast<uint8_t*>(S.insideW);
            B[t] = (uint8_t)inb;
            __syncthreads();
            {   // early discard like PS_Polygonizer.cpp:602-609: all corners inside or all outside
                const uint4* W4 = reinterpret_cast<const uint4*>(S.insideW);
                const uint4 a = W4[0], b = W4[1], c = W4[2], d = W4[3];
                const uint32_t o = a.x | a.y | a.z | a.w | b.x | b.y | b.z | b.w | c.x | c.y | c.z | c.w | d.x | d.y | d.z | d.w;
                const uint32_t n = a.x & a.y & a.z & a.w & b.x & b.y & b.z & b.w & c.x & c.y & c.z & c.w & d.x & d.y & d.z & d.w;
                mixed = (o != 0u) && (n != 0xffffffffu);
            }
            if (mixed) {
                // ---- crossing edges this MPU owns: per corner the x, y, z edge leaving it; bit 3*(8t + n) + axis of the edge mask
                const bool lastX = (r.x == G.hi[0] - 1), lastY = (r.y == G.hi[1] - 1), lastZ = (r.z == G.zOwnLast);
                const bool perMpu = (G.output == PSGPU_OUT_MPU);
                const bool ownX = perMpu || ci < 7 || lastX, ownY = perMpu || cj < 7 || lastY;
                const uint32_t zmask = (perMpu || lastZ) ? 0xffu : 0x7fu;  // edges starting at ck = 7 belong to the next MPU in z
                const uint32_t bx = (ci < 7) ? B[t + 1] : inb, by = (cj < 7) ? B[t + 8] : inb;
                const uint32_t xc = (ownY ? zmask : 0u) & (inb ^ bx);
                const uint32_t yc = (ownX ? zmask : 0u) & (inb ^ by);
                const uint32_t zc = ((ownX && ownY) ? 0x7fu : 0u) & (inb ^ (inb >> 1));
                const uint32_t ebits = S.spread3[xc] | (S.spread3[yc] << 1) | (S.spread3[zc] << 2);
                uint8_t* E = reinterpret_cast<uint8_t*>(S.edgeW);  // 24 bits at bit 24 t = bytes 3t .. 3t + 2
                E[3 * t] = (uint8_t)ebits; E[3 * t + 1] = (uint8_t)(ebits >> 8); E[3 * t + 2] = (uint8_t)(ebits >> 16);
                // ---- triangles: the 7 cells above (ci, cj), cube case from the four neighbouring columns' bytes
                if (ci < 7 && cj < 7) {
                    const uint32_t b00 = inb, b10 = bx, b01 = by, b11 = B[t + 9];
#pragma unroll
                    for (int cz = 0; cz < 7; cz++) {
                        const uint32_t cfg = ((b00 >> cz) & 3u) | (((b01 >> cz) & 3u) << 2) | (((b10 >> cz) & 3u) << 4) | (((b11 >> cz) & 3u) << 6);
                        myT += S.nTri[cfg];
                    }
                }
            }
        } else {
        // early discard like PS_Polygonizer.cpp:602-609: all corners inside or all outside
        const int anyIn = __syncthreads_or(inb != 0u);
        const int anyOut = __syncthreads_or(inb != ((1u << kPts) - 1u));
        mixed = anyIn && anyOut;
        if (mixed) {
            atomicOr(&S.insideW[(kPts * t) >> 5], inb << ((kPts * t) & 31));
            __syncthreads();
            // ---- crossing edges this MPU owns: 3 bits per corner (x, y, z edge leaving it)
            const bool lastX = (r.x == G.hi[0] - 1), lastY = (r.y == G.hi[1] - 1), lastZ = (r.z == G.zOwnLast);
            const bool perMpu = (G.output == PSGPU_OUT_MPU);
            const bool ownX = perMpu || ci < 7 || lastX, ownY = perMpu || cj < 7 || lastY;
            uint32_t ebits = 0;
#pragma unroll
            for (int n = 0; n < kPts; n++) {
                const int ck = k0 + n;
                const int idx = kPts * t + n;
                const uint32_t me = (inb >> n) & 1u;
                const bool ownZ = perMpu || ck < 7 || lastZ;
                if (ci < 7 && ownY && ownZ) { const int nb = idx + 8; if (((S.insideW[nb >> 5] >> (nb & 31)) & 1u) != me) ebits |= 1u << (n * 3 + 0); }
                if (cj < 7 && ownX && ownZ) { const int nb = idx + 64; if (((S.insideW[nb >> 5] >> (nb & 31)) & 1u) != me) ebits |= 1u << (n * 3 + 1); }
                if (ck < 7 && ownX && ownY) { const int nb = idx + 1; if (((S.insideW[nb >> 5] >> (nb & 31)) & 1u) != me) ebits |= 1u << (n * 3 + 2); }
            }
            if (ebits) {
                const int e0 = 3 * kPts * t, sh = e0 & 31;
                atomicOr(&S.edgeW[e0 >> 5], ebits << sh);
                if (sh + 3 * kPts > 32) atomicOr(&S.edgeW[(e0 >> 5) + 1], ebits >> (32 - sh));
            }
            // ---- triangles: cube case of cell (x, y, z) from the inside bits read as bytes (byte i + 8j = the 8 corners along k)
            const uint8_t* B = reinterpret_cast<const uint8_t*>(S.insideW);
            for (int c = t; c < 343; c += kThreads) {
                const int cx = c / 49, cy = (c - cx * 49) / 7, cz = c - cx * 49 - cy * 7;
                const uint32_t b00 = B[cy * 8 + cx], b10 = B[cy * 8 + cx + 1], b01 = B[(cy + 1) * 8 + cx], b11 = B[(cy + 1) * 8 + cx + 1];
                const uint32_t cfg = ((b00 >> cz) & 3u) | (((b01 >> cz) & 3u) << 2) | (((b10 >> cz) & 3u) << 4) | (((b11 >> cz) & 3u) << 6);
                myT += S.nTri[cfg];
            }
        }
        }
        if (mixed) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) myT += __shfl_xor_sync(0xffffffffu, myT, o);
            if (lane == 0) S.warpSum[t >> 5] = myT;
            __syncthreads();
#pragma unroll
            for (int w = 0; w < kThreads / 32; w++) nT += S.warpSum[w];
            // ---- exclusive popcount prefix over the 48 edge words (warp 0)
            if (t < 32) {
                const uint32_t w0 = S.edgeW[t], w1 = (t < 16) ? S.edgeW[32 + t] : 0u;
                uint32_t p0 = __popc(w0), p1 = __popc(w1);
                uint32_t s0 = p0, s1 = p1;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t a = __shfl_up_sync(0xffffffffu, s0, o), b = __shfl_up_sync(0xffffffffu, s1, o);
                    if (lane >= o) { s0 += a; s1 += b; }
                }
                const uint32_t tot0 = __shfl_sync(0xffffffffu, s0, 31), tot1 = __shfl_sync(0xffffffffu, s1, 15);
                out.edgeMask[(size_t)item * 48 + t] = w0;
                out.edgePrefix[(size_t)item * 48 + t] = (uint16_t)(s0 - p0);
                if (t < 16) {
                    out.edgeMask[(size_t)item * 48 + 32 + t] = w1;
                    out.edgePrefix[(size_t)item * 48 + 32 + t] = (uint16_t)(tot0 + s1 - p1);
                }
                if (t == 0) S.warpSum[7] = tot0 + tot1;
            }
            __syncthreads();
            nV = S.warpSum[7];
        }
        if (G.forceFv || nT > 0) {
#pragma unroll
            for (int q = 0; q < kPts / 4; q++) out.fv[(size_t)item * 128 + t * (kPts / 4) + q] = reinterpret_cast<float4*>(S.fv)[t * (kPts / 4) + q];
            // the 64 bytes tri_kernel builds its cube cases from (it used to re-derive them from the 2 KB field cache)
            if (nT > 0 && t < 16) out.inside[(size_t)item * 16 + t] = S.insideW[t];
        }
        if (t == 0) {
            out.vcount[item] = (nT > 0) ? nV : 0u;
            out.tcount[item] = nT;
            if (nT > 0) {
                const uint32_t slot = atomicAdd(&out.counters[1], 1u);
                out.mixedList[slot] = (uint32_t)item;
                if (G.output == PSGPU_OUT_MPU && (nV > 512u || nT > 512u)) atomicAdd(&out.counters[2], 1u);
            }
        }
        __syncthreads();
        item = S.next;
        stage ^= 1;
        __syncthreads();
    }
}

// ---- parity hooks: the full program at arbitrary points ------------------------------------------------
template <bool STRICT>
__global__ void __launch_bounds__(128) points_field_kernel(const float4* __restrict__ prog, const float* __restrict__ xyz, int64_t n,
                                                           float* __restrict__ f, int sem) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = i < n ? i : n - 1;
    float X[1] = {xyz[j * 3]}, Y[1] = {xyz[j * 3 + 1]}, Z[1] = {xyz[j * 3 + 2]}, F[1];
    eval_program<1, false, STRICT>(prog, X, Y, Z, F, nullptr, nullptr, sem);
    if (i < n) f[i] = F[0];
}

// field + colour at p and the forward-difference normal with NORMAL_DELTA (PS_Polygonizer.cpp:764-782, 1598-1622)
template <bool STRICT>
__global__ void __launch_bounds__(128) points_fcn_kernel(const float4* __restrict__ prog, const float* __restrict__ colorTab,
                                                         const float* __restrict__ xyz, int64_t n, float* __restrict__ f,
                                                         float* __restrict__ col, float* __restrict__ nrm, int sem) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = i < n ? i : n - 1;
    const float px = xyz[j * 3], py = xyz[j * 3 + 1], pz = xyz[j * 3 + 2];
    float X[4] = {px, px + PS_NORMAL_DELTA, px, px}, Y[4] = {py, py, py + PS_NORMAL_DELTA, py}, Z[4] = {pz, pz, pz, pz + PS_NORMAL_DELTA};
    float F[4], rgb[3];
    eval_program<4, true, STRICT>(prog, X, Y, Z, F, rgb, colorTab, sem);
    if (i < n) {
        const float dinv = -1.0f / PS_NORMAL_DELTA;
        float nx = (F[1] - F[0]) * dinv, ny = (F[2] - F[0]) * dinv, nz = (F[3] - F[0]) * dinv;
        const float l2 = nx * nx + ny * ny + nz * nz;
        if (l2 > 0.f) { const float inv = rsqrtf(l2); nx *= inv; ny *= inv; nz *= inv; }
        else if (sem == PSGPU_SEM_SCALAR) { nx = ny = nz = 1.0f; }  // normalizeXYZ of a zero vector, FW/PS_Vector.h:1072-1086
        f[i] = F[0];
        col[i * 3] = rgb[0]; col[i * 3 + 1] = rgb[1]; col[i * 3 + 2] = rgb[2];
        nrm[i * 3] = nx; nrm[i * 3 + 1] = ny; nrm[i * 3 + 2] = nz;
    }
}

// Shrink-wrap Newton root between two points, one edge per thread, in the uploaded program's semantics:
//  STRICT  FieldComputer::computeRootNewtonRaphson (PS_SimdPoly/include/PS_Polygonizer.cpp:1625-1681): starts at the end whose field is
//          nearer the target and CARRIES that field value; gradient() by forward differences (:1576-1596); g = (t - f) / |grad|^2
//  else    ComputeRootNewtonRaphsonVEC4 (ParsipHaptics/include/CompactBlobTree.cpp:1581-1622): field re-evaluated with the gradient,
//          g = 1 / (|grad|^2 + f^2) (its vec4 dot includes the field lane)
// evals = (i + 1) * 4 with i the loop counter at exit (iterations + 1 rounds reported when the loop runs out), as both return it.
template <bool STRICT>
__global__ void __launch_bounds__(128) points_newton_kernel(const float4* __restrict__ prog, const float* __restrict__ p1, const float* __restrict__ p2,
                                                            const float* __restrict__ f1, const float* __restrict__ f2, int64_t n, float target,
                                                            int iterations, float* __restrict__ root, float* __restrict__ rootField,
                                                            int32_t* __restrict__ evals, int sem) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = i < n ? i : n - 1;
    const bool first = fabsf(f1[j] - target) < fabsf(f2[j] - target);
    const float* s = first ? p1 : p2;
    float x = s[j * 3], y = s[j * 3 + 1], z = s[j * 3 + 2];
    float field = first ? f1[j] : f2[j];
    const float eps = PS_FIELDVALUE_EPSILON, inv = 1.0f / PS_FIELDVALUE_EPSILON;
    int it = 0;
    for (; it < iterations; it++) {
        float X[4] = {x, __fadd_rn(x, eps), x, x}, Y[4] = {y, y, __fadd_rn(y, eps), y}, Z[4] = {z, z, z, __fadd_rn(z, eps)}, F[4];
        eval_program<4, false, STRICT>(prog, X, Y, Z, F, nullptr, nullptr, sem);
        if (!STRICT) field = F[0];
        const float gx = __fmul_rn(__fsub_rn(F[1], field), inv), gy = __fmul_rn(__fsub_rn(F[2], field), inv), gz = __fmul_rn(__fsub_rn(F[3], field), inv);
        const float d = __fsub_rn(target, field);
        const float gg = __fadd_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)), __fmul_rn(gz, gz));
        if (STRICT) {
            const float g = __fdiv_rn(d, gg);
            x = __fadd_rn(x, __fmul_rn(gx, g)); y = __fadd_rn(y, __fmul_rn(gy, g)); z = __fadd_rn(z, __fmul_rn(gz, g));
        } else {
            const float g = __fdiv_rn(1.0f, __fadd_rn(gg, __fmul_rn(field, field)));
            x = __fadd_rn(x, __fmul_rn(__fmul_rn(d, gx), g)); y = __fadd_rn(y, __fmul_rn(__fmul_rn(d, gy), g)); z = __fadd_rn(z, __fmul_rn(__fmul_rn(d, gz), g));
        }
        float X1[1] = {x}, Y1[1] = {y}, Z1[1] = {z}, F1[1];
        eval_program<1, false, STRICT>(prog, X1, Y1, Z1, F1, nullptr, nullptr, sem);
        field = F1[0];
        if (fabsf(__fsub_rn(field, target)) < eps) break;
    }
    if (i < n) {
        root[i * 3] = x; root[i * 3 + 1] = y; root[i * 3 + 2] = z;
        rootField[i] = field;
        evals[i] = iterations <= 0 ? -1 : (it + 1) * 4;
    }
}

// exact SURVEY 8(d) bookkeeping over the final per-MPU programs: acc[0] += 512 * flops, acc[1] += 512 * prims,
// acc[2] += quads, acc[3] += 512 (corner evaluations)
__global__ void program_stats_kernel(const Region* __restrict__ cands, int nCand, const float4* __restrict__ pool,
                                     const uint32_t* __restrict__ vcount, unsigned long long* __restrict__ acc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long fl = 0, pr = 0, qd = 0, ev = 0, vf = 0;
    if (i < nCand) {
        const Region r = cands[i];
        const float4* p = pool + r.off;
        for (uint32_t pc = 0;;) {
            const uint32_t hdr = __float_as_uint(p[pc].x);
            if (PS_HDR_OPC(hdr) == OPC_END) break;
            fl += (unsigned long long)ps_algorithmic_flops(hdr);
            if (PS_HDR_OPC(hdr) < OPC_FIRST_NONPRIM) pr++;
            pc += ps_instr_quads(hdr);
        }
        qd = r.len; ev = 512;
        if (vcount) vf = fl * vcount[i];
        fl *= 512; pr *= 512;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        fl += __shfl_xor_sync(0xffffffffu, fl, o); pr += __shfl_xor_sync(0xffffffffu, pr, o);
        qd += __shfl_xor_sync(0xffffffffu, qd, o); ev += __shfl_xor_sync(0xffffffffu, ev, o);
        vf += __shfl_xor_sync(0xffffffffu, vf, o);
    }
    if ((threadIdx.x & 31) == 0 && ev) { atomicAdd(&acc[0], fl); atomicAdd(&acc[1], pr); atomicAdd(&acc[2], qd); atomicAdd(&acc[3], ev); atomicAdd(&acc[4], vf); }
}

// FP32 FMA-chain microbenchmark: 8 independent FFMA chains per thread, 2048 threads per SM
__global__ void __launch_bounds__(256) fp32_peak_kernel(float* out, int iters, float a, float b) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = (float)(threadIdx.x + i);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = fmaf(v[i], a, b);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += v[i];
    if (s == 12345.678f) out[0] = s;
}

}  // namespace

int field_kernel_smem_bytes(uint32_t maxProgQuads, uint32_t* capQuads) {
    uint32_t cap = maxProgQuads < 8 ? 8 : maxProgQuads;
    if (cap > kMaxProgCapQuads) cap = kMaxProgCapQuads;
    cap = (cap + 7u) & ~7u;
    *capQuads = cap;
    return (int)(sizeof(FieldSmemHeader) + 2u * cap * 16u);
}

void launch_field_kernel(const GridParams& G, const Region* cands, int nCand, const float4* pool, const FieldOutputs& out, uint32_t maxProgQuads,
                         bool hasWarps, int numSMs, cudaStream_t st) {
    if (nCand <= 0) return;
    uint32_t cap;
    int smem = field_kernel_smem_bytes(maxProgQuads, &cap);
    bool staged = (maxProgQuads <= cap);
    if (getenv("PSGPU_NO_STAGE")) staged = false;  // debugging aid: interpret straight from global memory
    if (!staged) { cap = 8; smem = (int)(sizeof(FieldSmemHeader) + 2u * cap * 16u); }
    const bool strict = (G.sem == PSGPU_SEM_SIMD);
    const bool four = !strict && hasWarps;
    using KernelT = void (*)(GridParams, const Region*, int, const float4*, FieldOutputs, uint32_t);
    KernelT k;
    if (strict && G.hasMasks) k = staged ? field_kernel<true, true, 8, true> : field_kernel<true, false, 8, true>;  // PSGPU_LIN_AS_SHIPPED programs
    else if (strict) k = staged ? field_kernel<true, true, 8> : field_kernel<true, false, 8>;
    else if (four) k = staged ? field_kernel<false, true, 4> : field_kernel<false, false, 4>;
    else k = staged ? field_kernel<false, true, 8> : field_kernel<false, false, 8>;
    const int threads = four ? 128 : 64;
    // per launch: the attribute is per device and function, a process may drive several devices (a microsecond of host time)
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int perSM = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k, threads, smem);
    if (perSM < 1) perSM = 1;
    int grid = numSMs * perSM;  // persistent CTAs: a whole number of waves on 148 SMs
    if (grid > nCand) grid = nCand;
    k<<<grid, threads, smem, st>>>(G, cands, nCand, pool, out, cap);
}

void launch_points_field(const float4* prog, const float* xyz, int64_t n, float* f, int sem, cudaStream_t st) {
    if (n <= 0) return;
    if (sem == PSGPU_SEM_SIMD) points_field_kernel<true><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(prog, xyz, n, f, sem);
    else points_field_kernel<false><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(prog, xyz, n, f, sem);
}

void launch_points_field_color_normal(const float4* prog, const float* colorTab, const float* xyz, int64_t n, float* f, float* col, float* nrm,
                                      int sem, cudaStream_t st) {
    if (n <= 0) return;
    if (sem == PSGPU_SEM_SIMD) points_fcn_kernel<true><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(prog, colorTab, xyz, n, f, col, nrm, sem);
    else points_fcn_kernel<false><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(prog, colorTab, xyz, n, f, col, nrm, sem);
}

void launch_points_newton(const float4* prog, const float* p1, const float* p2, const float* f1, const float* f2, int64_t n, float target, int iterations,
                          float* root, float* rootField, int32_t* evals, int sem, cudaStream_t st) {
    if (n <= 0) return;
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (sem == PSGPU_SEM_SIMD) points_newton_kernel<true><<<grid, 128, 0, st>>>(prog, p1, p2, f1, f2, n, target, iterations, root, rootField, evals, sem);
    else points_newton_kernel<false><<<grid, 128, 0, st>>>(prog, p1, p2, f1, f2, n, target, iterations, root, rootField, evals, sem);
}

void launch_program_stats(const Region* cands, int nCand, const float4* pool, const uint32_t* vcount, unsigned long long* acc, cudaStream_t st) {
    if (nCand <= 0) return;
    program_stats_kernel<<<(nCand + 127) / 128, 128, 0, st>>>(cands, nCand, pool, vcount, acc);
}

float run_fp32_peak(int numSMs, cudaStream_t st, float* ms_out) {
    float* d = nullptr;
    cudaMalloc(&d, 4);
    const int iters = 4096, blocks = numSMs * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0, st);
        fp32_peak_kernel<<<blocks, 256, 0, st>>>(d, iters, 1.0000001f, 1e-9f);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d);
    if (ms_out) *ms_out = best;
    const double flops = 2.0 * 64.0 * iters * 256.0 * blocks;
    return (float)(flops / (best * 1e-3) / 1e12);
}

}  // namespace psgpu
