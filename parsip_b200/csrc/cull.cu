// Per-MPU primitive culling (north-star subsystem 2): prunes the postfix program against a region of the
// MPU lattice, top-down over a 4x4x4-branching hierarchy of regions, and ends with one self-contained pruned
// program per surviving MPU in a device pool (which the field kernel stages into shared memory by TMA).
//
// The reference has only on/off analogues of this: the 8-corner zero test
// (PS_SimdPoly/include/PS_Polygonizer.cpp:483-540), the MPU-vs-primitive-octree scan
// (ParsipHaptics/include/CPolyParsipOptimized.cpp:160-183) and the per-packet operator-box prune
// (PS_Polygonizer.cpp:1228-1252). Here a primitive is dropped only when its Wyvill support (distance >= 1)
// provably misses the region, so the pruned program evaluates to exactly the same floats as the full one
// (zero-folding rules of SURVEY Q16), and - PSGPU_CULL_INTERVAL - a region is dropped when interval bounds of
// the field over the region exclude the iso value (no surface can cross it).
#include <cub/device/device_scan.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/transform_iterator.h>
#include <thrust/iterator/transform_output_iterator.h>
#include <thrust/iterator/zip_iterator.h>
#include <thrust/tuple.h>

#include "cull_walk.cuh"
#include "kernels.h"

namespace psgpu {

namespace {

__device__ __forceinline__ bool child_region(const CullLevel& L, const Region& parent, int c, int (&coord)[3], Box3& box) {
    coord[0] = parent.x * 4 + (c & 3); coord[1] = parent.y * 4 + ((c >> 2) & 3); coord[2] = parent.z * 4 + (c >> 4);
    const int s = L.childSize;
    int m0[3], m1[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        m0[a] = coord[a] * s;
        m1[a] = min(m0[a] + s, L.hi[a]);
        if (m0[a] < L.lo[a]) m0[a] = L.lo[a];
        if (m0[a] >= m1[a]) return false;
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        // covers both coordinate conventions (per-MPU origin and global index differ by an ulp or two)
        const float pad = 1e-5f * (fabsf(L.gridLo[a]) + (float)L.hi[a] * L.mpuSide) + 1e-6f;
        box.lo[a] = L.gridLo[a] + (float)m0[a] * L.mpuSide - pad;
        box.hi[a] = L.gridLo[a] + (float)m1[a] * L.mpuSide + pad;
    }
    return true;
}

__device__ __forceinline__ void publish_child(const CullLevel& L, const int (&coord)[3], uint32_t off, uint32_t len, uint32_t idx,
                                              Region* __restrict__ children, int32_t* __restrict__ candOfMpu, uint32_t* __restrict__ maxLen) {
    Region r;
    r.x = coord[0]; r.y = coord[1]; r.z = coord[2]; r.off = off; r.len = len;
    children[idx] = r;
    if (maxLen) atomicMax(maxLen, len);
    if (candOfMpu) {
        const int ny = L.hi[1] - L.lo[1], nz = L.hi[2] - L.lo[2];
        candOfMpu[((size_t)(coord[0] - L.lo[0]) * ny + (coord[1] - L.lo[1])) * nz + (coord[2] - L.lo[2])] = (int32_t)idx;
    }
}

__global__ void __launch_bounds__(128) prune_count_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                          const float4* __restrict__ pool, uint32_t* __restrict__ childLen,
                                                          uint32_t* __restrict__ maxLen) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nParents * 64) return;
    const Region parent = parents[tid >> 6];
    int coord[3];
    Box3 box;
    uint32_t len = 0;
    if (child_region(L, parent, tid & 63, coord, box))
        len = prune_walk<false>(pool + parent.off, box, L.cullMode, L.iso, nullptr, L.colorExact != 0);
    childLen[tid] = len;
    // the longest pruned program of the level (the host reads it back with the scan totals): one atomic per warp
    const uint32_t act = __activemask();
    const uint32_t m = __reduce_max_sync(act, len);
    if (m && (threadIdx.x & 31) == __ffs((int)act) - 1) atomicMax(maxLen, m);
}

__global__ void __launch_bounds__(128) prune_emit_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                         const float4* __restrict__ pool, const uint32_t* __restrict__ childLen,
                                                         const uint32_t* __restrict__ childOff, const uint32_t* __restrict__ childIdx,
                                                         float4* __restrict__ outPool, uint8_t* __restrict__ outFlag, uint32_t outBase,
                                                         Region* __restrict__ children, int32_t* __restrict__ candOfMpu,
                                                         uint32_t* __restrict__ maxLen) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nParents * 64) return;
    const uint32_t len = childLen[tid];
    if (len == 0) return;
    const Region parent = parents[tid >> 6];
    int coord[3];
    Box3 box;
    child_region(L, parent, tid & 63, coord, box);
    const uint32_t off = outBase + childOff[tid];
    prune_walk<true>(pool + parent.off, box, L.cullMode, L.iso, outPool + off, L.colorExact != 0, len, outFlag ? outFlag + off : nullptr);
    publish_child(L, coord, off, len, childIdx[tid], children, candOfMpu, maxLen);
}

// single-walk variant of the thread-per-MPU level: emit into a private scratch slot, compacted after the scans (one walk instead of
// count + emit; the slot stride is the longest parent program)
__global__ void __launch_bounds__(128) prune_fused_thread_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                                 const float4* __restrict__ pool, float4* __restrict__ scratch, uint32_t stride,
                                                                 uint32_t* __restrict__ childLen, uint32_t* __restrict__ maxLen) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nParents * 64) return;
    const Region parent = parents[tid >> 6];
    int coord[3];
    Box3 box;
    uint32_t len = 0;
    if (child_region(L, parent, tid & 63, coord, box))
        len = prune_walk<true>(pool + parent.off, box, L.cullMode, L.iso, scratch + (size_t)tid * stride, L.colorExact != 0, stride, nullptr);
    if (len > stride) len = 0xFFFFFFFFu;  // cannot happen (emitted <= source); poisons the total so the host notices
    childLen[tid] = len;
    if (len) atomicMax(maxLen, len);
}

// ---- warp-cooperative pruning --------------------------------------------------------------------------------------
// The coarse levels have few regions and long parent programs (level 0 walks the whole BlobTree), so one thread per region
// is bound by the latency of ~1000 dependent primitive bounds (ncu r01b: 0.7 ms per pass on 4 CTAs). Here a WARP prunes one
// region: 32 quad positions at a time, every lane that sits on an instruction header (flag byte) computes that primitive's
// bound, all-zero blend operands are dropped on the spot, and lane 0 feeds the few remaining instructions to the Pruner state
// machine in program order - the same walk, so the output is identical to prune_walk's.
constexpr int kCoopWarps = 4;
struct __align__(16) CoopEntry { uint32_t pc; uint32_t za; float flo, fhi; };  // pc bit 31 = primitive is all-zero here

template <bool EMIT>
__device__ uint32_t prune_walk_warp(const float4* __restrict__ src, const uint8_t* __restrict__ sflag, uint32_t srcLen, const Box3& region,
                                    int cullMode, float iso, float4* __restrict__ dst, uint8_t* dflag, bool colorExact, uint32_t limit,
                                    CoopEntry* ent, bool nonNeg = false) {
    const int lane = threadIdx.x & 31;
    const uint32_t FULL = 0xffffffffu;
    Pruner<EMIT> P;
    P.init(src, region, dst, dflag, colorExact, limit);
    Box3 box = region;
    uint32_t base = 0, zaTotal = 0, zaPrev = 0;
    for (;;) {
        const uint32_t p = base + lane;
        const bool isInstr = (p < srcLen) && (sflag[p] != 0);
        float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (isInstr) q0 = src[p];
        const uint32_t hdr = __float_as_uint(q0.x);
        const uint32_t opc = PS_HDR_OPC(hdr);
        // Primitive bounds are computed for the whole batch against the box at its start, so the batch ends BEFORE the first
        // primitive that follows an instruction changing the region box (warp, operator matrix, OPC_POPPOINT): lane 0 steps through
        // the changers themselves (a run of them costs one batch, not one each), and the next batch starts at that primitive with
        // the updated box. OPC_END closes the walk.
        const bool isPrim = isInstr && opc != OPC_END && opc < OPC_FIRST_NONPRIM;
        const uint32_t chMask = __ballot_sync(FULL, isInstr && ps_changes_box(opc));
        const uint32_t staleMask = __ballot_sync(FULL, isPrim && (chMask & ((1u << lane) - 1u)) != 0u);
        const uint32_t endMask = __ballot_sync(FULL, isInstr && opc == OPC_END);
        const int staleLane = staleMask ? (__ffs((int)staleMask) - 1) : 32;
        const int endLane = endMask ? (__ffs((int)endMask) - 1) : 32;
        const bool active = isInstr && lane < staleLane && lane <= endLane;
        bool pz = false, skip = false;
        float flo = 0.f, fhi = 0.f;
        uint32_t scale = 0;
        if (active && opc != OPC_END && opc < OPC_FIRST_NONPRIM) {
            pz = prim_bound(hdr, q0, src + p + 1, box, flo, fhi);
            skip = Pruner<EMIT>::skippable(hdr, pz, scale, nonNeg);
        }
        const bool isEntry = active && !skip;
        const uint32_t zs = (active && skip) ? scale : 0u;
        uint32_t incl = zs;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t entMask = __ballot_sync(FULL, isEntry);
        // "PUSH <all-zero primitive>" whose union (OPC_COMBINE, CMB_MAX, no colour-scale marker) is the very next surviving
        // instruction: the culled blend group of a union of blends. Found here by all lanes; lane 0 then steps over the pair (below).
        const uint32_t above = entMask & ~((2u << lane) - 1u);
        const uint32_t hdrNext = __shfl_sync(FULL, hdr, above ? (__ffs((int)above) - 1) : lane);
        const bool zeroPair = isEntry && isPrim && pz && PS_HDR_CMB(hdr) == CMB_PUSH && above != 0u && ps_is_plain_union(hdrNext, colorExact);
        if (isEntry) {
            CoopEntry e;
            e.pc = p | (pz ? 0x80000000u : 0u) | (zeroPair ? 0x40000000u : 0u); e.za = zaTotal + incl - zs; e.flo = flo; e.fhi = fhi;
            ent[__popc(entMask & ((1u << lane) - 1u))] = e;
        }
        zaTotal += __shfl_sync(FULL, incl, 31);
        __syncwarp();
        int endSeen = 0;
        if (lane == 0) {
            const int n = __popc(entMask);
            for (int i = 0; i < n; i++) {
                const CoopEntry e = ent[i];
                const uint32_t pc = e.pc & 0x3FFFFFFFu;
                P.skipped_scale(e.za - zaPrev);
                zaPrev = e.za;
                if ((e.pc & 0x40000000u) && P.zero_union_is_noop()) {  // max(T, 0) with T >= 0 or T all-zero: nothing to emit, no state to change
                    zaPrev = ent[i + 1].za;  // scale steps owed between the two belong to the zero entry: dropped
                    i++;
                    continue;
                }
                const float4 h0 = src[pc];
                const uint32_t h = __float_as_uint(h0.x), o = PS_HDR_OPC(h);
                if (o == OPC_END) { endSeen = 1; break; }
                if (o < OPC_FIRST_NONPRIM) P.step_prim(pc, h0, h, (e.pc >> 31) != 0u, e.flo, e.fhi);
                else P.step_op(pc, h0, h);
            }
        }
        __syncwarp();
        endSeen = __shfl_sync(FULL, endSeen, 0);
        if (endSeen) break;
        if (chMask) {
#pragma unroll
            for (int a = 0; a < 3; a++) { box.lo[a] = __shfl_sync(FULL, P.box.lo[a], 0); box.hi[a] = __shfl_sync(FULL, P.box.hi[a], 0); }
        }
        base += (uint32_t)staleLane;  // 32 when no primitive of the batch was left behind
        if (base >= srcLen) break;  // malformed program without OPC_END
    }
    uint32_t len = 0;
    if (lane == 0) len = P.finish(cullMode, iso);
    return __shfl_sync(FULL, len, 0);
}

__global__ void __launch_bounds__(kCoopWarps * 32) prune_count_warp_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                                           const float4* __restrict__ pool, const uint8_t* __restrict__ qflag,
                                                                           uint32_t* __restrict__ childLen, uint32_t* __restrict__ maxLen) {
    __shared__ CoopEntry sEnt[kCoopWarps][32];
    const int gw = blockIdx.x * kCoopWarps + (threadIdx.x >> 5);
    if (gw >= nParents * 64) return;
    const Region parent = parents[gw >> 6];
    int coord[3];
    Box3 box;
    uint32_t len = 0;
    if (child_region(L, parent, gw & 63, coord, box))
        len = prune_walk_warp<false>(pool + parent.off, qflag + parent.off, parent.len, box, L.cullMode, L.iso, nullptr, nullptr, L.colorExact != 0,
                                     0xFFFFFFFFu, sEnt[threadIdx.x >> 5], L.nonNeg != 0);
    if ((threadIdx.x & 31) == 0) {
        childLen[gw] = len;
        if (len) atomicMax(maxLen, len);
    }
}

__global__ void __launch_bounds__(kCoopWarps * 32) prune_emit_warp_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                                          const float4* __restrict__ pool, const uint8_t* __restrict__ qflag,
                                                                          const uint32_t* __restrict__ childLen, const uint32_t* __restrict__ childOff,
                                                                          const uint32_t* __restrict__ childIdx, float4* __restrict__ outPool,
                                                                          uint8_t* __restrict__ outFlag, uint32_t outBase, Region* __restrict__ children,
                                                                          int32_t* __restrict__ candOfMpu, uint32_t* __restrict__ maxLen) {
    __shared__ CoopEntry sEnt[kCoopWarps][32];
    const int gw = blockIdx.x * kCoopWarps + (threadIdx.x >> 5);
    if (gw >= nParents * 64) return;
    const uint32_t len = childLen[gw];
    if (len == 0) return;
    const Region parent = parents[gw >> 6];
    int coord[3];
    Box3 box;
    child_region(L, parent, gw & 63, coord, box);
    const uint32_t off = outBase + childOff[gw];
    prune_walk_warp<true>(pool + parent.off, qflag + parent.off, parent.len, box, L.cullMode, L.iso, outPool + off, outFlag ? outFlag + off : nullptr,
                          L.colorExact != 0, len, sEnt[threadIdx.x >> 5], L.nonNeg != 0);
    if ((threadIdx.x & 31) == 0) publish_child(L, coord, off, len, childIdx[gw], children, candOfMpu, maxLen);
}

// ---- single-walk variant of the warp levels: the walk emits straight into a private scratch slot of `stride` quads (a pruned
// list is never longer than its parent's), the scans run on the lengths it returns, and prune_compact_kernel moves the slots to
// their final pool offsets - one walk per region instead of count + emit.
__global__ void __launch_bounds__(kCoopWarps * 32) prune_fused_warp_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                                           const float4* __restrict__ pool, const uint8_t* __restrict__ qflag,
                                                                           float4* __restrict__ scratch, uint8_t* __restrict__ scratchFlag, uint32_t stride,
                                                                           uint32_t* __restrict__ childLen, uint32_t* __restrict__ maxLen) {
    __shared__ CoopEntry sEnt[kCoopWarps][32];
    const int gw = blockIdx.x * kCoopWarps + (threadIdx.x >> 5);
    if (gw >= nParents * 64) return;
    const Region parent = parents[gw >> 6];
    int coord[3];
    Box3 box;
    uint32_t len = 0;
    if (child_region(L, parent, gw & 63, coord, box))
        len = prune_walk_warp<true>(pool + parent.off, qflag + parent.off, parent.len, box, L.cullMode, L.iso, scratch + (size_t)gw * stride,
                                    scratchFlag + (size_t)gw * stride, L.colorExact != 0, stride, sEnt[threadIdx.x >> 5], L.nonNeg != 0);
    if ((threadIdx.x & 31) == 0) {
        if (len > stride) len = 0xFFFFFFFFu;  // cannot happen (emitted <= source); poisons the total so the host notices
        childLen[gw] = len;
        if (len) atomicMax(maxLen, len);
    }
}

// ---- CTA-cooperative pruning for the top level ----------------------------------------------------------------------
// Level 0 walks the WHOLE program for a few hundred regions (blob10k@1024^3 / 8 ranks: 200 regions x 17.5k quads): one warp per
// region leaves 1-2 warps per SM, each ~5000 cycles per 32-quad batch with nothing to overlap (ncu r01e: 1.38 ms). Here a CTA
// of kCtaWarps warps takes the region: kCtaWarps*32 quad positions per step, the surviving instructions of the step are
// compacted IN PROGRAM ORDER into shared memory (with their header quad, so the serial part never touches global memory for a
// header), and thread 0 feeds them to the same Pruner state machine - the output is identical to prune_walk's.
constexpr int kCtaWarps = 16;
struct __align__(16) CtaEntry { float4 h0; uint32_t pc, za; float flo, fhi; };
struct CtaWalkSmem {
    CtaEntry ent[kCtaWarps * 32];
    uint32_t cnt[kCtaWarps], za[kCtaWarps];
    uint32_t chg[kCtaWarps];  // per warp: ballot of box-changing instructions in this step
    uint32_t cut, stale, endSeen, len;
    float box[6];
};

template <bool EMIT, bool FINE = false>
__device__ uint32_t prune_walk_cta(const float4* __restrict__ src, const uint8_t* __restrict__ sflag, uint32_t srcLen, const Box3& region,
                                   int cullMode, float iso, float4* __restrict__ dst, uint8_t* dflag, bool colorExact, uint32_t limit,
                                   CtaWalkSmem& S, bool nonNeg = false) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t FULL = 0xffffffffu, NONE = 0xffffffffu;
    Pruner<EMIT> P;
    P.init(src, region, dst, dflag, colorExact, limit);
    Box3 box = region;
    uint32_t base = 0, zaTotal = 0, zaPrev = 0;
    for (;;) {
        if (tid == 0) { S.cut = NONE; S.stale = NONE; }
        __syncthreads();
        const uint32_t p = base + (uint32_t)tid;
        const bool isInstr = (p < srcLen) && (sflag[p] != 0);
        float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (isInstr) q0 = src[p];
        const uint32_t hdr = __float_as_uint(q0.x);
        const uint32_t opc = PS_HDR_OPC(hdr);
        // as in prune_walk_warp: the step ends before the first primitive that follows a box-changing instruction, and at OPC_END
        const bool isPrim = isInstr && opc != OPC_END && opc < OPC_FIRST_NONPRIM;
        const uint32_t chMask = __ballot_sync(FULL, isInstr && ps_changes_box(opc));
        const uint32_t endMask = __ballot_sync(FULL, isInstr && opc == OPC_END);
        if (lane == 0) S.chg[w] = chMask;
        if (endMask && lane == __ffs((int)endMask) - 1) atomicMin(&S.cut, p);
        __syncthreads();
        bool changerBefore = (chMask & ((1u << lane) - 1u)) != 0u;
#pragma unroll
        for (int i = 0; i < kCtaWarps; i++) changerBefore = changerBefore || (i < w && S.chg[i] != 0u);
        const uint32_t staleMask = __ballot_sync(FULL, isPrim && changerBefore);
        if (staleMask && lane == __ffs((int)staleMask) - 1) atomicMin(&S.stale, p);
        bool anyChanger = false;
#pragma unroll
        for (int i = 0; i < kCtaWarps; i++) anyChanger = anyChanger || S.chg[i] != 0u;
        __syncthreads();
        const uint32_t cut = S.cut, stale = S.stale;
        const bool active = isInstr && p < stale && p <= cut;
        bool pz = false, skip = false;
        float flo = 0.f, fhi = 0.f;
        uint32_t scale = 0;
        if (active && opc != OPC_END && opc < OPC_FIRST_NONPRIM) {
            pz = FINE ? prim_bound_tiled(hdr, q0, src + p + 1, box, flo, fhi, false) : prim_bound(hdr, q0, src + p + 1, box, flo, fhi);  // <= 64 tiles: with the exact box test finer ones drop nothing more and cost 7 ms per upload
            skip = Pruner<EMIT>::skippable(hdr, pz, scale, nonNeg);
        }
        const bool isEntry = active && !skip;
        const uint32_t zs = (active && skip) ? scale : 0u;
        uint32_t incl = zs;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t entMask = __ballot_sync(FULL, isEntry);
        if (lane == 31) S.za[w] = incl;
        if (lane == 0) S.cnt[w] = (uint32_t)__popc(entMask);
        __syncthreads();
        uint32_t eoff = 0, zoff = 0, etot = 0, ztot = 0;
#pragma unroll
        for (int i = 0; i < kCtaWarps; i++) {
            const uint32_t c = S.cnt[i], z = S.za[i];
            if (i < w) { eoff += c; zoff += z; }
            etot += c; ztot += z;
        }
        if (isEntry) {
            CtaEntry e;
            e.h0 = q0; e.pc = p | (pz ? 0x80000000u : 0u); e.za = zaTotal + zoff + incl - zs; e.flo = flo; e.fhi = fhi;
            S.ent[eoff + (uint32_t)__popc(entMask & ((1u << lane) - 1u))] = e;
        }
        zaTotal += ztot;
        __syncthreads();
        {   // culled blend group of a union of blends (see prune_walk_warp): flagged by all threads, stepped over by thread 0
            bool zeroPair = false;
            if ((uint32_t)tid + 1u < etot) {
                const uint32_t h1 = __float_as_uint(S.ent[tid].h0.x), h2 = __float_as_uint(S.ent[tid + 1].h0.x);
                zeroPair = (S.ent[tid].pc >> 31) != 0u && PS_HDR_OPC(h1) < OPC_FIRST_NONPRIM && PS_HDR_CMB(h1) == CMB_PUSH && ps_is_plain_union(h2, colorExact);
            }
            if (zeroPair) S.ent[tid].pc |= 0x40000000u;
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t endSeen = 0;
            for (uint32_t i = 0; i < etot; i++) {
                const CtaEntry e = S.ent[i];
                const uint32_t pc = e.pc & 0x3FFFFFFFu;
                P.skipped_scale(e.za - zaPrev);
                zaPrev = e.za;
                if ((e.pc & 0x40000000u) && P.zero_union_is_noop()) {
                    zaPrev = S.ent[i + 1].za;
                    i++;
                    continue;
                }
                const uint32_t h = __float_as_uint(e.h0.x), o = PS_HDR_OPC(h);
                if (o == OPC_END) { endSeen = 1; break; }
                if (o < OPC_FIRST_NONPRIM) {
                    const bool zero = (e.pc >> 31) != 0u;
                    if (zero && PS_HDR_CMB(h) == CMB_PUSH && i + 1 < etot) {
                        // an all-zero operand whose OPC_COMBINE is the very next surviving instruction (its blend partners were
                        // all skipped): fold the pair in one step. Scale steps owed between the two belong to the zero entry: dropped
                        const uint32_t h2 = __float_as_uint(S.ent[i + 1].h0.x);
                        if (PS_HDR_OPC(h2) == OPC_COMBINE) {
                            zaPrev = S.ent[i + 1].za;
                            P.step_zero_operand(h2);
                            i++;
                            continue;
                        }
                    }
                    P.step_prim(pc, e.h0, h, zero, e.flo, e.fhi);
                } else P.step_op(pc, e.h0, h);
            }
            S.endSeen = endSeen;
#pragma unroll
            for (int a = 0; a < 3; a++) { S.box[a] = P.box.lo[a]; S.box[3 + a] = P.box.hi[a]; }
        }
        __syncthreads();
        if (S.endSeen) break;
        if (anyChanger) {
#pragma unroll
            for (int a = 0; a < 3; a++) { box.lo[a] = S.box[a]; box.hi[a] = S.box[3 + a]; }
        }
        base = (stale != NONE) ? stale : base + (uint32_t)(kCtaWarps * 32);
        if (base >= srcLen) break;  // malformed program without OPC_END
    }
    if (tid == 0) S.len = P.finish(cullMode, iso);
    __syncthreads();
    return S.len;
}

__global__ void __launch_bounds__(kCtaWarps * 32) prune_fused_cta_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                                         const float4* __restrict__ pool, const uint8_t* __restrict__ qflag,
                                                                         float4* __restrict__ scratch, uint8_t* __restrict__ scratchFlag, uint32_t stride,
                                                                         uint32_t* __restrict__ childLen, uint32_t* __restrict__ maxLen) {
    __shared__ CtaWalkSmem S;
    const int gw = blockIdx.x;
    const Region parent = parents[gw >> 6];
    int coord[3];
    Box3 box;
    uint32_t len = 0;
    if (child_region(L, parent, gw & 63, coord, box))  // uniform over the CTA
        len = prune_walk_cta<true>(pool + parent.off, qflag + parent.off, parent.len, box, L.cullMode, L.iso, scratch + (size_t)gw * stride,
                                   scratchFlag + (size_t)gw * stride, L.colorExact != 0, stride, S, L.nonNeg != 0);
    if (threadIdx.x == 0) {
        if (len > stride) len = 0xFFFFFFFFu;
        childLen[gw] = len;
        if (len) atomicMax(maxLen, len);
    }
}

__global__ void __launch_bounds__(kCoopWarps * 32) prune_compact_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                                        const float4* __restrict__ scratch, const uint8_t* __restrict__ scratchFlag,
                                                                        uint32_t stride, const uint32_t* __restrict__ childLen,
                                                                        const uint32_t* __restrict__ childOff, const uint32_t* __restrict__ childIdx,
                                                                        float4* __restrict__ outPool, uint8_t* __restrict__ outFlag, uint32_t outBase,
                                                                        Region* __restrict__ children, int32_t* __restrict__ candOfMpu) {
    const int gw = blockIdx.x * kCoopWarps + (threadIdx.x >> 5);
    if (gw >= nParents * 64) return;
    const uint32_t len = childLen[gw];
    if (len == 0) return;
    const int lane = threadIdx.x & 31;
    const uint32_t off = outBase + childOff[gw];
    const float4* src = scratch + (size_t)gw * stride;
    const uint8_t* sf = scratchFlag + (size_t)gw * stride;
    for (uint32_t i = lane; i < len; i += 32) {
        outPool[off + i] = src[i];
        if (outFlag) outFlag[off + i] = sf[i];
    }
    if (lane == 0) {
        const Region parent = parents[gw >> 6];
        int coord[3];
        Box3 box;
        child_region(L, parent, gw & 63, coord, box);
        publish_child(L, coord, off, len, childIdx[gw], children, candOfMpu, nullptr);
    }
}

__global__ void flag_kernel(const uint32_t* __restrict__ len, uint32_t* __restrict__ flag, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = len[i] ? 1u : 0u;
}

// PSGPU_CULL_NONE: every MPU of the slab is a candidate running the full program
__global__ void all_mpus_kernel(CullLevel L, Region* __restrict__ out, uint32_t fullLen, int32_t* __restrict__ candOfMpu, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int ny = L.hi[1] - L.lo[1], nz = L.hi[2] - L.lo[2];
    Region r;
    r.x = L.lo[0] + i / (ny * nz); r.y = L.lo[1] + (i / nz) % ny; r.z = L.lo[2] + i % nz;
    r.off = 0; r.len = fullLen;
    out[i] = r;
    candOfMpu[i] = i;
}

// parity hook: one thread prunes the full program against an arbitrary box (psgpu_prune_region)
__global__ void prune_one_kernel(const float4* __restrict__ prog, Box3 box, int cullMode, float iso, int colorExact, float4* __restrict__ out,
                                 uint32_t cap, uint32_t* __restrict__ lenOut) {
    const uint32_t len = prune_walk<false>(prog, box, cullMode, iso, nullptr, colorExact != 0);
    *lenOut = len;
    if (len > 0 && len <= cap) prune_walk<true>(prog, box, cullMode, iso, out, colorExact != 0, len);
}

// the whole program against ONE box, with instruction-start flags: the slab program a rank's hierarchy starts from
// (context.cu: ensure_slab_program). A plan-time cost, cached per (program, slab) - but paid again after every upload, so it is the
// CTA-cooperative walk (512 bounds per step, thread 0 steps the survivors; one pass straight into the output): 15 ms -> 2 ms for
// a rank's eighth of blob10k, which a single thread walked twice (count, emit).
__global__ void __launch_bounds__(kCtaWarps * 32) prune_slab_kernel(const float4* __restrict__ prog, const uint8_t* __restrict__ progFlag, uint32_t progLen,
                                                                    Box3 box, float iso, int colorExact, float4* __restrict__ out,
                                                                    uint8_t* __restrict__ outFlag, uint32_t cap, uint32_t* __restrict__ lenOut) {
    __shared__ CtaWalkSmem S;
    const uint32_t len = prune_walk_cta<true, true>(prog, progFlag, progLen, box, PSGPU_CULL_SUPPORT, iso, out, outFlag, colorExact != 0, cap, S);
    if (threadIdx.x == 0) *lenOut = len;
}

}  // namespace

void launch_prune_slab(const float4* prog, const uint8_t* progFlag, uint32_t progLen, const float* lo, const float* hi, float iso, int colorExact,
                       float4* out, uint8_t* outFlag, uint32_t cap, uint32_t* lenOut, cudaStream_t st) {
    Box3 b;
    for (int a = 0; a < 3; a++) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
    prune_slab_kernel<<<1, kCtaWarps * 32, 0, st>>>(prog, progFlag, progLen, b, iso, colorExact, out, outFlag, cap, lenOut);
}

void launch_prune_one(const float4* prog, const float* lo, const float* hi, int cullMode, float iso, int colorExact, float4* out, uint32_t cap,
                      uint32_t* lenOut, cudaStream_t st) {
    Box3 b;
    for (int a = 0; a < 3; a++) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
    prune_one_kernel<<<1, 1, 0, st>>>(prog, b, cullMode, iso, colorExact, out, cap, lenOut);
}

void launch_all_mpus(const CullLevel& L, Region* out, uint32_t fullLen, int32_t* candOfMpu, int n, cudaStream_t st) {
    all_mpus_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, out, fullLen, candOfMpu, n);
}

// regions per level below which a warp per region pays (few regions, long parent programs)
// A coarse level (regions of several MPUs, long parent programs) takes a warp per region up to 65 536 regions; the last level (single MPUs,
// short programs: 190 000 of them on blob1k@512^3, where a warp each measured 1.57 ms against 0.61 ms) only when it is small.
// (One threshold for both was a cliff: a rank's slab that touches THREE layers of 16^3 regions instead of two has 300 x 64 = 19 200 regions
// at the 4^3 level, fell back to a thread per region there and culled in 1.9 ms instead of 1.3 ms.)
static bool use_warp_walk(const CullLevel& L, int nParents) { return (long long)nParents * 64 <= (L.childSize == 1 ? 16384 : 65536); }

void launch_prune_count(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint8_t* qflag, uint32_t* childLen,
                        uint32_t* maxLen, cudaStream_t st) {
    const int n = nParents * 64;
    if (use_warp_walk(L, nParents))
        prune_count_warp_kernel<<<(n + kCoopWarps - 1) / kCoopWarps, kCoopWarps * 32, 0, st>>>(L, parents, nParents, pool, qflag, childLen, maxLen);
    else prune_count_kernel<<<(n + 127) / 128, 128, 0, st>>>(L, parents, nParents, pool, childLen, maxLen);
}

void launch_flags(const uint32_t* len, uint32_t* flag, int n, cudaStream_t st) { flag_kernel<<<(n + 255) / 256, 256, 0, st>>>(len, flag, n); }

void launch_prune_emit(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint8_t* qflag, const uint32_t* childLen,
                       const uint32_t* childOff, const uint32_t* childIdx, float4* outPool, uint8_t* outFlag, uint32_t outBase, Region* children,
                       int32_t* candOfMpu, uint32_t* maxLen, cudaStream_t st) {
    const int n = nParents * 64;
    if (use_warp_walk(L, nParents))
        prune_emit_warp_kernel<<<(n + kCoopWarps - 1) / kCoopWarps, kCoopWarps * 32, 0, st>>>(L, parents, nParents, pool, qflag, childLen, childOff, childIdx,
                                                                                             outPool, outFlag, outBase, children, candOfMpu, maxLen);
    else
        prune_emit_kernel<<<(n + 127) / 128, 128, 0, st>>>(L, parents, nParents, pool, childLen, childOff, childIdx, outPool, outFlag, outBase, children,
                                                           candOfMpu, maxLen);
}

bool prune_level_is_fused(const CullLevel& L, int nParents) { return use_warp_walk(L, nParents); }

void launch_prune_fused(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint8_t* qflag, float4* scratch,
                        uint8_t* scratchFlag, uint32_t stride, uint32_t* childLen, uint32_t* maxLen, bool ctaPerRegion, cudaStream_t st) {
    const int n = nParents * 64;
    if (ctaPerRegion) {
        prune_fused_cta_kernel<<<n, kCtaWarps * 32, 0, st>>>(L, parents, nParents, pool, qflag, scratch, scratchFlag, stride, childLen, maxLen);
        return;
    }
    if (!use_warp_walk(L, nParents)) {
        prune_fused_thread_kernel<<<(n + 127) / 128, 128, 0, st>>>(L, parents, nParents, pool, scratch, stride, childLen, maxLen);
        return;
    }
    prune_fused_warp_kernel<<<(n + kCoopWarps - 1) / kCoopWarps, kCoopWarps * 32, 0, st>>>(L, parents, nParents, pool, qflag, scratch, scratchFlag, stride,
                                                                                          childLen, maxLen);
}

void launch_prune_compact(const CullLevel& L, const Region* parents, int nParents, const float4* scratch, const uint8_t* scratchFlag, uint32_t stride,
                          const uint32_t* childLen, const uint32_t* childOff, const uint32_t* childIdx, float4* outPool, uint8_t* outFlag,
                          uint32_t outBase, Region* children, int32_t* candOfMpu, cudaStream_t st) {
    const int n = nParents * 64;
    prune_compact_kernel<<<(n + kCoopWarps - 1) / kCoopWarps, kCoopWarps * 32, 0, st>>>(L, parents, nParents, scratch, scratchFlag, stride, childLen, childOff,
                                                                                       childIdx, outPool, outFlag, outBase, children, candOfMpu);
}

// ---- two exclusive sums in one pass: the pair (a[i], b[i]) travels as one 64-bit value (neither total reaches 2^32: the callers
// check the totals they read back), the results are split into two arrays again on the way out. n real elements; element n, the
// closing entry whose scan output is the total, is a constant 0 of the input iterator - nothing to memset. b == nullptr: b[i] = (a[i] != 0),
// i.e. the offsets AND the compacted indices of the non-empty entries from one array of lengths.
namespace {
struct PairIn {
    const uint32_t* a;
    const uint32_t* b;
    int n;
    __host__ __device__ unsigned long long operator()(int i) const {
        if (i >= n) return 0ull;
        const uint32_t x = a[i];
        const uint32_t y = b ? b[i] : (x ? 1u : 0u);
        return (unsigned long long)x | ((unsigned long long)y << 32);
    }
};
struct PairOut {
    __host__ __device__ thrust::tuple<uint32_t, uint32_t> operator()(unsigned long long v) const {
        return thrust::make_tuple((uint32_t)v, (uint32_t)(v >> 32));
    }
};
using PairInIt = thrust::transform_iterator<PairIn, thrust::counting_iterator<int>>;
using PairOutIt = thrust::transform_output_iterator<PairOut, thrust::zip_iterator<thrust::tuple<uint32_t*, uint32_t*>>>;
}  // namespace

size_t scan_temp_bytes(int n) {
    size_t bytes = 0, pair = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, n);
    PairInIt in(thrust::counting_iterator<int>(0), PairIn{nullptr, nullptr, 0});
    PairOutIt out(thrust::make_zip_iterator(thrust::make_tuple((uint32_t*)nullptr, (uint32_t*)nullptr)), PairOut{});
    cub::DeviceScan::ExclusiveSum(nullptr, pair, in, out, n);
    return bytes > pair ? bytes : pair;
}

void exclusive_scan_pair(void* temp, size_t tempBytes, const uint32_t* a, const uint32_t* b, uint32_t* outA, uint32_t* outB, int n, cudaStream_t st) {
    PairInIt in(thrust::counting_iterator<int>(0), PairIn{a, b, n});
    PairOutIt out(thrust::make_zip_iterator(thrust::make_tuple(outA, outB)), PairOut{});
    cub::DeviceScan::ExclusiveSum(temp, tempBytes, in, out, n + 1, st);
}

void exclusive_scan_u32(void* temp, size_t tempBytes, const uint32_t* in, uint32_t* out, int n, cudaStream_t st) {
    cub::DeviceScan::ExclusiveSum(temp, tempBytes, in, out, n, st);
}

}  // namespace psgpu
