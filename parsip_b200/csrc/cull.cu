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

__global__ void __launch_bounds__(128) prune_count_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                          const float4* __restrict__ pool, uint32_t* __restrict__ childLen) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nParents * 64) return;
    const Region parent = parents[tid >> 6];
    int coord[3];
    Box3 box;
    uint32_t len = 0;
    if (child_region(L, parent, tid & 63, coord, box))
        len = prune_walk<false>(pool + parent.off, box, L.cullMode, L.iso, nullptr, L.colorExact != 0);
    childLen[tid] = len;
}

__global__ void __launch_bounds__(128) prune_emit_kernel(CullLevel L, const Region* __restrict__ parents, int nParents,
                                                         const float4* __restrict__ pool, const uint32_t* __restrict__ childLen,
                                                         const uint32_t* __restrict__ childOff, const uint32_t* __restrict__ childIdx,
                                                         float4* __restrict__ outPool, uint32_t outBase, Region* __restrict__ children,
                                                         int32_t* __restrict__ candOfMpu, uint32_t* __restrict__ maxLen) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nParents * 64) return;
    const uint32_t len = childLen[tid];
    if (len == 0) return;
    const Region parent = parents[tid >> 6];
    int coord[3];
    Box3 box;
    child_region(L, parent, tid & 63, coord, box);
    const uint32_t off = outBase + childOff[tid];
    prune_walk<true>(pool + parent.off, box, L.cullMode, L.iso, outPool + off, L.colorExact != 0, len);
    const uint32_t idx = childIdx[tid];
    Region r;
    r.x = coord[0]; r.y = coord[1]; r.z = coord[2]; r.off = off; r.len = len;
    children[idx] = r;
    if (maxLen) atomicMax(maxLen, len);
    if (candOfMpu) {
        const int nx = L.hi[0] - L.lo[0], ny = L.hi[1] - L.lo[1], nz = L.hi[2] - L.lo[2];
        (void)nx;
        candOfMpu[((size_t)(coord[0] - L.lo[0]) * ny + (coord[1] - L.lo[1])) * nz + (coord[2] - L.lo[2])] = (int32_t)idx;
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

}  // namespace

void launch_prune_one(const float4* prog, const float* lo, const float* hi, int cullMode, float iso, int colorExact, float4* out, uint32_t cap,
                      uint32_t* lenOut, cudaStream_t st) {
    Box3 b;
    for (int a = 0; a < 3; a++) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
    prune_one_kernel<<<1, 1, 0, st>>>(prog, b, cullMode, iso, colorExact, out, cap, lenOut);
}

void launch_all_mpus(const CullLevel& L, Region* out, uint32_t fullLen, int32_t* candOfMpu, int n, cudaStream_t st) {
    all_mpus_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, out, fullLen, candOfMpu, n);
}

void launch_prune_count(const CullLevel& L, const Region* parents, int nParents, const float4* pool, uint32_t* childLen, cudaStream_t st) {
    const int n = nParents * 64;
    prune_count_kernel<<<(n + 127) / 128, 128, 0, st>>>(L, parents, nParents, pool, childLen);
}

void launch_flags(const uint32_t* len, uint32_t* flag, int n, cudaStream_t st) { flag_kernel<<<(n + 255) / 256, 256, 0, st>>>(len, flag, n); }

void launch_prune_emit(const CullLevel& L, const Region* parents, int nParents, const float4* pool, const uint32_t* childLen,
                       const uint32_t* childOff, const uint32_t* childIdx, float4* outPool, uint32_t outBase, Region* children,
                       int32_t* candOfMpu, uint32_t* maxLen, cudaStream_t st) {
    const int n = nParents * 64;
    prune_emit_kernel<<<(n + 127) / 128, 128, 0, st>>>(L, parents, nParents, pool, childLen, childOff, childIdx, outPool, outBase, children,
                                                       candOfMpu, maxLen);
}

size_t scan_temp_bytes(int n) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, n);
    return bytes;
}

void exclusive_scan_u32(void* temp, size_t tempBytes, const uint32_t* in, uint32_t* out, int n, cudaStream_t st) {
    cub::DeviceScan::ExclusiveSum(temp, tempBytes, in, out, n, st);
}

}  // namespace psgpu
