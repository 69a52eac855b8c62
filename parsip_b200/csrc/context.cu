// psgpu context: device memory, the cull -> field -> scan -> emit pipeline, and the C ABI of include/psgpu.h.
// Host orchestration of what Polygonize (PS_SimdPoly/include/PS_Polygonizer.cpp:315-385) and
// CParsipOptimized::setup/run (ParsipHaptics/include/CPolyParsipOptimized.cpp:330-410) do on the CPU:
// lay out the MPU lattice, then process every MPU in parallel. There is no CPU fallback in here.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/psgpu.h"
#include "host_program.h"
#include "kernels.h"

extern "C" void psgpu_lattice_dims(float cellsize, const float* lo, const float* hi, int32_t* nm);

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    // grows geometrically; contents are NOT preserved unless keep == true
    cudaError_t reserve(size_t bytes, bool keep = false, cudaStream_t st = 0) {
        if (bytes <= cap) return cudaSuccess;
        size_t ncap = std::max(bytes, cap + cap / 2);
        ncap = (ncap + 255) & ~(size_t)255;
        void* np = nullptr;
        cudaError_t e = cudaMalloc(&np, ncap);
        if (e != cudaSuccess) return e;
        if (keep && p && cap) cudaMemcpyAsync(np, p, cap, cudaMemcpyDeviceToDevice, st);
        if (p) { cudaStreamSynchronize(st); cudaFree(p); }
        p = np; cap = ncap;
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace

struct psgpu_ctx {
    int device = 0;
    int numSMs = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copyStream = nullptr;  // device -> host mesh copies that overlap the emission kernels (psgpu_polygonize_host)
    std::string err;
    // program
    uint32_t progQuads = 0;
    int nPrims = 0;
    int semantics = PSGPU_SEM_SCALAR;
    bool hasWarps = false;  // the program uses the point stack (warps / operator matrices)
    DevBuf pool;      // float4: [full program | level 0 programs | level 1 programs | ...]
    DevBuf qflag;     // one byte per pool quad: 1 = first quad of an instruction (read by the warp-cooperative pruning passes)
    DevBuf colors;
    // cull scratch
    DevBuf regionsA, regionsB, childLen, childOff, childFlag, childIdx, scanTemp, candOfMpu, scalars, cullScratch, cullScratchFlag;
    // field outputs
    DevBuf fv, vcount, tcount, vbase, tbase, edgeMask, edgePrefix, mixedList, mixedFlag, mixedIdx, counters, statsAcc;
    // mesh
    DevBuf pos, nrm, col, tri, mpuIjk, mpuVbase, mpuTbase, vperm;
    DevBuf ptsIn, ptsF, ptsCol, ptsNrm;
    uint32_t* hostScalars = nullptr;  // pinned readback
    void* hostStage = nullptr;        // pinned upload staging (program + colours)
    size_t hostStageCap = 0;
    void* hostMesh = nullptr;         // pinned arena behind psgpu_mesh_map_host
    size_t hostMeshCap = 0;
    cudaEvent_t ev[8];
    psgpu_mesh last{};
    psgpu_stats stats{};
    psgpu::Region* cands = nullptr;
    int nCand = 0;
};

#define CK(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t _e = (call);                                                                              \
        if (_e != cudaSuccess) {                                                                              \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(_e);                                    \
            return (_e == cudaErrorMemoryAllocation) ? PSGPU_NOT_ENOUGH_MEM : PSGPU_ERROR_CUDA;               \
        }                                                                                                     \
    } while (0)

extern "C" int psgpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" int psgpu_create(int device, psgpu_ctx** out) {
    if (!out) return PSGPU_PARAM_ERROR;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return PSGPU_ERROR_NO_DEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return PSGPU_ERROR_NO_DEVICE;
    psgpu_ctx* ctx = new psgpu_ctx();
    ctx->device = device;
    cudaDeviceGetAttribute(&ctx->numSMs, cudaDevAttrMultiProcessorCount, device);
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return PSGPU_ERROR_CUDA; }
    if (cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return PSGPU_ERROR_CUDA; }
    for (auto& e : ctx->ev) cudaEventCreate(&e);
    cudaMallocHost(&ctx->hostScalars, 64 * sizeof(uint32_t));
    *out = ctx;
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_destroy(psgpu_ctx* ctx) {
    if (!ctx) return PSGPU_PARAM_ERROR;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf* all[] = {&ctx->pool, &ctx->qflag, &ctx->colors, &ctx->regionsA, &ctx->regionsB, &ctx->childLen, &ctx->childOff, &ctx->childFlag, &ctx->childIdx,
                     &ctx->scanTemp, &ctx->candOfMpu, &ctx->scalars, &ctx->cullScratch, &ctx->cullScratchFlag, &ctx->fv, &ctx->vcount, &ctx->tcount, &ctx->vbase, &ctx->tbase,
                     &ctx->edgeMask, &ctx->edgePrefix, &ctx->mixedList, &ctx->mixedFlag, &ctx->mixedIdx, &ctx->counters, &ctx->statsAcc,
                     &ctx->pos, &ctx->nrm, &ctx->col, &ctx->tri, &ctx->mpuIjk, &ctx->mpuVbase, &ctx->mpuTbase, &ctx->vperm, &ctx->ptsIn, &ctx->ptsF,
                     &ctx->ptsCol, &ctx->ptsNrm};
    for (DevBuf* b : all) b->release();
    for (auto& e : ctx->ev) cudaEventDestroy(e);
    if (ctx->hostScalars) cudaFreeHost(ctx->hostScalars);
    if (ctx->hostStage) cudaFreeHost(ctx->hostStage);
    if (ctx->hostMesh) cudaFreeHost(ctx->hostMesh);
    cudaStreamDestroy(ctx->copyStream);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return PSGPU_SUCCESS;
}

extern "C" const char* psgpu_last_error(const psgpu_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int psgpu_upload_program(psgpu_ctx* ctx, const psgpu_program* p) {
    if (!ctx || !p || p->words.empty()) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    const size_t bytes = p->words.size() * 4;
    const size_t cbytes = p->colors.size() * 4;
    if (bytes + cbytes + bytes / 16 > ctx->hostStageCap) {  // host -> device copies go through pinned memory
        if (ctx->hostStage) cudaFreeHost(ctx->hostStage);
        ctx->hostStage = nullptr;
        ctx->hostStageCap = (bytes + cbytes + bytes / 16) * 2;
        CK(cudaMallocHost(&ctx->hostStage, ctx->hostStageCap));
    }
    memcpy(ctx->hostStage, p->words.data(), bytes);
    if (cbytes) memcpy((char*)ctx->hostStage + bytes, p->colors.data(), cbytes);
    // instruction-start flags of the full program, staged behind the colours
    const size_t nQuads = p->words.size() / 4;
    uint8_t* hflag = (uint8_t*)ctx->hostStage + bytes + cbytes;
    memset(hflag, 0, nQuads);
    for (size_t pc = 0; pc < nQuads;) {
        const uint32_t hdr = p->words[pc * 4];
        hflag[pc] = 1;
        if (PS_HDR_OPC(hdr) == OPC_END) break;
        pc += (size_t)ps_instr_quads(hdr);
    }
    CK(ctx->pool.reserve(std::max(bytes * 4, (size_t)64 << 20)));
    CK(ctx->qflag.reserve(ctx->pool.cap / 16));
    CK(cudaMemcpyAsync(ctx->pool.p, ctx->hostStage, bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->qflag.p, hflag, nQuads, cudaMemcpyHostToDevice, ctx->stream));
    CK(ctx->colors.reserve(std::max(cbytes, (size_t)256)));
    if (cbytes) CK(cudaMemcpyAsync(ctx->colors.p, (char*)ctx->hostStage + bytes, cbytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->progQuads = (uint32_t)(p->words.size() / 4);
    ctx->nPrims = p->nPrims;
    ctx->semantics = p->semantics;
    ctx->hasWarps = p->maxPointStack > 0;
    return PSGPU_SUCCESS;
}

namespace {

using namespace psgpu;

int fill_grid(psgpu_ctx* ctx, const psgpu_params* P, GridParams& G) {
    if (!(P->cellsize > 0.f)) return PSGPU_PARAM_ERROR;
    int32_t nm[3];
    psgpu_lattice_dims(P->cellsize, P->bbox_lo, P->bbox_hi, nm);
    if (nm[0] <= 0 || nm[1] <= 0 || nm[2] <= 0) return PSGPU_PARAM_ERROR;
    for (int a = 0; a < 3; a++) { G.gridLo[a] = P->bbox_lo[a]; G.dims[a] = nm[a]; G.lo[a] = 0; G.hi[a] = nm[a]; }
    if (P->z_end > 0) {
        G.lo[2] = std::max(0, P->z_begin);
        G.hi[2] = std::min(nm[2], P->z_end);
    }
    G.cellsize = P->cellsize;
    G.hasWarps = ctx->hasWarps ? 1 : 0;
    G.mpuSide = P->cellsize * (float)PS_CELLS_PER_MPU;
    G.coordMode = P->coord_mode; G.sem = ctx->semantics; G.output = P->output; G.rootMode = P->root_mode;
    G.rootIters = P->root_iterations > 0 ? P->root_iterations : (P->root_mode == PSGPU_ROOT_NEWTON_REF ? 8 : 4);
    G.insideGE = (ctx->semantics == PSGPU_SEM_SIMD) ? 1 : 0;
    G.wantColors = P->want_colors; G.wantNormals = P->want_normals;
    G.forceFv = 0;
    G.iso = (P->iso > 0.f) ? P->iso : PS_ISO_VALUE;
    return PSGPU_SUCCESS;
}

// Hierarchical pruning; leaves ctx->cands / ctx->nCand (device) and returns the longest candidate program.
int run_cull(psgpu_ctx* ctx, const GridParams& G, int cullMode, bool colorExact, uint32_t& maxLen, uint32_t& poolUsed, int& launches) {
    cudaStream_t st = ctx->stream;
    const int nx = G.hi[0] - G.lo[0], ny = G.hi[1] - G.lo[1], nz = G.hi[2] - G.lo[2];
    const size_t nMpu = (size_t)nx * ny * nz;
    CK(ctx->candOfMpu.reserve(nMpu * 4));
    CullLevel L;
    for (int a = 0; a < 3; a++) { L.lo[a] = G.lo[a]; L.hi[a] = G.hi[a]; L.gridLo[a] = G.gridLo[a]; }
    L.mpuSide = G.mpuSide; L.cullMode = cullMode; L.iso = G.iso; L.childSize = 1; L.colorExact = colorExact ? 1 : 0;
    poolUsed = ctx->progQuads;
    if (cullMode == PSGPU_CULL_NONE) {
        if (nMpu > (size_t)1 << 30) { ctx->err = "lattice too large"; return PSGPU_NOT_ENOUGH_MEM; }
        CK(ctx->regionsA.reserve(nMpu * sizeof(Region)));
        launch_all_mpus(L, ctx->regionsA.as<Region>(), ctx->progQuads, ctx->candOfMpu.as<int32_t>(), (int)nMpu, st);
        launches++;
        ctx->cands = ctx->regionsA.as<Region>();
        ctx->nCand = (int)nMpu;
        maxLen = ctx->progQuads;
        return PSGPU_SUCCESS;
    }
    CK(cudaMemsetAsync(ctx->candOfMpu.p, 0xFF, nMpu * 4, st));
    const int maxDim = std::max(nx, std::max(ny, nz));
    std::vector<int> sizes;
    if (maxDim > 64) sizes = {16, 4, 1};
    else if (maxDim > 4) sizes = {4, 1};
    else sizes = {1};
    if (sizes.size() == 3 && ctx->progQuads <= 4096) {
        // A short program does not pay for the 16^3 level: its walk is bound by the serial Pruner, about as long per region as the
        // 4^3 level's own (blob1k@512^3: cull 0.72 -> 0.64 ms without it; the 17.5k-quad blob10k needs it: 1.6 vs 5.6 ms). Only while
        // the 4^3 level still gets a warp per region (<= 256 top regions of 16^3 MPUs).
        const auto cdiv = [](int a, int b) { return (a + b - 1) / b; };
        if ((long long)cdiv(nx + 15, 16) * cdiv(ny + 15, 16) * cdiv(nz + 15, 16) <= 256) sizes = {4, 1};
    }
    // a tiny program (a dozen primitives) is cheaper to test once per MPU than to refine level by level (small@128^3: 0.27 -> 0.23 ms)
    if (ctx->progQuads <= 64) sizes = {1};
    if (const char* ov = getenv("PSGPU_CULL_LEVELS")) {  // tuning / debugging aid: "16,4,1", "4,1" or "1" (each level 4x the next)
        std::vector<int> o;
        for (const char* c = ov; *c;) { o.push_back(atoi(c)); while (*c && *c != ',') c++; if (*c == ',') c++; }
        bool ok = !o.empty() && o.back() == 1;
        for (size_t i = 0; ok && i + 1 < o.size(); i++) ok = o[i] == 4 * o[i + 1];
        if (ok) sizes = o;
    }
    // top-level parents: regions of 4 * sizes[0] MPUs carrying the full program
    std::vector<Region> top;
    {
        const int ps = sizes[0] * 4;
        for (int x = G.lo[0] / ps; x * ps < G.hi[0]; x++)
            for (int y = G.lo[1] / ps; y * ps < G.hi[1]; y++)
                for (int z = G.lo[2] / ps; z * ps < G.hi[2]; z++) top.push_back(Region{x, y, z, 0u, ctx->progQuads});
    }
    DevBuf* cur = &ctx->regionsA;
    DevBuf* nxt = &ctx->regionsB;
    CK(cur->reserve(top.size() * sizeof(Region)));
    CK(cudaMemcpyAsync(cur->p, top.data(), top.size() * sizeof(Region), cudaMemcpyHostToDevice, st));
    int nParents = (int)top.size();
    CK(ctx->scalars.reserve(64));
    maxLen = 0;
    uint32_t parentMaxLen = ctx->progQuads;  // longest program among the current level's parents
    for (size_t lv = 0; lv < sizes.size(); lv++) {
        L.childSize = sizes[lv];
        const bool last = (lv + 1 == sizes.size());
        const int n = nParents * 64;
        CK(ctx->childLen.reserve((size_t)(n + 1) * 4)); CK(ctx->childOff.reserve((size_t)(n + 1) * 4));
        CK(ctx->childFlag.reserve((size_t)(n + 1) * 4)); CK(ctx->childIdx.reserve((size_t)(n + 1) * 4));
        const size_t tb = scan_temp_bytes(n + 1);
        CK(ctx->scanTemp.reserve(tb));
        CK(cudaMemsetAsync(ctx->childLen.as<uint32_t>() + n, 0, 4, st));
        // coarse levels (few regions, long programs): ONE warp-cooperative walk per region into scratch slots, compacted after the scans;
        // the scratch bound keeps 10k-primitive trees within ~1 GB
        const size_t scratchQuads = (size_t)n * parentMaxLen;
        const bool fused = prune_level_is_fused(nParents) && scratchQuads <= ((size_t)64 << 20) && !getenv("PSGPU_CULL_TWO_PASS");
        uint32_t* levelMax = ctx->scalars.as<uint32_t>() + 1 + lv;
        if (fused) {
            CK(ctx->cullScratch.reserve(scratchQuads * 16)); CK(ctx->cullScratchFlag.reserve(scratchQuads));
            CK(cudaMemsetAsync(levelMax, 0, 4, st));
            // few regions with a long program (the top level): a whole CTA per region instead of a warp
            const char* ctaMin = getenv("PSGPU_CULL_CTA_MINLEN");
            const bool ctaPerRegion = n <= 4096 && parentMaxLen >= (uint32_t)(ctaMin ? atoi(ctaMin) : 1024);
            launch_prune_fused(L, cur->as<Region>(), nParents, ctx->pool.as<float4>(), ctx->qflag.as<uint8_t>(), ctx->cullScratch.as<float4>(),
                               ctx->cullScratchFlag.as<uint8_t>(), parentMaxLen, ctx->childLen.as<uint32_t>(), levelMax, ctaPerRegion, st);
        } else
            launch_prune_count(L, cur->as<Region>(), nParents, ctx->pool.as<float4>(), ctx->qflag.as<uint8_t>(), ctx->childLen.as<uint32_t>(), st);
        launch_flags(ctx->childLen.as<uint32_t>(), ctx->childFlag.as<uint32_t>(), n + 1, st);
        exclusive_scan_u32(ctx->scanTemp.p, tb, ctx->childLen.as<uint32_t>(), ctx->childOff.as<uint32_t>(), n + 1, st);
        exclusive_scan_u32(ctx->scanTemp.p, tb, ctx->childFlag.as<uint32_t>(), ctx->childIdx.as<uint32_t>(), n + 1, st);
        launches += 4;
        CK(cudaMemcpyAsync(&ctx->hostScalars[0], ctx->childOff.as<uint32_t>() + n, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&ctx->hostScalars[1], ctx->childIdx.as<uint32_t>() + n, 4, cudaMemcpyDeviceToHost, st));
        if (fused) CK(cudaMemcpyAsync(&ctx->hostScalars[3], levelMax, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        const uint32_t totalQuads = ctx->hostScalars[0], nChildren = ctx->hostScalars[1];
        if (fused && ctx->hostScalars[3] > parentMaxLen) { ctx->err = "pruned program longer than its parent"; return PSGPU_ERROR_CUDA; }
        if ((uint64_t)poolUsed + totalQuads >= ((uint64_t)1 << 32)) { ctx->err = "pruned programs exceed 2^32 quads"; return PSGPU_NOT_ENOUGH_MEM; }
        CK(ctx->pool.reserve(((size_t)poolUsed + totalQuads + 16) * 16, true, st));
        CK(ctx->qflag.reserve((size_t)poolUsed + totalQuads + 16, true, st));
        CK(nxt->reserve(std::max((size_t)nChildren, (size_t)1) * sizeof(Region)));
        if (last) CK(cudaMemsetAsync(ctx->scalars.p, 0, 4, st));
        if (nChildren > 0 && fused) {
            launch_prune_compact(L, cur->as<Region>(), nParents, ctx->cullScratch.as<float4>(), ctx->cullScratchFlag.as<uint8_t>(), parentMaxLen,
                                 ctx->childLen.as<uint32_t>(), ctx->childOff.as<uint32_t>(), ctx->childIdx.as<uint32_t>(), ctx->pool.as<float4>(),
                                 last ? nullptr : ctx->qflag.as<uint8_t>(), poolUsed, nxt->as<Region>(), last ? ctx->candOfMpu.as<int32_t>() : nullptr, st);
            if (last) CK(cudaMemcpyAsync(ctx->scalars.p, levelMax, 4, cudaMemcpyDeviceToDevice, st));
            launches++;
            parentMaxLen = ctx->hostScalars[3];
        } else if (nChildren > 0) {
            parentMaxLen = std::min(parentMaxLen, totalQuads);  // a bound: no per-level maximum is tracked on the two-pass path
            launch_prune_emit(L, cur->as<Region>(), nParents, ctx->pool.as<float4>(), ctx->qflag.as<uint8_t>(), ctx->childLen.as<uint32_t>(),
                              ctx->childOff.as<uint32_t>(), ctx->childIdx.as<uint32_t>(), ctx->pool.as<float4>(),
                              last ? nullptr : ctx->qflag.as<uint8_t>(), poolUsed, nxt->as<Region>(),
                              last ? ctx->candOfMpu.as<int32_t>() : nullptr, last ? ctx->scalars.as<uint32_t>() : nullptr, st);
            launches++;
        }
        poolUsed += totalQuads;
        std::swap(cur, nxt);
        nParents = (int)nChildren;
        if (nParents == 0) break;
    }
    ctx->cands = cur->as<Region>();
    ctx->nCand = nParents;
    if (nParents > 0) {
        CK(cudaMemcpyAsync(&ctx->hostScalars[2], ctx->scalars.p, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        maxLen = ctx->hostScalars[2];
    }
    return PSGPU_SUCCESS;
}

// caller-provided HOST buffers the mesh is streamed into while the emission kernels still run
struct HostSink {
    float *pos, *nrm, *col;
    uint32_t* tri;
    int64_t capV, capT;
};

int polygonize_impl(psgpu_ctx* ctx, const psgpu_params* P, psgpu_mesh* out, bool forceFv, const HostSink* sink = nullptr) {
    if (!ctx || !P || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    GridParams G;
    int rc = fill_grid(ctx, P, G);
    if (rc != PSGPU_SUCCESS) { ctx->err = "bad grid parameters"; return rc; }
    G.forceFv = forceFv ? 1 : 0;
    if (P->output == PSGPU_OUT_INDEXED && P->coord_mode != PSGPU_COORD_GLOBAL) {
        ctx->err = "PSGPU_OUT_INDEXED needs PSGPU_COORD_GLOBAL (shared corners must be bit-identical across MPUs)";
        return PSGPU_PARAM_ERROR;
    }
    int launches = 0;
    psgpu_stats& S = ctx->stats;
    S = psgpu_stats{};
    S.mpus_total = (int64_t)(G.hi[0] - G.lo[0]) * (G.hi[1] - G.lo[1]) * (G.hi[2] - G.lo[2]);
    S.program_quads = ctx->progQuads;
    CK(cudaEventRecord(ctx->ev[0], st));
    uint32_t maxLen = 0, poolUsed = 0;
    rc = run_cull(ctx, G, P->cull, ctx->semantics == PSGPU_SEM_SIMD && P->want_colors, maxLen, poolUsed, launches);
    if (rc != PSGPU_SUCCESS) return rc;
    CK(cudaEventRecord(ctx->ev[1], st));
    const int nCand = ctx->nCand;
    S.mpus_evaluated = nCand;
    S.pruned_quads = (int64_t)poolUsed - ctx->progQuads;

    psgpu_mesh M{};
    for (int a = 0; a < 3; a++) M.lattice[a] = G.dims[a];
    uint32_t nV = 0, nT = 0, nMixed = 0, nOver = 0;
    bool streamed = false;
    if (nCand > 0) {
        const size_t n1 = (size_t)nCand + 1;
        CK(ctx->fv.reserve((size_t)nCand * 2048));
        CK(ctx->vcount.reserve(n1 * 4)); CK(ctx->tcount.reserve(n1 * 4)); CK(ctx->vbase.reserve(n1 * 4)); CK(ctx->tbase.reserve(n1 * 4));
        CK(ctx->edgeMask.reserve((size_t)nCand * 192)); CK(ctx->edgePrefix.reserve((size_t)nCand * 96));
        CK(ctx->mixedList.reserve((size_t)nCand * 4));
        CK(ctx->counters.reserve(64));
        CK(cudaMemsetAsync(ctx->counters.p, 0, 64, st));
        CK(cudaMemsetAsync(ctx->vcount.as<uint32_t>() + nCand, 0, 4, st));
        CK(cudaMemsetAsync(ctx->tcount.as<uint32_t>() + nCand, 0, 4, st));
        FieldOutputs fo;
        fo.fv = ctx->fv.as<float4>(); fo.vcount = ctx->vcount.as<uint32_t>(); fo.tcount = ctx->tcount.as<uint32_t>();
        fo.edgeMask = ctx->edgeMask.as<uint32_t>(); fo.edgePrefix = ctx->edgePrefix.as<uint16_t>();
        fo.mixedList = ctx->mixedList.as<uint32_t>(); fo.counters = ctx->counters.as<uint32_t>();
        launch_field_kernel(G, ctx->cands, nCand, ctx->pool.as<float4>(), fo, maxLen, ctx->hasWarps, ctx->numSMs, st);
        launches++;
        CK(cudaEventRecord(ctx->ev[2], st));
        const size_t tb = scan_temp_bytes((int)n1);
        CK(ctx->scanTemp.reserve(tb));
        exclusive_scan_u32(ctx->scanTemp.p, tb, fo.vcount, ctx->vbase.as<uint32_t>(), (int)n1, st);
        exclusive_scan_u32(ctx->scanTemp.p, tb, fo.tcount, ctx->tbase.as<uint32_t>(), (int)n1, st);
        launches += 2;
        CK(cudaMemcpyAsync(&ctx->hostScalars[4], ctx->vbase.as<uint32_t>() + nCand, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&ctx->hostScalars[5], ctx->tbase.as<uint32_t>() + nCand, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&ctx->hostScalars[6], ctx->counters.as<uint32_t>(), 16, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        nV = ctx->hostScalars[4]; nT = ctx->hostScalars[5]; nMixed = ctx->hostScalars[7]; nOver = ctx->hostScalars[8];
        MeshOutputs mo{};
        CK(ctx->pos.reserve(std::max((size_t)nV, (size_t)1) * 12)); CK(ctx->nrm.reserve(std::max((size_t)nV, (size_t)1) * 12));
        CK(ctx->col.reserve(std::max((size_t)nV, (size_t)1) * 12)); CK(ctx->tri.reserve(std::max((size_t)nT, (size_t)1) * 12));
        mo.pos = ctx->pos.as<float>(); mo.nrm = ctx->nrm.as<float>(); mo.col = ctx->col.as<float>(); mo.tri = ctx->tri.as<uint32_t>();
        mo.vbase = ctx->vbase.as<uint32_t>(); mo.tbase = ctx->tbase.as<uint32_t>();
        if (P->output == PSGPU_OUT_MPU) {
            CK(ctx->mixedFlag.reserve(n1 * 4)); CK(ctx->mixedIdx.reserve(n1 * 4));
            launch_flags(fo.tcount, ctx->mixedFlag.as<uint32_t>(), (int)n1, st);
            exclusive_scan_u32(ctx->scanTemp.p, tb, ctx->mixedFlag.as<uint32_t>(), ctx->mixedIdx.as<uint32_t>(), (int)n1, st);
            launches += 2;
            CK(ctx->mpuIjk.reserve(((size_t)nMixed + 1) * 12)); CK(ctx->mpuVbase.reserve(((size_t)nMixed + 1) * 4));
            CK(ctx->mpuTbase.reserve(((size_t)nMixed + 1) * 4));
            mo.mixedIdx = ctx->mixedIdx.as<uint32_t>(); mo.mpuIjk = ctx->mpuIjk.as<int32_t>();
            mo.mpuVbase = ctx->mpuVbase.as<uint32_t>(); mo.mpuTbase = ctx->mpuTbase.as<uint32_t>();
            CK(ctx->vperm.reserve(std::max((size_t)nV, (size_t)1) * 2));
            mo.vperm = ctx->vperm.as<uint16_t>();
            // closing entries of the per-MPU base tables
            CK(cudaMemcpyAsync(ctx->mpuVbase.as<uint32_t>() + nMixed, ctx->vbase.as<uint32_t>() + nCand, 4, cudaMemcpyDeviceToDevice, st));
            CK(cudaMemcpyAsync(ctx->mpuTbase.as<uint32_t>() + nMixed, ctx->tbase.as<uint32_t>() + nCand, 4, cudaMemcpyDeviceToDevice, st));
        }
        if (sink && ((int64_t)nV > sink->capV || (int64_t)nT > sink->capT)) {
            ctx->err = "host buffers too small";
            M.n_vertices = nV; M.n_triangles = nT;
            if (out) *out = M;
            return PSGPU_NOT_ENOUGH_MEM;
        }
        CK(cudaEventRecord(ctx->ev[3], st));
        if (nMixed > 0) {
            launch_emit_kernels(G, ctx->cands, ctx->pool.as<float4>(), ctx->colors.as<float>(), fo, mo, ctx->candOfMpu.as<int32_t>(), (int)nMixed,
                                maxLen, ctx->numSMs, st, ctx->ev[5], ctx->ev[6]);
            launches += 2 + (G.wantColors ? 1 : 0);  // tri_kernel, vertex_kernel, color_kernel
        }
        CK(cudaEventRecord(ctx->ev[4], st));
        if (sink && nMixed > 0 && P->output == PSGPU_OUT_INDEXED) {
            // each array leaves for the host as soon as the kernel that writes it is done: triangles behind tri_kernel (while
            // vertex_kernel runs), positions + normals behind vertex_kernel (while color_kernel runs), colours last
            cudaStream_t cs = ctx->copyStream;
            CK(cudaStreamWaitEvent(cs, ctx->ev[5], 0));
            if (sink->tri && nT) CK(cudaMemcpyAsync(sink->tri, ctx->tri.p, (size_t)nT * 12, cudaMemcpyDeviceToHost, cs));
            CK(cudaStreamWaitEvent(cs, ctx->ev[6], 0));
            if (sink->pos && nV) CK(cudaMemcpyAsync(sink->pos, ctx->pos.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            if (sink->nrm && nV && P->want_normals) CK(cudaMemcpyAsync(sink->nrm, ctx->nrm.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            CK(cudaStreamWaitEvent(cs, ctx->ev[4], 0));
            if (sink->col && nV && P->want_colors) CK(cudaMemcpyAsync(sink->col, ctx->col.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            streamed = true;
        }
        if (P->collect_stats) {
            CK(ctx->statsAcc.reserve(64));
            CK(cudaMemsetAsync(ctx->statsAcc.p, 0, 64, st));
            launch_program_stats(ctx->cands, nCand, ctx->pool.as<float4>(), fo.vcount, ctx->statsAcc.as<unsigned long long>(), st);
            launches++;
            unsigned long long acc[5];
            CK(cudaMemcpyAsync(acc, ctx->statsAcc.p, 40, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            S.algorithmic_flops = (int64_t)acc[0]; S.prim_evals = (int64_t)acc[1]; S.vertex_eval_flops = (int64_t)acc[4];
        }
    } else {
        CK(cudaEventRecord(ctx->ev[2], st)); CK(cudaEventRecord(ctx->ev[3], st)); CK(cudaEventRecord(ctx->ev[4], st));
    }
    CK(cudaStreamSynchronize(st));
    if (streamed) CK(cudaStreamSynchronize(ctx->copyStream));
    {
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { ctx->err = std::string("kernel failure: ") + cudaGetErrorString(e); return PSGPU_ERROR_CUDA; }
    }
    cudaEventElapsedTime(&S.ms_cull, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&S.ms_field, ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&S.ms_scan, ctx->ev[2], ctx->ev[3]);
    cudaEventElapsedTime(&S.ms_emit, ctx->ev[3], ctx->ev[4]);
    cudaEventElapsedTime(&S.ms_total, ctx->ev[0], ctx->ev[4]);
    if (nMixed > 0) {
        cudaEventElapsedTime(&S.ms_tri, ctx->ev[3], ctx->ev[5]);
        cudaEventElapsedTime(&S.ms_vertex, ctx->ev[5], ctx->ev[6]);
        cudaEventElapsedTime(&S.ms_color, ctx->ev[6], ctx->ev[4]);
    }
    S.mpus_crossed = nMixed;
    const int evalsPerVertex = (P->root_mode == PSGPU_ROOT_FOURSAMPLE) ? 8 : 14;
    S.field_evals = (int64_t)nCand * 512 + (int64_t)nV * evalsPerVertex;
    S.gpu_launches = launches;
    M.n_vertices = nV; M.n_triangles = nT;
    M.pos = ctx->pos.as<float>(); M.nrm = ctx->nrm.as<float>(); M.col = ctx->col.as<float>(); M.tri = ctx->tri.as<uint32_t>();
    M.n_mpus_out = (P->output == PSGPU_OUT_MPU) ? nMixed : 0;
    M.mpu_ijk = ctx->mpuIjk.as<int32_t>(); M.mpu_vbase = ctx->mpuVbase.as<uint32_t>(); M.mpu_tbase = ctx->mpuTbase.as<uint32_t>();
    M.over_capacity = (int32_t)nOver;
    ctx->last = M;
    if (out) *out = M;
    if (P->output == PSGPU_OUT_MPU && nOver > 0) {
        ctx->err = "MPU mesh exceeds the reference's 512 vertex/triangle capacity (SURVEY Q10)";
        return PSGPU_NOT_ENOUGH_MEM;
    }
    if (sink && !streamed && (nV || nT)) return PSGPU_SUCCESS + 1;  // nothing was copied yet (per-MPU output): caller downloads
    return PSGPU_SUCCESS;
}

}  // namespace

extern "C" int psgpu_polygonize(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out) { return polygonize_impl(ctx, params, out, false); }

extern "C" int psgpu_mesh_download(psgpu_ctx* ctx, float* pos, float* nrm, float* col, uint32_t* tri, int32_t* mpu_ijk, uint32_t* mpu_vbase,
                                   uint32_t* mpu_tbase) {
    if (!ctx) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    const psgpu_mesh& M = ctx->last;
    cudaStream_t st = ctx->stream;
    if (pos && M.n_vertices) CK(cudaMemcpyAsync(pos, M.pos, (size_t)M.n_vertices * 12, cudaMemcpyDeviceToHost, st));
    if (nrm && M.n_vertices) CK(cudaMemcpyAsync(nrm, M.nrm, (size_t)M.n_vertices * 12, cudaMemcpyDeviceToHost, st));
    if (col && M.n_vertices) CK(cudaMemcpyAsync(col, M.col, (size_t)M.n_vertices * 12, cudaMemcpyDeviceToHost, st));
    if (tri && M.n_triangles) CK(cudaMemcpyAsync(tri, M.tri, (size_t)M.n_triangles * 12, cudaMemcpyDeviceToHost, st));
    if (M.n_mpus_out) {
        if (mpu_ijk) CK(cudaMemcpyAsync(mpu_ijk, M.mpu_ijk, (size_t)M.n_mpus_out * 12, cudaMemcpyDeviceToHost, st));
        if (mpu_vbase) CK(cudaMemcpyAsync(mpu_vbase, M.mpu_vbase, ((size_t)M.n_mpus_out + 1) * 4, cudaMemcpyDeviceToHost, st));
        if (mpu_tbase) CK(cudaMemcpyAsync(mpu_tbase, M.mpu_tbase, ((size_t)M.n_mpus_out + 1) * 4, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    return PSGPU_SUCCESS;
}

// The last mesh in a context-owned PINNED arena: one asynchronous copy per array at full link rate (a copy into pageable
// caller memory is staged by the driver chunk by chunk), pointers valid until the next polygonize / map on this context.
extern "C" int psgpu_mesh_map_host(psgpu_ctx* ctx, const float** pos, const float** nrm, const float** col, const uint32_t** tri,
                                   const int32_t** mpu_ijk, const uint32_t** mpu_vbase, const uint32_t** mpu_tbase) {
    if (!ctx) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    const psgpu_mesh& M = ctx->last;
    cudaStream_t st = ctx->stream;
    auto pad = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t vB = pad((size_t)M.n_vertices * 12), tB = pad((size_t)M.n_triangles * 12);
    const size_t iB = pad((size_t)M.n_mpus_out * 12), bB = pad(((size_t)M.n_mpus_out + 1) * 4);
    const size_t need = (pos ? vB : 0) + (nrm ? vB : 0) + (col ? vB : 0) + (tri ? tB : 0) + (mpu_ijk ? iB : 0) + (mpu_vbase ? bB : 0) + (mpu_tbase ? bB : 0) + 256;
    if (need > ctx->hostMeshCap) {
        if (ctx->hostMesh) cudaFreeHost(ctx->hostMesh);
        ctx->hostMesh = nullptr;
        ctx->hostMeshCap = 0;
        const size_t cap = need + need / 4;
        CK(cudaMallocHost(&ctx->hostMesh, cap));
        ctx->hostMeshCap = cap;
    }
    char* cur = (char*)ctx->hostMesh;
    auto take = [&](const void* dev, size_t bytes, size_t padded) -> const void* {
        char* dst = cur;
        cur += padded;
        if (dev && bytes) cudaMemcpyAsync(dst, dev, bytes, cudaMemcpyDeviceToHost, st);
        return dst;
    };
    if (pos) *pos = (const float*)take(M.pos, (size_t)M.n_vertices * 12, vB);
    if (nrm) *nrm = (const float*)take(M.nrm, (size_t)M.n_vertices * 12, vB);
    if (col) *col = (const float*)take(M.col, (size_t)M.n_vertices * 12, vB);
    if (tri) *tri = (const uint32_t*)take(M.tri, (size_t)M.n_triangles * 12, tB);
    const bool tables = M.n_mpus_out > 0;
    if (mpu_ijk) *mpu_ijk = (const int32_t*)take(tables ? M.mpu_ijk : nullptr, (size_t)M.n_mpus_out * 12, iB);
    if (mpu_vbase) *mpu_vbase = (const uint32_t*)take(tables ? M.mpu_vbase : nullptr, ((size_t)M.n_mpus_out + 1) * 4, bB);
    if (mpu_tbase) *mpu_tbase = (const uint32_t*)take(tables ? M.mpu_tbase : nullptr, ((size_t)M.n_mpus_out + 1) * 4, bB);
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_polygonize_host(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, float* pos, float* nrm, float* col,
                                     uint32_t* tri, int64_t cap_vertices, int64_t cap_triangles) {
    psgpu_mesh M{};
    const HostSink sink{pos, nrm, col, tri, cap_vertices, cap_triangles};
    int rc = polygonize_impl(ctx, params, &M, false, &sink);
    if (out) *out = M;
    if (rc == PSGPU_SUCCESS + 1) return psgpu_mesh_download(ctx, pos, nrm, col, tri, nullptr, nullptr, nullptr);
    return rc;
}

extern "C" int psgpu_field_eval(psgpu_ctx* ctx, const float* xyz, int64_t n, float* f_out) {
    if (!ctx || !xyz || !f_out || n < 0 || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    if (n == 0) return PSGPU_SUCCESS;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(ctx->ptsIn.reserve((size_t)n * 12)); CK(ctx->ptsF.reserve((size_t)n * 4));
    CK(cudaMemcpyAsync(ctx->ptsIn.p, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    psgpu::launch_points_field(ctx->pool.as<float4>(), ctx->ptsIn.as<float>(), n, ctx->ptsF.as<float>(), ctx->semantics, st);
    CK(cudaMemcpyAsync(f_out, ctx->ptsF.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_field_color_normal(psgpu_ctx* ctx, const float* xyz, int64_t n, float* f, float* col, float* nrm) {
    if (!ctx || !xyz || !f || !col || !nrm || n < 0 || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    if (n == 0) return PSGPU_SUCCESS;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(ctx->ptsIn.reserve((size_t)n * 12)); CK(ctx->ptsF.reserve((size_t)n * 4));
    CK(ctx->ptsCol.reserve((size_t)n * 12)); CK(ctx->ptsNrm.reserve((size_t)n * 12));
    CK(cudaMemcpyAsync(ctx->ptsIn.p, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    psgpu::launch_points_field_color_normal(ctx->pool.as<float4>(), ctx->colors.as<float>(), ctx->ptsIn.as<float>(), n, ctx->ptsF.as<float>(),
                                            ctx->ptsCol.as<float>(), ctx->ptsNrm.as<float>(), ctx->semantics, st);
    CK(cudaMemcpyAsync(f, ctx->ptsF.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(col, ctx->ptsCol.p, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(nrm, ctx->ptsNrm.p, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_mpu_cache(psgpu_ctx* ctx, const psgpu_params* params, const int32_t* ijk, float* fv512, uint8_t* case343) {
    if (!ctx || !params || !ijk || !fv512 || !case343) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    psgpu::GridParams G;
    int rc = fill_grid(ctx, params, G);
    if (rc != PSGPU_SUCCESS) return rc;
    G.forceFv = 1;
    G.output = PSGPU_OUT_MPU;
    psgpu::Region r{ijk[0], ijk[1], ijk[2], 0u, ctx->progQuads};
    CK(ctx->regionsA.reserve(sizeof(r)));
    CK(cudaMemcpyAsync(ctx->regionsA.p, &r, sizeof(r), cudaMemcpyHostToDevice, st));
    CK(ctx->fv.reserve(2048)); CK(ctx->vcount.reserve(8)); CK(ctx->tcount.reserve(8));
    CK(ctx->edgeMask.reserve(192)); CK(ctx->edgePrefix.reserve(96)); CK(ctx->mixedList.reserve(4)); CK(ctx->counters.reserve(64));
    CK(cudaMemsetAsync(ctx->counters.p, 0, 64, st));
    psgpu::FieldOutputs fo;
    fo.fv = ctx->fv.as<float4>(); fo.vcount = ctx->vcount.as<uint32_t>(); fo.tcount = ctx->tcount.as<uint32_t>();
    fo.edgeMask = ctx->edgeMask.as<uint32_t>(); fo.edgePrefix = ctx->edgePrefix.as<uint16_t>();
    fo.mixedList = ctx->mixedList.as<uint32_t>(); fo.counters = ctx->counters.as<uint32_t>();
    psgpu::launch_field_kernel(G, ctx->regionsA.as<psgpu::Region>(), 1, ctx->pool.as<float4>(), fo, ctx->progQuads, ctx->hasWarps, ctx->numSMs, st);
    CK(cudaMemcpyAsync(fv512, ctx->fv.p, 2048, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    for (int i = 0; i < 7; i++) for (int j = 0; j < 7; j++) for (int k = 0; k < 7; k++) {
        int cfg = 0;
        for (int b = 0; b < 8; b++) {
            const float f = fv512[(i + ((b >> 2) & 1)) * 8 + (j + ((b >> 1) & 1)) * 64 + (k + (b & 1))];
            const bool in = G.insideGE ? (f >= G.iso) : (f > G.iso);
            if (in) cfg |= 1 << b;
        }
        case343[i * 49 + j * 7 + k] = (uint8_t)cfg;
    }
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_mpu_caches_culled(psgpu_ctx* ctx, const psgpu_params* params, int32_t* ijk_out, float* fv_out, int64_t cap_mpus, int64_t* n_out) {
    if (!ctx || !params || !n_out) return PSGPU_PARAM_ERROR;
    int rc = polygonize_impl(ctx, params, nullptr, true);
    if (rc != PSGPU_SUCCESS && rc != PSGPU_NOT_ENOUGH_MEM) return rc;
    *n_out = ctx->nCand;
    if (ctx->nCand > cap_mpus || !ijk_out || !fv_out) return PSGPU_SUCCESS;
    std::vector<psgpu::Region> regs((size_t)ctx->nCand);
    if (ctx->nCand > 0) {
        CK(cudaMemcpyAsync(regs.data(), ctx->cands, regs.size() * sizeof(psgpu::Region), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaMemcpyAsync(fv_out, ctx->fv.p, (size_t)ctx->nCand * 2048, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    for (int i = 0; i < ctx->nCand; i++) { ijk_out[i * 3] = regs[i].x; ijk_out[i * 3 + 1] = regs[i].y; ijk_out[i * 3 + 2] = regs[i].z; }
    return PSGPU_SUCCESS;
}

// debugging aid (not in psgpu.h): pruned program of candidate `idx` of the last polygonize
extern "C" int psgpu_debug_candidate_program(psgpu_ctx* ctx, int idx, uint32_t* quads_out, uint32_t cap, uint32_t* len_out) {
    if (!ctx || idx < 0 || idx >= ctx->nCand) return PSGPU_PARAM_ERROR;
    psgpu::Region r;
    CK(cudaMemcpy(&r, ctx->cands + idx, sizeof(r), cudaMemcpyDeviceToHost));
    *len_out = r.len;
    if (r.len <= cap) CK(cudaMemcpy(quads_out, ctx->pool.as<float4>() + r.off, (size_t)r.len * 16, cudaMemcpyDeviceToHost));
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_prune_region(psgpu_ctx* ctx, const float* lo3, const float* hi3, int cull, float iso, int color_exact, uint32_t* quads_out,
                                  uint32_t cap_quads, uint32_t* len_out) {
    if (!ctx || !lo3 || !hi3 || !quads_out || !len_out || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CK(ctx->ptsIn.reserve((size_t)cap_quads * 16 + 16));
    CK(ctx->scalars.reserve(64));
    psgpu::launch_prune_one(ctx->pool.as<float4>(), lo3, hi3, cull, iso > 0.f ? iso : PS_ISO_VALUE, color_exact, ctx->ptsIn.as<float4>(), cap_quads,
                            ctx->scalars.as<uint32_t>(), st);
    CK(cudaMemcpyAsync(&ctx->hostScalars[10], ctx->scalars.p, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    *len_out = ctx->hostScalars[10];
    if (*len_out > 0 && *len_out <= cap_quads) {
        CK(cudaMemcpyAsync(quads_out, ctx->ptsIn.p, (size_t)*len_out * 16, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_get_stats(const psgpu_ctx* ctx, psgpu_stats* out) {
    if (!ctx || !out) return PSGPU_PARAM_ERROR;
    *out = ctx->stats;
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_measure_fp32_peak(psgpu_ctx* ctx, float* tflops_out, float* sm_clock_mhz_out) {
    if (!ctx || !tflops_out) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    float ms = 0.f;
    *tflops_out = psgpu::run_fp32_peak(ctx->numSMs, ctx->stream, &ms);
    if (sm_clock_mhz_out) {
        int khz = 0;
        cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device);
        *sm_clock_mhz_out = khz / 1000.0f;
    }
    CK(cudaGetLastError());
    return PSGPU_SUCCESS;
}
