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
#include <mutex>
#include <string>
#include <vector>

#include "../../include/psgpu.h"
#include "host_program.h"
#include "nccl_dl.h"
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
        if (keep && p && cap) {
            e = cudaMemcpyAsync(np, p, cap, cudaMemcpyDeviceToDevice, st);
            if (e != cudaSuccess) { cudaFree(np); return e; }
        }
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
    cudaStream_t commStream = nullptr;  // NCCL exchange of a stitched run, beside the vertex / colour kernels (highest priority)
    std::string err;
    // program
    uint32_t progQuads = 0;
    int nPrims = 0;
    int semantics = PSGPU_SEM_SCALAR;
    bool hasWarps = false;  // the program uses the point stack (warps / operator matrices)
    bool nonNeg = false;    // no difference operator anywhere in the program: every value on the stack is >= 0
    bool hasMasks = false;  // the program contains OPC_BOXMASK (linearized with PSGPU_LIN_AS_SHIPPED)
    DevBuf pool;      // float4: [full program | level 0 programs | level 1 programs | ...]
    DevBuf qflag;     // one byte per pool quad: 1 = first quad of an instruction (read by the warp-cooperative pruning passes)
    DevBuf colors;
    // cull scratch
    DevBuf regionsA, regionsB, childLen, childOff, childIdx, scanTemp, candOfMpu, scalars, cullScratch, cullScratchFlag;
    // field outputs
    DevBuf fv, inside, vcount, tcount, vbase, tbase, edgeMask, edgePrefix, mixedList, mixedFlag, mixedIdx, counters, statsAcc;
    // mesh
    DevBuf pos, nrm, col, tri, mpuIjk, mpuVbase, mpuTbase, vperm;
    DevBuf ptsIn, ptsF, ptsCol, ptsNrm;
    uint32_t* hostScalars = nullptr;  // pinned readback; [20 ..) = the ranks' {vertices, triangles} of a stitched run
    static constexpr int kHostScalars = 1024;
    void* hostStage = nullptr;        // pinned upload staging (program + colours)
    size_t hostStageCap = 0;
    void* hostMesh = nullptr;         // pinned arena behind psgpu_mesh_map_host
    size_t hostMeshCap = 0;
    cudaEvent_t ev[10];
    psgpu_mesh last{};
    psgpu_stats stats{};
    psgpu::Region* cands = nullptr;
    int nCand = 0;
    // multi-GPU (psgpu_comm_init): one context = one rank
    void* comm = nullptr;
    int nranks = 1, rank = 0;
    DevBuf shardCounts, haloSend, haloRecv;
    std::vector<uint64_t> rankCounts;  // {vertices, triangles} of every rank after the last sharded polygonize
    // slab program: the uploaded program pruned once against this rank's slab box, cached behind the full program in the pool
    struct SlabProg { bool valid = false; int zb = 0, ze = 0; float lo[3] = {0, 0, 0}, side = 0.f; int nx = 0, ny = 0; int colorExact = 0; uint32_t len = 0; } slab;
    uint32_t rootOff = 0, rootLen = 0;  // program the cull hierarchy starts from in the current polygonize
    int polysSinceUpload = 0;           // polygonize calls since the last psgpu_upload_program (ensure_slab_program)
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
    {
        int prLo = 0, prHi = 0;
        cudaDeviceGetStreamPriorityRange(&prLo, &prHi);
        if (cudaStreamCreateWithPriority(&ctx->commStream, cudaStreamNonBlocking, prHi) != cudaSuccess) { delete ctx; return PSGPU_ERROR_CUDA; }
    }
    bool ok = true;
    for (auto& e : ctx->ev) { e = nullptr; ok = ok && cudaEventCreate(&e) == cudaSuccess; }
    ok = ok && cudaMallocHost(&ctx->hostScalars, psgpu_ctx::kHostScalars * sizeof(uint32_t)) == cudaSuccess;
    if (!ok) { psgpu_destroy(ctx); return PSGPU_ERROR_CUDA; }
    *out = ctx;
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_destroy(psgpu_ctx* ctx) {
    if (!ctx) return PSGPU_PARAM_ERROR;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf* all[] = {&ctx->pool, &ctx->qflag, &ctx->colors, &ctx->regionsA, &ctx->regionsB, &ctx->childLen, &ctx->childOff, &ctx->childIdx,
                     &ctx->scanTemp, &ctx->candOfMpu, &ctx->scalars, &ctx->cullScratch, &ctx->cullScratchFlag, &ctx->fv, &ctx->inside, &ctx->vcount, &ctx->tcount, &ctx->vbase, &ctx->tbase,
                     &ctx->edgeMask, &ctx->edgePrefix, &ctx->mixedList, &ctx->mixedFlag, &ctx->mixedIdx, &ctx->counters, &ctx->statsAcc,
                     &ctx->pos, &ctx->nrm, &ctx->col, &ctx->tri, &ctx->mpuIjk, &ctx->mpuVbase, &ctx->mpuTbase, &ctx->vperm, &ctx->ptsIn, &ctx->ptsF,
                     &ctx->ptsCol, &ctx->ptsNrm, &ctx->shardCounts, &ctx->haloSend, &ctx->haloRecv};
    if (ctx->comm) { psgpu::nccl_api().CommDestroy(ctx->comm); ctx->comm = nullptr; }
    for (DevBuf* b : all) b->release();
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    if (ctx->hostScalars) cudaFreeHost(ctx->hostScalars);
    if (ctx->hostStage) cudaFreeHost(ctx->hostStage);
    if (ctx->hostMesh) cudaFreeHost(ctx->hostMesh);
    cudaStreamDestroy(ctx->copyStream);
    if (ctx->commStream) cudaStreamDestroy(ctx->commStream);
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
    ctx->slab.valid = false;
    ctx->polysSinceUpload = 0;
    ctx->nPrims = p->nPrims;
    ctx->semantics = p->semantics;
    ctx->hasWarps = p->maxPointStack > 0;
    {   // differences (min(a, 1 - b), a * (1 - b) and their reversed forms) are the only operators that can leave a negative value
        bool neg = false, masks = false;
        const uint32_t* w = p->words.data();
        const size_t nq = p->words.size() / 4;
        for (size_t pc = 0; pc < nq;) {
            const uint32_t hdr = w[pc * 4], opc = PS_HDR_OPC(hdr), cmb = PS_HDR_CMB(hdr);
            if (opc == OPC_END) break;
            if ((opc < OPC_FIRST_NONPRIM || opc == OPC_COMBINE) && (cmb == CMB_DIF || cmb == CMB_SDIF || cmb == CMB_RDIF || cmb == CMB_RSDIF)) neg = true;
            if (opc == OPC_BOXMASK) masks = true;
            if (opc == OPC_DIF0 || opc == OPC_RICCI_SIMD) neg = true;  // fast_pow's quotient changes sign beyond its pole
            pc += (size_t)ps_instr_quads(hdr);
        }
        ctx->nonNeg = !neg;
        ctx->hasMasks = masks;
    }
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
    G.hasMasks = ctx->hasMasks ? 1 : 0;
    G.mpuSide = P->cellsize * (float)PS_CELLS_PER_MPU;
    G.coordMode = P->coord_mode; G.sem = ctx->semantics; G.output = P->output; G.rootMode = P->root_mode;
    G.rootIters = P->root_iterations > 0 ? P->root_iterations : (P->root_mode == PSGPU_ROOT_NEWTON_REF ? 8 : 4);
    G.insideGE = (ctx->semantics == PSGPU_SEM_SIMD) ? 1 : 0;
    G.wantColors = P->want_colors; G.wantNormals = P->want_normals;
    G.forceFv = 0;
    G.zOwnLast = G.hi[2] - 1;
    G.vertexOffset = 0u;
    G.haloVertexOffset = 0u;
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
    L.mpuSide = G.mpuSide; L.cullMode = cullMode; L.iso = G.iso; L.childSize = 1; L.colorExact = colorExact ? 1 : 0; L.nonNeg = ctx->nonNeg ? 1 : 0;
    poolUsed = ctx->rootOff + ctx->rootLen;  // [full program | slab program] stay; the levels are rebuilt behind them
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
    if (sizes.size() == 3 && ctx->rootLen <= 4096) {
        // A short program does not pay for the 16^3 level: its walk is bound by the serial Pruner, about as long per region as the
        // 4^3 level's own (blob1k@512^3: cull 0.72 -> 0.64 ms without it; the 17.5k-quad blob10k needs it: 1.6 vs 5.6 ms). Only while
        // the 4^3 level still gets a warp per region (<= 256 top regions of 16^3 MPUs).
        const auto cdiv = [](int a, int b) { return (a + b - 1) / b; };
        if ((long long)cdiv(nx + 15, 16) * cdiv(ny + 15, 16) * cdiv(nz + 15, 16) <= 256) sizes = {4, 1};
    }
    // a tiny program (a dozen primitives) is cheaper to test once per MPU than to refine level by level (small@128^3: 0.27 -> 0.23 ms)
    if (ctx->rootLen <= 64) sizes = {1};
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
                for (int z = G.lo[2] / ps; z * ps < G.hi[2]; z++) top.push_back(Region{x, y, z, ctx->rootOff, ctx->rootLen});
    }
    DevBuf* cur = &ctx->regionsA;
    DevBuf* nxt = &ctx->regionsB;
    CK(cur->reserve(top.size() * sizeof(Region)));
    CK(cudaMemcpyAsync(cur->p, top.data(), top.size() * sizeof(Region), cudaMemcpyHostToDevice, st));
    int nParents = (int)top.size();
    CK(ctx->scalars.reserve(64));
    CK(cudaMemsetAsync(ctx->scalars.p, 0, 64, st));  // [0] longest candidate program (two-pass emit), [1 + level] the levels' maxima
    maxLen = 0;
    uint32_t parentMaxLen = ctx->rootLen;  // longest program among the current level's parents
    for (size_t lv = 0; lv < sizes.size(); lv++) {
        L.childSize = sizes[lv];
        const bool last = (lv + 1 == sizes.size());
        const int n = nParents * 64;
        CK(ctx->childLen.reserve((size_t)(n + 1) * 4)); CK(ctx->childOff.reserve((size_t)(n + 1) * 4));
        CK(ctx->childIdx.reserve((size_t)(n + 1) * 4));
        const size_t tb = scan_temp_bytes(n + 1);
        CK(ctx->scanTemp.reserve(tb));
        // coarse levels (few regions, long programs): ONE warp-cooperative walk per region into scratch slots, compacted after the scans;
        // the scratch bound keeps 10k-primitive trees within ~1 GB
        const size_t scratchQuads = (size_t)n * parentMaxLen;
        // The thread-per-MPU level too when it is the last one, its slots fit 4 GB and the program has warps: there a walk is dominated by
        // the warp-box arithmetic and walking twice costs more than the slot copy (csgwarp@512^3: cull 1.34 -> 1.22 ms); for warp-free
        // programs count + emit stays (blob1k@512^3: 0.67 vs 0.70 ms fused). No flags at the last level: the field kernel never reads them.
        const bool warpLevel = prune_level_is_fused(L, nParents);
        const bool fused = !getenv("PSGPU_CULL_TWO_PASS") &&
                           (warpLevel ? scratchQuads <= ((size_t)64 << 20) : (last && ctx->hasWarps && scratchQuads <= ((size_t)256 << 20)));
        uint32_t* levelMax = ctx->scalars.as<uint32_t>() + 1 + lv;
        if (fused) {
            CK(ctx->cullScratch.reserve(scratchQuads * 16)); CK(ctx->cullScratchFlag.reserve(warpLevel ? scratchQuads : 16));
            // few regions with a long program (the top level): a whole CTA per region instead of a warp
            const char* ctaMin = getenv("PSGPU_CULL_CTA_MINLEN");
            const char* ctaMaxN = getenv("PSGPU_CULL_CTA_MAXN");
            const bool ctaPerRegion = n <= (ctaMaxN ? atoi(ctaMaxN) : 4096) && parentMaxLen >= (uint32_t)(ctaMin ? atoi(ctaMin) : 1024);
            launch_prune_fused(L, cur->as<Region>(), nParents, ctx->pool.as<float4>(), ctx->qflag.as<uint8_t>(), ctx->cullScratch.as<float4>(),
                               ctx->cullScratchFlag.as<uint8_t>(), parentMaxLen, ctx->childLen.as<uint32_t>(), levelMax, ctaPerRegion, st);
        } else
            launch_prune_count(L, cur->as<Region>(), nParents, ctx->pool.as<float4>(), ctx->qflag.as<uint8_t>(), ctx->childLen.as<uint32_t>(), levelMax, st);
        // (two-pass levels: the longest pruned program is known once the lengths are - it comes back with the scan totals below, and
        //  the emit pass and the field kernel follow without another host round trip)
        // pool offsets (sums of the lengths) and compacted indices (sums of the flags length != 0) of the children in one scan
        exclusive_scan_pair(ctx->scanTemp.p, tb, ctx->childLen.as<uint32_t>(), nullptr, ctx->childOff.as<uint32_t>(), ctx->childIdx.as<uint32_t>(), n, st);
        launches += 2;
        CK(cudaMemcpyAsync(&ctx->hostScalars[0], ctx->childOff.as<uint32_t>() + n, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&ctx->hostScalars[1], ctx->childIdx.as<uint32_t>() + n, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&ctx->hostScalars[3], levelMax, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        const uint32_t totalQuads = ctx->hostScalars[0], nChildren = ctx->hostScalars[1];
        if (fused && ctx->hostScalars[3] > parentMaxLen) { ctx->err = "pruned program longer than its parent"; return PSGPU_ERROR_CUDA; }
        if ((uint64_t)poolUsed + totalQuads >= ((uint64_t)1 << 32)) { ctx->err = "pruned programs exceed 2^32 quads"; return PSGPU_NOT_ENOUGH_MEM; }
        CK(ctx->pool.reserve(((size_t)poolUsed + totalQuads + 16) * 16, true, st));
        CK(ctx->qflag.reserve((size_t)poolUsed + totalQuads + 16, true, st));
        CK(nxt->reserve(std::max((size_t)nChildren, (size_t)1) * sizeof(Region)));
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
        if (last) maxLen = ctx->hostScalars[3];
        if (nParents == 0) break;
    }
    ctx->cands = cur->as<Region>();
    ctx->nCand = nParents;
    if (nParents == 0) maxLen = 0;
    return PSGPU_SUCCESS;
}

// caller-provided HOST buffers the mesh is streamed into while the emission kernels still run
struct HostSink {
    float *pos, *nrm, *col;
    uint32_t* tri;
    int64_t capV, capT;
};

#define NK(call)                                                                                                       \
    do {                                                                                                               \
        int _r = (call);                                                                                               \
        if (_r != 0) { ctx->err = std::string(#call) + ": " + psgpu::nccl_api().GetErrorString(_r); return PSGPU_ERROR_CUDA; } \
    } while (0)

// The uploaded program pruned ONCE against the slab's box (support culling, colour markers as the levels leave them): with R
// ranks every rank's hierarchy then walks ~1/R of the primitives instead of all of them at its top level (SCALE_r01: cull 0.64 ms
// at N = 1 but 1.30 ms per rank at N = 8). One CTA-cooperative walk of the whole program (prune_slab_kernel) - milliseconds for 10k
// primitives - so it is a plan-time cost like the upload: cached behind the full program until the program or the slab changes, and only taken for
// programs long enough to pay back (>= 2048 quads) on a slab that is a proper part of the lattice.
int ensure_slab_program(psgpu_ctx* ctx, const GridParams& G, bool colorExact) {
    ctx->rootOff = 0; ctx->rootLen = ctx->progQuads;
    const bool partial = (G.hi[2] - G.lo[2]) < G.dims[2];
    if (!partial || ctx->progQuads < 2048 || getenv("PSGPU_NO_SLAB_PROGRAM")) { ctx->slab.valid = false; return PSGPU_SUCCESS; }
    psgpu_ctx::SlabProg& S = ctx->slab;
    // The prune is a plan-time cost (4 ms for a rank's eighth of blob10k) that saves ~0.1 ms of culling per polygonize: it is only
    // built once the uploaded program has proved to be REUSED - from its third polygonize on. A caller that uploads a new program for
    // every frame (animation, the end-to-end path of bench.py) never pays it; PSGPU_SLAB_PROGRAM_NOW=1 builds it at once.
    if (++ctx->polysSinceUpload < 3 && !S.valid && !getenv("PSGPU_SLAB_PROGRAM_NOW")) return PSGPU_SUCCESS;
    const bool hit = S.valid && S.zb == G.lo[2] && S.ze == G.hi[2] && S.side == G.mpuSide && S.nx == G.dims[0] && S.ny == G.dims[1] &&
                     S.lo[0] == G.gridLo[0] && S.lo[1] == G.gridLo[1] && S.lo[2] == G.gridLo[2] && S.colorExact == (colorExact ? 1 : 0);
    if (!hit) {
        cudaStream_t st = ctx->stream;
        float lo[3], hi[3];
        for (int a = 0; a < 3; a++) {
            const float pad = 1.0e-3f * G.mpuSide;  // child boxes are computed as gridLo + index * side: keep them inside
            lo[a] = G.gridLo[a] + (float)G.lo[a] * G.mpuSide - pad;
            hi[a] = G.gridLo[a] + (float)G.hi[a] * G.mpuSide + pad;
        }
        const uint32_t cap = ctx->progQuads;
        CK(ctx->pool.reserve(((size_t)ctx->progQuads * 2 + 16) * 16, true, st));
        CK(ctx->qflag.reserve((size_t)ctx->progQuads * 2 + 16, true, st));
        CK(ctx->scalars.reserve(64));
        launch_prune_slab(ctx->pool.as<float4>(), ctx->qflag.as<uint8_t>(), ctx->progQuads, lo, hi, PS_ISO_VALUE, colorExact ? 1 : 0, ctx->pool.as<float4>() + ctx->progQuads,
                          ctx->qflag.as<uint8_t>() + ctx->progQuads, cap, ctx->scalars.as<uint32_t>() + 12, st);
        CK(cudaMemcpyAsync(&ctx->hostScalars[12], ctx->scalars.as<uint32_t>() + 12, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        CK(cudaGetLastError());
        S.valid = true; S.zb = G.lo[2]; S.ze = G.hi[2]; S.side = G.mpuSide; S.nx = G.dims[0]; S.ny = G.dims[1];
        for (int a = 0; a < 3; a++) S.lo[a] = G.gridLo[a];
        S.colorExact = colorExact ? 1 : 0;
        S.len = ctx->hostScalars[12];
        if (S.len > cap) { ctx->err = "slab program longer than the program"; S.valid = false; return PSGPU_ERROR_CUDA; }
    }
    ctx->rootOff = ctx->progQuads; ctx->rootLen = S.len;  // len 0: nothing of the program reaches this slab
    return PSGPU_SUCCESS;
}

// stitching state of one sharded polygonize (psgpu_polygonize_sharded)
struct ShardCtl {
    bool stitch;
    psgpu_shard* info;
};

// Count exchange + halo exchange of a stitched z-slab run, on the context's COMMUNICATION stream, enqueued once the scans are through
// and before the vertex / colour kernels are launched on the compute stream - those need nothing from the other ranks, so a rank that
// reaches this point early computes while it waits for its neighbours instead of idling in the collective:
//   1. ncclAllGather of {vertices, triangles} (16 bytes per rank) -> every rank's global vertex / triangle offset;
//   2. every rank but the first sends the table of its FIRST MPU layer (rank-local vertex base + crossing-edge mask / prefix per MPU)
//      to the rank below, which resolves the edges of its closing z-plane through it (tri_kernel) - the lower-z rank does not emit
//      those vertices, so the ranks' meshes concatenate to the single-GPU mesh (same vertex count, same triangle soup).
// shard_exchange_finish waits for both (host side: the offsets become kernel parameters of tri_kernel) and leaves ev[9] for the
// compute stream to wait on.
int shard_exchange_begin(psgpu_ctx* ctx, const GridParams& G, uint32_t nV, uint32_t nT, const FieldOutputs* fo, MeshOutputs& mo) {
    psgpu::NcclApi& N = psgpu::nccl_api();
    cudaStream_t cs = ctx->commStream;  // (the host has read the scans' totals: everything the halo table is built from is complete)
    const int R = ctx->nranks, me = ctx->rank;
    if (20 + 4 * R > psgpu_ctx::kHostScalars) { ctx->err = "too many ranks"; return PSGPU_PARAM_ERROR; }
    CK(cudaEventRecord(ctx->ev[8], cs));
    CK(ctx->shardCounts.reserve((size_t)(2 + 2 * R) * 8));
    uint64_t* hsend = reinterpret_cast<uint64_t*>(&ctx->hostScalars[16]);
    hsend[0] = nV; hsend[1] = nT;
    uint64_t* dsend = ctx->shardCounts.as<uint64_t>();
    uint64_t* drecv = dsend + 2;
    CK(cudaMemcpyAsync(dsend, hsend, 16, cudaMemcpyHostToDevice, cs));
    NK(N.AllGather(dsend, drecv, 2, psgpu::NcclApi::kUint64, ctx->comm, cs));
    CK(cudaMemcpyAsync(&ctx->hostScalars[20], drecv, (size_t)2 * R * 8, cudaMemcpyDeviceToHost, cs));  // pinned: does not block the host
    const int nx = G.hi[0] - G.lo[0], ny = G.hi[1] - G.lo[1];
    const size_t nxy = (size_t)nx * ny;
    const size_t bytes = nxy * (4 + 192 + 96);
    auto carve = [&](DevBuf& b, uint32_t*& vb, uint32_t*& mk, uint16_t*& pf) {
        vb = b.as<uint32_t>(); mk = vb + nxy; pf = reinterpret_cast<uint16_t*>(mk + nxy * 48);
    };
    uint32_t *sv = nullptr, *sm = nullptr, *rv = nullptr, *rm = nullptr;
    uint16_t *sp = nullptr, *rp = nullptr;
    if (me > 0) {
        CK(ctx->haloSend.reserve(bytes));
        carve(ctx->haloSend, sv, sm, sp);
        if (fo) launch_halo_build(G, ctx->candOfMpu.as<int32_t>(), *fo, ctx->vbase.as<uint32_t>(), sv, sm, sp, (int)nxy, cs);
        else CK(cudaMemsetAsync(sv, 0xFF, nxy * 4, cs));  // no candidates in this slab: no surface anywhere in the layer
    }
    if (me < R - 1) { CK(ctx->haloRecv.reserve(bytes)); carve(ctx->haloRecv, rv, rm, rp); }
    NK(N.GroupStart());
    if (me > 0) NK(N.Send(sv, bytes, psgpu::NcclApi::kUint8, me - 1, ctx->comm, cs));
    if (me < R - 1) NK(N.Recv(rv, bytes, psgpu::NcclApi::kUint8, me + 1, ctx->comm, cs));
    NK(N.GroupEnd());
    mo.haloVbase = rv; mo.haloMask = rm; mo.haloPrefix = rp;
    CK(cudaEventRecord(ctx->ev[9], cs));
    return PSGPU_SUCCESS;
}

int shard_exchange_finish(psgpu_ctx* ctx, GridParams& G, psgpu_shard* info) {
    const int R = ctx->nranks, me = ctx->rank;
    CK(cudaStreamSynchronize(ctx->commStream));
    const uint64_t* got = reinterpret_cast<const uint64_t*>(&ctx->hostScalars[20]);
    ctx->rankCounts.assign(got, got + (size_t)2 * R);
    uint64_t vOff = 0, tOff = 0, vTot = 0, tTot = 0;
    for (int r = 0; r < R; r++) {
        if (r < me) { vOff += ctx->rankCounts[2 * r]; tOff += ctx->rankCounts[2 * r + 1]; }
        vTot += ctx->rankCounts[2 * r]; tTot += ctx->rankCounts[2 * r + 1];
    }
    if (vTot >= ((uint64_t)1 << 32)) { ctx->err = "stitched mesh exceeds 2^32 vertices"; return PSGPU_NOT_ENOUGH_MEM; }
    G.vertexOffset = (uint32_t)vOff;
    G.haloVertexOffset = (uint32_t)(vOff + ctx->rankCounts[2 * me]);  // = the vertex offset of the rank above
    if (info) {
        info->nranks = R; info->rank = me; info->z_begin = G.lo[2]; info->z_end = G.hi[2];
        info->vertex_offset = (int64_t)vOff; info->triangle_offset = (int64_t)tOff;
        info->total_vertices = (int64_t)vTot; info->total_triangles = (int64_t)tTot;
        info->stitched = 1;
    }
    return PSGPU_SUCCESS;
}

int polygonize_impl(psgpu_ctx* ctx, const psgpu_params* P, psgpu_mesh* out, bool forceFv, const HostSink* sink = nullptr, ShardCtl* shard = nullptr) {
    if (!ctx || !P || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    GridParams G;
    int rc = fill_grid(ctx, P, G);
    if (rc != PSGPU_SUCCESS) { ctx->err = "bad grid parameters"; return rc; }
    G.forceFv = forceFv ? 1 : 0;
    const bool stitch = shard && shard->stitch && ctx->nranks > 1;
    if (stitch) {
        if (P->output != PSGPU_OUT_INDEXED) { ctx->err = "sharded stitching needs PSGPU_OUT_INDEXED"; return PSGPU_PARAM_ERROR; }
        G.zOwnLast = G.dims[2] - 1;  // only the lattice's last layer owns a closing z-plane
    }
    if (P->output == PSGPU_OUT_INDEXED && P->coord_mode != PSGPU_COORD_GLOBAL) {
        ctx->err = "PSGPU_OUT_INDEXED needs PSGPU_COORD_GLOBAL (shared corners must be bit-identical across MPUs)";
        return PSGPU_PARAM_ERROR;
    }
    int launches = 0;
    psgpu_stats& S = ctx->stats;
    S = psgpu_stats{};
    S.mpus_total = (int64_t)(G.hi[0] - G.lo[0]) * (G.hi[1] - G.lo[1]) * (G.hi[2] - G.lo[2]);
    S.program_quads = ctx->progQuads;
    const bool colorExact = ctx->semantics == PSGPU_SEM_SIMD && P->want_colors;
    rc = ensure_slab_program(ctx, G, colorExact);
    if (rc != PSGPU_SUCCESS) return rc;
    CK(cudaEventRecord(ctx->ev[0], st));
    uint32_t maxLen = 0, poolUsed = 0;
    if (ctx->rootLen == 0) { ctx->nCand = 0; poolUsed = ctx->rootOff; }
    else {
        rc = run_cull(ctx, G, P->cull, colorExact, maxLen, poolUsed, launches);
        if (rc != PSGPU_SUCCESS) return rc;
    }
    CK(cudaEventRecord(ctx->ev[1], st));
    const int nCand = ctx->nCand;
    S.mpus_evaluated = nCand;
    S.pruned_quads = (int64_t)poolUsed - (ctx->rootOff + ctx->rootLen);

    psgpu_mesh M{};
    for (int a = 0; a < 3; a++) M.lattice[a] = G.dims[a];
    uint32_t nV = 0, nT = 0, nMixed = 0, nOver = 0;
    bool streamed = false;
    if (nCand > 0) {
        const size_t n1 = (size_t)nCand + 1;
        CK(ctx->fv.reserve((size_t)nCand * 2048)); CK(ctx->inside.reserve((size_t)nCand * 64));
        CK(ctx->vcount.reserve(n1 * 4)); CK(ctx->tcount.reserve(n1 * 4)); CK(ctx->vbase.reserve(n1 * 4)); CK(ctx->tbase.reserve(n1 * 4));
        CK(ctx->edgeMask.reserve((size_t)nCand * 192)); CK(ctx->edgePrefix.reserve((size_t)nCand * 96));
        CK(ctx->mixedList.reserve((size_t)nCand * 4));
        CK(ctx->counters.reserve(64));
        CK(cudaMemsetAsync(ctx->counters.p, 0, 64, st));
        // (the closing entries vcount[nCand] = tcount[nCand] = 0 of the scans below are written by field_kernel)
        FieldOutputs fo;
        fo.fv = ctx->fv.as<float4>(); fo.inside = ctx->inside.as<uint32_t>(); fo.vcount = ctx->vcount.as<uint32_t>(); fo.tcount = ctx->tcount.as<uint32_t>();
        fo.edgeMask = ctx->edgeMask.as<uint32_t>(); fo.edgePrefix = ctx->edgePrefix.as<uint16_t>();
        fo.mixedList = ctx->mixedList.as<uint32_t>(); fo.counters = ctx->counters.as<uint32_t>();
        launch_field_kernel(G, ctx->cands, nCand, ctx->pool.as<float4>(), fo, maxLen, ctx->hasWarps, ctx->numSMs, st);
        launches++;
        CK(cudaEventRecord(ctx->ev[2], st));
        const size_t tb = scan_temp_bytes((int)n1);
        CK(ctx->scanTemp.reserve(tb));
        exclusive_scan_pair(ctx->scanTemp.p, tb, fo.vcount, fo.tcount, ctx->vbase.as<uint32_t>(), ctx->tbase.as<uint32_t>(), nCand, st);  // both bases in one scan
        launches += 1;
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
        if (stitch) {  // enqueued on the communication stream; nothing on the compute stream waits for it before tri_kernel
            rc = shard_exchange_begin(ctx, G, nV, nT, &fo, mo);
            if (rc != PSGPU_SUCCESS) return rc;
            launches++;
        }
        if (sink && ((int64_t)nV > sink->capV || (int64_t)nT > sink->capT)) {
            if (stitch) shard_exchange_finish(ctx, G, shard->info);  // the other ranks are in the same collectives
            ctx->err = "host buffers too small";
            M.n_vertices = nV; M.n_triangles = nT;
            if (out) *out = M;
            return PSGPU_NOT_ENOUGH_MEM;
        }
        CK(cudaEventRecord(ctx->ev[3], st));
        // A stitched run: vertex and colour kernels first (they need nothing from the other ranks), tri_kernel - global vertex
        // offsets, halo table of the rank above - once the exchange is through; what a rank would wait for its neighbours is
        // spent computing. Otherwise tri_kernel first: its output is ready soonest and leaves for the host while the others run.
        const bool lateTri = stitch;
        if (lateTri) {
            if (nMixed > 0)
                launch_emit_kernels(G, ctx->cands, ctx->pool.as<float4>(), ctx->colors.as<float>(), fo, mo, ctx->candOfMpu.as<int32_t>(), (int)nMixed,
                                    maxLen, ctx->numSMs, st, nullptr, ctx->ev[6], PS_EMIT_VERTEX_COLOR);
            CK(cudaEventRecord(ctx->ev[7], st));
            rc = shard_exchange_finish(ctx, G, shard->info);
            if (rc != PSGPU_SUCCESS) return rc;
            CK(cudaStreamWaitEvent(st, ctx->ev[9], 0));
            if (nMixed > 0)
                launch_emit_kernels(G, ctx->cands, ctx->pool.as<float4>(), ctx->colors.as<float>(), fo, mo, ctx->candOfMpu.as<int32_t>(), (int)nMixed,
                                    maxLen, ctx->numSMs, st, ctx->ev[5], nullptr, PS_EMIT_TRI);
        } else if (nMixed > 0)
            launch_emit_kernels(G, ctx->cands, ctx->pool.as<float4>(), ctx->colors.as<float>(), fo, mo, ctx->candOfMpu.as<int32_t>(), (int)nMixed,
                                maxLen, ctx->numSMs, st, ctx->ev[5], ctx->ev[6]);
        if (nMixed > 0) launches += 2 + (G.wantColors ? 1 : 0);  // tri_kernel, vertex_kernel, color_kernel
        CK(cudaEventRecord(ctx->ev[4], st));
        if (sink && nMixed > 0 && P->output == PSGPU_OUT_INDEXED) {
            // each array leaves for the host as soon as the kernel that writes it is done: triangles behind tri_kernel (while
            // vertex_kernel runs), positions + normals behind vertex_kernel (while color_kernel runs), colours last; in a stitched
            // run (tri_kernel last) the triangles go last
            cudaStream_t cs = ctx->copyStream;
            const bool nrmLate = normals_in_color_pass(G);  // then the colour pass writes them
            if (!lateTri) {
                CK(cudaStreamWaitEvent(cs, ctx->ev[5], 0));
                if (sink->tri && nT) CK(cudaMemcpyAsync(sink->tri, ctx->tri.p, (size_t)nT * 12, cudaMemcpyDeviceToHost, cs));
            }
            CK(cudaStreamWaitEvent(cs, ctx->ev[6], 0));
            if (sink->pos && nV) CK(cudaMemcpyAsync(sink->pos, ctx->pos.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            if (sink->nrm && nV && P->want_normals && !nrmLate) CK(cudaMemcpyAsync(sink->nrm, ctx->nrm.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            CK(cudaStreamWaitEvent(cs, lateTri ? ctx->ev[7] : ctx->ev[4], 0));
            if (sink->nrm && nV && P->want_normals && nrmLate) CK(cudaMemcpyAsync(sink->nrm, ctx->nrm.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            if (sink->col && nV && P->want_colors) CK(cudaMemcpyAsync(sink->col, ctx->col.p, (size_t)nV * 12, cudaMemcpyDeviceToHost, cs));
            if (lateTri) {
                CK(cudaStreamWaitEvent(cs, ctx->ev[5], 0));
                if (sink->tri && nT) CK(cudaMemcpyAsync(sink->tri, ctx->tri.p, (size_t)nT * 12, cudaMemcpyDeviceToHost, cs));
            }
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
        CK(cudaEventRecord(ctx->ev[2], st));
        if (stitch) {  // an empty slab still takes part in the collectives
            MeshOutputs mo{};
            CK(cudaEventRecord(ctx->ev[7], st));
            rc = shard_exchange_begin(ctx, G, 0u, 0u, nullptr, mo);
            if (rc == PSGPU_SUCCESS) rc = shard_exchange_finish(ctx, G, shard->info);
            if (rc != PSGPU_SUCCESS) return rc;
        }
        CK(cudaEventRecord(ctx->ev[3], st)); CK(cudaEventRecord(ctx->ev[4], st));
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
    if (stitch) {
        // what the compute stream waited for the exchange: from the end of the colour kernel (ev[7]) to the end of the exchange on
        // the communication stream (ev[9]), 0 when the exchange was through by then. ms_total - ms_exchange = the rank's own work.
        float w = 0.f;
        cudaEventElapsedTime(&w, ctx->ev[7], ctx->ev[9]);
        S.ms_exchange = w > 0.f ? w : 0.f;
    }
    if (nMixed > 0 && stitch) {  // vertex, colour, then tri_kernel
        cudaEventElapsedTime(&S.ms_vertex, ctx->ev[3], ctx->ev[6]);
        cudaEventElapsedTime(&S.ms_color, ctx->ev[6], ctx->ev[7]);
        cudaEventElapsedTime(&S.ms_tri, ctx->ev[7], ctx->ev[5]);
        S.ms_tri = S.ms_tri > S.ms_exchange ? S.ms_tri - S.ms_exchange : 0.f;
    } else if (nMixed > 0) {
        cudaEventElapsedTime(&S.ms_tri, ctx->ev[3], ctx->ev[5]);
        cudaEventElapsedTime(&S.ms_vertex, ctx->ev[5], ctx->ev[6]);
        cudaEventElapsedTime(&S.ms_color, ctx->ev[6], ctx->ev[4]);
    }
    if (shard && shard->info) shard->info->ms_exchange = S.ms_exchange;
    if (nMixed > 0 && P->output == PSGPU_OUT_INDEXED) {
        // edges whose owner did not report the crossing (tri_kernel): never expected, never silent
        CK(cudaMemcpy(&ctx->hostScalars[13], ctx->counters.as<uint32_t>() + 6, 4, cudaMemcpyDeviceToHost));
        S.unresolved_edges = ctx->hostScalars[13];
        if (S.unresolved_edges) {
            ctx->err = "indexed mesh: " + std::to_string(S.unresolved_edges) + " shared edges were not found in their owner's crossing mask";
            return PSGPU_ERROR_CUDA;
        }
    }
    S.mpus_crossed = nMixed;
    // 4-sample: 4 root samples + 4 gradient samples; section search: 4 samples per round + 4; reference Newton: ~2 iterations of 5 + 4
    const int evalsPerVertex = (P->root_mode == PSGPU_ROOT_FOURSAMPLE) ? 8 : (P->root_mode == PSGPU_ROOT_BISECT ? 4 * G.rootIters + 4 : 14);
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

// ---- multi-GPU: one context per rank, NCCL over NVLink --------------------------------------------------------------------------
extern "C" int psgpu_comm_get_unique_id(void* id128) {
    if (!id128) return PSGPU_PARAM_ERROR;
    psgpu::NcclApi& N = psgpu::nccl_api();
    if (!N.load()) return PSGPU_ERROR_NO_DEVICE;
    psgpu::NcclApi::unique_id id;
    if (N.GetUniqueId(&id) != 0) return PSGPU_ERROR_CUDA;
    memcpy(id128, &id, sizeof(id));
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_comm_init(psgpu_ctx* ctx, int nranks, int rank, const void* id128) {
    if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return PSGPU_PARAM_ERROR;
    psgpu::NcclApi& N = psgpu::nccl_api();
    if (!N.load()) { ctx->err = "libnccl.so.2 not found"; return PSGPU_ERROR_NO_DEVICE; }
    CK(cudaSetDevice(ctx->device));
    if (ctx->comm) { N.CommDestroy(ctx->comm); ctx->comm = nullptr; }
    psgpu::NcclApi::unique_id id;
    memcpy(&id, id128, sizeof(id));
    NK(N.CommInitRank(&ctx->comm, nranks, id, rank));
    ctx->nranks = nranks; ctx->rank = rank;
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_comm_destroy(psgpu_ctx* ctx) {
    if (!ctx) return PSGPU_PARAM_ERROR;
    if (ctx->comm) { psgpu::nccl_api().CommDestroy(ctx->comm); ctx->comm = nullptr; }
    ctx->nranks = 1; ctx->rank = 0;
    return PSGPU_SUCCESS;
}

// [begin, end) MPU layers of `rank`: equal slabs, or - with one weight per layer - the cut points that balance the weight prefix
// sum (parsip_b200/sharding.py: slab_bounds is the same rule; computed identically on every rank, no communication)
extern "C" int psgpu_slab_bounds(int n_layers, int nranks, int rank, const float* layer_weights, int32_t* z_begin, int32_t* z_end) {
    if (n_layers < 0 || nranks <= 0 || rank < 0 || rank >= nranks || !z_begin || !z_end) return PSGPU_PARAM_ERROR;
    double total = 0.0;
    if (layer_weights) for (int i = 0; i < n_layers; i++) total += layer_weights[i];
    if (!layer_weights || !(total > 0.0)) {
        const int base = n_layers / nranks, extra = n_layers % nranks;
        *z_begin = rank * base + std::min(rank, extra);
        *z_end = *z_begin + base + (rank < extra ? 1 : 0);
        return PSGPU_SUCCESS;
    }
    std::vector<int> cuts(1, 0);
    double acc = 0.0;
    int layer = 0;
    for (int r = 1; r < nranks; r++) {
        const double target = total * r / nranks;
        while (layer < n_layers && acc + 0.5 * layer_weights[layer] < target) { acc += layer_weights[layer]; layer++; }
        cuts.push_back(layer);
    }
    cuts.push_back(n_layers);
    *z_begin = cuts[rank]; *z_end = cuts[rank + 1];
    return PSGPU_SUCCESS;
}

static int polygonize_sharded_impl(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, psgpu_shard* shard, const HostSink* sink) {
    if (!ctx || !params) return PSGPU_PARAM_ERROR;
    if (shard) *shard = psgpu_shard{};
    psgpu_params P = *params;
    if (ctx->nranks > 1 && !ctx->comm) { ctx->err = "psgpu_comm_init first"; return PSGPU_PARAM_ERROR; }
    if (P.z_end <= 0) {  // no explicit slab: equal shares of the MPU layers
        int32_t nm[3];
        psgpu_lattice_dims(P.cellsize, P.bbox_lo, P.bbox_hi, nm);
        if (nm[2] < ctx->nranks) { ctx->err = "fewer MPU layers than ranks"; return PSGPU_PARAM_ERROR; }
        psgpu_slab_bounds(nm[2], ctx->nranks, ctx->rank, nullptr, &P.z_begin, &P.z_end);
    }
    if (P.z_end <= P.z_begin) { ctx->err = "empty slab"; return PSGPU_PARAM_ERROR; }
    ShardCtl sc{true, shard};
    const int rc = polygonize_impl(ctx, &P, out, false, sink, &sc);
    if (shard && ctx->nranks <= 1 && rc == PSGPU_SUCCESS) {
        shard->nranks = 1; shard->rank = 0; shard->z_begin = P.z_begin; shard->z_end = P.z_end;
        shard->total_vertices = ctx->last.n_vertices; shard->total_triangles = ctx->last.n_triangles;
    }
    return rc;
}

extern "C" int psgpu_polygonize_sharded(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, psgpu_shard* shard) {
    return polygonize_sharded_impl(ctx, params, out, shard, nullptr);
}

// the same with this rank's share streamed into HOST buffers behind the emission kernels (like psgpu_polygonize_host)
extern "C" int psgpu_polygonize_sharded_host(psgpu_ctx* ctx, const psgpu_params* params, psgpu_mesh* out, psgpu_shard* shard, float* pos,
                                             float* nrm, float* col, uint32_t* tri, int64_t cap_vertices, int64_t cap_triangles) {
    psgpu_mesh M{};
    const HostSink sink{pos, nrm, col, tri, cap_vertices, cap_triangles};
    const int rc = polygonize_sharded_impl(ctx, params, &M, shard, &sink);
    if (out) *out = M;
    return rc;
}

// Assembles the stitched mesh on `root`: DEVICE buffers of total_vertices * 3 floats (pos, nrm, col; NULL skips one) and
// total_triangles * 3 indices, written at every rank's offset; other ranks pass NULL. Grouped ncclSend / ncclRecv over NVLink.
extern "C" int psgpu_mesh_gather(psgpu_ctx* ctx, int root, float* pos, float* nrm, float* col, uint32_t* tri) {
    if (!ctx || root < 0 || root >= ctx->nranks) return PSGPU_PARAM_ERROR;
    if ((int)ctx->rankCounts.size() != 2 * ctx->nranks) { ctx->err = "psgpu_polygonize_sharded first"; return PSGPU_PARAM_ERROR; }
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    psgpu::NcclApi& N = psgpu::nccl_api();
    const int R = ctx->nranks, me = ctx->rank;
    const psgpu_mesh& M = ctx->last;
    if (R == 1 || me == root) {
        uint64_t vo = 0, to = 0;
        if (R > 1) NK(N.GroupStart());
        for (int r = 0; r < R; r++) {
            const uint64_t nv = ctx->rankCounts.empty() ? (uint64_t)M.n_vertices : ctx->rankCounts[2 * r];
            const uint64_t nt = ctx->rankCounts.empty() ? (uint64_t)M.n_triangles : ctx->rankCounts[2 * r + 1];
            if (r == me) {
                if (pos && nv) CK(cudaMemcpyAsync(pos + 3 * vo, M.pos, nv * 12, cudaMemcpyDeviceToDevice, st));
                if (nrm && nv) CK(cudaMemcpyAsync(nrm + 3 * vo, M.nrm, nv * 12, cudaMemcpyDeviceToDevice, st));
                if (col && nv) CK(cudaMemcpyAsync(col + 3 * vo, M.col, nv * 12, cudaMemcpyDeviceToDevice, st));
                if (tri && nt) CK(cudaMemcpyAsync(tri + 3 * to, M.tri, nt * 12, cudaMemcpyDeviceToDevice, st));
            } else {
                if (pos && nv) NK(N.Recv(pos + 3 * vo, nv * 3, psgpu::NcclApi::kFloat32, r, ctx->comm, st));
                if (nrm && nv) NK(N.Recv(nrm + 3 * vo, nv * 3, psgpu::NcclApi::kFloat32, r, ctx->comm, st));
                if (col && nv) NK(N.Recv(col + 3 * vo, nv * 3, psgpu::NcclApi::kFloat32, r, ctx->comm, st));
                if (tri && nt) NK(N.Recv(tri + 3 * to, nt * 3, psgpu::NcclApi::kUint32, r, ctx->comm, st));
            }
            vo += nv; to += nt;
        }
        if (R > 1) NK(N.GroupEnd());
    } else {
        // which arrays the root wants is part of the call contract: every rank passes the same NULL pattern through `pos/nrm/col/tri`
        // being non-NULL on the root; the senders mirror it with the flags below (non-NULL = send)
        const uint64_t nv = (uint64_t)M.n_vertices, nt = (uint64_t)M.n_triangles;
        NK(N.GroupStart());
        if (pos && nv) NK(N.Send(M.pos, nv * 3, psgpu::NcclApi::kFloat32, root, ctx->comm, st));
        if (nrm && nv) NK(N.Send(M.nrm, nv * 3, psgpu::NcclApi::kFloat32, root, ctx->comm, st));
        if (col && nv) NK(N.Send(M.col, nv * 3, psgpu::NcclApi::kFloat32, root, ctx->comm, st));
        if (tri && nt) NK(N.Send(M.tri, nt * 3, psgpu::NcclApi::kUint32, root, ctx->comm, st));
        NK(N.GroupEnd());
    }
    CK(cudaStreamSynchronize(st));
    return PSGPU_SUCCESS;
}

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

extern "C" int psgpu_newton_roots(psgpu_ctx* ctx, const float* p1, const float* p2, const float* f1, const float* f2, int64_t n, float target_field,
                                  int iterations, float* root_out, float* field_out, int32_t* evals_out) {
    if (!ctx || !p1 || !p2 || !f1 || !f2 || !root_out || !field_out || !evals_out || n < 0 || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    if (n == 0) return PSGPU_SUCCESS;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    // in: p1 | p2 | f1 | f2 (8 floats per pair); out: root | field | evals (5 words per pair)
    CK(ctx->ptsIn.reserve((size_t)n * 32)); CK(ctx->ptsNrm.reserve((size_t)n * 20));
    float* in = ctx->ptsIn.as<float>();
    float* out = ctx->ptsNrm.as<float>();
    CK(cudaMemcpyAsync(in, p1, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(in + n * 3, p2, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(in + n * 6, f1, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(in + n * 7, f2, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    psgpu::launch_points_newton(ctx->pool.as<float4>(), in, in + n * 3, in + n * 6, in + n * 7, n, target_field, iterations, out, out + n * 3,
                                reinterpret_cast<int32_t*>(out + n * 4), ctx->semantics, st);
    CK(cudaMemcpyAsync(root_out, out, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(field_out, out + n * 3, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(evals_out, out + n * 4, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
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
    CK(ctx->fv.reserve(2048)); CK(ctx->inside.reserve(64)); CK(ctx->vcount.reserve(8)); CK(ctx->tcount.reserve(8));
    CK(ctx->edgeMask.reserve(192)); CK(ctx->edgePrefix.reserve(96)); CK(ctx->mixedList.reserve(4)); CK(ctx->counters.reserve(64));
    CK(cudaMemsetAsync(ctx->counters.p, 0, 64, st));
    psgpu::FieldOutputs fo;
    fo.fv = ctx->fv.as<float4>(); fo.inside = ctx->inside.as<uint32_t>(); fo.vcount = ctx->vcount.as<uint32_t>(); fo.tcount = ctx->tcount.as<uint32_t>();
    fo.edgeMask = ctx->edgeMask.as<uint32_t>(); fo.edgePrefix = ctx->edgePrefix.as<uint16_t>();
    fo.mixedList = ctx->mixedList.as<uint32_t>(); fo.counters = ctx->counters.as<uint32_t>();
    psgpu::launch_field_kernel(G, ctx->regionsA.as<psgpu::Region>(), 1, ctx->pool.as<float4>(), fo, ctx->progQuads, ctx->hasWarps, ctx->numSMs, st);
    CK(ctx->ptsCol.reserve(512));
    psgpu::launch_cube_cases(G, ctx->fv.as<float>(), ctx->ptsCol.as<uint8_t>(), st);  // the device classifier, not a host re-statement
    CK(cudaMemcpyAsync(fv512, ctx->fv.p, 2048, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(case343, ctx->ptsCol.p, 343, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
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

namespace {
__global__ void layer_work_kernel(const psgpu::Region* __restrict__ cands, const uint32_t* __restrict__ tcount, int nCand, int nLayers,
                                  int32_t* __restrict__ evaluated, int32_t* __restrict__ crossed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nCand) return;
    const int z = cands[i].z;
    if (z < 0 || z >= nLayers) return;
    atomicAdd(&evaluated[z], 1);
    if (tcount[i] != 0u) atomicAdd(&crossed[z], 1);
}
}  // namespace

// Per-MPU-layer work of the LAST polygonize on this context (layers of the global lattice; zero outside the context's slab): MPUs
// evaluated and MPUs that produced triangles. Summed over ranks it is the histogram psgpu_slab_bounds balances the next frame with.
extern "C" int psgpu_layer_work(psgpu_ctx* ctx, int32_t* evaluated, int32_t* crossed, int32_t n_layers) {
    if (!ctx || !evaluated || !crossed || n_layers <= 0) return PSGPU_PARAM_ERROR;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    memset(evaluated, 0, sizeof(int32_t) * (size_t)n_layers); memset(crossed, 0, sizeof(int32_t) * (size_t)n_layers);
    if (ctx->nCand <= 0) return PSGPU_SUCCESS;
    CK(ctx->ptsF.reserve((size_t)n_layers * 8));
    int32_t* d = ctx->ptsF.as<int32_t>();
    CK(cudaMemsetAsync(d, 0, (size_t)n_layers * 8, st));
    layer_work_kernel<<<(ctx->nCand + 255) / 256, 256, 0, st>>>(ctx->cands, ctx->tcount.as<uint32_t>(), ctx->nCand, n_layers, d, d + n_layers);
    CK(cudaMemcpyAsync(evaluated, d, (size_t)n_layers * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(crossed, d + n_layers, (size_t)n_layers * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
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

// ---------------------------------------------------------------------------------------------------------------------
// psgpu_field_eval_noalloc: psgpu_field_eval for callers that must not allocate (code under CUDA graph capture, tight
// serving / haptics loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// The device staging psgpu_field_eval keeps in growable context buffers (ptsIn / ptsF, cudaMalloc on first
// use and on growth) lives here in ONE file-scope device array of fixed size:
//     [ xyz: 3 * PSGPU_NOALLOC_MAX_POINTS floats | f: PSGPU_NOALLOC_MAX_POINTS floats ]  = 1 MiB,
// held by every GPU that loads the library until the process exits. All calls on a device share it and it keeps the last
// call's contents. Nothing in it has to start at zero: a call writes xyz[0 .. 3n) before the kernel reads it and the kernel
// writes f[0 .. n) before the copy-out reads it (the kernel clamps its read index to n - 1), so no clear is issued.
// Calls that overlap on the GPU (two contexts = two streams on one device, or two host threads) RACE on the array: the
// caller serialises them. In HOST code `g_noallocScratch` is a host-side stand-in, not a device address - the pointer the
// kernel gets comes from cudaGetSymbolAddress, looked up once per device with that device current.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_noallocScratch[4 * PSGPU_NOALLOC_MAX_POINTS];

namespace {
constexpr int kNoallocMaxDevices = 64;
float* g_noallocAddr[kNoallocMaxDevices];  // device address of g_noallocScratch per device ordinal, nullptr = not looked up yet
std::mutex g_noallocLock;
}  // namespace

extern "C" int psgpu_field_eval_noalloc(psgpu_ctx* ctx, const float* xyz, int64_t n, float* f_out) {
    if (!ctx || !xyz || !f_out || n < 0 || ctx->progQuads == 0) return PSGPU_PARAM_ERROR;
    if (n > (int64_t)PSGPU_NOALLOC_MAX_POINTS) {
        ctx->err = "psgpu_field_eval_noalloc: " + std::to_string((long long)n) + " points exceed the static scratch of " +
                   std::to_string((long long)PSGPU_NOALLOC_MAX_POINTS) + " points (PSGPU_NOALLOC_MAX_POINTS); split the batch or use psgpu_field_eval";
        return PSGPU_NOT_ENOUGH_MEM;
    }
    if (ctx->device < 0 || ctx->device >= kNoallocMaxDevices) { ctx->err = "psgpu_field_eval_noalloc: device ordinal out of range"; return PSGPU_PARAM_ERROR; }
    if (n == 0) return PSGPU_SUCCESS;
    CK(cudaSetDevice(ctx->device));
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> guard(g_noallocLock);
        if (!g_noallocAddr[ctx->device]) {  // set-up, once per device, with that device current
            void* p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_noallocScratch));
            if (!p) { ctx->err = "cudaGetSymbolAddress(g_noallocScratch) returned a null address"; return PSGPU_ERROR_CUDA; }
            g_noallocAddr[ctx->device] = static_cast<float*>(p);
        }
        scratch = g_noallocAddr[ctx->device];
    }
    cudaStream_t st = ctx->stream;
    float* dXyz = scratch;
    float* dF = scratch + (size_t)3 * PSGPU_NOALLOC_MAX_POINTS;
    CK(cudaMemcpyAsync(dXyz, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    psgpu::launch_points_field(ctx->pool.as<float4>(), dXyz, n, dF, ctx->semantics, st);  // the kernel psgpu_field_eval launches
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(f_out, dF, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return PSGPU_SUCCESS;
}
