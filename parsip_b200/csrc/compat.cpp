// Implementation of the PS::SIMDPOLY facade (include/PS_SimdPolyCompat.h) on top of the psgpu C ABI.
// Reference behaviour restated: PrepareBBoxes PS_SimdPoly/include/PS_Polygonizer.cpp:55-309,
// CountMPUNeeded :388-412, Polygonize :315-385 (MPU order x outer / z inner, bboxLo = lo + idx*side,
// MPU-local U16 indices), PrintThreadResults :414-428, SimdPoly ParsipHaptics/include/PS_HighPerformanceRender.cpp:33-376.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <functional>
#include <thread>
#include <mutex>
#include <vector>

#include "../../include/PS_SimdPolyCompat.h"
#include "../../include/psgpu.h"

namespace PS {
namespace SIMDPOLY {

namespace {

std::mutex g_lock;       // the reference's Polygonize is not re-entrant either (file-scope statics, cpp:18-19)
psgpu_ctx* g_ctx = nullptr;
U32 g_processed = 0, g_crossed = 0;

int typeFromRef(int t) {
    switch (t) {
    case bntPrimPoint: return PSGPU_PRIM_POINT;
    case bntPrimLine: return PSGPU_PRIM_LINE;
    case bntPrimCylinder: return PSGPU_PRIM_CYLINDER;
    case bntPrimDisc: return PSGPU_PRIM_DISC;
    case bntPrimRing: return PSGPU_PRIM_RING;
    case bntPrimCube: return PSGPU_PRIM_CUBE;
    case bntPrimNull: return PSGPU_PRIM_NULL;
    case bntOpUnion: return PSGPU_OP_UNION;
    case bntOpIntersect: return PSGPU_OP_INTERSECT;
    case bntOpDif: return PSGPU_OP_DIF;
    case bntOpSmoothDif: return PSGPU_OP_SMOOTHDIF;
    case bntOpBlend: return PSGPU_OP_BLEND;
    case bntOpRicciBlend: return PSGPU_OP_RICCI;
    case bntOpWarpTwist: return PSGPU_OP_WARP_TWIST;
    case bntOpWarpTaper: return PSGPU_OP_WARP_TAPER;
    case bntOpWarpBend: return PSGPU_OP_WARP_BEND;
    case bntOpWarpShear: return PSGPU_OP_WARP_SHEAR;
    }
    return -1;
}
int typeToRef(int t) {
    switch (t) {
    case PSGPU_PRIM_POINT: return bntPrimPoint;
    case PSGPU_PRIM_LINE: return bntPrimLine;
    case PSGPU_PRIM_CYLINDER: return bntPrimCylinder;
    case PSGPU_PRIM_DISC: return bntPrimDisc;
    case PSGPU_PRIM_RING: return bntPrimRing;
    case PSGPU_PRIM_CUBE: return bntPrimCube;
    case PSGPU_PRIM_NULL: return bntPrimNull;
    case PSGPU_OP_UNION: return bntOpUnion;
    case PSGPU_OP_INTERSECT: return bntOpIntersect;
    case PSGPU_OP_DIF: return bntOpDif;
    case PSGPU_OP_SMOOTHDIF: return bntOpSmoothDif;
    case PSGPU_OP_BLEND: return bntOpBlend;
    case PSGPU_OP_RICCI: return bntOpRicciBlend;
    case PSGPU_OP_WARP_TWIST: return bntOpWarpTwist;
    case PSGPU_OP_WARP_TAPER: return bntOpWarpTaper;
    case PSGPU_OP_WARP_BEND: return bntOpWarpBend;
    case PSGPU_OP_WARP_SHEAR: return bntOpWarpShear;
    }
    return -1;
}

// SOA binary tree -> flat psgpu_tree (ops first: node id = op id; prims follow: node id = ctOps + prim id)
struct FlatTree {
    std::vector<int32_t> type, first, nk, child;
    std::vector<float> pos, dir, res, color, xform;
    std::vector<uint8_t> hasxf;
    std::vector<float> boxLo, boxHi;  // the caller's operator boxes (SOABlobOps.vBox*, written by PrepareBBoxes)
    psgpu_tree t{};
};

int buildTree(const SOABlobPrims& P, const SOABlobPrimMatrices& M, const SOABlobOps& O, FlatTree& F) {
    const int nO = (int)O.ctOps, nP = (int)P.ctPrims;
    if (nP <= 0 || nP > MAX_TREE_NODES || nO > MAX_TREE_NODES) return RET_PARAM_ERROR;
    int n = nO + nP;
    int root = 0;
    int blendAll = -1;
    if (nO == 0 && nP > 1) { n += 1; blendAll = nO + nP; root = blendAll; }  // no operators: sum of all primitives (cpp:1356-1368)
    F.type.assign(n, 0); F.first.assign(n, 0); F.nk.assign(n, 0);
    F.pos.assign((size_t)n * 3, 0.f); F.dir.assign((size_t)n * 3, 0.f); F.res.assign((size_t)n * 4, 0.f); F.color.assign((size_t)n * 3, 0.f);
    F.xform.assign((size_t)n * 12, 0.f); F.hasxf.assign(n, 0);
    F.boxLo.assign((size_t)n * 3, 0.f); F.boxHi.assign((size_t)n * 3, 0.f);
    for (int i = 0; i < nO; i++) {
        F.boxLo[i * 3] = O.vBoxLoX[i]; F.boxLo[i * 3 + 1] = O.vBoxLoY[i]; F.boxLo[i * 3 + 2] = O.vBoxLoZ[i];
        F.boxHi[i * 3] = O.vBoxHiX[i]; F.boxHi[i * 3 + 1] = O.vBoxHiY[i]; F.boxHi[i * 3 + 2] = O.vBoxHiZ[i];
        const int ty = typeFromRef(O.opType[i]);
        if (ty < 0) return RET_INVALID_BVH;
        F.type[i] = ty;
        F.first[i] = (int)F.child.size();
        F.nk[i] = 2;
        const int lop = (O.opChildKind[i] & 2) >> 1, rop = O.opChildKind[i] & 1;
        F.child.push_back(lop ? O.opLeftChild[i] : nO + O.opLeftChild[i]);
        F.child.push_back(rop ? O.opRightChild[i] : nO + O.opRightChild[i]);
        F.res[i * 4] = O.resX[i]; F.res[i * 4 + 1] = O.resY[i]; F.res[i * 4 + 2] = O.resZ[i]; F.res[i * 4 + 3] = O.resW[i];
    }
    for (int i = 0; i < nP; i++) {
        const int id = nO + i;
        const int ty = typeFromRef(P.skeletType[i]);
        if (ty < 0) return PSGPU_ERROR_UNSUPPORTED_NODE;
        F.type[id] = ty;
        F.pos[id * 3] = P.posX[i]; F.pos[id * 3 + 1] = P.posY[i]; F.pos[id * 3 + 2] = P.posZ[i];
        F.dir[id * 3] = P.dirX[i]; F.dir[id * 3 + 1] = P.dirY[i]; F.dir[id * 3 + 2] = P.dirZ[i];
        F.res[id * 4] = P.resX[i]; F.res[id * 4 + 1] = P.resY[i]; F.res[id * 4 + 2] = P.resZ[i];
        F.color[id * 3] = P.colorX[i]; F.color[id * 3 + 1] = P.colorY[i]; F.color[id * 3 + 2] = P.colorZ[i];
        const int im = P.idxMatrix[i];
        if (im != 0) {
            if (im >= (int)M.count) return RET_PARAM_ERROR;
            memcpy(&F.xform[(size_t)id * 12], &M.matrix[im * PRIM_MATRIX_STRIDE], 12 * sizeof(float));
            F.hasxf[id] = 1;
        }
    }
    if (blendAll >= 0) {
        F.type[blendAll] = PSGPU_OP_BLEND;
        F.first[blendAll] = (int)F.child.size();
        F.nk[blendAll] = nP;
        for (int i = 0; i < nP; i++) F.child.push_back(nO + i);
    } else if (nO == 0) root = 0;
    if (F.child.empty()) F.child.push_back(0);
    F.t.n_nodes = n; F.t.root = root;
    F.t.type = F.type.data(); F.t.first_child = F.first.data(); F.t.n_children = F.nk.data(); F.t.child = F.child.data();
    F.t.pos = F.pos.data(); F.t.dir = F.dir.data(); F.t.res = F.res.data(); F.t.color = F.color.data();
    F.t.xform = F.xform.data(); F.t.has_xform = F.hasxf.data();
    F.t.op_box_lo = F.boxLo.data(); F.t.op_box_hi = F.boxHi.data();
    return RET_SUCCESS;
}

inline int nOpsOf(const SOABlobOps& O) { return (int)O.ctOps; }

int ensureCtx() {
    if (g_ctx) return RET_SUCCESS;
    return psgpu_create(0, &g_ctx);
}

int uploadScene(const SOABlobPrims& P, const SOABlobPrimMatrices& M, const SOABlobOps& O) {
    FlatTree F;
    int rc = buildTree(P, M, O, F);
    if (rc != RET_SUCCESS) return rc;
    rc = ensureCtx();
    if (rc != RET_SUCCESS) return rc;
    psgpu_program* prog = nullptr;
    // exact association of the binary tree is kept (no PSGPU_LIN_REASSOCIATE): with <= 128 nodes the stack fits.
    // PSGPU_FACADE_AS_SHIPPED=1: also the reference's prune of operators deeper than 3 levels by the caller's operator boxes
    // (PSGPU_LIN_AS_SHIPPED, psgpu.h) - field values, positions and indices of the SHIPPED reference; default: every contribution kept
    const char* shipped = getenv("PSGPU_FACADE_AS_SHIPPED");
    const int base = (shipped && atoi(shipped) != 0 && nOpsOf(O) > 0) ? PSGPU_LIN_AS_SHIPPED : 0;
    rc = psgpu_linearize(&F.t, PSGPU_SEM_SIMD, base, &prog);
    if (rc == PSGPU_NOT_ENOUGH_MEM && base == 0) rc = psgpu_linearize(&F.t, PSGPU_SEM_SIMD, PSGPU_LIN_REASSOCIATE, &prog);
    if (rc != RET_SUCCESS) return rc;
    rc = psgpu_upload_program(g_ctx, prog);
    psgpu_program_free(prog);
    return rc;
}

}  // namespace

U32 CountMPUNeeded(float cellsize, const svec3f& lo, const svec3f& hi) {
    const float l[3] = {lo.x, lo.y, lo.z}, h[3] = {hi.x, hi.y, hi.z};
    return psgpu_count_mpus(cellsize, l, h);
}

int PrepareBBoxes(float cellsize, SOABlobPrims& P, SOABlobBoxMatrices& B, SOABlobOps& O) {
    (void)cellsize;
    if (P.ctPrims == 0) return RET_PARAM_ERROR;
    const float iso = ISO_DIST + 5.0f * MIN_CELL_SIZE;  // ComputeWyvillIsoDistance(ISO_VALUE) + 5*MIN_CELL_SIZE, cpp:64-65
    svec3f bl(0, 0, 0), bh(0, 0, 0);
    for (U32 i = 0; i < P.ctPrims; i++) {
        const float p[3] = {P.posX[i], P.posY[i], P.posZ[i]}, d[3] = {P.dirX[i], P.dirY[i], P.dirZ[i]};
        float lo[3] = {P.vPrimBoxLoX[i], P.vPrimBoxLoY[i], P.vPrimBoxLoZ[i]}, hi[3] = {P.vPrimBoxHiX[i], P.vPrimBoxHiY[i], P.vPrimBoxHiZ[i]};
        switch (P.skeletType[i]) {
        case bntPrimPoint:
            for (int a = 0; a < 3; a++) { lo[a] = p[a] - iso; hi[a] = p[a] + iso; }
            break;
        case bntPrimLine:
            for (int a = 0; a < 3; a++) { const float e = iso * 1.0f + (3.0f * iso) * (d[a] - p[a]); lo[a] = p[a] - e; hi[a] = d[a] + e; }
            break;
        case bntPrimRing: case bntPrimDisc: {
            const float radius = P.resX[i] + iso;
            for (int a = 0; a < 3; a++) { const float e = (radius + iso) * (1.0f - d[a]) + iso * d[a]; lo[a] = p[a] - e; hi[a] = p[a] + e; }
        } break;
        case bntPrimCylinder:
            for (int a = 0; a < 3; a++) {
                const float s1 = p[a] + P.resY[i] * d[a];
                const float e = (iso + P.resX[i]) * 1.0f + (0.5f * iso) * d[a];
                lo[a] = p[a] - e; hi[a] = s1 + e;
            }
            break;
        case bntPrimCube: {
            const float side = P.resX[i] + iso;
            for (int a = 0; a < 3; a++) { lo[a] = p[a] - side; hi[a] = p[a] + side; }
        } break;
        default: break;
        }
        const int im = P.idxMatrix[i];
        if (im != 0 && im < (int)B.count) {  // cpp:200-218: the two corners through the forward 4x4; mat4Transform reads the
            // 16 floats as e[col][row] with the translation in e[3][0..2] (PS_MATRIX4.h:199-209)
            const float* m = &B.matrix[im * BOX_MATRIX_STRIDE];
            float l2[3], h2[3];
            for (int a = 0; a < 3; a++) {
                l2[a] = (m[0 * 4 + a] * lo[0] + m[1 * 4 + a] * lo[1] + m[2 * 4 + a] * lo[2]) + m[3 * 4 + a];
                h2[a] = (m[0 * 4 + a] * hi[0] + m[1 * 4 + a] * hi[1] + m[2 * 4 + a] * hi[2]) + m[3 * 4 + a];
            }
            for (int a = 0; a < 3; a++) { lo[a] = l2[a]; hi[a] = h2[a]; }
        }
        P.vPrimBoxLoX[i] = lo[0]; P.vPrimBoxLoY[i] = lo[1]; P.vPrimBoxLoZ[i] = lo[2];
        P.vPrimBoxHiX[i] = hi[0]; P.vPrimBoxHiY[i] = hi[1]; P.vPrimBoxHiZ[i] = hi[2];
        if (i == 0) { bl = svec3f(lo[0], lo[1], lo[2]); bh = svec3f(hi[0], hi[1], hi[2]); }
        else {
            bl = svec3f(std::fmin(bl.x, lo[0]), std::fmin(bl.y, lo[1]), std::fmin(bl.z, lo[2]));
            bh = svec3f(std::fmax(bh.x, hi[0]), std::fmax(bh.y, hi[1]), std::fmax(bh.z, hi[2]));
        }
    }
    P.bboxLo = bl; P.bboxHi = bh;
    if (O.ctOps > 0) {  // cpp:238-307: operator box = union of children, post-order
        std::vector<int> stk; std::vector<U8> done(O.ctOps, 0);
        stk.push_back(0);
        while (!stk.empty()) {
            const int op = stk.back(), lc = O.opLeftChild[op], rc = O.opRightChild[op];
            const int lop = (O.opChildKind[op] & 2) >> 1, rop = O.opChildKind[op] & 1;
            if (!((lop && !done[lc]) || (rop && !done[rc]))) {
                stk.pop_back();
                const float ll[3] = {lop ? O.vBoxLoX[lc] : P.vPrimBoxLoX[lc], lop ? O.vBoxLoY[lc] : P.vPrimBoxLoY[lc], lop ? O.vBoxLoZ[lc] : P.vPrimBoxLoZ[lc]};
                const float lh[3] = {lop ? O.vBoxHiX[lc] : P.vPrimBoxHiX[lc], lop ? O.vBoxHiY[lc] : P.vPrimBoxHiY[lc], lop ? O.vBoxHiZ[lc] : P.vPrimBoxHiZ[lc]};
                const float rl[3] = {rop ? O.vBoxLoX[rc] : P.vPrimBoxLoX[rc], rop ? O.vBoxLoY[rc] : P.vPrimBoxLoY[rc], rop ? O.vBoxLoZ[rc] : P.vPrimBoxLoZ[rc]};
                const float rh[3] = {rop ? O.vBoxHiX[rc] : P.vPrimBoxHiX[rc], rop ? O.vBoxHiY[rc] : P.vPrimBoxHiY[rc], rop ? O.vBoxHiZ[rc] : P.vPrimBoxHiZ[rc]};
                O.vBoxLoX[op] = std::fmin(ll[0], rl[0]); O.vBoxLoY[op] = std::fmin(ll[1], rl[1]); O.vBoxLoZ[op] = std::fmin(ll[2], rl[2]);
                O.vBoxHiX[op] = std::fmax(lh[0], rh[0]); O.vBoxHiY[op] = std::fmax(lh[1], rh[1]); O.vBoxHiZ[op] = std::fmax(lh[2], rh[2]);
                done[op] = 1;
            } else {
                if (lop && !done[lc]) stk.push_back(lc);
                if (rop && !done[rc]) stk.push_back(rc);
            }
        }
    }
    return RET_SUCCESS;
}

int Polygonize(float cellsize, const SOABlobPrims& P, const SOABlobPrimMatrices& M, const SOABlobOps& O, PolyMPUs& out, MPUSTATS* stats) {
    if (P.ctPrims == 0) return RET_PARAM_ERROR;
    std::lock_guard<std::mutex> guard(g_lock);
    const bool timing = getenv("PSGPU_FACADE_TIMING") != nullptr;  // stage times of this call on stderr
    const auto tStart = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (timing) fprintf(stderr, "[PS_SimdPolyCompat] %-10s %8.3f ms\n", what,
                            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tStart).count());
    };
    int rc = uploadScene(P, M, O);
    if (rc != RET_SUCCESS) return rc;
    lap("upload");
    psgpu_params prm;
    memset(&prm, 0, sizeof(prm));
    prm.cellsize = cellsize;
    prm.bbox_lo[0] = P.bboxLo.x; prm.bbox_lo[1] = P.bboxLo.y; prm.bbox_lo[2] = P.bboxLo.z;
    prm.bbox_hi[0] = P.bboxHi.x; prm.bbox_hi[1] = P.bboxHi.y; prm.bbox_hi[2] = P.bboxHi.z;
    prm.root_mode = PSGPU_ROOT_FOURSAMPLE; prm.coord_mode = PSGPU_COORD_MPU_LOCAL; prm.output = PSGPU_OUT_MPU;
    prm.cull = PSGPU_CULL_INTERVAL; prm.want_colors = 1; prm.want_normals = 1;
    psgpu_mesh mesh;
    rc = psgpu_polygonize(g_ctx, &prm, &mesh);
    if (rc != RET_SUCCESS) return rc;
    lap("polygonize");
    // MPU lattice exactly as cpp:328-371
    const float side = cellsize * (float)(GRID_DIM - 1);
    const int nx = mesh.lattice[0], ny = mesh.lattice[1], nz = mesh.lattice[2];
    const long long lattice = (long long)nx * ny * nz;
    if (lattice >= MAX_MPU_COUNT) printf("WARNING! MAX MPU Count Reached! Total = %lld MAX = %d\n", lattice, MAX_MPU_COUNT);
    out.ctMPUs = (U32)std::min<long long>(lattice, MAX_MPU_COUNT);
    const size_t nm = (size_t)mesh.n_mpus_out;
    const float *pos = nullptr, *nrm = nullptr, *col = nullptr;
    const uint32_t *tri = nullptr, *vb = nullptr, *tb = nullptr;
    const int32_t* ijk = nullptr;
    rc = psgpu_mesh_map_host(g_ctx, &pos, &nrm, &col, &tri, &ijk, &vb, &tb);
    if (rc != RET_SUCCESS) return rc;
    lap("download");
    psgpu_stats st;
    psgpu_get_stats(g_ctx, &st);
    // The writer touches ~3 cache lines of every MPU struct (21.5 KB apart) and copies the surface MPUs' arrays: memory-latency
    // work, split over a few host threads when the lattice is large (the reference fills PolyMPUs from all its TBB workers)
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const int nThreads = (int)std::min<size_t>(std::min(hw, 8u), 1 + (out.ctMPUs + 4 * nm) / 8192);
    std::atomic<int> overflow(0);
    auto work = [&](int w) {
        const size_t m0 = (size_t)out.ctMPUs * w / nThreads, m1 = (size_t)out.ctMPUs * (w + 1) / nThreads;
        for (size_t q = m0; q < m1; q++) {  // q = (i * ny + j) * nz + k: x outer, z inner
            const int k = (int)(q % nz), j = (int)((q / nz) % ny), i = (int)(q / ((size_t)nz * ny));
            MPU& m = out.vMPUs[q];
            m.bboxLo = svec3f(P.bboxLo.x + (float)i * side, P.bboxLo.y + (float)j * side, P.bboxLo.z + (float)k * side);
            m.ctVertices = 0; m.ctTriangles = 0; m.ctFieldEvals = 0;
            if (stats) { stats[q].idxThread = -1; stats[q].bIntersected = 0; stats[q].msStart = 0; stats[q].msEnd = st.ms_total; }
        }
        // ranks follow the candidate order, so both base tables are monotone and closed by the totals; a thread takes a
        // contiguous range of ranks (their MPU structs may lie in another thread's lattice range: disjoint fields)
        for (size_t r = nm * w / nThreads; r < nm * (w + 1) / nThreads; r++) {
            const uint32_t t0 = tb[r], t1 = tb[r + 1], v0 = vb[r], v1 = vb[r + 1];
            const long long idx = ((long long)ijk[r * 3] * ny + ijk[r * 3 + 1]) * nz + ijk[r * 3 + 2];
            if (idx < 0 || idx >= (long long)out.ctMPUs) continue;
            MPU& m = out.vMPUs[idx];
            const uint32_t cv = v1 - v0, ct = t1 - t0;
            if (cv > MAX_MPU_VERTEX_COUNT || ct > MAX_MPU_TRIANGLE_COUNT) { overflow = 1; continue; }
            memcpy(m.vPos, &pos[(size_t)v0 * 3], (size_t)cv * 12);
            memcpy(m.vNorm, &nrm[(size_t)v0 * 3], (size_t)cv * 12);
            memcpy(m.vColor, &col[(size_t)v0 * 3], (size_t)cv * 12);
            const uint32_t* src = &tri[(size_t)t0 * 3];
            for (uint32_t q = 0; q < ct * 3; q++) m.triangles[q] = (U16)src[q];
        }
    };
    auto counts = [&](int w) {  // after every thread's lattice pass: the counters of the surface MPUs
        for (size_t r = nm * w / nThreads; r < nm * (w + 1) / nThreads; r++) {
            const long long idx = ((long long)ijk[r * 3] * ny + ijk[r * 3 + 1]) * nz + ijk[r * 3 + 2];
            if (idx < 0 || idx >= (long long)out.ctMPUs) continue;
            MPU& m = out.vMPUs[idx];
            m.ctVertices = (U16)(vb[r + 1] - vb[r]); m.ctTriangles = (U16)(tb[r + 1] - tb[r]);
            m.ctFieldEvals = 128;  // SIMD evaluations of the cache phase, the reference's unit (cpp:594)
            if (stats) { stats[idx].idxThread = (int)r; stats[idx].bIntersected = 1; }
        }
    };
    auto fan = [&](const std::function<void(int)>& f) {
        if (nThreads <= 1) { f(0); return; }
        std::vector<std::thread> pool;
        for (int w = 1; w < nThreads; w++) pool.emplace_back(f, w);
        f(0);
        for (auto& t : pool) t.join();
    };
    fan(work);
    if (overflow) return RET_NOT_ENOUGH_MEM;  // the reference overruns the MPU's arrays here
    fan(counts);
    g_processed = out.ctMPUs; g_crossed = (U32)nm;
    lap("fill");
    return RET_SUCCESS;
}

void PrintThreadResults(int ctAttempts, U32* lpThreadProcessed, U32* lpThreadCrossed) {
    if (ctAttempts <= 0) ctAttempts = 1;
    if (lpThreadProcessed) lpThreadProcessed[0] = g_processed / ctAttempts;
    if (lpThreadCrossed) lpThreadCrossed[0] = g_crossed / ctAttempts;
    printf("Thread#  %d, Processed MPUs %d, Crossed MPUs %d \n", 1, g_processed / ctAttempts, g_crossed / ctAttempts);
    g_processed = g_crossed = 0;
}

int FieldValues(const SOABlobPrims& P, const SOABlobPrimMatrices& M, const SOABlobOps& O, const float* xyz, int n, float* outField) {
    std::lock_guard<std::mutex> guard(g_lock);
    int rc = uploadScene(P, M, O);
    if (rc != RET_SUCCESS) return rc;
    return psgpu_field_eval(g_ctx, xyz, n, outField);
}

int RootsNewtonRaphson(const SOABlobPrims& P, const SOABlobPrimMatrices& M, const SOABlobOps& O, const float* p1, const float* p2, const float* fp1,
                       const float* fp2, int n, float* outRoot, float* outField, int* outEvals, float target_field, int iterations) {
    std::lock_guard<std::mutex> guard(g_lock);
    int rc = uploadScene(P, M, O);
    if (rc != RET_SUCCESS) return rc;
    return psgpu_newton_roots(g_ctx, p1, p2, fp1, fp2, n, target_field, iterations, outRoot, outField, outEvals);
}

}  // namespace SIMDPOLY
}  // namespace PS

// ---- SimdPoly wrapper -----------------------------------------------------------------------------------
using namespace PS::SIMDPOLY;

SimdPoly::SimdPoly() : m_polyMPUs(nullptr) { reset(); }
SimdPoly::~SimdPoly() { delete m_polyMPUs; }

void SimdPoly::reset() {  // PS_HighPerformanceRender.cpp:33-40
    m_blobPrims.ctPrims = 0; m_blobOps.ctOps = 0; m_blobPrimMatrices.count = 0; m_blobBoxMatrices.count = 0;
    if (m_polyMPUs) m_polyMPUs->ctMPUs = 0;
}

// Pre-order walk: operators and primitives numbered in visit order, exactly 2 children per operator
// (PS_HighPerformanceRender.cpp:42-364; errors -1 prim overflow, -2 operator overflow, -3 non-binary operator).
int SimdPoly::linearizeBlobTree(const psgpu_tree* t) {
    reset();
    if (!t || t->n_nodes <= 0) return RET_PARAM_ERROR;
    static const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    m_blobPrimMatrices.count = 1; m_blobBoxMatrices.count = 1;
    memcpy(m_blobPrimMatrices.matrix, ident, 12 * sizeof(float));
    memcpy(m_blobBoxMatrices.matrix, ident, 16 * sizeof(float));
    struct Walker {
        SimdPoly* self; const psgpu_tree* t; SOABlobPrims& P; SOABlobOps& O; SOABlobPrimMatrices& M; int err;
        int walk(int n, int& isOp) {
            const int ty = t->type[n];
            isOp = !PSGPU_IS_PRIM(ty);
            if (isOp) {
                if (O.ctOps >= MAX_TREE_NODES) { err = -2; return -2; }
                const int id = (int)O.ctOps++;
                O.opType[id] = (U8)PS::SIMDPOLY::typeToRef(ty);
                if (t->n_children[n] != 2) { err = -3; return -3; }
                int k0, k1;
                const int c0 = walk(t->child[t->first_child[n]], k0);
                if (err) return err;
                const int c1 = walk(t->child[t->first_child[n] + 1], k1);
                if (err) return err;
                O.opLeftChild[id] = (U8)c0; O.opRightChild[id] = (U8)c1; O.opChildKind[id] = (U8)(k0 * 2 + k1);
                O.resX[id] = t->res[n * 4]; O.resY[id] = t->res[n * 4 + 1]; O.resZ[id] = t->res[n * 4 + 2]; O.resW[id] = t->res[n * 4 + 3];
                return id;
            }
            if (P.ctPrims >= MAX_TREE_NODES) { err = -1; return -1; }
            const int id = (int)P.ctPrims++;
            P.skeletType[id] = (U8)PS::SIMDPOLY::typeToRef(ty);
            P.posX[id] = t->pos[n * 3]; P.posY[id] = t->pos[n * 3 + 1]; P.posZ[id] = t->pos[n * 3 + 2];
            P.dirX[id] = t->dir[n * 3]; P.dirY[id] = t->dir[n * 3 + 1]; P.dirZ[id] = t->dir[n * 3 + 2];
            P.resX[id] = t->res[n * 4]; P.resY[id] = t->res[n * 4 + 1]; P.resZ[id] = t->res[n * 4 + 2];
            P.colorX[id] = t->color[n * 3]; P.colorY[id] = t->color[n * 3 + 1]; P.colorZ[id] = t->color[n * 3 + 2];
            P.idxMatrix[id] = 0;
            if (t->xform && t->has_xform && t->has_xform[n]) {
                const int im = (int)M.count++;
                P.idxMatrix[id] = (U8)im;
                memcpy(&M.matrix[im * PRIM_MATRIX_STRIDE], &t->xform[n * 12], 12 * sizeof(float));
            }
            return id;
        }
    } w{this, t, m_blobPrims, m_blobOps, m_blobPrimMatrices, 0};
    int isOp;
    const int id = w.walk(t->root, isOp);
    if (w.err) return w.err;
    m_blobBoxMatrices.count = m_blobPrimMatrices.count;
    int rc = PrepareBBoxes(DEFAULT_CELL_SIZE, m_blobPrims, m_blobBoxMatrices, m_blobOps);
    if (rc != RET_SUCCESS) return rc;
    return id;
}

int SimdPoly::run(float cellsize) {  // PS_HighPerformanceRender.cpp:373-376
    if (!m_polyMPUs) m_polyMPUs = new PolyMPUs();
    return PS::SIMDPOLY::Polygonize(cellsize, m_blobPrims, m_blobPrimMatrices, m_blobOps, *m_polyMPUs);
}
