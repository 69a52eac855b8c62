// TEST INFRASTRUCTURE ONLY. C wrapper around the UNMODIFIED reference SIMD polygonizer
// (/root/reference/Parsip100/PS_SimdPoly/include/PS_Polygonizer.{h,cpp}), compiled in place by
// oracle/Makefile into oracle/_ref/. Nothing here is shipped or measured as product; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
//
// It exposes exactly the reference's public entry points (PS_Polygonizer.h:268-326, 384-393):
// PrepareBBoxes, CountMPUNeeded, Polygonize, FieldComputer::{fieldValue, fieldValueAndColor, normal},
// plus a streamed driver that runs the reference's own CMPUProcessor (PS_Polygonizer.cpp:441-470)
// over caller-chosen MPU sub-blocks so that grids beyond MAX_MPU_COUNT (24000) can be timed
// without the 516 MB static PolyMPUs.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "PS_Polygonizer.h"
#include "../include/psgpu_types.h"

using namespace PS;
using namespace PS::SIMDPOLY;

namespace PS { namespace SIMDPOLY {
typedef tbb::enumerable_thread_specific<FieldComputer> ThreadFieldComputer_;
extern ThreadFieldComputer_ g_threadFieldComputer;   // PS_Polygonizer.cpp:19
} }

struct psref_desc {
    int32_t n_prims, n_ops, n_mats;
    const int32_t* prim_type;   // psgpu_node_type
    const float* prim_pos;      // n*3
    const float* prim_dir;      // n*3
    const float* prim_res;      // n*3
    const float* prim_color;    // n*3
    const int32_t* prim_mat;    // matrix slot, 0 = identity
    const int32_t* op_type;     // psgpu_node_type
    const int32_t* op_left;
    const int32_t* op_right;
    const int32_t* op_kind;     // bit1: left is op, bit0: right is op
    const float* op_res;        // n*4
    const float* mats_bwd;      // n_mats*12 (slot 0 = identity)
    const float* mats_fwd;      // n_mats*16 (slot 0 = identity)
};

struct psref_scene {
    SOABlobPrims* prims;
    SOABlobOps* ops;
    SOABlobPrimMatrices* primMats;
    SOABlobBoxMatrices* boxMats;
    PolyMPUs* poly;          // lazily allocated (516 MB of untouched pages)
    FieldComputer* fc;
    // streamed results
    std::vector<float> sPos, sNrm, sCol;
    std::vector<uint32_t> sTri;
    std::vector<int32_t> sMpuIdx;   // lattice index (i,j,k) of every MPU that produced triangles
    std::vector<uint32_t> sMpuCounts; // (nv, nt) per such MPU
};

template <typename T> static T* alloc_aligned() {
    void* p = 0;
    if (posix_memalign(&p, 64, sizeof(T)) != 0) return 0;
    memset(p, 0, sizeof(T));
    return (T*)p;
}

static int map_type(int t) {
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

extern "C" {

void psref_build_info(int64_t* out) {
    out[0] = MAX_TREE_NODES;
    out[1] = MAX_MPU_COUNT;
    out[2] = (int64_t)sizeof(MPU);
    out[3] = (int64_t)sizeof(PolyMPUs);
    out[4] = PS_SIMD_FLEN;
    out[5] = GRID_DIM;
    out[6] = (int64_t)sizeof(SOABlobPrims);
    out[7] = (int64_t)sizeof(SOABlobOps);
    out[8] = (int64_t)sizeof(SOABlobPrimMatrices);
    out[9] = (int64_t)sizeof(SOABlobBoxMatrices);
    out[10] = MAX_MPU_VERTEX_COUNT;
    out[11] = MAX_MPU_TRIANGLE_COUNT;
}

void psref_set_threads(int n) { tbbshim::set_num_threads(n); }

psref_scene* psref_scene_create(const psref_desc* d) {
    if (d->n_prims <= 0 || d->n_prims > MAX_TREE_NODES || d->n_ops > MAX_TREE_NODES ||
        d->n_mats > MAX_MATRIX_COUNT) return 0;
    psref_scene* s = new psref_scene();
    s->prims = alloc_aligned<SOABlobPrims>();
    s->ops = alloc_aligned<SOABlobOps>();
    s->primMats = alloc_aligned<SOABlobPrimMatrices>();
    s->boxMats = alloc_aligned<SOABlobBoxMatrices>();
    s->poly = 0;
    s->fc = 0;
    for (int i = 0; i < d->n_prims; i++) {
        int t = map_type(d->prim_type[i]);
        if (t < 0) { return 0; }
        s->prims->skeletType[i] = (U8)t;
        s->prims->idxMatrix[i] = d->prim_mat ? d->prim_mat[i] : 0;
        s->prims->posX[i] = d->prim_pos[i*3+0]; s->prims->posY[i] = d->prim_pos[i*3+1]; s->prims->posZ[i] = d->prim_pos[i*3+2];
        s->prims->dirX[i] = d->prim_dir[i*3+0]; s->prims->dirY[i] = d->prim_dir[i*3+1]; s->prims->dirZ[i] = d->prim_dir[i*3+2];
        s->prims->resX[i] = d->prim_res[i*3+0]; s->prims->resY[i] = d->prim_res[i*3+1]; s->prims->resZ[i] = d->prim_res[i*3+2];
        s->prims->colorX[i] = d->prim_color[i*3+0]; s->prims->colorY[i] = d->prim_color[i*3+1]; s->prims->colorZ[i] = d->prim_color[i*3+2];
    }
    s->prims->ctPrims = (U32)d->n_prims;
    for (int i = 0; i < d->n_ops; i++) {
        int t = map_type(d->op_type[i]);
        if (t < 0) { return 0; }
        s->ops->opType[i] = (U8)t;
        s->ops->opLeftChild[i] = d->op_left[i];
        s->ops->opRightChild[i] = d->op_right[i];
        s->ops->opChildKind[i] = (U8)d->op_kind[i];
        s->ops->resX[i] = d->op_res[i*4+0]; s->ops->resY[i] = d->op_res[i*4+1];
        s->ops->resZ[i] = d->op_res[i*4+2]; s->ops->resW[i] = d->op_res[i*4+3];
    }
    s->ops->ctOps = (U32)d->n_ops;
    int nm = d->n_mats > 0 ? d->n_mats : 1;
    static const float ident12[12] = {1,0,0,0, 0,1,0,0, 0,0,1,0};
    static const float ident16[16] = {1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1};
    memcpy(s->primMats->matrix, ident12, sizeof(ident12));
    memcpy(s->boxMats->matrix, ident16, sizeof(ident16));
    for (int m = 0; m < d->n_mats; m++) {
        memcpy(&s->primMats->matrix[m*PRIM_MATRIX_STRIDE], &d->mats_bwd[m*12], 12*sizeof(float));
        if (d->mats_fwd) memcpy(&s->boxMats->matrix[m*BOX_MATRIX_STRIDE], &d->mats_fwd[m*16], 16*sizeof(float));
    }
    s->primMats->count = (U32)nm;
    s->boxMats->count = (U32)nm;
    return s;
}

void psref_scene_destroy(psref_scene* s) {
    if (!s) return;
    free(s->prims); free(s->ops); free(s->primMats); free(s->boxMats);
    if (s->poly) free(s->poly);
    if (s->fc) { s->fc->~FieldComputer(); free(s->fc); }
    delete s;
}

int psref_prepare_bboxes(psref_scene* s, float cellsize) {
    return PrepareBBoxes(cellsize, *s->prims, *s->boxMats, *s->ops);
}

void psref_get_bbox(const psref_scene* s, float* lo, float* hi) {
    lo[0] = s->prims->bboxLo.x; lo[1] = s->prims->bboxLo.y; lo[2] = s->prims->bboxLo.z;
    hi[0] = s->prims->bboxHi.x; hi[1] = s->prims->bboxHi.y; hi[2] = s->prims->bboxHi.z;
}
void psref_set_bbox(psref_scene* s, const float* lo, const float* hi) {
    s->prims->bboxLo = svec3f(lo[0], lo[1], lo[2]);
    s->prims->bboxHi = svec3f(hi[0], hi[1], hi[2]);
}
void psref_get_prim_boxes(const psref_scene* s, float* out6) {
    for (U32 i = 0; i < s->prims->ctPrims; i++) {
        out6[i*6+0] = s->prims->vPrimBoxLoX[i]; out6[i*6+1] = s->prims->vPrimBoxLoY[i]; out6[i*6+2] = s->prims->vPrimBoxLoZ[i];
        out6[i*6+3] = s->prims->vPrimBoxHiX[i]; out6[i*6+4] = s->prims->vPrimBoxHiY[i]; out6[i*6+5] = s->prims->vPrimBoxHiZ[i];
    }
}
void psref_get_op_boxes(const psref_scene* s, float* out6) {
    for (U32 i = 0; i < s->ops->ctOps; i++) {
        out6[i*6+0] = s->ops->vBoxLoX[i]; out6[i*6+1] = s->ops->vBoxLoY[i]; out6[i*6+2] = s->ops->vBoxLoZ[i];
        out6[i*6+3] = s->ops->vBoxHiX[i]; out6[i*6+4] = s->ops->vBoxHiY[i]; out6[i*6+5] = s->ops->vBoxHiZ[i];
    }
}
// Pads every operator box by `pad` (harness-side neutralisation of the depth>3 op-box prune).
void psref_pad_op_boxes(psref_scene* s, float pad) {
    for (U32 i = 0; i < s->ops->ctOps; i++) {
        s->ops->vBoxLoX[i] -= pad; s->ops->vBoxLoY[i] -= pad; s->ops->vBoxLoZ[i] -= pad;
        s->ops->vBoxHiX[i] += pad; s->ops->vBoxHiY[i] += pad; s->ops->vBoxHiZ[i] += pad;
    }
}

uint32_t psref_count_mpus(const psref_scene* s, float cellsize) {
    return CountMPUNeeded(cellsize, s->prims->bboxLo, s->prims->bboxHi);
}

static FieldComputer* get_fc(psref_scene* s) {
    if (s->fc) { s->fc->~FieldComputer(); free(s->fc); s->fc = 0; }
    void* p = 0;
    if (posix_memalign(&p, 64, sizeof(FieldComputer)) != 0) return 0;
    s->fc = new (p) FieldComputer();
    memcpy(&s->fc->m_blobPrims, s->prims, sizeof(SOABlobPrims));
    memcpy(&s->fc->m_blobPrimMatrices, s->primMats, sizeof(SOABlobPrimMatrices));
    memcpy(&s->fc->m_blobOps, s->ops, sizeof(SOABlobOps));
    return s->fc;
}

// FieldComputer::fieldValue at n points, PS_SIMD_FLEN per call (PS_Polygonizer.cpp:1184-1376).
// Points are grouped in input order; the last group is padded with its last point.
void psref_field(psref_scene* s, int64_t n, const float* xyz, float* out) {
    FieldComputer* fc = get_fc(s);
    const int L = PS_SIMD_FLEN;
    float PS_SIMD_ALIGN(ax[PS_SIMD_FLEN]); float PS_SIMD_ALIGN(ay[PS_SIMD_FLEN]); float PS_SIMD_ALIGN(az[PS_SIMD_FLEN]);
    for (int64_t b = 0; b < n; b += L) {
        for (int l = 0; l < L; l++) {
            int64_t i = b + l < n ? b + l : n - 1;
            ax[l] = xyz[i*3]; ay[l] = xyz[i*3+1]; az[l] = xyz[i*3+2];
        }
        Float_ X(ax), Y(ay), Z(az), F;
        fc->fieldValue(X, Y, Z, F);
        for (int l = 0; l < L && b + l < n; l++) out[b + l] = F[l];
    }
}

// What process_cells_simd does per new vertex (PS_Polygonizer.cpp:764-782): broadcast the point to all
// lanes, fieldValueAndColor, then normal with NORMAL_DELTA; lane 0 is kept.
void psref_field_color_normal(psref_scene* s, int64_t n, const float* xyz, float* f, float* col, float* nrm) {
    FieldComputer* fc = get_fc(s);
    for (int64_t i = 0; i < n; i++) {
        Float_ X(xyz[i*3]), Y(xyz[i*3+1]), Z(xyz[i*3+2]), F, cx, cy, cz, nx, ny, nz;
        fc->fieldValueAndColor(X, Y, Z, F, cx, cy, cz);
        fc->normal(X, Y, Z, F, nx, ny, nz, NORMAL_DELTA);
        f[i] = F[0];
        col[i*3] = cx[0]; col[i*3+1] = cy[0]; col[i*3+2] = cz[0];
        nrm[i*3] = nx[0]; nrm[i*3+1] = ny[0]; nrm[i*3+2] = nz[0];
    }
}

// FieldComputer::computeRootNewtonRaphson (PS_Polygonizer.cpp:1625-1681), called as declared; evals = its return value.
void psref_newton(psref_scene* s, int64_t n, const float* p1, const float* p2, const float* fp1, const float* fp2, float target,
                  int iterations, float* root, float* root_field, int32_t* evals) {
    FieldComputer* fc = get_fc(s);
    for (int64_t i = 0; i < n; i++) {
        svec3f a(p1[i*3], p1[i*3+1], p1[i*3+2]), b(p2[i*3], p2[i*3+1], p2[i*3+2]), out(0.0f, 0.0f, 0.0f);
        float of = 0.0f;
        evals[i] = fc->computeRootNewtonRaphson(a, b, fp1[i], fp2[i], out, of, target, iterations);
        root[i*3] = out.x; root[i*3+1] = out.y; root[i*3+2] = out.z;
        root_field[i] = of;
    }
}

// The reference's Polygonize (PS_Polygonizer.cpp:315-385), whole grid, into its own PolyMPUs.
// stats: [0] ctMPUs [1] MPUs with triangles [2] vertices [3] triangles [4] sum ctFieldEvals (SIMD evals)
int psref_polygonize(psref_scene* s, float cellsize, int64_t* stats) {
    if (!s->poly) {
        void* p = 0;
        if (posix_memalign(&p, 64, sizeof(PolyMPUs)) != 0) return RET_NOT_ENOUGH_MEM;
        s->poly = (PolyMPUs*)p;
    }
    // Polygonize only writes ctVertices/ctTriangles/ctFieldEvals for MPUs that survive its corner
    // test (cpp:518-519 returns before the init at :546-548), so the caller must zero them.
    for (size_t i = 0; i < MAX_MPU_COUNT; i++) {
        s->poly->vMPUs[i].ctVertices = 0; s->poly->vMPUs[i].ctTriangles = 0; s->poly->vMPUs[i].ctFieldEvals = 0;
    }
    int ret = Polygonize(cellsize, *s->prims, *s->primMats, *s->ops, *s->poly, NULL);
    if (stats) {
        int64_t crossed = 0, nv = 0, nt = 0, fe = 0;
        for (U32 i = 0; i < s->poly->ctMPUs; i++) {
            const MPU& m = s->poly->vMPUs[i];
            if (m.ctTriangles > 0) crossed++;
            nv += m.ctVertices; nt += m.ctTriangles; fe += m.ctFieldEvals;
        }
        stats[0] = s->poly->ctMPUs; stats[1] = crossed; stats[2] = nv; stats[3] = nt; stats[4] = fe;
    }
    return ret;
}

// Per-MPU access to the result of psref_polygonize. info: nv, nt, ctFieldEvals ; lo: bboxLo
void psref_mpu_info(const psref_scene* s, uint32_t idx, int32_t* info, float* lo) {
    const MPU& m = s->poly->vMPUs[idx];
    info[0] = m.ctVertices; info[1] = m.ctTriangles; info[2] = (int32_t)m.ctFieldEvals;
    lo[0] = m.bboxLo.x; lo[1] = m.bboxLo.y; lo[2] = m.bboxLo.z;
}
void psref_mpu_copy(const psref_scene* s, uint32_t idx, float* pos, float* nrm, float* col, uint16_t* tri) {
    const MPU& m = s->poly->vMPUs[idx];
    memcpy(pos, m.vPos, m.ctVertices * 3 * sizeof(float));
    memcpy(nrm, m.vNorm, m.ctVertices * 3 * sizeof(float));
    memcpy(col, m.vColor, m.ctVertices * 3 * sizeof(float));
    memcpy(tri, m.triangles, m.ctTriangles * 3 * sizeof(uint16_t));
}
// Concatenated mesh (vertices in MPU order, indices rebased to u32). Buffers sized from stats[2], stats[3].
void psref_mesh_concat(const psref_scene* s, float* pos, float* nrm, float* col, uint32_t* tri) {
    size_t vb = 0, tb = 0;
    for (U32 i = 0; i < s->poly->ctMPUs; i++) {
        const MPU& m = s->poly->vMPUs[i];
        memcpy(pos + vb*3, m.vPos, m.ctVertices * 3 * sizeof(float));
        memcpy(nrm + vb*3, m.vNorm, m.ctVertices * 3 * sizeof(float));
        memcpy(col + vb*3, m.vColor, m.ctVertices * 3 * sizeof(float));
        for (size_t t = 0; t < (size_t)m.ctTriangles * 3; t++) tri[tb*3 + t] = (uint32_t)(m.triangles[t] + vb);
        vb += m.ctVertices; tb += m.ctTriangles;
    }
}

// Streamed driver: the reference's CMPUProcessor + ThreadStartSetup + parallel_for exactly as
// Polygonize wires them (cpp:377-382), over the lattice sub-block [b0,e0) of the MPU grid that
// Polygonize would build from the scene bbox (cpp:328-371; same origin expression lo + idx*side).
// keep_mesh != 0 appends every produced MPU mesh to the scene's stream buffers.
// stats: [0] MPUs processed [1] with triangles [2] vertices [3] triangles [4] SIMD evals
//        [5..7] lattice dims
int psref_polygonize_stream(psref_scene* s, float cellsize, const int32_t* b0, const int32_t* e0,
                            int keep_mesh, int64_t* stats) {
    if (s->prims->ctPrims == 0) return RET_PARAM_ERROR;
    const int cellsPerMPU = GRID_DIM - 1;
    const float side = cellsize * (float)cellsPerMPU;
    svec3f lo = s->prims->bboxLo, hi = s->prims->bboxHi;
    svec3f all = vsub3f(hi, lo);
    int nc[3] = { (int)ceil(all.x / cellsize), (int)ceil(all.y / cellsize), (int)ceil(all.z / cellsize) };
    int nm[3];
    for (int a = 0; a < 3; a++) { nm[a] = nc[a] / cellsPerMPU; if (nc[a] % cellsPerMPU != 0) nm[a]++; }
    int B[3], E[3];
    for (int a = 0; a < 3; a++) {
        B[a] = b0 ? b0[a] : 0; E[a] = e0 ? e0[a] : nm[a];
        if (B[a] < 0) B[a] = 0; if (E[a] > nm[a]) E[a] = nm[a];
    }
    s->sPos.clear(); s->sNrm.clear(); s->sCol.clear(); s->sTri.clear(); s->sMpuIdx.clear(); s->sMpuCounts.clear();
    const size_t CHUNK = 8192;
    MPU* chunk = (MPU*)malloc(sizeof(MPU) * CHUNK);
    std::vector<int32_t> cidx(CHUNK * 3);
    if (!chunk) return RET_NOT_ENOUGH_MEM;
    int64_t nproc = 0, ncross = 0, nv = 0, nt = 0, fe = 0;
    ThreadStartSetup startSetup(s->prims, s->primMats, s->ops);
    size_t fill = 0;
    for (int i = B[0]; i < E[0]; i++) for (int j = B[1]; j < E[1]; j++) for (int k = B[2]; k <= E[2]; k++) {
        bool last = (i == E[0]-1 && j == E[1]-1 && k == E[2]);
        if (k < E[2]) {
            MPU& m = chunk[fill];
            m.bboxLo = vadd3f(lo, svec3f((float)i * side, (float)j * side, (float)k * side));
            m.ctVertices = 0; m.ctTriangles = 0; m.ctFieldEvals = 0;
            cidx[fill*3] = i; cidx[fill*3+1] = j; cidx[fill*3+2] = k;
            fill++;
        }
        if (fill == CHUNK || (last && fill > 0)) {
            CMPUProcessor body(cellsize, fill, chunk, NULL);
            tbb::parallel_for(blocked_range<size_t>(0, fill), body, tbb::auto_partitioner());
            for (size_t q = 0; q < fill; q++) {
                const MPU& m = chunk[q];
                nproc++; fe += m.ctFieldEvals;
                if (m.ctTriangles > 0) {
                    ncross++;
                    if (keep_mesh) {
                        uint32_t vb = (uint32_t)(s->sPos.size() / 3);
                        s->sPos.insert(s->sPos.end(), m.vPos, m.vPos + m.ctVertices*3);
                        s->sNrm.insert(s->sNrm.end(), m.vNorm, m.vNorm + m.ctVertices*3);
                        s->sCol.insert(s->sCol.end(), m.vColor, m.vColor + m.ctVertices*3);
                        for (int t = 0; t < m.ctTriangles*3; t++) s->sTri.push_back(vb + m.triangles[t]);
                        s->sMpuIdx.push_back(cidx[q*3]); s->sMpuIdx.push_back(cidx[q*3+1]); s->sMpuIdx.push_back(cidx[q*3+2]);
                        s->sMpuCounts.push_back(m.ctVertices); s->sMpuCounts.push_back(m.ctTriangles);
                    }
                }
                nv += m.ctVertices; nt += m.ctTriangles;
            }
            fill = 0;
        }
    }
    g_threadFieldComputer.clear();
    free(chunk);
    if (stats) {
        stats[0] = nproc; stats[1] = ncross; stats[2] = nv; stats[3] = nt; stats[4] = fe;
        stats[5] = nm[0]; stats[6] = nm[1]; stats[7] = nm[2];
    }
    return RET_SUCCESS;
}

int64_t psref_stream_counts(const psref_scene* s, int64_t* out) {
    out[0] = (int64_t)(s->sPos.size() / 3); out[1] = (int64_t)(s->sTri.size() / 3); out[2] = (int64_t)(s->sMpuIdx.size() / 3);
    return out[0];
}
void psref_stream_copy(const psref_scene* s, float* pos, float* nrm, float* col, uint32_t* tri,
                       int32_t* mpuIdx, uint32_t* mpuCounts) {
    if (pos) memcpy(pos, s->sPos.data(), s->sPos.size() * sizeof(float));
    if (nrm) memcpy(nrm, s->sNrm.data(), s->sNrm.size() * sizeof(float));
    if (col) memcpy(col, s->sCol.data(), s->sCol.size() * sizeof(float));
    if (tri) memcpy(tri, s->sTri.data(), s->sTri.size() * sizeof(uint32_t));
    if (mpuIdx) memcpy(mpuIdx, s->sMpuIdx.data(), s->sMpuIdx.size() * sizeof(int32_t));
    if (mpuCounts) memcpy(mpuCounts, s->sMpuCounts.data(), s->sMpuCounts.size() * sizeof(uint32_t));
}

}  // extern "C"
