// TEST INFRASTRUCTURE: a caller of the C++ facade (include/PS_SimdPolyCompat.h) written the way a ParsipCmd-style client of
// the reference's PS_SimdPoly is (fill the SoA tables, PrepareBBoxes, Polygonize into a caller-owned PolyMPUs, read the
// per-MPU meshes) and exposed to the Python tests through a few extern "C" probes. Linked against libpsgpu.so only.
#include <cstdlib>
#include <cstring>

#include "../../include/PS_SimdPolyCompat.h"
#include "../../include/psgpu_types.h"

using namespace PS::SIMDPOLY;

struct fac_desc {
    int32_t n_prims, n_ops, n_mats;
    const int32_t* prim_type;  // psgpu_node_type
    const float *prim_pos, *prim_dir, *prim_res, *prim_color;
    const int32_t* prim_mat;
    const int32_t *op_type, *op_left, *op_right, *op_kind;
    const float* op_res;
    const float *mats_bwd, *mats_fwd;
};

struct fac_scene {
    SOABlobPrims prims;
    SOABlobOps ops;
    SOABlobPrimMatrices primMats;
    SOABlobBoxMatrices boxMats;
    PolyMPUs* poly;
};

// standalone BlobNodeType numbering of the compat header (== PS_Polygonizer.h:86-90)
static int ref_type(int t) {
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
    default: return -1;
    }
}

extern "C" {

fac_scene* fac_scene_create(const fac_desc* d) {
    if (d->n_prims <= 0 || d->n_prims > MAX_TREE_NODES || d->n_ops > MAX_TREE_NODES || d->n_mats > MAX_MATRIX_COUNT) return 0;
    fac_scene* s = (fac_scene*)calloc(1, sizeof(fac_scene));
    for (int i = 0; i < d->n_prims; i++) {
        const int t = ref_type(d->prim_type[i]);
        if (t < 0) { free(s); return 0; }
        s->prims.skeletType[i] = (U8)t;
        s->prims.idxMatrix[i] = (U8)(d->prim_mat ? d->prim_mat[i] : 0);
        s->prims.posX[i] = d->prim_pos[i * 3]; s->prims.posY[i] = d->prim_pos[i * 3 + 1]; s->prims.posZ[i] = d->prim_pos[i * 3 + 2];
        s->prims.dirX[i] = d->prim_dir[i * 3]; s->prims.dirY[i] = d->prim_dir[i * 3 + 1]; s->prims.dirZ[i] = d->prim_dir[i * 3 + 2];
        s->prims.resX[i] = d->prim_res[i * 3]; s->prims.resY[i] = d->prim_res[i * 3 + 1]; s->prims.resZ[i] = d->prim_res[i * 3 + 2];
        s->prims.colorX[i] = d->prim_color[i * 3]; s->prims.colorY[i] = d->prim_color[i * 3 + 1]; s->prims.colorZ[i] = d->prim_color[i * 3 + 2];
    }
    s->prims.ctPrims = (U32)d->n_prims;
    for (int i = 0; i < d->n_ops; i++) {
        const int t = ref_type(d->op_type[i]);
        if (t < 0) { free(s); return 0; }
        s->ops.opType[i] = (U8)t;
        s->ops.opLeftChild[i] = (U8)d->op_left[i]; s->ops.opRightChild[i] = (U8)d->op_right[i]; s->ops.opChildKind[i] = (U8)d->op_kind[i];
        s->ops.resX[i] = d->op_res[i * 4]; s->ops.resY[i] = d->op_res[i * 4 + 1]; s->ops.resZ[i] = d->op_res[i * 4 + 2]; s->ops.resW[i] = d->op_res[i * 4 + 3];
    }
    s->ops.ctOps = (U32)d->n_ops;
    static const float id12[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    static const float id16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    memcpy(s->primMats.matrix, id12, sizeof(id12));
    memcpy(s->boxMats.matrix, id16, sizeof(id16));
    for (int m = 0; m < d->n_mats; m++) {
        memcpy(&s->primMats.matrix[m * PRIM_MATRIX_STRIDE], &d->mats_bwd[m * 12], 12 * sizeof(float));
        if (d->mats_fwd) memcpy(&s->boxMats.matrix[m * BOX_MATRIX_STRIDE], &d->mats_fwd[m * 16], 16 * sizeof(float));
    }
    s->primMats.count = s->boxMats.count = (U32)(d->n_mats > 0 ? d->n_mats : 1);
    return s;
}

void fac_scene_destroy(fac_scene* s) {
    if (!s) return;
    free(s->poly);
    free(s);
}

int fac_prepare_bboxes(fac_scene* s, float cellsize) { return PrepareBBoxes(cellsize, s->prims, s->boxMats, s->ops); }

void fac_get_bbox(const fac_scene* s, float* lo, float* hi) {
    lo[0] = s->prims.bboxLo.x; lo[1] = s->prims.bboxLo.y; lo[2] = s->prims.bboxLo.z;
    hi[0] = s->prims.bboxHi.x; hi[1] = s->prims.bboxHi.y; hi[2] = s->prims.bboxHi.z;
}

void fac_get_boxes(const fac_scene* s, float* prim6, float* op6) {
    for (U32 i = 0; i < s->prims.ctPrims; i++) {
        float* o = prim6 + i * 6;
        o[0] = s->prims.vPrimBoxLoX[i]; o[1] = s->prims.vPrimBoxLoY[i]; o[2] = s->prims.vPrimBoxLoZ[i];
        o[3] = s->prims.vPrimBoxHiX[i]; o[4] = s->prims.vPrimBoxHiY[i]; o[5] = s->prims.vPrimBoxHiZ[i];
    }
    for (U32 i = 0; i < s->ops.ctOps; i++) {
        float* o = op6 + i * 6;
        o[0] = s->ops.vBoxLoX[i]; o[1] = s->ops.vBoxLoY[i]; o[2] = s->ops.vBoxLoZ[i];
        o[3] = s->ops.vBoxHiX[i]; o[4] = s->ops.vBoxHiY[i]; o[5] = s->ops.vBoxHiZ[i];
    }
}

uint32_t fac_count_mpus(const fac_scene* s, float cellsize) { return CountMPUNeeded(cellsize, s->prims.bboxLo, s->prims.bboxHi); }

int fac_field(fac_scene* s, int n, const float* xyz, float* out) { return FieldValues(s->prims, s->primMats, s->ops, xyz, n, out); }
int fac_newton(fac_scene* s, int n, const float* p1, const float* p2, const float* f1, const float* f2, float target, int iters, float* root, float* field,
               int* evals) {
    return RootsNewtonRaphson(s->prims, s->primMats, s->ops, p1, p2, f1, f2, n, root, field, evals, target, iters);
}

// stats: ctMPUs, crossed, vertices, triangles
int fac_polygonize(fac_scene* s, float cellsize, int64_t* stats) {
    if (!s->poly) s->poly = (PolyMPUs*)malloc(sizeof(PolyMPUs));  // 516 MB of untouched pages, like the reference's callers
    if (!s->poly) return RET_NOT_ENOUGH_MEM;
    const int ret = Polygonize(cellsize, s->prims, s->primMats, s->ops, *s->poly, 0);
    int64_t crossed = 0, nv = 0, nt = 0;
    if (ret == RET_SUCCESS)
        for (U32 i = 0; i < s->poly->ctMPUs; i++) {
            const MPU& m = s->poly->vMPUs[i];
            if (m.ctTriangles > 0) crossed++;
            nv += m.ctVertices; nt += m.ctTriangles;
        }
    stats[0] = ret == RET_SUCCESS ? s->poly->ctMPUs : 0; stats[1] = crossed; stats[2] = nv; stats[3] = nt;
    return ret;
}

// concatenated PolyMPUs content: vertices in MPU order, MPU-local U16 indices rebased to u32
void fac_mesh_concat(const fac_scene* s, float* pos, float* nrm, float* col, uint32_t* tri, float* mpuLo) {
    size_t vb = 0, tb = 0;
    for (U32 i = 0; i < s->poly->ctMPUs; i++) {
        const MPU& m = s->poly->vMPUs[i];
        if (mpuLo) { mpuLo[i * 3] = m.bboxLo.x; mpuLo[i * 3 + 1] = m.bboxLo.y; mpuLo[i * 3 + 2] = m.bboxLo.z; }
        memcpy(pos + vb * 3, m.vPos, (size_t)m.ctVertices * 12);
        memcpy(nrm + vb * 3, m.vNorm, (size_t)m.ctVertices * 12);
        memcpy(col + vb * 3, m.vColor, (size_t)m.ctVertices * 12);
        for (int q = 0; q < m.ctTriangles * 3; q++) tri[tb * 3 + q] = (uint32_t)m.triangles[q] + (uint32_t)vb;
        vb += m.ctVertices; tb += m.ctTriangles;
    }
}

// the SimdPoly wrapper class on a flat tree: returns run()'s code, stats as above
int fac_simdpoly_run(const psgpu_tree* tree, float cellsize, int64_t* stats) {
    SimdPoly sp;
    const int id = sp.linearizeBlobTree(tree);
    if (id < 0) return id;
    const int ret = sp.run(cellsize);
    int64_t crossed = 0, nv = 0, nt = 0;
    if (ret == RET_SUCCESS) {
        const PolyMPUs& p = sp.polyMPUs();
        for (U32 i = 0; i < p.ctMPUs; i++) { if (p.vMPUs[i].ctTriangles) crossed++; nv += p.vMPUs[i].ctVertices; nt += p.vMPUs[i].ctTriangles; }
        stats[0] = p.ctMPUs;
    }
    stats[1] = crossed; stats[2] = nv; stats[3] = nt;
    return ret;
}

}  // extern "C"
