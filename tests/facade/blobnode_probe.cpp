// TEST INFRASTRUCTURE: a translation unit as a Parsip maintainer would write it - the reference's scene API headers
// (Parsip100/PS_BlobTree/include) first, then the drop-in facade and the CBlobNode* walker - to prove that
// SimdPoly::linearizeBlobTree(CBlobNode*) (reference signature, ParsipHaptics/include/PS_HighPerformanceRender.h:27) compiles
// against the real classes and fills the same tables as the flat-tree route. Built only where /root/reference exists
// (oracle/Makefile, next to oracle/_ref/libcompactref.so which provides the scene API's object code and the tree builder).
#include "PS_BlobTree/include/BlobTreeLibraryAll.h"
#include "PS_BlobTree/include/CQuadricPoint.h"
#include "PS_BlobTree/include/CInstance.h"
#include "PS_BlobTree/include/CNullPrimitive.h"

#include "PS_SimdPolyCompat.h"
#include "psgpu_blobtree_walk.h"
#include "PS_FrameWork/include/PS_MeshVV.h"
#include "psgpu_meshvv_export.h"

#include <cstring>

using namespace PS::BLOBTREE;
using namespace PS::SIMDPOLY;

extern "C" {

/* linearize through the CBlobNode* overload; copies out what the reference's linearizer would have produced */
int bnp_linearize(void* cblobnode_root, int32_t* counts3, uint8_t* prim_type, float* prim_pos, float* prim_dir, float* prim_res,
                  float* prim_color, uint8_t* prim_mat, uint8_t* op_type, uint8_t* op_left, uint8_t* op_right, uint8_t* op_kind, float* op_res,
                  float* mats) {
    SimdPoly* sp = new SimdPoly();
    const int rc = sp->linearizeBlobTree(reinterpret_cast<CBlobNode*>(cblobnode_root));
    const SOABlobPrims& P = sp->blobPrims();
    const SOABlobOps& O = sp->blobOps();
    const SOABlobPrimMatrices& M = sp->blobPrimMatrices();
    counts3[0] = (int32_t)P.ctPrims; counts3[1] = (int32_t)O.ctOps; counts3[2] = (int32_t)M.count;
    if (rc >= 0) {
        for (U32 i = 0; i < P.ctPrims; i++) {
            prim_type[i] = P.skeletType[i]; prim_mat[i] = P.idxMatrix[i];
            prim_pos[3 * i] = P.posX[i]; prim_pos[3 * i + 1] = P.posY[i]; prim_pos[3 * i + 2] = P.posZ[i];
            prim_dir[3 * i] = P.dirX[i]; prim_dir[3 * i + 1] = P.dirY[i]; prim_dir[3 * i + 2] = P.dirZ[i];
            prim_res[3 * i] = P.resX[i]; prim_res[3 * i + 1] = P.resY[i]; prim_res[3 * i + 2] = P.resZ[i];
            prim_color[3 * i] = P.colorX[i]; prim_color[3 * i + 1] = P.colorY[i]; prim_color[3 * i + 2] = P.colorZ[i];
        }
        for (U32 i = 0; i < O.ctOps; i++) {
            op_type[i] = O.opType[i]; op_left[i] = O.opLeftChild[i]; op_right[i] = O.opRightChild[i]; op_kind[i] = O.opChildKind[i];
            op_res[4 * i] = O.resX[i]; op_res[4 * i + 1] = O.resY[i]; op_res[4 * i + 2] = O.resZ[i]; op_res[4 * i + 3] = O.resW[i];
        }
        memcpy(mats, M.matrix, sizeof(float) * PRIM_MATRIX_STRIDE * (size_t)M.count);
    }
    delete sp;
    return rc;
}

/* linearize + run on the GPU through the facade; stats = {ret, mpus, vertices, triangles} */
int bnp_run(void* cblobnode_root, float cellsize, int64_t* stats4) {
    SimdPoly* sp = new SimdPoly();
    int rc = sp->linearizeBlobTree(reinterpret_cast<CBlobNode*>(cblobnode_root));
    stats4[0] = rc; stats4[1] = stats4[2] = stats4[3] = 0;
    if (rc >= 0) {
        rc = sp->run(cellsize);
        stats4[0] = rc;
        if (rc == RET_SUCCESS) {
            const PolyMPUs& pm = sp->polyMPUs();
            stats4[1] = pm.ctMPUs;
            for (U32 i = 0; i < pm.ctMPUs; i++) { stats4[2] += pm.vMPUs[i].ctVertices; stats4[3] += pm.vMPUs[i].ctTriangles; }
        }
    }
    delete sp;
    return rc;
}

/* f4: CBlobNode* -> walker -> scalar-semantics program -> GPU polygonize (reference-Newton per-MPU output or the native indexed mesh)
 * -> psgpu::export_mesh into a reference CMeshVV. Two-call protocol: arrays NULL = counts only. counts2 = {vertices, faces}. */
int bnp_export_mesh(void* cblobnode_root, const float* lo3, const float* hi3, float cellsize, int per_mpu, int64_t* counts2, float* pos, float* nrm,
                    float* col4, uint32_t* faces) {
    static PS::CMeshVV* mesh = 0;
    if (!pos) {
        psgpu::FlatBlobTree flat;
        if (flat.walk(reinterpret_cast<CBlobNode*>(cblobnode_root)) != 0) return -1;
        const psgpu_tree t = flat.view();
        psgpu_program* prog = 0;
        if (psgpu_linearize(&t, PSGPU_SEM_SCALAR, PSGPU_LIN_REASSOCIATE, &prog) != PSGPU_SUCCESS) return -2;
        psgpu_ctx* ctx = 0;
        if (psgpu_create(0, &ctx) != PSGPU_SUCCESS) { psgpu_program_free(prog); return -3; }
        int rc = psgpu_upload_program(ctx, prog);
        psgpu_params P;
        memset(&P, 0, sizeof(P));
        P.cellsize = cellsize;
        for (int a = 0; a < 3; a++) { P.bbox_lo[a] = lo3[a]; P.bbox_hi[a] = hi3[a]; }
        P.root_mode = per_mpu ? PSGPU_ROOT_NEWTON_REF : PSGPU_ROOT_BISECT;
        P.coord_mode = per_mpu ? PSGPU_COORD_MPU_LOCAL : PSGPU_COORD_GLOBAL;
        P.output = per_mpu ? PSGPU_OUT_MPU : PSGPU_OUT_INDEXED;
        P.cull = PSGPU_CULL_INTERVAL;
        P.want_colors = P.want_normals = 1;
        psgpu_mesh M;
        if (rc == PSGPU_SUCCESS) rc = psgpu_polygonize(ctx, &P, &M);
        delete mesh;
        mesh = new PS::CMeshVV();
        const bool ok = rc == PSGPU_SUCCESS && psgpu::export_mesh(ctx, M, mesh);
        psgpu_destroy(ctx);
        psgpu_program_free(prog);
        counts2[0] = mesh->countVertices(); counts2[1] = mesh->countFaces();
        return ok ? 1 : 0;
    }
    if (!mesh) return -1;
    memcpy(pos, mesh->m_lstVertices.data(), mesh->m_lstVertices.size() * 4);
    memcpy(nrm, mesh->m_lstNormals.data(), mesh->m_lstNormals.size() * 4);
    memcpy(col4, mesh->m_lstColors.data(), mesh->m_lstColors.size() * 4);
    memcpy(faces, mesh->m_lstFaces.data(), mesh->m_lstFaces.size() * 4);
    return 1;
}

}  // extern "C"
