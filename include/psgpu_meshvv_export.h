/* psgpu mesh -> PS::CMeshVV: the hand-over CParsipOptimized::exportMesh makes to the GUI
 * (Parsip100/ParsipHaptics/include/CPolyParsipOptimized.cpp:594-613: copyFrom the first MPU's mesh, appendFrom the others, i.e.
 * per-MPU vertex blocks concatenated and face indices rebased by the running vertex count, PS_FrameWork/include/PS_MeshVV.cpp:2347-2400).
 *
 * Compiled INSIDE the reference tree: include after PS_FrameWork/include/PS_MeshVV.h and psgpu.h. The mesh keeps the layout the
 * reference's polygonizer sets up (CPolyParsipOptimized.cpp:19-20: TRIANGLES, 3 floats per vertex, 4 per colour, no texture channels).
 * Works for both outputs of psgpu_polygonize: PSGPU_OUT_MPU (MPU-local indices, rebased here exactly as appendFrom does) and
 * PSGPU_OUT_INDEXED (one shared vertex array, indices copied). Colours get alpha 1 (the product carries RGB). */
#ifndef PSGPU_MESHVV_EXPORT_H
#define PSGPU_MESHVV_EXPORT_H

#include <stdint.h>
#include <vector>
#include "psgpu.h"

namespace psgpu {

/* Fills *out from the LAST polygonize of ctx (m = the psgpu_mesh it returned). Returns what exportMesh returns: true iff the mesh
 * has faces. On a download error the mesh is left empty and false is returned (psgpu_last_error(ctx) has the reason). */
inline bool export_mesh(psgpu_ctx* ctx, const psgpu_mesh& m, PS::CMeshVV* out) {
    if (!ctx || !out) return false;
    out->removeAll();
    out->initMesh(PS::CMeshVV::TRIANGLES, 3, 4, 0, 0);
    if (m.n_triangles <= 0 || m.n_vertices <= 0) return false;
    const size_t nv = (size_t)m.n_vertices, nt = (size_t)m.n_triangles, nm = (size_t)m.n_mpus_out;
    const bool perMpu = nm > 0 && m.mpu_vbase && m.mpu_tbase;
    std::vector<float> rgb(nv * 3);
    std::vector<uint32_t> tri(nt * 3), vb(perMpu ? nm + 1 : 0), tb(perMpu ? nm + 1 : 0);
    out->m_lstVertices.resize(nv * 3);
    out->m_lstNormals.resize(nv * 3);
    const int rc = psgpu_mesh_download(ctx, &out->m_lstVertices[0], &out->m_lstNormals[0], &rgb[0], &tri[0], 0, perMpu ? &vb[0] : 0,
                                       perMpu ? &tb[0] : 0);
    if (rc != PSGPU_SUCCESS) { out->removeAll(); return false; }
    out->m_lstColors.resize(nv * 4);
    for (size_t v = 0; v < nv; v++) {
        out->m_lstColors[v * 4] = rgb[v * 3]; out->m_lstColors[v * 4 + 1] = rgb[v * 3 + 1]; out->m_lstColors[v * 4 + 2] = rgb[v * 3 + 2];
        out->m_lstColors[v * 4 + 3] = 1.0f;
    }
    out->m_lstFaces.resize(nt * 3);
    if (perMpu) {
        for (size_t u = 0; u < nm; u++)
            for (size_t i = (size_t)tb[u] * 3; i < (size_t)tb[u + 1] * 3; i++) out->m_lstFaces[i] = tri[i] + vb[u];
    } else {
        for (size_t i = 0; i < nt * 3; i++) out->m_lstFaces[i] = tri[i];
    }
    return out->countFaces() > 0;
}

}  // namespace psgpu
#endif
