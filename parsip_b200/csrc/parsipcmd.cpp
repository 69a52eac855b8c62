// parsipcmd: command-line polygonizer over libpsgpu.so - the role BASELINE configs[0] gives "ParsipCmd" (whose sources the reference
// repository does not ship): load a .scene file, run the GUI's pre-passes (flattenTransformations, setOctreeFromBlobTree), linearize,
// polygonize on the GPU, print the counters the reference logs per run (Distrib/ParsipHaptics_Release.csv: primitives, operators,
// cell size, polygonization time, MPUs total / intersected, mesh faces / vertices), optionally write the mesh as Wavefront OBJ.
//
//   parsipcmd scene.scene [cellsize] [--reps N] [--newton] [--obj out.obj] [--device D]
//
// --newton reproduces the reference's scalar polygonizer (Newton roots, per-MPU vertices, CPolyParsipOptimized.cpp:130-327);
// the default is the native indexed mesh (bracketed roots, vertices shared between cells and MPUs).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/psgpu.h"

static int fail(const char* what, int rc, psgpu_ctx* ctx = nullptr) {
    fprintf(stderr, "parsipcmd: %s failed (%d)%s%s\n", what, rc, ctx ? ": " : "", ctx ? psgpu_last_error(ctx) : "");
    return 1;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: parsipcmd scene.scene [cellsize=0.13] [--reps N] [--newton] [--obj out.obj] [--device D]\n");
        return 2;
    }
    const char* scene = argv[1];
    float cellsize = 0.13f;  // the cell size of the reference's recorded runs (ParsipHaptics_Release.csv)
    int reps = 5, device = 0;
    bool newton = false;
    const char* obj = nullptr;
    for (int i = 2; i < argc; i++) {
        if (!strcmp(argv[i], "--reps") && i + 1 < argc) reps = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--obj") && i + 1 < argc) obj = argv[++i];
        else if (!strcmp(argv[i], "--newton")) newton = true;
        else cellsize = (float)atof(argv[i]);
    }
    psgpu_tree_buf *loaded = nullptr, *flat = nullptr;
    int rc = psgpu_scene_load(scene, -1, &loaded);
    if (rc != PSGPU_SUCCESS) return fail("psgpu_scene_load", rc);
    rc = psgpu_tree_flatten_transforms(psgpu_tree_view(loaded), &flat);
    if (rc != PSGPU_SUCCESS) return fail("psgpu_tree_flatten_transforms", rc);
    const psgpu_tree* tree = psgpu_tree_view(flat);
    std::vector<float> lo((size_t)tree->n_nodes * 3), hi((size_t)tree->n_nodes * 3);
    rc = psgpu_tree_octrees(tree, 1, lo.data(), hi.data());
    if (rc != PSGPU_SUCCESS) return fail("psgpu_tree_octrees", rc);
    int nPrims = 0, nOps = 0;
    for (int n = 0; n < tree->n_nodes; n++) (PSGPU_IS_PRIM(tree->type[n]) ? nPrims : nOps)++;
    psgpu_program* prog = nullptr;
    rc = psgpu_linearize(tree, PSGPU_SEM_SCALAR, PSGPU_LIN_REASSOCIATE, &prog);
    if (rc != PSGPU_SUCCESS) return fail("psgpu_linearize", rc);
    psgpu_ctx* ctx = nullptr;
    rc = psgpu_create(device, &ctx);
    if (rc != PSGPU_SUCCESS) return fail("psgpu_create (no usable CUDA device; there is no CPU fallback)", rc);
    rc = psgpu_upload_program(ctx, prog);
    if (rc != PSGPU_SUCCESS) return fail("psgpu_upload_program", rc, ctx);
    psgpu_params P;
    memset(&P, 0, sizeof(P));
    P.cellsize = cellsize;
    const int root = tree->root;
    for (int a = 0; a < 3; a++) { P.bbox_lo[a] = lo[(size_t)root * 3 + a]; P.bbox_hi[a] = hi[(size_t)root * 3 + a]; }
    P.root_mode = newton ? PSGPU_ROOT_NEWTON_REF : PSGPU_ROOT_BISECT;
    P.coord_mode = newton ? PSGPU_COORD_MPU_LOCAL : PSGPU_COORD_GLOBAL;
    P.output = newton ? PSGPU_OUT_MPU : PSGPU_OUT_INDEXED;
    P.cull = PSGPU_CULL_INTERVAL;
    P.want_colors = P.want_normals = 1;
    psgpu_mesh M;
    psgpu_stats S;
    std::vector<float> ms;
    for (int r = 0; r < std::max(1, reps); r++) {
        rc = psgpu_polygonize(ctx, &P, &M);
        if (rc != PSGPU_SUCCESS && !(newton && rc == PSGPU_NOT_ENOUGH_MEM && M.over_capacity > 0)) return fail("psgpu_polygonize", rc, ctx);
        psgpu_get_stats(ctx, &S);
        ms.push_back(S.ms_total);
    }
    std::sort(ms.begin(), ms.end());
    printf("scene,Blob Primitives,Blob Operators,Poly Cell Size,Total Poly Time ms (median of %d),Total FieldEvaluations,MPUs Total,MPUs Evaluated,"
           "MPUs Intersected,Cells Total,Mesh Faces,Mesh Vertices,mode\n", (int)ms.size());
    printf("%s,%d,%d,%.4f,%.3f,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%s\n", scene, nPrims, nOps, cellsize, ms[ms.size() / 2], (long long)S.field_evals,
           (long long)S.mpus_total, (long long)S.mpus_evaluated, (long long)S.mpus_crossed, (long long)S.mpus_total * 343, (long long)M.n_triangles,
           (long long)M.n_vertices, newton ? "reference Newton roots, per-MPU vertices" : "native indexed mesh");
    if (obj) {
        std::vector<float> pos((size_t)M.n_vertices * 3), nrm((size_t)M.n_vertices * 3);
        std::vector<uint32_t> tri((size_t)M.n_triangles * 3), vb((size_t)M.n_mpus_out + 1), tb((size_t)M.n_mpus_out + 1);
        rc = psgpu_mesh_download(ctx, pos.data(), nrm.data(), nullptr, tri.data(), nullptr, newton ? vb.data() : nullptr, newton ? tb.data() : nullptr);
        if (rc != PSGPU_SUCCESS) return fail("psgpu_mesh_download", rc, ctx);
        FILE* f = fopen(obj, "w");
        if (!f) return fail("fopen", 0);
        for (int64_t v = 0; v < M.n_vertices; v++) fprintf(f, "v %.6f %.6f %.6f\n", pos[(size_t)v * 3], pos[(size_t)v * 3 + 1], pos[(size_t)v * 3 + 2]);
        for (int64_t v = 0; v < M.n_vertices; v++) fprintf(f, "vn %.5f %.5f %.5f\n", nrm[(size_t)v * 3], nrm[(size_t)v * 3 + 1], nrm[(size_t)v * 3 + 2]);
        if (!newton) {
            for (int64_t t = 0; t < M.n_triangles; t++)
                fprintf(f, "f %u//%u %u//%u %u//%u\n", tri[(size_t)t * 3] + 1, tri[(size_t)t * 3] + 1, tri[(size_t)t * 3 + 1] + 1, tri[(size_t)t * 3 + 1] + 1,
                        tri[(size_t)t * 3 + 2] + 1, tri[(size_t)t * 3 + 2] + 1);
        } else {  // per-MPU output: indices are MPU-local, rebase by the MPU's vertex base (CParsipOptimized::exportMesh -> CMeshVV::appendFrom)
            for (int64_t m = 0; m < M.n_mpus_out; m++)
                for (uint32_t t = tb[(size_t)m]; t < tb[(size_t)m + 1]; t++) {
                    const uint32_t a = tri[(size_t)t * 3] + vb[(size_t)m] + 1, b = tri[(size_t)t * 3 + 1] + vb[(size_t)m] + 1, c = tri[(size_t)t * 3 + 2] + vb[(size_t)m] + 1;
                    fprintf(f, "f %u//%u %u//%u %u//%u\n", a, a, b, b, c, c);
                }
        }
        fclose(f);
    }
    psgpu_destroy(ctx);
    psgpu_program_free(prog);
    psgpu_tree_free(flat);
    psgpu_tree_free(loaded);
    return 0;
}
