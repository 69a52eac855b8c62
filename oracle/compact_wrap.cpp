// TEST INFRASTRUCTURE ONLY: C wrapper around the reference's OWN scalar path, compiled by oracle/Makefile together with the
// unmodified reference sources where they lie (nothing is copied into this repo):
//   Parsip100/PS_BlobTree/include/*.cpp            scene API (CBlobNode and the node classes)
//   Parsip100/ParsipHaptics/include/CompactBlobTree.cpp      COMPACTBLOBTREE: convert / fieldvalue / normal / baseColor / Newton
//   Parsip100/ParsipHaptics/include/CPolyParsipOptimized.cpp CParsipOptimized: the scalar marching-cubes polygonizer
//   Parsip100/PS_FrameWork/include/{PS_AppConfig,PS_String,PS_Property,PS_ErrorManager,PS_GeometryFuncs,PS_Octree,PS_Matrix,
//                                    PS_Quaternion,PS_FileDirectory,PS_DateTime,PS_PerfLogger}.cpp
// Third-party pieces the reference needs and this image lacks are shimmed under oracle/shims (Loki, QString, the GL half of
// _parsDebug.h, the container half of CMeshVV) and oracle/tbbshim; none of them holds arithmetic.
//
// The wrapper (1) builds a tree with the reference classes from a flat description, (2) walks it back with the PRODUCT header
// include/psgpu_blobtree_walk.h (the CBlobNode* -> psgpu_tree binding under test), (3) exposes the reference evaluator:
// COMPACTBLOBTREE::fieldvalue / normal / baseColor, ComputeRootNewtonRaphsonVEC4, CBlobNode::fieldValue, CParsipOptimized.
#include "PS_BlobTree/include/BlobTreeLibraryAll.h"
#include "PS_BlobTree/include/CQuadricPoint.h"
#include "PS_BlobTree/include/CInstance.h"
#include "PS_BlobTree/include/CNullPrimitive.h"
#include "CompactBlobTree.h"
#include "CPolyParsipOptimized.h"

#include "../include/psgpu_blobtree_walk.h"

#include <cstring>
#include <vector>

using namespace PS;
using namespace PS::BLOBTREE;

extern "C" {

typedef struct cref_desc {
    int32_t n_nodes;
    const int32_t* type;        /* psgpu_node_type */
    const int32_t* first_child;
    const int32_t* n_children;
    const int32_t* child;
    const float* pos;
    const float* dir;
    const float* res;
    const float* color;
    const float* srt;           /* [n*10] scale xyz, quaternion xyzw, translate xyz; NULL = none */
    const uint8_t* has_srt;
} cref_desc;

struct cref_h {
    CBlobNode* root;
    COMPACTBLOBTREE* cpt;
    psgpu::FlatBlobTree flat;
    int walk_rc;
    int convert_rc;
    CParsipOptimized* poly;
    cref_h() : root(0), cpt(0), walk_rc(0), convert_rc(0), poly(0) {}
};

static MajorAxices axis_of(float v) { return (MajorAxices)(int)v; }

static CBlobNode* make_node(const cref_desc* d, int i, std::vector<CBlobNode*>& made) {
    const float* p = d->pos + 3 * i;
    const float* q = d->dir + 3 * i;
    const float* r = d->res + 4 * i;
    const vec3f pos(p[0], p[1], p[2]), dir(q[0], q[1], q[2]);
    CBlobNode* n = 0;
    switch (d->type[i]) {
    case PSGPU_PRIM_POINT: n = new CSkeletonPrimitive(new CSkeletonPoint(pos)); break;
    case PSGPU_PRIM_LINE: n = new CSkeletonPrimitive(new CSkeletonLine(pos, dir)); break;
    case PSGPU_PRIM_CYLINDER: n = new CSkeletonPrimitive(new CSkeletonCylinder(pos, dir, r[0], r[1])); break;
    case PSGPU_PRIM_DISC: n = new CSkeletonPrimitive(new CSkeletonDisc(pos, dir, r[0])); break;
    case PSGPU_PRIM_RING: n = new CSkeletonPrimitive(new CSkeletonRing(pos, dir, r[0])); break;
    case PSGPU_PRIM_CUBE: n = new CSkeletonPrimitive(new CSkeletonCube(pos, r[0])); break;
    case PSGPU_PRIM_NULL: n = new CNullPrimitive(); break;
    case PSGPU_PRIM_QUADRICPOINT: n = new CQuadricPoint(pos, r[0], r[1]); break;
    case PSGPU_PRIM_INSTANCE: n = new CInstance(); break; /* origin patched once every node exists */
    case PSGPU_OP_UNION: n = new CUnion(); break;
    case PSGPU_OP_INTERSECT: n = new CIntersection(); break;
    case PSGPU_OP_DIF: n = new CDifference(); break;
    case PSGPU_OP_SMOOTHDIF: n = new CSmoothDifference(); break;
    case PSGPU_OP_BLEND: n = new CBlend(); break;
    case PSGPU_OP_RICCI: n = new CRicciBlend(r[0]); break;
    case PSGPU_OP_WARP_TWIST: n = new CWarpTwist(r[0], axis_of(r[1])); break;
    case PSGPU_OP_WARP_TAPER: n = new CWarpTaper(r[0], axis_of(r[2]), axis_of(r[1])); break; /* ctor order: (factor, taper, along) */
    case PSGPU_OP_WARP_BEND: n = new CWarpBend(r[0], r[1], r[2], r[3]); break;
    case PSGPU_OP_WARP_SHEAR: n = new CWarpShear(r[0], axis_of(r[1]), axis_of(r[2])); break;
    default: return 0;
    }
    n->setID(i);
    if (PSGPU_IS_PRIM(d->type[i])) {
        const float* c = d->color + 3 * i;
        CMaterial m;
        m.diffused = vec4f(c[0], c[1], c[2], 1.0f);
        n->setMaterial(m);
    }
    if (d->srt && d->has_srt && d->has_srt[i]) {
        const float* s = d->srt + 10 * i;
        n->getTransform().set(vec3f(s[0], s[1], s[2]), CQuaternion(s[3], s[4], s[5], s[6]), vec3f(s[7], s[8], s[9]));
    }
    made[(size_t)i] = n;
    for (int k = 0; k < d->n_children[i]; k++) {
        CBlobNode* c = make_node(d, d->child[d->first_child[i] + k], made);
        if (!c) return 0;
        n->addChild(c);
    }
    return n;
}

cref_h* cref_create(const cref_desc* d) {
    cref_h* h = new cref_h();
    std::vector<CBlobNode*> made((size_t)d->n_nodes, (CBlobNode*)0);
    h->root = make_node(d, 0, made);
    if (!h->root) { delete h; return 0; }
    for (int i = 0; i < d->n_nodes; i++)
        if (d->type[i] == PSGPU_PRIM_INSTANCE && made[(size_t)i])
            reinterpret_cast<CInstance*>(made[(size_t)i])->setOriginalNode(made[(size_t)(int)d->res[4 * i]]);
    h->cpt = new COMPACTBLOBTREE();
    h->convert_rc = h->cpt->convert(h->root);
    h->walk_rc = h->flat.walk(h->root);
    return h;
}

void cref_destroy(cref_h* h) {
    if (!h) return;
    delete h->poly;
    delete h->cpt;
    delete h->root;
    delete h;
}

void* cref_root(const cref_h* h) { return h->root; } /* CBlobNode*, for tests/facade/blobnode_probe.cpp */
int cref_convert_rc(const cref_h* h) { return h->convert_rc; }
int cref_walk_rc(const cref_h* h) { return h->walk_rc; }
void cref_counts(const cref_h* h, int32_t* out4) {
    out4[0] = h->cpt->countPrimitives(); out4[1] = h->cpt->countOperators();
    out4[2] = (int32_t)h->flat.type.size(); out4[3] = (int32_t)h->flat.child.size();
}

/* the product walker's output (include/psgpu_blobtree_walk.h) */
void cref_walk_copy(const cref_h* h, int32_t* type, int32_t* first_child, int32_t* n_children, int32_t* child, float* pos, float* dir,
                    float* res, float* color, float* xform, uint8_t* has_xform, float* oct_lo, float* oct_hi) {
    const psgpu::FlatBlobTree& f = h->flat;
    const size_t N = f.type.size();
    memcpy(type, f.type.data(), N * 4); memcpy(first_child, f.first_child.data(), N * 4); memcpy(n_children, f.n_children.data(), N * 4);
    memcpy(child, f.child.data(), f.child.size() * 4);
    memcpy(pos, f.pos.data(), N * 12); memcpy(dir, f.dir.data(), N * 12); memcpy(res, f.res.data(), N * 16);
    memcpy(color, f.color.data(), N * 12); memcpy(xform, f.xform.data(), N * 48); memcpy(has_xform, f.has_xform.data(), N);
    memcpy(oct_lo, f.oct_lo.data(), N * 12); memcpy(oct_hi, f.oct_hi.data(), N * 12);
}
void cref_walk_srt(const cref_h* h, float* srt10n, int32_t* ids) {
    memcpy(srt10n, h->flat.srt.data(), h->flat.srt.size() * 4); memcpy(ids, h->flat.org_id.data(), h->flat.org_id.size() * 4);
}

void cref_octree(const cref_h* h, float* lo3, float* hi3) {
    COctree o = h->root->getOctree();
    lo3[0] = o.lower.x; lo3[1] = o.lower.y; lo3[2] = o.lower.z;
    hi3[0] = o.upper.x; hi3[1] = o.upper.y; hi3[2] = o.upper.z;
}

/* COMPACTBLOBTREE::fieldvalue (CompactBlobTree.cpp:476-487) */
void cref_field(cref_h* h, int64_t n, const float* xyz, float* out) {
    for (int64_t i = 0; i < n; i++) out[i] = h->cpt->fieldvalue(vec4f(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 0.0f));
}

/* what doMarchingCubes does at a vertex (CPolyParsipOptimized.cpp:283-302): field with per-node stores, normal, baseColor */
void cref_field_color_normal(cref_h* h, int64_t n, const float* xyz, float* f, float* col4, float* nrm) {
    std::vector<float> sp((size_t)h->cpt->countPrimitives() + 1), so((size_t)h->cpt->countOperators() + 1);
    for (int64_t i = 0; i < n; i++) {
        const vec4f p(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 0.0f);
        const float fp = h->cpt->fieldvalue(p, so.data(), sp.data());
        const vec4f nn = h->cpt->normal(p, fp, NORMAL_DELTA);
        const vec4f c = h->cpt->baseColor(p, so.data(), sp.data());
        f[i] = fp;
        nrm[3 * i] = nn.x; nrm[3 * i + 1] = nn.y; nrm[3 * i + 2] = nn.z;
        col4[4 * i] = c.x; col4[4 * i + 1] = c.y; col4[4 * i + 2] = c.z; col4[4 * i + 3] = c.w;
    }
}

/* ComputeRootNewtonRaphsonVEC4 (CompactBlobTree.cpp:1581-1622) */
void cref_newton(cref_h* h, int64_t n, const float* p1, const float* p2, const float* f1, const float* f2, float iso, int iterations,
                 float* out_xyz, float* out_f, int32_t* out_evals) {
    std::vector<float> sp((size_t)h->cpt->countPrimitives() + 1), so((size_t)h->cpt->countOperators() + 1);
    for (int64_t i = 0; i < n; i++) {
        vec4f o; float of = 0.0f;
        out_evals[i] = ComputeRootNewtonRaphsonVEC4(h->cpt, so.data(), sp.data(), vec4f(p1[3 * i], p1[3 * i + 1], p1[3 * i + 2], 0.0f),
                                                    vec4f(p2[3 * i], p2[3 * i + 1], p2[3 * i + 2], 0.0f), f1[i], f2[i], o, of, iso, iterations);
        out_xyz[3 * i] = o.x; out_xyz[3 * i + 1] = o.y; out_xyz[3 * i + 2] = o.z;
        out_f[i] = of;
    }
}

/* the scene API's own recursive evaluation, CBlobNode::fieldValue (PS_BlobTree/include/CBlobTree.h) */
void cref_scene_field(cref_h* h, int64_t n, const float* xyz, float* out) {
    for (int64_t i = 0; i < n; i++) out[i] = h->root->fieldValue(vec3f(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
}

/* ---- pre-linearization passes and .scene output with the reference's own classes. The recursions below restate
 * CLayer::recursive_FlattenTransformations / recursive_RecomputeAllOctrees (ParsipHaptics/include/CLayerManager.cpp:604-645; that
 * file needs Qt) around the reference's CAffineTransformation::add, CMatrix and COctree, which do the arithmetic. */
static void ref_flatten(CBlobNode* node, const CAffineTransformation& transformBranch) {
    if (node == NULL) return;
    if (node->isOperator() && (!node->isUnary())) {
        CAffineTransformation t = transformBranch;
        t.add(node->getTransform());
        node->getTransform().init();
        for (size_t i = 0; i < node->countChildren(); i++) ref_flatten(node->getChild(i), t);
    } else {
        CAffineTransformation t = transformBranch;
        t.add(node->getTransform());
        node->getTransform().set(t);
    }
}

static void ref_recompute_octrees(CBlobNode* node, const CMatrix& mtxBranch) {
    if (node == NULL) return;
    CMatrix mtxCurrent = mtxBranch;
    mtxCurrent.multiply(node->getTransform().getForwardMatrix());
    if (node->isOperator()) {
        for (size_t i = 0; i < node->countChildren(); i++) ref_recompute_octrees(node->getChild(i), mtxCurrent);
        node->computeOctree();
    } else {
        COctree oct = node->computeOctree();
        oct.transform(mtxCurrent);
        node->setOctree(oct.lower, oct.upper);
    }
}

void cref_flatten_transforms(cref_h* h) {
    CAffineTransformation t;
    ref_flatten(h->root, t);
    h->convert_rc = h->cpt->convert(h->root);
    h->walk_rc = h->flat.walk(h->root);
}

void cref_recompute_octrees(cref_h* h) {
    CMatrix m;
    ref_recompute_octrees(h->root, m);
    h->convert_rc = h->cpt->convert(h->root);
    h->walk_rc = h->flat.walk(h->root);
}

/* CBlobNode::saveScript (recursive, PS_BlobTree/include/CBlobTree.cpp:40-118 + the node classes) into an INI file */
int cref_save_scene(cref_h* h, const char* path) {
    CSketchConfig* cfg = new CSketchConfig(DAnsiStr(path), CAppConfig::fmWrite);
    const bool ok = h->root->saveScript(cfg);
    delete cfg;
    return ok ? 1 : 0;
}

/* CParsipOptimized::setup + run (CPolyParsipOptimized.cpp:330-410) on a caller-chosen box */
int cref_polygonize(cref_h* h, const float* lo3, const float* hi3, float cellsize, float iso, int nthreads) {
    delete h->poly;
    h->poly = new CParsipOptimized();
    COctree oct;
    oct.lower = vec3f(lo3[0], lo3[1], lo3[2]);
    oct.upper = vec3f(hi3[0], hi3[1], hi3[2]);
    tbb::shim::thread_override() = nthreads;
    h->poly->setup(h->cpt, oct, 0, cellsize, iso);
    h->poly->run();
    tbb::shim::thread_override() = 0;
    return (int)h->poly->countMPUs();
}

void cref_result_counts(const cref_h* h, int64_t* out4) {
    size_t v = 0, t = 0;
    h->poly->statsMeshInfo(v, t);
    out4[0] = (int64_t)h->poly->countMPUs(); out4[1] = (int64_t)v; out4[2] = (int64_t)t; out4[3] = (int64_t)h->poly->statsTotalFieldEvals();
}

/* MPUs that kept surface, in list order (removeExtraPUs drops the rest); triangles rebased to the concatenated vertex array */
void cref_result_copy(const cref_h* h, float* pos, float* nrm, float* col4, uint32_t* tri, float* mpu_origin, int32_t* mpu_counts) {
    size_t vb = 0, tb = 0;
    for (size_t m = 0; m < h->poly->countMPUs(); m++) {
        CSIMDMPU* u = h->poly->getMPU((int)m);
        const CMeshVV* me = u->lpOutputMesh;
        const size_t nv = me->countVertices(), nt = me->countFaces();
        const vec3f o = u->getOrigin();
        mpu_origin[3 * m] = o.x; mpu_origin[3 * m + 1] = o.y; mpu_origin[3 * m + 2] = o.z;
        mpu_counts[2 * m] = (int32_t)nv; mpu_counts[2 * m + 1] = (int32_t)nt;
        memcpy(pos + 3 * vb, me->m_lstVertices.data(), nv * 12);
        memcpy(nrm + 3 * vb, me->m_lstNormals.data(), nv * 12);
        memcpy(col4 + 4 * vb, me->m_lstColors.data(), nv * 16);
        for (size_t i = 0; i < nt * 3; i++) tri[3 * tb + i] = me->m_lstFaces[i] + (uint32_t)vb;
        vb += nv; tb += nt;
    }
}

/* CParsipOptimized::exportMesh (CPolyParsipOptimized.cpp:594-613): the single CMeshVV the GUI hands on */
int cref_export_mesh(const cref_h* h, float* pos, float* nrm, float* col4, uint32_t* tri, int64_t* counts2) {
    CMeshVV out;
    const bool ok = h->poly->exportMesh(&out);
    counts2[0] = out.countVertices(); counts2[1] = out.countFaces();
    if (pos) {
        memcpy(pos, out.m_lstVertices.data(), out.m_lstVertices.size() * 4);
        memcpy(nrm, out.m_lstNormals.data(), out.m_lstNormals.size() * 4);
        memcpy(col4, out.m_lstColors.data(), out.m_lstColors.size() * 4);
        memcpy(tri, out.m_lstFaces.data(), out.m_lstFaces.size() * 4);
    }
    return ok ? 1 : 0;
}

}  // extern "C"
