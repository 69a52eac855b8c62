/* CBlobNode* -> psgpu_tree walker: the piece a Parsip maintainer compiles INSIDE the reference tree.
 *
 * Include AFTER the reference scene API (Parsip100/PS_BlobTree/include/BlobTreeLibraryAll.h, or CBlobTree.h plus the
 * node classes in use): this header only reads the public getters the reference's own two linearizers read
 *   - SimdPoly::linearizeBlobTree(CBlobNode*)       ParsipHaptics/include/PS_HighPerformanceRender.cpp:42-371
 *   - COMPACTBLOBTREE::convert(CBlobNode*, int)     ParsipHaptics/include/CompactBlobTree.cpp:93-408
 * and fills the flat n-ary description of include/psgpu_types.h. No field arithmetic happens here.
 *
 * Unlike SimdPoly::linearizeBlobTree the walk accepts n-ary operators (the native linearizer folds them), keeps operator
 * matrices (the compact tree applies them, CompactBlobTree.cpp:685-694) and numbers nodes in pre-order.
 * tests/test_compactref.py builds trees with the reference classes, walks them with this header and checks the result
 * against the harness-built flat tree and, through it, the GPU path against the reference's own evaluator. */
#ifndef PSGPU_BLOBTREE_WALK_H
#define PSGPU_BLOBTREE_WALK_H

#ifndef BLOBTREE_DEFINITIONS
#error "include the reference scene API (PS_BlobTree/include/CBlobTree.h and the node classes) before psgpu_blobtree_walk.h"
#endif

#include <stdint.h>
#include <map>
#include <vector>
#include "psgpu_types.h"

namespace psgpu {

/* BlobNodeType of PS_BlobTree/include/_constSettings.h:26-38 -> psgpu_node_type; -1 = not lowered */
inline int node_type_from_reference(BlobNodeType t) {
    switch (t) {
    case bntPrimPoint: return PSGPU_PRIM_POINT;
    case bntPrimLine: return PSGPU_PRIM_LINE;
    case bntPrimCylinder: return PSGPU_PRIM_CYLINDER;
    case bntPrimDisc: return PSGPU_PRIM_DISC;
    case bntPrimRing: return PSGPU_PRIM_RING;
    case bntPrimCube: return PSGPU_PRIM_CUBE;
    case bntPrimNull: return PSGPU_PRIM_NULL;
    case bntPrimQuadricPoint: return PSGPU_PRIM_QUADRICPOINT;
    case bntPrimInstance: return PSGPU_PRIM_INSTANCE;
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
    default: return -1; /* triangle (a stub in the reference), polygon, half plane, PCM, gradient blend, cache, texture */
    }
}

/* Owns the arrays a psgpu_tree points to. */
struct FlatBlobTree {
    std::vector<int32_t> type, first_child, n_children, child, org_id;
    std::vector<float> pos, dir, res, color, xform, oct_lo, oct_hi, srt;
    std::vector<uint8_t> has_xform, has_srt;
    int unsupported_type;  /* reference BlobNodeType that stopped the walk, -1 if none */

    FlatBlobTree() : unsupported_type(-1) {}

    /* returns the root index (0) or PSGPU_ERROR_UNSUPPORTED_NODE / PSGPU_PARAM_ERROR */
    int walk(PS::BLOBTREE::CBlobNode* root) {
        using PS::BLOBTREE::CBlobNode;
        clear();
        if (!root) return PSGPU_PARAM_ERROR;
        /* pass 1: pre-order numbering with an explicit stack (ConvertToBinaryTree makes 10k-deep chains) */
        std::vector<CBlobNode*> nodes, stack;
        std::map<CBlobNode*, int32_t> index;
        stack.push_back(root);
        while (!stack.empty()) {
            CBlobNode* n = stack.back();
            stack.pop_back();
            index[n] = (int32_t)nodes.size();
            nodes.push_back(n);
            for (size_t i = n->countChildren(); i-- > 0;) stack.push_back(n->getChild((int)i));
        }
        /* pass 2: one record per node */
        const size_t N = nodes.size();
        type.resize(N); first_child.resize(N); n_children.resize(N); org_id.resize(N); has_xform.resize(N);
        pos.assign(N * 3, 0.f); dir.assign(N * 3, 0.f); res.assign(N * 4, 0.f); color.assign(N * 3, 0.f);
        xform.assign(N * 12, 0.f); oct_lo.assign(N * 3, 0.f); oct_hi.assign(N * 3, 0.f); srt.assign(N * 10, 0.f); has_srt.assign(N, 1);
        for (size_t i = 0; i < N; i++) {
            CBlobNode* n = nodes[i];
            const int t = node_type_from_reference(n->getNodeType());
            if (t < 0) { unsupported_type = (int)n->getNodeType(); return PSGPU_ERROR_UNSUPPORTED_NODE; }
            type[i] = t;
            org_id[i] = n->getID();
            first_child[i] = (int32_t)child.size();
            n_children[i] = (int32_t)n->countChildren();
            for (size_t k = 0; k < n->countChildren(); k++) child.push_back(index[n->getChild((int)k)]);
            /* backward (world -> local) matrix rows 0-2, row-major: PS_HighPerformanceRender.cpp:196-207 */
            PS::CMatrix back = n->getTransform().getBackwardMatrix();
            has_xform[i] = back.isIdentity() ? 0 : 1;
            float row[4];
            for (int r = 0; r < 3; r++) { back.getRow(row, r); for (int c = 0; c < 4; c++) xform[i * 12 + r * 4 + c] = row[c]; }
            {   // the transform as the scene API holds it (what flattenTransformations composes and saveScript writes)
                const vec3f sc = n->getTransform().getScale(), tr = n->getTransform().getTranslate();
                const vec4f ro = n->getTransform().getRotationVec4();
                float* q = &srt[i * 10];
                q[0] = sc.x; q[1] = sc.y; q[2] = sc.z; q[3] = ro.x; q[4] = ro.y; q[5] = ro.z; q[6] = ro.w; q[7] = tr.x; q[8] = tr.y; q[9] = tr.z;
            }
            PS::COctree oct = n->getOctree();
            put3(oct_lo, i, oct.lower); put3(oct_hi, i, oct.upper);
            if (PSGPU_IS_PRIM(t)) {
                vec4f d = n->getMaterial().diffused;
                color[i * 3 + 0] = d.x; color[i * 3 + 1] = d.y; color[i * 3 + 2] = d.z;
                fill_primitive(n, t, i);
            } else {
                fill_operator(n, t, i);
            }
        }
        if (child.empty()) child.push_back(0);
        /* instances name their origin by the reference node id (CompactBlobTree.cpp:385-392): res[0] = flat index of that node */
        for (size_t i = 0; i < N; i++) {
            if (type[i] != PSGPU_PRIM_INSTANCE) continue;
            PS::BLOBTREE::CInstance* inst = reinterpret_cast<PS::BLOBTREE::CInstance*>(nodes[i]);
            std::map<CBlobNode*, int32_t>::const_iterator it = index.find(inst->getOriginalNode());
            if (it == index.end()) { unsupported_type = (int)bntPrimInstance; return PSGPU_ERROR_UNSUPPORTED_NODE; }
            res[i * 4 + 0] = (float)it->second;
        }
        return 0;
    }

    psgpu_tree view() const {
        psgpu_tree t = psgpu_tree();
        t.n_nodes = (int32_t)type.size(); t.root = 0;
        t.type = type.data(); t.first_child = first_child.data(); t.n_children = n_children.data(); t.child = child.data();
        t.pos = pos.data(); t.dir = dir.data(); t.res = res.data(); t.color = color.data();
        t.xform = xform.data(); t.has_xform = has_xform.data();
        t.srt = srt.data(); t.has_srt = has_srt.data(); t.node_id = org_id.data();
        return t;
    }

    void clear() {
        type.clear(); first_child.clear(); n_children.clear(); child.clear(); org_id.clear();
        pos.clear(); dir.clear(); res.clear(); color.clear(); xform.clear(); oct_lo.clear(); oct_hi.clear(); has_xform.clear(); srt.clear(); has_srt.clear();
        unsupported_type = -1;
    }

private:
    static void put3(std::vector<float>& a, size_t i, const vec3f& v) { a[i * 3] = v.x; a[i * 3 + 1] = v.y; a[i * 3 + 2] = v.z; }

    /* same getters as PS_HighPerformanceRender.cpp:218-346 / CompactBlobTree.cpp:275-393 */
    void fill_primitive(PS::BLOBTREE::CBlobNode* n, int t, size_t i) {
        using namespace PS::BLOBTREE;
        if (t == PSGPU_PRIM_NULL || t == PSGPU_PRIM_INSTANCE) return;
        if (t == PSGPU_PRIM_QUADRICPOINT) {
            CQuadricPoint* q = reinterpret_cast<CQuadricPoint*>(n);
            put3(pos, i, q->getPosition());
            res[i * 4 + 0] = q->getFieldRadius();
            res[i * 4 + 1] = q->getFieldScale();
            return;
        }
        CSkeletonPrimitive* sprim = reinterpret_cast<CSkeletonPrimitive*>(n);
        CSkeleton* sk = sprim->getSkeleton();
        switch (t) {
        case PSGPU_PRIM_POINT: put3(pos, i, reinterpret_cast<CSkeletonPoint*>(sk)->getPosition()); break;
        case PSGPU_PRIM_LINE: {
            CSkeletonLine* l = reinterpret_cast<CSkeletonLine*>(sk);
            put3(pos, i, l->getStartPosition());
            put3(dir, i, l->getEndPosition()); /* dir = END point, as in SOABlobPrims */
        } break;
        case PSGPU_PRIM_CYLINDER: {
            CSkeletonCylinder* c = reinterpret_cast<CSkeletonCylinder*>(sk);
            put3(pos, i, c->getPosition()); put3(dir, i, c->getDirection());
            res[i * 4 + 0] = c->getRadius(); res[i * 4 + 1] = c->getHeight();
        } break;
        case PSGPU_PRIM_DISC: {
            CSkeletonDisc* d = reinterpret_cast<CSkeletonDisc*>(sk);
            put3(pos, i, d->getPosition()); put3(dir, i, d->getDirection());
            const float r = d->getRadius();
            res[i * 4 + 0] = r; res[i * 4 + 1] = r * r;
        } break;
        case PSGPU_PRIM_RING: {
            CSkeletonRing* g = reinterpret_cast<CSkeletonRing*>(sk);
            put3(pos, i, g->getPosition()); put3(dir, i, g->getDirection());
            const float r = g->getRadius();
            res[i * 4 + 0] = r; res[i * 4 + 1] = r * r;
        } break;
        case PSGPU_PRIM_CUBE: {
            CSkeletonCube* c = reinterpret_cast<CSkeletonCube*>(sk);
            put3(pos, i, c->getPosition());
            res[i * 4 + 0] = c->getSide();
        } break;
        default: break;
        }
    }

    /* PS_HighPerformanceRender.cpp:108-160 / CompactBlobTree.cpp:159-241 */
    void fill_operator(PS::BLOBTREE::CBlobNode* n, int t, size_t i) {
        using namespace PS::BLOBTREE;
        float* r = &res[i * 4];
        switch (t) {
        case PSGPU_OP_RICCI: {
            const float p = reinterpret_cast<CRicciBlend*>(n)->getN();
            r[0] = p;
            if (p != 0.0f) r[1] = 1.0f / p;
        } break;
        case PSGPU_OP_WARP_TWIST: {
            CWarpTwist* w = reinterpret_cast<CWarpTwist*>(n);
            r[0] = w->getWarpFactor(); r[1] = (float)w->getMajorAxis();
        } break;
        case PSGPU_OP_WARP_TAPER: {
            CWarpTaper* w = reinterpret_cast<CWarpTaper*>(n);
            r[0] = w->getWarpFactor(); r[1] = (float)w->getAxisAlong(); r[2] = (float)w->getAxisTaper();
        } break;
        case PSGPU_OP_WARP_BEND: {
            CWarpBend* w = reinterpret_cast<CWarpBend*>(n);
            r[0] = w->getBendRate(); r[1] = w->getBendCenter(); r[2] = w->getBendRegion().left; r[3] = w->getBendRegion().right;
        } break;
        case PSGPU_OP_WARP_SHEAR: {
            CWarpShear* w = reinterpret_cast<CWarpShear*>(n);
            r[0] = w->getWarpFactor(); r[1] = (float)w->getAxisAlong(); r[2] = (float)w->getAxisDependent();
        } break;
        default: break;
        }
    }
};

}  // namespace psgpu

/* The reference signature of ParsipHaptics/include/PS_HighPerformanceRender.h:27, declared in PS_SimdPolyCompat.h and
 * defined here because only a TU that sees the scene API can walk a CBlobNode. Return values are the reference's:
 * the root id (>= 0) or PS_ERROR_PRIM_OVERFLOW -1 / OPERATOR_OVERFLOW -2 / NON_BINARY_OP -3 (cpp:10-12). */
#ifdef PS_SIMDPOLY_COMPAT_H
inline int SimdPoly::linearizeBlobTree(PS::BLOBTREE::CBlobNode* root) {
    psgpu::FlatBlobTree flat;
    const int rc = flat.walk(root);
    if (rc < 0) return rc;
    const psgpu_tree t = flat.view();
    return linearizeBlobTree(&t);
}
#endif

#endif
