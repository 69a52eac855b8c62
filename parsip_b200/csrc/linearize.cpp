// Host-side BlobTree linearizer and bounding boxes (no GPU code).
//
// Replaces SimdPoly::linearizeBlobTree (ParsipHaptics/include/PS_HighPerformanceRender.cpp:42-371) and
// COMPACTBLOBTREE::convert (ParsipHaptics/include/CompactBlobTree.cpp:25-408): instead of SOA / AoS node
// tables walked by an explicit-stack or recursive evaluator, the tree is flattened once into the postfix
// quad stream of program.h. Also restates PrepareBBoxes / CountMPUNeeded
// (PS_SimdPoly/include/PS_Polygonizer.cpp:55-309, 388-412), which are host-only in the reference too.
#include <cmath>
#include <cstdio>
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/psgpu.h"
#include "host_program.h"
#include "program.h"

namespace {

struct Emitter {
    const psgpu_tree* t;
    int semantics;
    int flags;
    psgpu_program* p;
    int depth = 0, pdepth = 0;
    int err = PSGPU_SUCCESS;
    std::string msg;

    void quad(uint32_t h, float a = 0.f, float b = 0.f, float c = 0.f) {
        p->words.push_back(h);
        uint32_t u;
        memcpy(&u, &a, 4); p->words.push_back(u);
        memcpy(&u, &b, 4); p->words.push_back(u);
        memcpy(&u, &c, 4); p->words.push_back(u);
    }
    void quadf(float a, float b, float c, float d) {
        float v[4] = {a, b, c, d};
        for (int i = 0; i < 4; i++) { uint32_t u; memcpy(&u, &v[i], 4); p->words.push_back(u); }
    }
    void push() { depth++; if (depth > p->maxValueStack) p->maxValueStack = depth; }
    void pop() { depth--; }
    void ppush() { pdepth++; if (pdepth > p->maxPointStack) p->maxPointStack = pdepth; }

    void rows(const float* m) { quadf(m[0], m[1], m[2], m[3]); quadf(m[4], m[5], m[6], m[7]); quadf(m[8], m[9], m[10], m[11]); }

    // An OPERATOR's own backward matrix in the compact tree is applied as row . (x, y, z, w) with the w lane of the incoming
    // vec4: 0 from COMPACTBLOBTREE::fieldvalue and after any warp (Vector4d() zero-initialises w, FW/PS_Vector.h:1004-1007),
    // 1 right after another operator's matrix (CompactBlobTree.cpp:685-694). So the translation column only acts below an
    // affine-only parent. The lane is a static property of the node's ancestors: w == 0 drops the column here, at no cost to
    // the interpreter. PSGPU_LIN_FULL_AFFINE applies the whole matrix instead (what the scene API's authors meant).
    void opAffine(int n, int w) {
        const float* m = &t->xform[n * 12];
        quad(PS_HDR(OPC_AFFINE, 0, false, 0));
        if (w || (flags & PSGPU_LIN_FULL_AFFINE)) rows(m);
        else { quadf(m[0], m[1], m[2], 0.f); quadf(m[4], m[5], m[6], 0.f); quadf(m[8], m[9], m[10], 0.f); }
        p->nInstr++; ppush();
    }

    bool hasXf(int n) const { return t->xform && t->has_xform && t->has_xform[n]; }

    void emitPrim(int n, uint32_t cmb) {
        const float* pos = &t->pos[n * 3];
        const float* dir = &t->dir[n * 3];
        const float* res = &t->res[n * 4];
        const bool xf = hasXf(n);
        uint32_t id = (uint32_t)p->nPrims;
        if (id >= PS_MAX_PRIMS) { err = PSGPU_NOT_ENOUGH_MEM; msg = "too many primitives"; return; }
        p->nPrims++;
        p->colors.push_back(t->color[n * 3]); p->colors.push_back(t->color[n * 3 + 1]); p->colors.push_back(t->color[n * 3 + 2]);
        p->nInstr++;
        if (cmb == CMB_PUSH) push();
        switch (t->type[n]) {
        case PSGPU_PRIM_POINT:
            quad(PS_HDR(OPC_POINT, cmb, xf, id), pos[0], pos[1], pos[2]);
            break;
        case PSGPU_PRIM_LINE: {
            // pos = start, dir = END point (PS_HighPerformanceRender.cpp:218-236)
            float dx = dir[0] - pos[0], dy = dir[1] - pos[1], dz = dir[2] - pos[2];
            float dd = dx * dx + dy * dy + dz * dz;  // same order as PS_Polygonizer.cpp:993
            if (semantics == PSGPU_SEM_SIMD) {
                quad(PS_HDR(OPC_LINE_INF, cmb, xf, id), pos[0], pos[1], pos[2]);
                quadf(dx, dy, dz, dd);
            } else {
                quad(PS_HDR(OPC_SEGMENT, cmb, xf, id), pos[0], pos[1], pos[2]);
                quadf(dx, dy, dz, dd > 0.f ? 1.0f / dd : 0.f);  // degenerate segment = its start point (PS_GeometryFuncs.cpp:161-165)
            }
        } break;
        case PSGPU_PRIM_CYLINDER:
            quad(PS_HDR(OPC_CYLINDER, cmb, xf, id), pos[0], pos[1], pos[2]);
            quadf(dir[0], dir[1], dir[2], res[0]);
            quadf(res[1], 0.f, 0.f, 0.f);
            break;
        case PSGPU_PRIM_CUBE:
            quad(PS_HDR(OPC_CUBE, cmb, xf, id), pos[0], pos[1], pos[2]);
            quadf(res[0], 0.f, 0.f, 0.f);
            break;
        case PSGPU_PRIM_DISC:
            quad(PS_HDR(OPC_DISC, cmb, xf, id), pos[0], pos[1], pos[2]);
            quadf(dir[0], dir[1], dir[2], res[0]);
            break;
        case PSGPU_PRIM_RING:
            quad(PS_HDR(OPC_RING, cmb, xf, id), pos[0], pos[1], pos[2]);
            quadf(dir[0], dir[1], dir[2], res[0]);
            break;
        case PSGPU_PRIM_NULL:
            quad(PS_HDR(OPC_NULLPRIM, cmb, false, id));
            break;
        case PSGPU_PRIM_QUADRICPOINT: {
            if (semantics == PSGPU_SEM_SIMD) { err = PSGPU_ERROR_UNSUPPORTED_NODE; msg = "quadric points exist only in the scalar semantics"; return; }
            // the culling pass folds operands that are identically zero, which needs fields >= 0
            if (!(res[0] > 0.f) || !(res[1] >= 0.f)) { err = PSGPU_PARAM_ERROR; msg = "quadric point needs radius > 0 and scale >= 0"; return; }
            quad(PS_HDR(OPC_QUADRIC, cmb, xf, id), pos[0], pos[1], pos[2]);
            quadf(1.0f / (res[0] * res[0]), res[1], res[0], 0.f);
        } break;
        default:
            err = PSGPU_ERROR_UNSUPPORTED_NODE;
            msg = "primitive type " + std::to_string(t->type[n]) + " is not lowered to the GPU (SURVEY Q13)";
            return;
        }
        if (xf && t->type[n] != PSGPU_PRIM_NULL) rows(&t->xform[n * 12]);
    }

    // SimdPoly::linearizeBlobTree never reads an operator's transform (PS_HighPerformanceRender.cpp:70-160): its callers run
    // flattenTransformations first (CLayerManager.cpp:604-641). Refuse instead of silently evaluating a different surface.
    void simdOpXform() {
        err = PSGPU_ERROR_UNSUPPORTED_NODE;
        msg = "SIMD semantics ignore operator transforms: flatten them into the primitives first (psgpu_tree_flatten_transforms)";
    }

    void combine(uint32_t cmb) {
        if (cmb == CMB_PUSH) return;
        quad(PS_HDR(OPC_COMBINE, cmb, false, 0));
        p->nInstr++;
        pop();
    }

    static uint32_t combineOf(int type) {
        switch (type) {
        case PSGPU_OP_BLEND: return CMB_ADD;
        case PSGPU_OP_UNION: return CMB_MAX;
        case PSGPU_OP_INTERSECT: return CMB_MIN;
        case PSGPU_OP_DIF: return CMB_DIF;
        case PSGPU_OP_SMOOTHDIF: return CMB_SDIF;
        }
        return CMB_PUSH;
    }

    // children of an associative operator with same-type operator children spliced in
    void gatherKids(int n, int type, std::vector<int>& out) {
        const int nk = t->n_children[n];
        for (int i = 0; i < nk; i++) {
            int c = t->child[t->first_child[n] + i];
            bool assoc = (type == PSGPU_OP_UNION || type == PSGPU_OP_INTERSECT) ||
                         (type == PSGPU_OP_BLEND && (flags & PSGPU_LIN_REASSOCIATE));
            if (asShipped()) assoc = false;  // every operator keeps its own value on the stack: it may be masked (OPC_BOXMASK)
            if (assoc && t->type[c] == type && !hasXf(c) && t->n_children[c] > 0) gatherKids(c, type, out);
            else out.push_back(c);
        }
    }

    // CInstance (BT/CInstance.h, CompactBlobTree.cpp:1069-1078): the origin node evaluated at the instance's local point
    // (row . (p, 1), so the whole matrix acts), with w = 1 below it. There is no call instruction on the device: the origin's
    // sub-tree is emitted again in place. Colours: the copy is coloured at the instance's own point (the reference looks the
    // origin's colour up from the stores of the origin's OWN evaluation elsewhere in the tree, CompactBlobTree.cpp:1109-1122).
    int instanceDepth = 0;
    void emitInstance(int n, uint32_t cmb) {
        if (semantics == PSGPU_SEM_SIMD) { err = PSGPU_ERROR_UNSUPPORTED_NODE; msg = "instances exist only in the scalar semantics"; return; }
        const int origin = (int)t->res[n * 4];
        if (origin < 0 || origin >= t->n_nodes || origin == n || instanceDepth > 8) {
            err = PSGPU_PARAM_ERROR; msg = "instance " + std::to_string(n) + ": origin index out of range or instancing too deep"; return;
        }
        const bool xf = hasXf(n);
        if (xf) {
            quad(PS_HDR(OPC_AFFINE, 0, false, 0)); rows(&t->xform[n * 12]);
            p->nInstr++; ppush();
            if (pdepth > PS_MAX_POINT_STACK) { err = PSGPU_NOT_ENOUGH_MEM; msg = "instance nesting deeper than PS_MAX_POINT_STACK"; return; }
        }
        instanceDepth++;
        if (xf) { emit(origin, CMB_PUSH, 1); quad(PS_HDR(OPC_POPPOINT, 0, false, 0)); p->nInstr++; pdepth--; combine(cmb); }
        else emit(origin, cmb, 1);
        instanceDepth--;
    }

    bool asShipped() const { return semantics == PSGPU_SEM_SIMD && (flags & PSGPU_LIN_AS_SHIPPED) != 0; }
    std::vector<int> heightMemo;
    int height(int n) {  // of the operator tree below n (primitives: 0); iterative, chains can be thousands of links deep
        if (heightMemo.empty()) heightMemo.assign((size_t)t->n_nodes, -1);
        if (heightMemo[n] >= 0) return heightMemo[n];
        std::vector<int> st{n};
        while (!st.empty()) {
            const int v = st.back();
            bool ready = true;
            int h = 0;
            for (int i = 0; i < t->n_children[v]; i++) {
                const int c = t->child[t->first_child[v] + i];
                if (heightMemo[c] < 0) { if (ready) ready = false; st.push_back(c); }
                else h = std::max(h, heightMemo[c] + 1);
            }
            if (ready) { heightMemo[v] = PSGPU_IS_PRIM(t->type[v]) ? 0 : h; st.pop_back(); }
        }
        return heightMemo[n];
    }
    int opDepth = -1;  // depth of the operator being emitted in the reference's traversal (root operator = 0)
    // the reference's prune of an operator deeper than 3 levels (PS_Polygonizer.cpp:1228-1252): its value, alone on the stack top, is
    // masked by its box. Bottom-up masking gives what the reference's top-down skip gives: a pruned ancestor's 0 overrides its sub-tree
    void boxMask(int n) {
        quad(PS_HDR(OPC_BOXMASK, 0, false, 0));
        quadf(t->op_box_lo[n * 3], t->op_box_lo[n * 3 + 1], t->op_box_lo[n * 3 + 2], 0.f);
        quadf(t->op_box_hi[n * 3], t->op_box_hi[n * 3 + 1], t->op_box_hi[n * 3 + 2], 0.f);
        p->nInstr++;
    }

    // Emits the sub-tree `n`; its value meets the current stack top with `cmb` (CMB_PUSH: new entry).
    void emit(int n, uint32_t cmb, int w = 0) {
        if (err != PSGPU_SUCCESS) return;
        const int type = t->type[n];
        if (asShipped() && !PSGPU_IS_PRIM(type)) {
            if (PSGPU_IS_WARP(type) || cmb != CMB_PUSH) { err = PSGPU_ERROR_UNSUPPORTED_NODE; msg = "PSGPU_LIN_AS_SHIPPED: binary operators without warps only"; return; }
            opDepth++;
            emitOp(n, cmb, w);
            if (err == PSGPU_SUCCESS && opDepth > 3) boxMask(n);
            opDepth--;
            return;
        }
        emitOp(n, cmb, w);
    }
    void emitOp(int n, uint32_t cmb, int w) {
        if (err != PSGPU_SUCCESS) return;
        const int type = t->type[n];
        if (type == PSGPU_PRIM_INSTANCE) { emitInstance(n, cmb); return; }
        if (PSGPU_IS_PRIM(type)) { emitPrim(n, cmb); return; }
        const int nk = t->n_children[n];
        if (nk == 0) {  // an operator without children contributes nothing (CompactBlobTree.cpp:1138-1139)
            quad(PS_HDR(OPC_NULLPRIM, cmb, false, 0)); p->nInstr++;
            if (cmb == CMB_PUSH) push();
            return;
        }
        if (PSGPU_IS_WARP(type)) {
            const int c0 = t->child[t->first_child[n]];
            if (semantics == PSGPU_SEM_SIMD) {  // PS_Polygonizer.cpp:1318-1337: warps return the left child unchanged (Q4)
                if (hasXf(n)) { simdOpXform(); return; }
                emit(c0, cmb);
                return;
            }
            int pops = 0;
            if (hasXf(n)) {  // operator's own matrix, then the warp (CompactBlobTree.cpp:685-746)
                opAffine(n, w); pops++;
            }
            const float* r = &t->res[n * 4];
            switch (type) {
            case PSGPU_OP_WARP_TWIST: quad(PS_HDR(OPC_TWIST, 0, false, 0), r[0], r[1]); break;
            case PSGPU_OP_WARP_TAPER: quad(PS_HDR(OPC_TAPER, 0, false, 0), r[0], r[1], r[2]); break;
            case PSGPU_OP_WARP_BEND: quad(PS_HDR(OPC_BEND, 0, false, 0), r[0], r[1], r[2]); quadf(r[3], 1.0f / r[0], 0.f, 0.f); break;
            case PSGPU_OP_WARP_SHEAR: quad(PS_HDR(OPC_SHEAR, 0, false, 0), r[0], r[1], r[2]); break;
            }
            p->nInstr++; ppush(); pops++;
            if (pdepth > PS_MAX_POINT_STACK) { err = PSGPU_NOT_ENOUGH_MEM; msg = "warp nesting deeper than PS_MAX_POINT_STACK"; return; }
            emit(c0, CMB_PUSH, 0);  // every warp returns a fresh vec4: w == 0 below it
            for (int i = 0; i < pops; i++) { quad(PS_HDR(OPC_POPPOINT, 0, false, 0)); p->nInstr++; pdepth--; }
            combine(cmb);
            return;
        }
        int pops = 0;
        if (hasXf(n)) {  // un-flattened operator transform (Q14): same treatment as the compact tree
            if (semantics == PSGPU_SEM_SIMD) { simdOpXform(); return; }
            opAffine(n, w); pops++; w = 1;
            if (pdepth > PS_MAX_POINT_STACK) { err = PSGPU_NOT_ENOUGH_MEM; msg = "affine nesting deeper than PS_MAX_POINT_STACK"; return; }
        }
        if (type == PSGPU_OP_RICCI) {
            const float nexp = t->res[n * 4], inv = (nexp != 0.f) ? 1.0f / nexp : 0.f;
            if (semantics == PSGPU_SEM_SIMD) {
                if (nk != 2) { err = PSGPU_INVALID_BVH; msg = "SIMD semantics need binary operators (PS_HighPerformanceRender.cpp:81-86)"; return; }
                emit(t->child[t->first_child[n]], CMB_PUSH, w);
                emit(t->child[t->first_child[n] + 1], CMB_PUSH, w);
                quad(PS_HDR(OPC_RICCI_SIMD, 0, false, 0), inv); p->nInstr++; pop();
            } else {
                for (int i = 0; i < nk; i++) {
                    emit(t->child[t->first_child[n] + i], CMB_PUSH, w);
                    quad(PS_HDR(OPC_POW, 1, false, 0), nexp); p->nInstr++;
                    if (i > 0) combine(CMB_RADD);
                }
                quad(PS_HDR(OPC_POW, 0, false, 0), inv); p->nInstr++;
            }
        } else {
            const uint32_t x = combineOf(type);
            if (x == CMB_PUSH) { err = PSGPU_ERROR_UNSUPPORTED_NODE; msg = "operator type " + std::to_string(type); return; }
            if (semantics == PSGPU_SEM_SIMD && nk != 2) {
                err = PSGPU_INVALID_BVH; msg = "SIMD semantics need binary operators (PS_HighPerformanceRender.cpp:81-86)"; return;
            }
            std::vector<int> kids;
            gatherKids(n, type, kids);
            if (semantics == PSGPU_SEM_SIMD && kids.size() == 2 && PSGPU_IS_PRIM(t->type[kids[0]]) && !PSGPU_IS_PRIM(t->type[kids[1]])) {
                // op(primitive, sub-tree): the shape of ConvertToBinaryTree's right-leaning chains (CLayerManager.cpp:131-176).
                // Evaluate the sub-tree first and fold the primitive in as the LEFT operand: one live stack entry instead of
                // one per chain link, and bit-identical (IEEE add / max / min commute; the differences have reverse codes).
                emit(kids[1], CMB_PUSH, w);
                emitPrim(kids[0], x == CMB_DIF ? (uint32_t)CMB_RDIF : (x == CMB_SDIF ? (uint32_t)CMB_RSDIF : x));
                for (int i = 0; i < pops; i++) { quad(PS_HDR(OPC_POPPOINT, 0, false, 0)); p->nInstr++; pdepth--; }
                combine(cmb);
                return;
            }
            if (asShipped() && kids.size() == 2 && !PSGPU_IS_PRIM(t->type[kids[0]]) && !PSGPU_IS_PRIM(t->type[kids[1]]) &&
                height(kids[1]) > height(kids[0])) {
                // op(sub-tree, taller sub-tree): nothing is gathered in this mode, so a right-leaning chain of operators (what
                // ConvertToBinaryTree makes of an n-ary union) would keep one stack entry per link. The taller side goes first and the
                // combine takes its reverse code (same operands, same bits): a chain needs two entries.
                emit(kids[1], CMB_PUSH, w);
                emit(kids[0], CMB_PUSH, w);
                combine(x == CMB_DIF ? (uint32_t)CMB_RDIF : (x == CMB_SDIF ? (uint32_t)CMB_RSDIF : x));
                for (int i = 0; i < pops; i++) { quad(PS_HDR(OPC_POPPOINT, 0, false, 0)); p->nInstr++; pdepth--; }
                combine(cmb);
                return;
            }
            emit(kids[0], CMB_PUSH, w);
            for (size_t i = 1; i < kids.size(); i++) {
                if (t->type[kids[i]] == PSGPU_PRIM_INSTANCE) emit(kids[i], x, w);
                else if (PSGPU_IS_PRIM(t->type[kids[i]])) emitPrim(kids[i], x);
                else { emit(kids[i], CMB_PUSH, w); combine(x); }
                if (err != PSGPU_SUCCESS) return;
            }
        }
        for (int i = 0; i < pops; i++) { quad(PS_HDR(OPC_POPPOINT, 0, false, 0)); p->nInstr++; pdepth--; }
        combine(cmb);
    }
};

}  // namespace

extern "C" int psgpu_linearize(const psgpu_tree* tree, int semantics, int flags, psgpu_program** out) {
    if (!tree || !out || tree->n_nodes <= 0 || tree->root < 0 || tree->root >= tree->n_nodes) return PSGPU_PARAM_ERROR;
    psgpu_program* p = new psgpu_program();
    p->semantics = semantics;
    Emitter e{tree, semantics, flags, p};
    if (semantics == PSGPU_SEM_SIMD && (flags & PSGPU_LIN_AS_SHIPPED) && (!tree->op_box_lo || !tree->op_box_hi)) {
        fprintf(stderr, "psgpu_linearize: PSGPU_LIN_AS_SHIPPED needs psgpu_tree.op_box_lo / op_box_hi\n");
        delete p;
        *out = nullptr;
        return PSGPU_PARAM_ERROR;
    }
    e.emit(tree->root, CMB_PUSH);
    if (e.err == PSGPU_SUCCESS && p->maxValueStack > PS_MAX_VALUE_STACK) {
        e.err = PSGPU_NOT_ENOUGH_MEM;
        e.msg = "value stack depth " + std::to_string(p->maxValueStack) + " exceeds PS_MAX_VALUE_STACK; linearize with PSGPU_LIN_REASSOCIATE";
    }
    if (e.err != PSGPU_SUCCESS) {
        fprintf(stderr, "psgpu_linearize: %s\n", e.msg.c_str());
        delete p;
        *out = nullptr;
        return e.err;
    }
    e.quad(PS_HDR(OPC_END, 0, false, 0));
    *out = p;
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_program_free(psgpu_program* p) { delete p; return PSGPU_SUCCESS; }

extern "C" int psgpu_program_info(const psgpu_program* p, int64_t* info) {
    if (!p || !info) return PSGPU_PARAM_ERROR;
    info[0] = (int64_t)p->words.size() / 4; info[1] = p->nPrims; info[2] = p->maxValueStack;
    info[3] = p->maxPointStack; info[4] = p->nInstr;
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_program_copy(const psgpu_program* p, uint32_t* out) {
    if (!p || !out) return PSGPU_PARAM_ERROR;
    memcpy(out, p->words.data(), p->words.size() * 4);
    return PSGPU_SUCCESS;
}

extern "C" int psgpu_program_colors(const psgpu_program* p, float* out) {
    if (!p || !out) return PSGPU_PARAM_ERROR;
    memcpy(out, p->colors.data(), p->colors.size() * 4);
    return PSGPU_SUCCESS;
}

// ---------------------------------------------------------------------------------------------------
// bounding boxes
// ---------------------------------------------------------------------------------------------------
namespace {

const float kIsoPad = 0.45420206f /* ISO_DIST, PS_Polygonizer.h:59 */ + 5.0f * 0.01f /* 5*MIN_CELL_SIZE, .cpp:65 */;

struct Box { float lo[3], hi[3]; bool valid = false; };

void boxUnion(Box& a, const Box& b) {
    if (!b.valid) return;
    if (!a.valid) { a = b; return; }
    for (int i = 0; i < 3; i++) { a.lo[i] = std::fmin(a.lo[i], b.lo[i]); a.hi[i] = std::fmax(a.hi[i], b.hi[i]); }
}

// primitive boxes: PS_Polygonizer.cpp:71-195
Box primBox(const psgpu_tree* t, int n) {
    Box b;
    const float* p = &t->pos[n * 3];
    const float* d = &t->dir[n * 3];
    const float* r = &t->res[n * 4];
    const float iso = kIsoPad;
    b.valid = true;
    switch (t->type[n]) {
    case PSGPU_PRIM_POINT:
        for (int a = 0; a < 3; a++) { b.lo[a] = p[a] - iso; b.hi[a] = p[a] + iso; }
        break;
    case PSGPU_PRIM_LINE:
        for (int a = 0; a < 3; a++) {
            float expand = iso * 1.0f + (3.0f * iso) * (d[a] - p[a]);
            b.lo[a] = p[a] - expand; b.hi[a] = d[a] + expand;
        }
        break;
    case PSGPU_PRIM_RING: case PSGPU_PRIM_DISC: {
        float radius = r[0] + iso;
        for (int a = 0; a < 3; a++) {
            float expand = (radius + iso) * (1.0f - d[a]) + iso * d[a];
            b.lo[a] = p[a] - expand; b.hi[a] = p[a] + expand;
        }
    } break;
    case PSGPU_PRIM_CYLINDER:
        for (int a = 0; a < 3; a++) {
            float s1 = p[a] + r[1] * d[a];
            float expand = (iso + r[0]) * 1.0f + (0.5f * iso) * d[a];
            b.lo[a] = p[a] - expand; b.hi[a] = s1 + expand;
        }
        break;
    case PSGPU_PRIM_CUBE: {
        float side = r[0] + iso;
        for (int a = 0; a < 3; a++) { b.lo[a] = p[a] - side; b.hi[a] = p[a] + side; }
    } break;
    default: b.valid = false; break;
    }
    return b;
}

// inverse of a backward 3x4 (world->local) affine, in double
bool invertAffine(const float* m, double* inv /*3x4*/) {
    double a[3][3] = {{m[0], m[1], m[2]}, {m[4], m[5], m[6]}, {m[8], m[9], m[10]}};
    double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                 a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    if (std::fabs(det) < 1e-30) return false;
    double id = 1.0 / det, r[3][3];
    r[0][0] = (a[1][1] * a[2][2] - a[1][2] * a[2][1]) * id; r[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) * id; r[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) * id;
    r[1][0] = (a[1][2] * a[2][0] - a[1][0] * a[2][2]) * id; r[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) * id; r[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) * id;
    r[2][0] = (a[1][0] * a[2][1] - a[1][1] * a[2][0]) * id; r[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) * id; r[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) * id;
    double tr[3] = {m[3], m[7], m[11]};
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) inv[i * 4 + j] = r[i][j];
        inv[i * 4 + 3] = -(r[i][0] * tr[0] + r[i][1] * tr[1] + r[i][2] * tr[2]);
    }
    return true;
}

// image of all 8 corners (the reference maps 2 corners only, which is wrong under rotation)
Box xformBox(const Box& b, const double* fwd) {
    Box o; o.valid = b.valid;
    if (!b.valid) return o;
    for (int c = 0; c < 8; c++) {
        double v[3] = {(c & 1) ? b.hi[0] : b.lo[0], (c & 2) ? b.hi[1] : b.lo[1], (c & 4) ? b.hi[2] : b.lo[2]};
        for (int i = 0; i < 3; i++) {
            float w = (float)(fwd[i * 4] * v[0] + fwd[i * 4 + 1] * v[1] + fwd[i * 4 + 2] * v[2] + fwd[i * 4 + 3]);
            if (c == 0) { o.lo[i] = o.hi[i] = w; } else { o.lo[i] = std::fmin(o.lo[i], w); o.hi[i] = std::fmax(o.hi[i], w); }
        }
    }
    return o;
}

// forward warps in double, CompactBlobTree.cpp:1315-1536
void warpFwd(int type, const float* r, const double* p, double* o) {
    o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
    switch (type) {
    case PSGPU_OP_WARP_TWIST: {
        int ax = (int)r[1]; double th = p[ax] * r[0], c = std::cos(th), s = std::sin(th);
        int u = ax == 0 ? 1 : 0, v = ax == 2 ? 1 : 2;
        o[u] = p[u] * c - p[v] * s; o[v] = p[u] * s + p[v] * c;
    } break;
    case PSGPU_OP_WARP_TAPER: {
        int al = (int)r[1], tp = (int)r[2], tgt;
        if (al == 0) tgt = (tp == 2) ? 2 : 1; else if (al == 1) tgt = (tp == 2) ? 2 : 0; else tgt = (tp == 2) ? 1 : 0;
        o[tgt] = p[tgt] * (1.0 + p[al] * r[0]);
    } break;
    case PSGPU_OP_WARP_BEND: {
        double k = r[0], y0 = r[1], l = r[2], rr = r[3], kd = 1.0 / k;
        double yh = p[1] <= l ? l : (p[1] >= rr ? rr : p[1]);
        double th = k * (yh - y0), c = std::cos(th), s = std::sin(th), dy = p[1] - yh;
        o[1] = -s * (p[2] - kd) + y0 + c * dy; o[2] = c * (p[2] - kd) + kd + s * dy;
    } break;
    case PSGPU_OP_WARP_SHEAR: {
        int al = (int)r[1], dp = (int)r[2];
        if (al == 1) o[1] = p[1] + r[0] * (dp == 2 ? p[2] : p[0]);
        else if (al == 2) o[2] = p[2] + r[0] * (dp == 1 ? p[1] : p[0]);
        else o[0] = p[0] + r[0] * (dp == 2 ? p[2] : p[1]);
    } break;
    }
}

// World-space box of {x : warp(x) in b}: the pre-image of a 9^3 lattice of b found by fixed-point iteration
// (x <- x + (q - w(x))), padded. Replaces the reference's "child box +- 0.6" heuristic
// (BT/CWarpBend.h:150-155, CWarpTaper.h:206-211, CWarpTwist.h:133-138); like it, this bounds the
// iso-surface for moderate warps and is a heuristic for extreme ones.
Box unwarpBox(const Box& b, int type, const float* r) {
    Box o; o.valid = b.valid;
    if (!b.valid) return o;
    const int S = 9;
    bool first = true;
    for (int i = 0; i < S; i++) for (int j = 0; j < S; j++) for (int k = 0; k < S; k++) {
        double q[3] = {b.lo[0] + (b.hi[0] - b.lo[0]) * i / (S - 1.0), b.lo[1] + (b.hi[1] - b.lo[1]) * j / (S - 1.0),
                       b.lo[2] + (b.hi[2] - b.lo[2]) * k / (S - 1.0)};
        double x[3] = {q[0], q[1], q[2]}, w[3];
        for (int it = 0; it < 40; it++) {
            warpFwd(type, r, x, w);
            double e = 0;
            for (int a = 0; a < 3; a++) { x[a] += q[a] - w[a]; e += std::fabs(q[a] - w[a]); }
            if (e < 1e-7) break;
        }
        for (int a = 0; a < 3; a++) {
            if (!std::isfinite(x[a])) continue;
            if (first) { o.lo[a] = o.hi[a] = (float)x[a]; }
            else { o.lo[a] = std::fmin(o.lo[a], (float)x[a]); o.hi[a] = std::fmax(o.hi[a], (float)x[a]); }
        }
        first = false;
    }
    for (int a = 0; a < 3; a++) { o.lo[a] -= 0.1f; o.hi[a] += 0.1f; }
    return o;
}

Box treeBox(const psgpu_tree* t, int n) {
    Box b;
    const int type = t->type[n];
    const bool xf = t->xform && t->has_xform && t->has_xform[n];
    if (PSGPU_IS_PRIM(type)) b = primBox(t, n);
    else {
        for (int i = 0; i < t->n_children[n]; i++) boxUnion(b, treeBox(t, t->child[t->first_child[n] + i]));
        if (PSGPU_IS_WARP(type)) b = unwarpBox(b, type, &t->res[n * 4]);
    }
    if (xf && b.valid) {
        double fwd[12];
        if (invertAffine(&t->xform[n * 12], fwd)) b = xformBox(b, fwd);
    }
    return b;
}

}  // namespace

extern "C" int psgpu_prepare_bboxes(const psgpu_tree* tree, float cellsize, float* lo, float* hi) {
    (void)cellsize;
    if (!tree || tree->n_nodes <= 0) return PSGPU_PARAM_ERROR;
    Box b = treeBox(tree, tree->root);
    if (!b.valid) return PSGPU_PARAM_ERROR;
    for (int a = 0; a < 3; a++) { lo[a] = b.lo[a]; hi[a] = b.hi[a]; }
    return PSGPU_SUCCESS;
}

// CountMPUNeeded, PS_Polygonizer.cpp:388-412
extern "C" void psgpu_lattice_dims(float cellsize, const float* lo, const float* hi, int32_t* nm) {
    for (int a = 0; a < 3; a++) {
        float side = hi[a] - lo[a];
        int cells = (int)std::ceil(side / cellsize);
        nm[a] = cells / PS_CELLS_PER_MPU;
        if (cells % PS_CELLS_PER_MPU != 0) nm[a]++;
    }
}

extern "C" uint32_t psgpu_count_mpus(float cellsize, const float* lo, const float* hi) {
    int32_t nm[3];
    psgpu_lattice_dims(cellsize, lo, hi, nm);
    return (uint32_t)(nm[0] * nm[1] * nm[2]);
}
