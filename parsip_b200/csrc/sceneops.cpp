// Host-side scene passes and file I/O on the flat BlobTree, Qt / Loki-free restatements of
//   CLayer::flattenTransformations / recursive_FlattenTransformations   ParsipHaptics/include/CLayerManager.cpp:604-624
//   CLayer::setOctreeFromBlobTree / recursive_RecomputeAllOctrees       ParsipHaptics/include/CLayerManager.cpp:421-427, 626-645
//   CBlobNode::getOctree / computeOctree of every node class            PS_BlobTree/include/C*.h (cited at each case)
//   the .scene reader CLayer::recursive_ReadBlobNode                    ParsipHaptics/include/CLayerManager.cpp:684-800
//   the .scene writer CBlobNode::saveGenericInfoScript + saveScript     PS_BlobTree/include/CBlobTree.cpp:40-118 and the node headers
// Nothing here touches the device.
#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/psgpu.h"
#include "tree_buf.h"

namespace {

const float kIso = 0.5f;             // ISO_VALUE, PS_BlobTree/include/_constSettings.h
const float kWarpExpand = 0.6f;      // BOUNDING_OCTREE_EXPANSION_FACTOR, _constSettings.h:4

struct Srt { float s[3], q[4], t[3]; };
Srt srt_identity() { return Srt{{1, 1, 1}, {0, 0, 0, 1}, {0, 0, 0}}; }
Srt srt_of(const psgpu_tree* t, int n) {
    if (!t->srt || (t->has_srt && !t->has_srt[n])) return srt_identity();
    const float* p = t->srt + (size_t)n * 10;
    return Srt{{p[0], p[1], p[2]}, {p[3], p[4], p[5], p[6]}, {p[7], p[8], p[9]}};
}
bool srt_is_identity(const Srt& a) {
    return a.s[0] == 1 && a.s[1] == 1 && a.s[2] == 1 && a.q[0] == 0 && a.q[1] == 0 && a.q[2] == 0 && a.q[3] == 1 && a.t[0] == 0 && a.t[1] == 0 && a.t[2] == 0;
}

// CAffineTransformation::add (PS_AffineTransformation.h:46-52): scales multiply, quaternions multiply (CQuaternion::multiply,
// PS_Quaternion.h:231-240, then Normalize), translations ADD - not a matrix product; reproduced as it is.
Srt srt_add(const Srt& a, const Srt& r) {
    Srt o;
    for (int i = 0; i < 3; i++) { o.s[i] = a.s[i] * r.s[i]; o.t[i] = a.t[i] + r.t[i]; }
    const float ax = a.q[0], ay = a.q[1], az = a.q[2], aw = a.q[3], bx = r.q[0], by = r.q[1], bz = r.q[2], bw = r.q[3];
    float x = bw * ax + bx * aw + by * az - bz * ay;
    float y = bw * ay + by * aw + bz * ax - bx * az;
    float z = bw * az + bz * aw + bx * ay - by * ax;
    float w = bw * aw - bx * ax - by * ay - bz * az;
    const float sq = x * x + y * y + z * z + w * w;
    if (sq > 0.0f) { const float d = (float)(1.0f / sqrt(sq)); x *= d; y *= d; z *= d; w *= d; }
    o.q[0] = x; o.q[1] = y; o.q[2] = z; o.q[3] = w;
    return o;
}

// forward 3x4 (column-vector form): p' = R (S p) + t, the matrix setReverseTransformations builds (PS_AffineTransformation.h:160-181)
void srt_forward(const Srt& a, double F[3][4]) {
    const double x = a.q[0], y = a.q[1], z = a.q[2], w = a.q[3];
    const double R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)},
                            {2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)},
                            {2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)}};
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) F[r][c] = R[r][c] * a.s[c]; F[r][3] = a.t[r]; }
}
// backward = inverse of forward. The reference inverts in float32 by cofactors (CMatrix::invert); here in double and rounded
// once - entries agree to ~1e-7 (tests/test_scene_passes.py compares with matrices read back from the reference's objects).
void srt_backward(const Srt& a, float B[12]) {
    double F[3][4];
    srt_forward(a, F);
    const double m[3][3] = {{F[0][0], F[0][1], F[0][2]}, {F[1][0], F[1][1], F[1][2]}, {F[2][0], F[2][1], F[2][2]}};
    const double det = m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
                       m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
    const double id = (det != 0.0) ? 1.0 / det : 1.0;
    double inv[3][3];
    inv[0][0] = (m[1][1] * m[2][2] - m[1][2] * m[2][1]) * id; inv[0][1] = (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * id; inv[0][2] = (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * id;
    inv[1][0] = (m[1][2] * m[2][0] - m[1][0] * m[2][2]) * id; inv[1][1] = (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * id; inv[1][2] = (m[0][2] * m[1][0] - m[0][0] * m[1][2]) * id;
    inv[2][0] = (m[1][0] * m[2][1] - m[1][1] * m[2][0]) * id; inv[2][1] = (m[0][1] * m[2][0] - m[0][0] * m[2][1]) * id; inv[2][2] = (m[0][0] * m[1][1] - m[0][1] * m[1][0]) * id;
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) B[r * 4 + c] = (float)inv[r][c];
        B[r * 4 + 3] = (float)(-(inv[r][0] * F[0][3] + inv[r][1] * F[1][3] + inv[r][2] * F[2][3]));
    }
}

void copy_node(const psgpu_tree* in, int n, psgpu_tree_buf* o, const Srt& tr) {
    o->type.push_back(in->type[n]);
    o->ids.push_back(in->node_id ? in->node_id[n] : n);
    for (int a = 0; a < 3; a++) { o->pos.push_back(in->pos[n * 3 + a]); o->dir.push_back(in->dir[n * 3 + a]); o->color.push_back(in->color[n * 3 + a]); }
    for (int a = 0; a < 4; a++) o->res.push_back(in->res[n * 4 + a]);
    const float* s = &tr.s[0];
    o->srt.insert(o->srt.end(), s, s + 3); o->srt.insert(o->srt.end(), tr.q, tr.q + 4); o->srt.insert(o->srt.end(), tr.t, tr.t + 3);
    const bool ident = srt_is_identity(tr);
    o->hassrt.push_back(ident ? 0 : 1);
    float B[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    if (!ident) srt_backward(tr, B);
    o->xform.insert(o->xform.end(), B, B + 12);
    o->hasxf.push_back(ident ? 0 : 1);
}

bool is_unary(int type) { return PSGPU_IS_WARP(type); }  // CBlobNode::isUnary: node types from bntOpCache on (CBlobTree.h:80)

}  // namespace

// recursive_FlattenTransformations: a non-unary operator hands branch + own transform down and becomes identity; a primitive or a
// unary operator (warp) takes branch + own and keeps it - the walk stops there, a warp's subtree keeps its own transforms.
extern "C" int psgpu_tree_flatten_transforms(const psgpu_tree* in, psgpu_tree_buf** outp) {
    if (!in || !outp || in->n_nodes <= 0 || !in->srt) return PSGPU_PARAM_ERROR;  // the reference composes SRT triples: they are required
    psgpu_tree_buf* o = new psgpu_tree_buf();
    const int N = in->n_nodes;
    std::vector<Srt> eff((size_t)N);
    std::vector<char> visited((size_t)N, 0);
    struct Item { int n; Srt branch; bool apply; };
    std::vector<Item> st;
    st.push_back(Item{in->root, srt_identity(), true});
    while (!st.empty()) {
        const Item it = st.back();
        st.pop_back();
        const int n = it.n, ty = in->type[n];
        visited[(size_t)n] = 1;
        if (!it.apply) { eff[(size_t)n] = srt_of(in, n); }
        else if (!PSGPU_IS_PRIM(ty) && !is_unary(ty)) {
            const Srt t = srt_add(it.branch, srt_of(in, n));
            eff[(size_t)n] = srt_identity();
            for (int i = 0; i < in->n_children[n]; i++) st.push_back(Item{in->child[in->first_child[n] + i], t, true});
            continue;
        } else eff[(size_t)n] = srt_add(it.branch, srt_of(in, n));
        for (int i = 0; i < in->n_children[n]; i++) st.push_back(Item{in->child[in->first_child[n] + i], srt_identity(), false});
    }
    for (int n = 0; n < N; n++) {
        copy_node(in, n, o, visited[(size_t)n] ? eff[(size_t)n] : srt_of(in, n));
        o->first.push_back(in->first_child[n]); o->nk.push_back(in->n_children[n]);
    }
    int nChild = 0;
    for (int n = 0; n < N; n++) nChild = std::max(nChild, in->first_child[n] + in->n_children[n]);
    o->child.assign(in->child, in->child + std::max(nChild, 1));
    o->finish(in->root);
    *outp = o;
    return PSGPU_SUCCESS;
}

namespace {
struct Box { float lo[3], hi[3]; };

// CSkeleton*::bound(), the scene API's heuristic iso-surface bounds (padded by ISO_VALUE)
Box prim_bound_box(const psgpu_tree* t, int n) {
    const float* p = t->pos + n * 3;
    const float* d = t->dir + n * 3;
    const float* r = t->res + n * 4;
    Box b;
    switch (t->type[n]) {
    case PSGPU_PRIM_POINT:  // CSkeletonPoint.h:59-64
        for (int a = 0; a < 3; a++) { b.lo[a] = p[a] - kIso; b.hi[a] = p[a] + kIso; }
        break;
    case PSGPU_PRIM_LINE:  // CSkeletonLine.h:80-85: expand = ISO + 3 ISO (end - start); box(start - expand, end + expand) (Q12)
        for (int a = 0; a < 3; a++) { const float e = kIso + 3.0f * kIso * (d[a] - p[a]); b.lo[a] = p[a] - e; b.hi[a] = d[a] + e; }
        break;
    case PSGPU_PRIM_CYLINDER:  // CSkeletonCylinder.h:103-110
        for (int a = 0; a < 3; a++) {
            const float s1 = p[a] + r[1] * d[a];
            const float e = (kIso + r[0]) * 1.0f + 0.5f * kIso * d[a];
            b.lo[a] = p[a] - e; b.hi[a] = s1 + e;
        }
        break;
    case PSGPU_PRIM_DISC: case PSGPU_PRIM_RING:  // CSkeletonDisc.h:93-100, CSkeletonRing.h:107-114
        for (int a = 0; a < 3; a++) {
            const float radius = r[0] + kIso;
            const float e = radius * (1.0f - d[a]) + kIso * d[a];
            b.lo[a] = p[a] - e; b.hi[a] = p[a] + e;
        }
        break;
    case PSGPU_PRIM_CUBE:  // CSkeletonCube.h:175-182
        for (int a = 0; a < 3; a++) { const float ss = r[0] + kIso; b.lo[a] = p[a] - ss; b.hi[a] = p[a] + ss; }
        break;
    case PSGPU_PRIM_QUADRICPOINT:  // CQuadricPoint.h:59-63
        for (int a = 0; a < 3; a++) { b.lo[a] = p[a] - r[0]; b.hi[a] = p[a] + r[0]; }
        break;
    default:  // CNullPrimitive.h:25-30
        for (int a = 0; a < 3; a++) b.lo[a] = b.hi[a] = 0.0f;
        break;
    }
    return b;
}
}  // namespace

// Octree (bounding box) of every node. recompute = 0: what CBlobNode::getOctree() returns on a fresh tree (lazy computeOctree:
// primitive boxes untransformed). recompute = 1: after CLayer::setOctreeFromBlobTree, i.e. recursive_RecomputeAllOctrees -
// primitive boxes mapped by the accumulated FORWARD matrices (their two corners only, COctree::transform, Q12), operator boxes
// rebuilt from their children: union (blend / union / Ricci), first child only (difference, smooth difference), box intersection,
// child + 0.6 (warps). The root's box is the grid domain the GUI hands to the polygonizer.
extern "C" int psgpu_tree_octrees(const psgpu_tree* t, int recompute, float* lo3n, float* hi3n) {
    if (!t || !lo3n || !hi3n || t->n_nodes <= 0) return PSGPU_PARAM_ERROR;
    const int N = t->n_nodes;
    std::vector<Box> box((size_t)N);
    std::vector<char> done((size_t)N, 0);
    if (recompute) {
        // the recompute walk starts from nodes that already hold their lazily computed boxes (COMPACTBLOBTREE::convert and every
        // getOctree() before it cached them): an instance met before its origin is recomputed sees the origin's FRESH box
        std::vector<float> flo((size_t)N * 3), fhi((size_t)N * 3);
        const int rc = psgpu_tree_octrees(t, 0, flo.data(), fhi.data());
        if (rc != PSGPU_SUCCESS) return rc;
        for (int n = 0; n < N; n++) for (int a = 0; a < 3; a++) { box[(size_t)n].lo[a] = flo[(size_t)n * 3 + a]; box[(size_t)n].hi[a] = fhi[(size_t)n * 3 + a]; }
    }
    // accumulated forward matrices (row-major 4x4, column-vector convention): mtxCurrent = mtxBranch * node.forward
    struct M4 { double m[4][4]; };
    auto ident = [] { M4 r; for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) r.m[i][j] = (i == j); return r; };
    std::vector<M4> acc((size_t)N, ident());
    struct Frame { int n; int next; };
    std::vector<Frame> st;
    st.push_back(Frame{t->root, 0});
    acc[(size_t)t->root] = ident();
    auto fwd_of = [&](int n) {
        M4 f = ident();
        if (t->srt && (!t->has_srt || t->has_srt[n])) { double F[3][4]; srt_forward(srt_of(t, n), F); for (int r = 0; r < 3; r++) for (int c = 0; c < 4; c++) f.m[r][c] = F[r][c]; }
        return f;
    };
    auto mul = [](const M4& a, const M4& b) { M4 r; for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { double s = 0; for (int k = 0; k < 4; k++) s += a.m[i][k] * b.m[k][j]; r.m[i][j] = s; } return r; };
    {
        const M4 f = fwd_of(t->root);
        acc[(size_t)t->root] = f;
    }
    while (!st.empty()) {
        Frame& fr = st.back();
        const int n = fr.n, ty = t->type[n];
        if (PSGPU_IS_PRIM(ty) && ty != PSGPU_PRIM_INSTANCE) {
            Box b = prim_bound_box(t, n);
            if (recompute) {  // oct.transform(mtxCurrent): the two corners mapped, then min / max
                const M4& m = acc[(size_t)n];
                float a[3], c[3];
                for (int r = 0; r < 3; r++) {
                    a[r] = (float)(m.m[r][0] * b.lo[0] + m.m[r][1] * b.lo[1] + m.m[r][2] * b.lo[2] + m.m[r][3]);
                    c[r] = (float)(m.m[r][0] * b.hi[0] + m.m[r][1] * b.hi[1] + m.m[r][2] * b.hi[2] + m.m[r][3]);
                }
                for (int r = 0; r < 3; r++) { b.lo[r] = std::min(a[r], c[r]); b.hi[r] = std::max(a[r], c[r]); }
            }
            box[(size_t)n] = b; done[(size_t)n] = 1; st.pop_back();
            continue;
        }
        if (ty == PSGPU_PRIM_INSTANCE) {  // CInstance.h:45-55: the origin's octree as it stands, not mapped into the instance's space
            const int origin = (int)t->res[n * 4];
            Box b;
            if (origin >= 0 && origin < N && (recompute || done[(size_t)origin])) b = box[(size_t)origin];
            else for (int a = 0; a < 3; a++) b.lo[a] = b.hi[a] = 0.0f;
            if (recompute) {  // ... but recursive_RecomputeAllOctrees maps every primitive's box, instances included
                const M4& m = acc[(size_t)n];
                float a0[3], c0[3];
                for (int r = 0; r < 3; r++) {
                    a0[r] = (float)(m.m[r][0] * b.lo[0] + m.m[r][1] * b.lo[1] + m.m[r][2] * b.lo[2] + m.m[r][3]);
                    c0[r] = (float)(m.m[r][0] * b.hi[0] + m.m[r][1] * b.hi[1] + m.m[r][2] * b.hi[2] + m.m[r][3]);
                }
                for (int r = 0; r < 3; r++) { b.lo[r] = std::min(a0[r], c0[r]); b.hi[r] = std::max(a0[r], c0[r]); }
            }
            box[(size_t)n] = b; done[(size_t)n] = 1; st.pop_back();
            continue;
        }
        if (fr.next < t->n_children[n]) {
            const int c = t->child[t->first_child[n] + fr.next];
            fr.next++;
            // mtxCurrent = mtxBranch.multiply(node forward) in the reference's row-vector CMatrix: the branch acts first, then the node
            acc[(size_t)c] = mul(fwd_of(c), acc[(size_t)n]);
            st.push_back(Frame{c, 0});
            continue;
        }
        Box b;
        for (int a = 0; a < 3; a++) b.lo[a] = b.hi[a] = 0.0f;
        const int nk = t->n_children[n];
        if (nk > 0) {
            b = box[(size_t)t->child[t->first_child[n]]];
            for (int i = 1; i < nk; i++) {
                const Box& k = box[(size_t)t->child[t->first_child[n] + i]];
                if (ty == PSGPU_OP_UNION || ty == PSGPU_OP_BLEND || ty == PSGPU_OP_RICCI) {  // CUnion.h:82-89, CBlend.h:92-99, CRicciBlend.h:140-146
                    for (int a = 0; a < 3; a++) { b.lo[a] = std::min(b.lo[a], k.lo[a]); b.hi[a] = std::max(b.hi[a], k.hi[a]); }
                } else if (ty == PSGPU_OP_INTERSECT) {  // COctree::csgIntersection, PS_Octree.cpp:362-380: per axis, only where the ranges overlap
                    for (int a = 0; a < 3; a++)
                        if (b.lo[a] <= k.hi[a] && b.hi[a] >= k.lo[a]) { b.lo[a] = std::max(b.lo[a], k.lo[a]); b.hi[a] = std::min(b.hi[a], k.hi[a]); }
                }  // difference / smooth difference: first child only (CDifference.h:81-85, CSmoothDifference.h:114-118)
            }
            if (PSGPU_IS_WARP(ty)) for (int a = 0; a < 3; a++) { b.lo[a] -= kWarpExpand; b.hi[a] += kWarpExpand; }  // CWarp*.h computeOctree
        }
        box[(size_t)n] = b; done[(size_t)n] = 1; st.pop_back();
    }
    for (int n = 0; n < N; n++) for (int a = 0; a < 3; a++) { lo3n[n * 3 + a] = box[(size_t)n].lo[a]; hi3n[n * 3 + a] = box[(size_t)n].hi[a]; }
    return PSGPU_SUCCESS;
}

// ---------------------------------------------------------------------------------------------------------------------------------
// .scene files: an INI with one [BLOBNODE id] section per node
// ---------------------------------------------------------------------------------------------------------------------------------
namespace {

typedef std::map<std::string, std::string> Section;

std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && isspace((unsigned char)s[a])) a++;
    while (b > a && isspace((unsigned char)s[b - 1])) b--;
    return s.substr(a, b - a);
}
std::string upper(std::string s) { for (char& c : s) c = (char)toupper((unsigned char)c); return s; }

std::vector<double> numbers(const std::string& v) {
    std::vector<double> out;
    std::string s = v;
    for (char& c : s) if (c == '(' || c == ')' || c == ',') c = ' ';
    std::istringstream is(s);
    double d;
    while (is >> d) out.push_back(d);
    return out;
}

int op_type_of(const std::string& name) {  // getName() of the operator classes
    static const std::map<std::string, int> m = {
        {"UNION", PSGPU_OP_UNION}, {"INTERSECTION", PSGPU_OP_INTERSECT}, {"DIFFERENCE", PSGPU_OP_DIF}, {"SMOOTH DIFFERENCE", PSGPU_OP_SMOOTHDIF},
        {"BLEND", PSGPU_OP_BLEND}, {"RICCI BLEND", PSGPU_OP_RICCI}, {"TWIST", PSGPU_OP_WARP_TWIST}, {"TAPER", PSGPU_OP_WARP_TAPER},
        {"BEND", PSGPU_OP_WARP_BEND}, {"SHEAR", PSGPU_OP_WARP_SHEAR}};
    const auto it = m.find(upper(name));
    return it == m.end() ? -1 : it->second;
}
int prim_type_of(const std::string& name) {  // getName() of the skeletons / non-skeletal primitives
    static const std::map<std::string, int> m = {
        {"POINT", PSGPU_PRIM_POINT}, {"LINE", PSGPU_PRIM_LINE}, {"CYLINDER", PSGPU_PRIM_CYLINDER}, {"DISC", PSGPU_PRIM_DISC}, {"RING", PSGPU_PRIM_RING},
        {"CUBE", PSGPU_PRIM_CUBE}, {"NULL", PSGPU_PRIM_NULL}, {"QUADRICPOINT", PSGPU_PRIM_QUADRICPOINT}, {"INSTANCE", PSGPU_PRIM_INSTANCE}};
    const auto it = m.find(upper(name));
    return it == m.end() ? -1 : it->second;
}
const char* type_name(int t) {
    switch (t) {
    case PSGPU_PRIM_POINT: return "POINT"; case PSGPU_PRIM_LINE: return "LINE"; case PSGPU_PRIM_CYLINDER: return "CYLINDER";
    case PSGPU_PRIM_DISC: return "DISC"; case PSGPU_PRIM_RING: return "RING"; case PSGPU_PRIM_CUBE: return "CUBE"; case PSGPU_PRIM_NULL: return "NULL";
    case PSGPU_PRIM_QUADRICPOINT: return "QUADRICPOINT"; case PSGPU_PRIM_INSTANCE: return "INSTANCE";
    case PSGPU_OP_UNION: return "UNION"; case PSGPU_OP_INTERSECT: return "INTERSECTION"; case PSGPU_OP_DIF: return "DIFFERENCE";
    case PSGPU_OP_SMOOTHDIF: return "SMOOTH DIFFERENCE"; case PSGPU_OP_BLEND: return "BLEND"; case PSGPU_OP_RICCI: return "RICCI BLEND";
    case PSGPU_OP_WARP_TWIST: return "TWIST"; case PSGPU_OP_WARP_TAPER: return "TAPER"; case PSGPU_OP_WARP_BEND: return "BEND";
    case PSGPU_OP_WARP_SHEAR: return "SHEAR";
    }
    return "?";
}

struct Loader {
    std::map<int, Section> sec;
    psgpu_tree_buf* o;
    std::map<int, int> indexOfId;
    std::vector<std::pair<int, int>> instances;  // (flat index, origin node id)
    std::string err;

    static std::vector<double> vec(const Section& s, const char* key, size_t n, double dflt) {
        const auto it = s.find(key);
        std::vector<double> v = it == s.end() ? std::vector<double>() : numbers(it->second);
        v.resize(n, dflt);
        return v;
    }
    static double num(const Section& s, const char* key, double dflt) {
        const auto it = s.find(key);
        if (it == s.end()) return dflt;
        const std::vector<double> v = numbers(it->second);
        return v.empty() ? dflt : v[0];
    }
    static std::string str(const Section& s, const char* key) { const auto it = s.find(key); return it == s.end() ? std::string() : trim(it->second); }

    // recursive_ReadBlobNode, pre-order; returns the flat index or -1
    int read(int id, int depth) {
        const auto sit = sec.find(id);
        if (sit == sec.end()) { err = "missing section BLOBNODE " + std::to_string(id); return -1; }
        if (depth > 100000) { err = "tree too deep"; return -1; }
        const Section& s = sit->second;
        const bool isOp = num(s, "IsOperator", 0) != 0;
        // the shipped sample (Distrib/train_corrected.scene) still says SkeletonType where the reader asks for PrimitiveType
        const std::string tn = isOp ? str(s, "OperatorType") : (str(s, "PrimitiveType").empty() ? str(s, "SkeletonType") : str(s, "PrimitiveType"));
        const int ty = isOp ? op_type_of(tn) : prim_type_of(tn);
        if (ty < 0) { err = "BLOBNODE " + std::to_string(id) + ": node kind '" + tn + "' is not lowered to the GPU"; return -1; }
        const int me = (int)o->type.size();
        indexOfId[id] = me;
        o->type.push_back(ty); o->ids.push_back(id);
        o->kids.emplace_back();
        float pos[3] = {0, 0, 0}, dir[3] = {0, 0, 0}, res[4] = {0, 0, 0, 0}, col[3] = {0, 0, 0};
        const std::vector<double> sc = vec(s, "AffineScale", 3, 1.0), ro = vec(s, "AffineRotate", 4, 0.0), tr = vec(s, "AffineTranslate", 3, 0.0);
        Srt t;
        for (int a = 0; a < 3; a++) { t.s[a] = (float)sc[a]; t.t[a] = (float)tr[a]; }
        const bool noRot = s.find("AffineRotate") == s.end();
        for (int a = 0; a < 4; a++) t.q[a] = noRot ? (a == 3 ? 1.0f : 0.0f) : (float)ro[a];
        auto v3 = [&](const char* key, float* dst) { const std::vector<double> v = vec(s, key, 3, 0.0); for (int a = 0; a < 3; a++) dst[a] = (float)v[a]; };
        switch (ty) {  // loadScript of the skeletons / operators (keys as their saveScript writes them)
        case PSGPU_PRIM_POINT: v3("position", pos); break;
        case PSGPU_PRIM_LINE: v3("start", pos); v3("end", dir); break;
        case PSGPU_PRIM_CYLINDER: {
            v3("position", pos); v3("direction", dir);
            const float l = std::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);  // CSkeletonCylinder normalises its direction
            if (l > 0.f) for (int a = 0; a < 3; a++) dir[a] /= l;
            res[0] = (float)num(s, "radius", 0); res[1] = (float)num(s, "height", 0);
        } break;
        case PSGPU_PRIM_DISC: case PSGPU_PRIM_RING: {
            v3("position", pos); v3("direction", dir);
            res[0] = (float)num(s, "radius", 0); res[1] = res[0] * res[0];
        } break;
        case PSGPU_PRIM_CUBE: v3("position", pos); res[0] = (float)num(s, "side", 0); break;
        case PSGPU_PRIM_QUADRICPOINT: v3("position", pos); res[0] = (float)num(s, "radius", 1); res[1] = (float)num(s, "scale", 1); break;
        case PSGPU_PRIM_INSTANCE: instances.push_back(std::make_pair(me, (int)num(s, "OriginalNodeIndex", -1))); break;
        case PSGPU_OP_RICCI: res[0] = (float)num(s, "power", 1); res[1] = res[0] != 0.f ? 1.0f / res[0] : 0.f; break;
        case PSGPU_OP_WARP_TWIST: res[0] = (float)num(s, "factor", 0); res[1] = (float)num(s, "axis", 0); break;
        case PSGPU_OP_WARP_TAPER: res[0] = (float)num(s, "factor", 0); res[1] = (float)num(s, "base axis", 0); res[2] = (float)num(s, "taper axis", 0); break;
        case PSGPU_OP_WARP_BEND: res[0] = (float)num(s, "rate", 0); res[1] = (float)num(s, "center", 0); res[2] = (float)num(s, "left bound", 0); res[3] = (float)num(s, "right bound", 0); break;
        case PSGPU_OP_WARP_SHEAR: res[0] = (float)num(s, "factor", 0); res[1] = (float)num(s, "base axis", 0); res[2] = (float)num(s, "shear axis", 0); break;
        default: break;
        }
        if (!isOp) { const std::vector<double> d = vec(s, "MtrlDiffused", 4, 0.0); for (int a = 0; a < 3; a++) col[a] = (float)d[a]; }
        o->pos.insert(o->pos.end(), pos, pos + 3); o->dir.insert(o->dir.end(), dir, dir + 3); o->res.insert(o->res.end(), res, res + 4);
        o->color.insert(o->color.end(), col, col + 3);
        o->srt.insert(o->srt.end(), t.s, t.s + 3); o->srt.insert(o->srt.end(), t.q, t.q + 4); o->srt.insert(o->srt.end(), t.t, t.t + 3);
        const bool ident = srt_is_identity(t);
        o->hassrt.push_back(ident ? 0 : 1);
        float B[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
        if (!ident) srt_backward(t, B);
        o->xform.insert(o->xform.end(), B, B + 12); o->hasxf.push_back(ident ? 0 : 1);
        if (isOp) {
            std::vector<int> kids;
            if (num(s, "ChildrenIDsUseRange", 0) != 0) {
                const std::vector<double> r = vec(s, "ChildrenIDsRange", 2, 0.0);
                for (int i = (int)r[0]; i <= (int)r[1]; i++) kids.push_back(i);
            }
            if (kids.empty()) { for (double d : numbers(str(s, "ChildrenIDs"))) kids.push_back((int)d); kids.resize((size_t)std::min((double)kids.size(), num(s, "ChildrenCount", (double)kids.size()))); }
            for (int k : kids) {
                const int c = read(k, depth + 1);
                if (c < 0) return -1;
                o->kids[(size_t)me].push_back(c);
            }
        }
        return me;
    }
};

}  // namespace

extern "C" int psgpu_scene_load(const char* path, int root_id, psgpu_tree_buf** outp) {
    if (!path || !outp) return PSGPU_PARAM_ERROR;
    *outp = nullptr;
    std::ifstream f(path);
    if (!f) return PSGPU_PARAM_ERROR;
    Loader L;
    std::string line;
    int cur = -1;
    int firstId = -1;
    while (std::getline(f, line)) {
        line.erase(std::remove(line.begin(), line.end(), '\0'), line.end());  // writeString leaves the DAnsiStr terminator in the file
        line = trim(line);
        if (line.empty() || line[0] == ';' || line[0] == '#') continue;
        if (line[0] == '[') {
            const size_t e = line.find(']');
            const std::string name = upper(trim(line.substr(1, e == std::string::npos ? std::string::npos : e - 1)));
            cur = -1;
            if (name.compare(0, 8, "BLOBNODE") == 0) { cur = atoi(name.c_str() + 8); if (firstId < 0) firstId = cur; L.sec[cur]; }
            continue;
        }
        const size_t eq = line.find('=');
        if (cur >= 0 && eq != std::string::npos) L.sec[cur][trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
    }
    if (L.sec.empty()) return PSGPU_PARAM_ERROR;
    psgpu_tree_buf* o = new psgpu_tree_buf();
    L.o = o;
    const int root = L.read(root_id >= 0 ? root_id : firstId, 0);
    if (root < 0) { fprintf(stderr, "psgpu_scene_load: %s\n", L.err.c_str()); delete o; return PSGPU_ERROR_UNSUPPORTED_NODE; }
    for (const auto& in : L.instances) {
        const auto it = L.indexOfId.find(in.second);
        if (it == L.indexOfId.end()) { fprintf(stderr, "psgpu_scene_load: instance origin %d not found\n", in.second); delete o; return PSGPU_PARAM_ERROR; }
        o->res[(size_t)in.first * 4] = (float)it->second;
    }
    const int N = (int)o->type.size();
    o->first.resize((size_t)N); o->nk.resize((size_t)N);
    for (int n = 0; n < N; n++) {
        o->first[(size_t)n] = (int32_t)o->child.size();
        o->nk[(size_t)n] = (int32_t)o->kids[(size_t)n].size();
        o->child.insert(o->child.end(), o->kids[(size_t)n].begin(), o->kids[(size_t)n].end());
    }
    if (o->child.empty()) o->child.push_back(0);
    o->kids.clear();
    o->finish(root);
    *outp = o;
    return PSGPU_SUCCESS;
}

// Writes the tree the way CBlobNode::saveScript does: one section per node, %f numbers, children listed by node id (a range when
// more than two primitive children carry consecutive ids, CBlobTree.cpp:75-101). Ambient / specular are not part of the flat tree:
// written as the diffuse colour scaled like CMaterial's presets (0.5 / 1.2 clipped), shininess 32.
extern "C" int psgpu_scene_save(const psgpu_tree* t, const char* path) {
    if (!t || !path || t->n_nodes <= 0) return PSGPU_PARAM_ERROR;
    FILE* f = fopen(path, "w");
    if (!f) return PSGPU_PARAM_ERROR;
    auto id_of = [&](int n) { return t->node_id ? t->node_id[n] : n; };
    std::vector<int> st(1, t->root);
    while (!st.empty()) {
        const int n = st.back();
        st.pop_back();
        const int ty = t->type[n];
        const bool isOp = !PSGPU_IS_PRIM(ty);
        const Srt s = srt_of(t, n);
        fprintf(f, "[BLOBNODE %d] \nIsOperator=%d \n", id_of(n), isOp ? 1 : 0);
        fprintf(f, "AffineScale=(%f, %f, %f) \nAffineRotate=(%f, %f, %f, %f) \nAffineTranslate=(%f, %f, %f) \n", s.s[0], s.s[1], s.s[2], s.q[0], s.q[1],
                s.q[2], s.q[3], s.t[0], s.t[1], s.t[2]);
        const float* p = t->pos + n * 3;
        const float* d = t->dir + n * 3;
        const float* r = t->res + n * 4;
        if (isOp) {
            const int nk = t->n_children[n];
            fprintf(f, "OperatorType=%s \nChildrenCount=%d \n", type_name(ty), nk);
            bool range = nk > 2, allPrims = true;
            for (int i = 0; i < nk; i++) {
                const int c = t->child[t->first_child[n] + i];
                allPrims = allPrims && PSGPU_IS_PRIM(t->type[c]);
                if (i > 0 && id_of(c) != id_of(t->child[t->first_child[n] + i - 1]) + 1) range = false;
            }
            range = range && allPrims;
            fprintf(f, "ChildrenIDsUseRange=%d \n", range ? 1 : 0);
            if (range) fprintf(f, "ChildrenIDsRange=(%d, %d) \n", id_of(t->child[t->first_child[n]]), id_of(t->child[t->first_child[n] + nk - 1]));
            else {
                fprintf(f, "ChildrenIDs=(");
                for (int i = 0; i < nk; i++) fprintf(f, "%d%s", id_of(t->child[t->first_child[n] + i]), i + 1 < nk ? ", " : "");
                fprintf(f, ") \n");
            }
            switch (ty) {
            case PSGPU_OP_RICCI: fprintf(f, "power=%f \n", r[0]); break;
            case PSGPU_OP_WARP_TWIST: fprintf(f, "factor=%f \naxis=%d \n", r[0], (int)r[1]); break;
            case PSGPU_OP_WARP_TAPER: fprintf(f, "factor=%f \nbase axis=%d \ntaper axis=%d \n", r[0], (int)r[1], (int)r[2]); break;
            case PSGPU_OP_WARP_BEND: fprintf(f, "rate=%f \ncenter=%f \nleft bound=%f \nright bound=%f \n", r[0], r[1], r[2], r[3]); break;
            case PSGPU_OP_WARP_SHEAR: fprintf(f, "factor=%f \nbase axis=%d \nshear axis=%d \n", r[0], (int)r[1], (int)r[2]); break;
            default: break;
            }
            for (int i = nk; i-- > 0;) st.push_back(t->child[t->first_child[n] + i]);
        } else {
            const float* c = t->color + n * 3;
            auto cl = [](float v) { return v > 1.0f ? 1.0f : v; };
            fprintf(f, "MtrlAmbient=(%f, %f, %f, %f) \n", 0.5f * c[0], 0.5f * c[1], 0.5f * c[2], 0.5f);
            fprintf(f, "MtrlDiffused=(%f, %f, %f, %f) \n", c[0], c[1], c[2], 1.0f);
            fprintf(f, "MtrlSpecular=(%f, %f, %f, %f) \n", cl(1.2f * c[0] + 0.4f), cl(1.2f * c[1] + 0.4f), cl(1.2f * c[2] + 0.4f), 1.0f);
            fprintf(f, "MtrlShininess=%f \nPrimitiveType=%s \n", 32.0f, type_name(ty));
            switch (ty) {
            case PSGPU_PRIM_POINT: fprintf(f, "position=(%f, %f, %f) \n", p[0], p[1], p[2]); break;
            case PSGPU_PRIM_LINE: fprintf(f, "start=(%f, %f, %f) \nend=(%f, %f, %f) \n", p[0], p[1], p[2], d[0], d[1], d[2]); break;
            case PSGPU_PRIM_CYLINDER: fprintf(f, "position=(%f, %f, %f) \ndirection=(%f, %f, %f) \nradius=%f \nheight=%f \n", p[0], p[1], p[2], d[0], d[1], d[2], r[0], r[1]); break;
            case PSGPU_PRIM_DISC: case PSGPU_PRIM_RING: fprintf(f, "position=(%f, %f, %f) \ndirection=(%f, %f, %f) \nradius=%f \n", p[0], p[1], p[2], d[0], d[1], d[2], r[0]); break;
            case PSGPU_PRIM_CUBE: fprintf(f, "position=(%f, %f, %f) \nside=%f \n", p[0], p[1], p[2], r[0]); break;
            case PSGPU_PRIM_QUADRICPOINT: fprintf(f, "position=(%f, %f, %f) \nscale=%f \nradius=%f \n", p[0], p[1], p[2], r[1], r[0]); break;
            case PSGPU_PRIM_INSTANCE: {
                const int origin = (int)r[0];
                fprintf(f, "OriginalNodeIndex=%d \nOriginalNodeIsOp=%d \nOriginalNodeType=%s \n", id_of(origin), PSGPU_IS_PRIM(t->type[origin]) ? 0 : 1, type_name(t->type[origin]));
            } break;
            default: break;
            }
        }
    }
    fclose(f);
    return PSGPU_SUCCESS;
}
