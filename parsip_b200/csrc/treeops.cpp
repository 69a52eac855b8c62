// Pre-linearization passes on the flat BlobTree, host side, Qt/Loki-free.
//
// psgpu_tree_binarize restates ConvertToBinaryTree::run (Parsip100/ParsipHaptics/include/CLayerManager.cpp:81-233):
//  * an operator with more than two children becomes the right-leaning chain op(c0, op(c1, ... op(c_{n-2}, c_{n-1}))) of
//    clones of itself (:140-176) - unless split_always is 0 and all its children are primitives, in which case it is kept
//    n-ary (:136-139, :178-186);
//  * an operator with one child that is not unary (a warp) is padded with a NULL primitive when pad_with_null is set
//    (:196-213);
//  * node ids are handed out in CLONE order (cloneNode bumps the id dispenser, :42-52): the root's clone first, then for every
//    operator child i: [the next chain link, i > 0] the child's clone, then the child's whole subtree. The output numbers its
//    nodes in exactly that order, so node index == the id the reference's converted tree would carry.
// Nothing here touches the device.
#include <cstring>
#include <vector>

#include "../../include/psgpu.h"

#include "tree_buf.h"

namespace {

struct Binarizer {
    const psgpu_tree* in;
    psgpu_tree_buf* out;
    bool splitAlways, padNull;
    int err = PSGPU_SUCCESS;

    int clone(int n) {  // copy of node n without children; n < 0: a NULL primitive
        const int id = (int)out->type.size();
        static const float ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
        out->type.push_back(n < 0 ? (int32_t)PSGPU_PRIM_NULL : in->type[n]);
        for (int a = 0; a < 3; a++) {
            out->pos.push_back(n < 0 ? 0.f : in->pos[n * 3 + a]);
            out->dir.push_back(n < 0 ? 0.f : in->dir[n * 3 + a]);
            out->color.push_back(n < 0 ? 0.f : in->color[n * 3 + a]);
        }
        for (int a = 0; a < 4; a++) out->res.push_back(n < 0 ? 0.f : in->res[n * 4 + a]);
        const bool xf = n >= 0 && in->xform && in->has_xform && in->has_xform[n];
        const float* m = xf ? in->xform + (size_t)n * 12 : ident;
        out->xform.insert(out->xform.end(), m, m + 12);
        out->hasxf.push_back(xf ? 1 : 0);
        out->kids.emplace_back();
        return id;
    }
    bool allChildrenPrims(int n) const {
        for (int i = 0; i < in->n_children[n]; i++)
            if (!PSGPU_IS_PRIM(in->type[in->child[in->first_child[n] + i]])) return false;
        return true;
    }
    // run(node, clonned) of the reference; explicit stack of (node, clone, next child, chain link so far)
    void run(int root, int rootClone) {
        struct Frame { int node, cur, i; };
        std::vector<Frame> st;
        st.push_back({root, rootClone, 0});
        while (!st.empty()) {
            Frame& f = st.back();
            const int n = f.node;
            if (PSGPU_IS_PRIM(in->type[n])) { st.pop_back(); continue; }
            const int nk = in->n_children[n];
            const int32_t* kid = in->child + in->first_child[n];
            if (nk > 2 && (splitAlways || !allChildrenPrims(n))) {
                if (f.i == 0) out->converted++;
                if (f.i >= nk) { st.pop_back(); continue; }
                const int i = f.i++;
                if (i > 0 && i < nk - 1) {  // the next link of the chain: a fresh clone of the operator, child of the previous link
                    const int link = clone(n);
                    out->kids[f.cur].push_back(link);
                    f.cur = link;
                }
                const int c = clone(kid[i]);
                out->kids[f.cur].push_back(c);
                const Frame next{kid[i], c, 0};
                st.push_back(next);  // invalidates f
            } else if (nk > 2) {  // kept n-ary: children cloned, not visited (they are primitives)
                for (int i = 0; i < nk; i++) { const int c = clone(kid[i]); out->kids[f.cur].push_back(c); }  // clone() grows kids
                st.pop_back();
            } else {
                if (f.i >= 2) { st.pop_back(); continue; }
                const int i = f.i++;
                if (i < nk) {
                    const int c = clone(kid[i]);
                    out->kids[f.cur].push_back(c);
                    const Frame next{kid[i], c, 0};
                    st.push_back(next);
                } else if (!PSGPU_IS_WARP(in->type[n]) && padNull) {
                    const int c = clone(-1);
                    out->kids[f.cur].push_back(c);
                    out->converted++;
                }
            }
        }
    }
};

bool valid(const psgpu_tree* t) {
    if (!t || t->n_nodes <= 0 || t->root < 0 || t->root >= t->n_nodes || !t->type || !t->first_child || !t->n_children || !t->child || !t->pos ||
        !t->dir || !t->res || !t->color)
        return false;
    for (int n = 0; n < t->n_nodes; n++) {
        if (t->n_children[n] < 0 || t->first_child[n] < 0) return false;
        if (PSGPU_IS_PRIM(t->type[n]) && t->n_children[n] != 0) return false;
    }
    return true;
}

}  // namespace

extern "C" int psgpu_tree_binarize(const psgpu_tree* in, int split_always, int pad_with_null, psgpu_tree_buf** outp, int32_t* n_converted) {
    if (!outp || !valid(in)) return PSGPU_PARAM_ERROR;
    for (int n = 0; n < in->n_nodes; n++)
        for (int i = 0; i < in->n_children[n]; i++) {
            const int c = in->child[in->first_child[n] + i];
            if (c < 0 || c >= in->n_nodes) return PSGPU_PARAM_ERROR;
        }
    psgpu_tree_buf* out = new psgpu_tree_buf();
    Binarizer B{in, out, split_always != 0, pad_with_null != 0};
    const int rc = B.clone(in->root);
    B.run(in->root, rc);
    const int N = (int)out->type.size();
    out->first.resize(N); out->nk.resize(N);
    for (int n = 0; n < N; n++) {
        out->first[n] = (int32_t)out->child.size();
        out->nk[n] = (int32_t)out->kids[n].size();
        out->child.insert(out->child.end(), out->kids[n].begin(), out->kids[n].end());
    }
    if (out->child.empty()) out->child.push_back(0);
    out->kids.clear(); out->kids.shrink_to_fit();
    out->finish(0);
    if (n_converted) *n_converted = out->converted;
    *outp = out;
    return PSGPU_SUCCESS;
}

extern "C" const psgpu_tree* psgpu_tree_view(const psgpu_tree_buf* b) { return b ? &b->view : nullptr; }

extern "C" int psgpu_tree_free(psgpu_tree_buf* b) {
    delete b;
    return PSGPU_SUCCESS;
}
