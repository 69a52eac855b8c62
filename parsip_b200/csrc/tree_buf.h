// Owner of the arrays behind a psgpu_tree that the library hands out (psgpu_tree_binarize, psgpu_tree_flatten_transforms,
// psgpu_scene_load); opaque `psgpu_tree_buf` of include/psgpu.h.
#ifndef PSGPU_TREE_BUF_H
#define PSGPU_TREE_BUF_H
#include <cstdint>
#include <vector>

#include "../../include/psgpu_types.h"

struct psgpu_tree_buf {
    std::vector<int32_t> type, first, nk, child, ids;
    std::vector<float> pos, dir, res, color, xform, srt;
    std::vector<uint8_t> hasxf, hassrt;
    std::vector<std::vector<int32_t>> kids;  // while building
    psgpu_tree view;
    int converted = 0;

    void finish(int root) {
        view = psgpu_tree();
        view.n_nodes = (int32_t)type.size(); view.root = root;
        view.type = type.data(); view.first_child = first.data(); view.n_children = nk.data(); view.child = child.data();
        view.pos = pos.data(); view.dir = dir.data(); view.res = res.data(); view.color = color.data();
        view.xform = xform.data(); view.has_xform = hasxf.data();
        view.srt = srt.size() == type.size() * 10 ? srt.data() : nullptr;
        view.has_srt = hassrt.size() == type.size() ? hassrt.data() : nullptr;
        view.node_id = ids.size() == type.size() ? ids.data() : nullptr;
    }
};
#endif
