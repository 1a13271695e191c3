/* stand-in: linked as oracle/_ref/lc_view/kd_tree.h so that the reference's vrl_lightcut.h (reached through the same
 * directory of links) finds THIS instead of its own kd_tree.h -- dead code (only the never-called buildTreeStable names
 * it) that g++ rejects.  Four declarations, written here, NOT reference code. */
#pragma once
#include <vector>
namespace mitsuba {
template <typename Float, typename Spectrum, class Node> class KdNode;
template <class Node> struct KdTree {
    explicit KdTree(float) {}
    void constructTree(std::vector<Node *> &, size_t) {}
    Node *queryNN(Node *) { return nullptr; }
    void invalidateNodesAndUpdate(Node *, Node *, Node *) {}
};
}
