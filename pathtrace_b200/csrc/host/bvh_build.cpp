// bvh_build.cpp -- host BVH builder: binned-SAH BVH2 over instance boxes, one instance per leaf, flattened
// depth-first into 64-byte PtNode records (two 32-byte child halves), then collapsed into 128-byte 4-wide nodes.
//
// The reference builds a random-axis median-split pointer tree (bvh.h:133-175); its topology has no effect on
// which hit is closest (every primitive is reachable through boxes that contain it), so this builder is free to
// pick a better tree.  What it must keep is the instance boxes themselves (instance::instance, primitive.h:266-297),
// which instance_bounds() reproduces.
#include "bvh_build.h"

#include <algorithm>
#include <cmath>
#include <limits>

namespace b200pt {

namespace {

struct Box {
    float lo[3], hi[3];
    void reset()
    {
        for (int k = 0; k < 3; k++) {
            lo[k] = std::numeric_limits<float>::max();
            hi[k] = -std::numeric_limits<float>::max();
        }
    }
    void grow(const Box &b)
    {
        for (int k = 0; k < 3; k++) {
            lo[k] = std::min(lo[k], b.lo[k]);
            hi[k] = std::max(hi[k], b.hi[k]);
        }
    }
    void grow_point(const float p[3])
    {
        for (int k = 0; k < 3; k++) {
            lo[k] = std::min(lo[k], p[k]);
            hi[k] = std::max(hi[k], p[k]);
        }
    }
    double area() const
    {
        double dx = (double)hi[0] - lo[0], dy = (double)hi[1] - lo[1], dz = (double)hi[2] - lo[2];
        if (dx < 0 || dy < 0 || dz < 0)
            return 0.0;
        return 2.0 * (dx * dy + dy * dz + dz * dx);
    }
};

struct BuildCtx {
    const std::vector<Aabb> *boxes;
    std::vector<Box> prim;
    std::vector<float> centroid; // 3 per prim
    std::vector<int> order;
    std::vector<pt::PtNode> nodes;
    int max_depth = 0;
};

constexpr int kBins = 32;

// builds the subtree over order[begin,end); returns its ref (node index or ~instance) and box
int build_range(BuildCtx &c, int begin, int end, Box &out_box, int depth)
{
    c.max_depth = std::max(c.max_depth, depth);
    const int n = end - begin;
    if (n == 1) {
        out_box = c.prim[c.order[begin]];
        return ~c.order[begin];
    }
    Box bounds, cbounds;
    bounds.reset();
    cbounds.reset();
    for (int i = begin; i < end; i++) {
        bounds.grow(c.prim[c.order[i]]);
        cbounds.grow_point(&c.centroid[3 * c.order[i]]);
    }
    int mid = -1;
    // SAH may produce long chains (nested shells); past this depth fall back to balanced median splits so that
    // total depth stays below the traversal stack (PT_STACK_DEPTH = 48): 24 + log2(n)
    if (n > 2 && depth < 24) {
        // as many bins as there are boxes, at most kBins: half of all nodes hold four boxes or fewer
        const int bins = std::min(kBins, std::max(4, n));
        double best_cost = std::numeric_limits<double>::infinity();
        int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; axis++) {
            float lo = cbounds.lo[axis], hi = cbounds.hi[axis];
            if (!(hi > lo))
                continue;
            Box bin_box[kBins];
            int bin_count[kBins];
            for (int b = 0; b < bins; b++) {
                bin_box[b].reset();
                bin_count[b] = 0;
            }
            const double scale = bins / ((double)hi - lo);
            for (int i = begin; i < end; i++) {
                int p = c.order[i];
                int b = (int)(((double)c.centroid[3 * p + axis] - lo) * scale);
                b = std::max(0, std::min(bins - 1, b));
                bin_box[b].grow(c.prim[p]);
                bin_count[b]++;
            }
            double right_area[kBins];
            int right_count[kBins];
            Box acc;
            acc.reset();
            int cnt = 0;
            for (int b = bins - 1; b > 0; b--) {
                acc.grow(bin_box[b]);
                cnt += bin_count[b];
                right_area[b] = acc.area();
                right_count[b] = cnt;
            }
            acc.reset();
            cnt = 0;
            for (int b = 0; b < bins - 1; b++) {
                acc.grow(bin_box[b]);
                cnt += bin_count[b];
                if (cnt == 0 || right_count[b + 1] == 0)
                    continue;
                double cost = acc.area() * cnt + right_area[b + 1] * right_count[b + 1];
                if (cost < best_cost) {
                    best_cost = cost;
                    best_axis = axis;
                    best_bin = b;
                }
            }
        }
        if (best_axis >= 0) {
            float lo = cbounds.lo[best_axis], hi = cbounds.hi[best_axis];
            const double scale = bins / ((double)hi - lo);
            auto it = std::partition(c.order.begin() + begin, c.order.begin() + end, [&](int p) {
                int b = (int)(((double)c.centroid[3 * p + best_axis] - lo) * scale);
                b = std::max(0, std::min(bins - 1, b));
                return b <= best_bin;
            });
            mid = (int)(it - c.order.begin());
        }
    }
    if (mid <= begin || mid >= end) {
        // median split on the widest centroid axis (also the n == 2 case)
        int axis = 0;
        float w = -1.f;
        for (int k = 0; k < 3; k++) {
            float e = cbounds.hi[k] - cbounds.lo[k];
            if (e > w) {
                w = e;
                axis = k;
            }
        }
        mid = begin + n / 2;
        std::nth_element(c.order.begin() + begin, c.order.begin() + mid, c.order.begin() + end,
                         [&](int a, int b) { return c.centroid[3 * a + axis] < c.centroid[3 * b + axis]; });
    }
    const int me = (int)c.nodes.size();
    c.nodes.push_back(pt::PtNode{});
    Box b0, b1;
    int r0 = build_range(c, begin, mid, b0, depth + 1);
    int r1 = build_range(c, mid, end, b1, depth + 1);
    pt::PtNode &node = c.nodes[me];
    for (int k = 0; k < 3; k++) {
        node.lo0[k] = b0.lo[k];
        node.hi0[k] = b0.hi[k];
        node.lo1[k] = b1.lo[k];
        node.hi1[k] = b1.hi[k];
    }
    node.ref0 = r0;
    node.ref1 = r1;
    node.pad0 = node.pad1 = 0;
    out_box = bounds;
    return me;
}

// Collapse of the binary tree: a node adopts its grandchildren in place of the inner child with the largest box,
// until it has four children or only leaves.  Every instance stays reachable through boxes that contain it, which
// is all the closest-hit result depends on.
struct Child {
    int ref;
    float lo[3], hi[3];
    double area() const
    {
        double dx = (double)hi[0] - lo[0], dy = (double)hi[1] - lo[1], dz = (double)hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
};

void children_of(const pt::PtNode &n, Child out[2])
{
    out[0].ref = n.ref0;
    out[1].ref = n.ref1;
    for (int k = 0; k < 3; k++) {
        out[0].lo[k] = n.lo0[k];
        out[0].hi[k] = n.hi0[k];
        out[1].lo[k] = n.lo1[k];
        out[1].hi[k] = n.hi1[k];
    }
}

// `stack`: the most entries a traversal below this node can have pending (its other children, then the same again
// below the child it descends into)
int collapse_node(const std::vector<pt::PtNode> &bin, int node, std::vector<pt::PtNode4> &out, int depth, int &max_depth,
                  int &stack)
{
    max_depth = std::max(max_depth, depth + 1);
    Child ch[4];
    int n = 2;
    children_of(bin[node], ch);
    while (n < 4) {
        int pick = -1;
        for (int i = 0; i < n; i++)
            if (ch[i].ref >= 0 && (pick < 0 || ch[i].area() > ch[pick].area()))
                pick = i;
        if (pick < 0)
            break;
        Child two[2];
        children_of(bin[ch[pick].ref], two);
        ch[pick] = two[0];
        ch[n++] = two[1];
    }
    const int me = (int)out.size();
    out.push_back(pt::PtNode4{});
    int refs[4];
    int below = 0;
    for (int i = 0; i < 4; i++) {
        refs[i] = i >= n ? PT_NO_CHILD : ch[i].ref;
        if (i < n && ch[i].ref >= 0) {
            int st = 0;
            refs[i] = collapse_node(bin, ch[i].ref, out, depth + 1, max_depth, st);
            below = std::max(below, st);
        }
    }
    stack = n - 1 + below;
    pt::PtNode4 &q = out[me];
    for (int i = 0; i < 4; i++) {
        const bool used = i < n;
        q.lox[i] = used ? ch[i].lo[0] : 0.f;
        q.loy[i] = used ? ch[i].lo[1] : 0.f;
        q.loz[i] = used ? ch[i].lo[2] : 0.f;
        q.hix[i] = used ? ch[i].hi[0] : 0.f;
        q.hiy[i] = used ? ch[i].hi[1] : 0.f;
        q.hiz[i] = used ? ch[i].hi[2] : 0.f;
        q.ref[i] = refs[i];
        q.pad[i] = 0;
    }
    return me;
}

} // namespace

BuiltBvh build_bvh(const std::vector<Aabb> &boxes)
{
    BuildCtx c;
    c.boxes = &boxes;
    const int n = (int)boxes.size();
    c.prim.resize(n);
    c.centroid.resize(3 * (size_t)n);
    c.order.resize(n);
    for (int i = 0; i < n; i++) {
        for (int k = 0; k < 3; k++) {
            c.prim[i].lo[k] = boxes[i].lo[k];
            c.prim[i].hi[k] = boxes[i].hi[k];
            c.centroid[3 * i + k] = 0.5f * boxes[i].lo[k] + 0.5f * boxes[i].hi[k];
        }
        c.order[i] = i;
    }
    BuiltBvh out;
    Box root;
    root.reset();
    if (n == 1) {
        // a root must be an inner node: both halves point at the only instance (intersecting it twice is
        // idempotent -- a medium draws the same Philox variate both times)
        pt::PtNode node{};
        for (int k = 0; k < 3; k++) {
            node.lo0[k] = node.lo1[k] = boxes[0].lo[k];
            node.hi0[k] = node.hi1[k] = boxes[0].hi[k];
        }
        node.ref0 = node.ref1 = ~0;
        c.nodes.push_back(node);
        root = c.prim[0];
        c.max_depth = 1;
    } else {
        c.nodes.reserve((size_t)n);
        build_range(c, 0, n, root, 0);
    }
    out.nodes = std::move(c.nodes);
    out.n_leaves = n;
    out.max_depth = c.max_depth;
    out.nodes4.reserve(out.nodes.size() / 2 + 1);
    collapse_node(out.nodes, 0, out.nodes4, 0, out.max_depth4, out.max_stack4);
    for (int k = 0; k < 3; k++) {
        out.bounds[k] = root.lo[k];
        out.bounds[3 + k] = root.hi[k];
    }
    return out;
}

} // namespace b200pt
