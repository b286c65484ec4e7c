/* host_bvh.cpp -- see host_bvh.h */
#include "host_bvh.h"
#include <algorithm>
#include <cmath>
#include <limits>
#include <cstdlib>

namespace {
struct Box {
    float lo[3], hi[3];
    void clear() { for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<float>::infinity(); hi[k] = -lo[k]; } }
    void add(const float *p) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    void add(const Box &b) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
};
struct Builder {
    const float *v; std::vector<Box> tb; std::vector<float> cen; std::vector<uint32_t> &order; std::vector<RTBvhNode> &nodes;
    float margin; int depth; uint32_t leaf;
    Builder(const float *v_, std::vector<uint32_t> &o, std::vector<RTBvhNode> &n) : v(v_), order(o), nodes(n), margin(0), depth(0), leaf(RT_BVH_LEAF) {}
    Box bounds(uint32_t b, uint32_t e) const { Box r; r.clear(); for (uint32_t k = b; k < e; k++) r.add(tb[order[k]]); return r; }
    /* builds [b, e): returns its reference and its box */
    uint32_t build(uint32_t b, uint32_t e, Box &box, int d) {
        depth = std::max(depth, d);
        box = bounds(b, e);
        if (e - b <= leaf) return RT_BVH_LEAF_FLAG | ((e - b - 1u) << 27) | b;
        Box cb; cb.clear();
        for (uint32_t k = b; k < e; k++) cb.add(&cen[3 * (size_t)order[k]]);
        int axis = 0;
        for (int k = 1; k < 3; k++) if (cb.hi[k] - cb.lo[k] > cb.hi[axis] - cb.lo[axis]) axis = k;
        const uint32_t mid = b + (e - b) / 2;
        std::nth_element(order.begin() + b, order.begin() + mid, order.begin() + e, [&](uint32_t x, uint32_t y) {
            const float cx = cen[3 * (size_t)x + axis], cy = cen[3 * (size_t)y + axis];
            return cx < cy || (cx == cy && x < y);
        });
        const uint32_t me = (uint32_t)nodes.size();
        nodes.push_back(RTBvhNode());
        Box b0, b1;
        const uint32_t r0 = build(b, mid, b0, d + 1), r1 = build(mid, e, b1, d + 1);
        RTBvhNode &n = nodes[me];
        for (int k = 0; k < 3; k++) {
            n.lo0[k] = b0.lo[k] - margin; n.hi0[k] = b0.hi[k] + margin;
            n.lo1[k] = b1.lo[k] - margin; n.hi1[k] = b1.hi[k] + margin;
        }
        n.ref0 = r0; n.ref1 = r1; n.pad0 = n.pad1 = 0;
        return me;
    }
};
}

int rtBuildBvh(const float *vertices, uint32_t T, std::vector<RTBvhNode> &nodes, std::vector<uint32_t> &order, float *marginOut) {
    nodes.clear(); order.resize(T);
    Builder B(vertices, order, nodes);
    if (const char *e = getenv("RT_BVH_LEAF")) { const int v = atoi(e); if (v >= 1 && v <= 16) B.leaf = (uint32_t)v; }   /* tuning: tools/time_bvh.py */
    B.tb.resize(T); B.cen.resize(3 * (size_t)T);
    Box scene; scene.clear();
    for (uint32_t i = 0; i < T; i++) {
        order[i] = i;
        Box &b = B.tb[i]; b.clear();
        for (int a = 0; a < 3; a++) b.add(vertices + ((size_t)i * 3 + a) * 4);
        for (int k = 0; k < 3; k++) B.cen[3 * (size_t)i + k] = 0.5f * b.lo[k] + 0.5f * b.hi[k];
        scene.add(b);
    }
    float ext = 0.0f;
    for (int k = 0; k < 3; k++) if (scene.hi[k] - scene.lo[k] > ext) ext = scene.hi[k] - scene.lo[k];
    B.margin = ext * (1.0f / 1024.0f);
    if (marginOut) *marginOut = B.margin;
    nodes.reserve(T / 2 + 2);
    Box rb;
    if (T <= B.leaf) {
        /* one leaf: a root whose second child is an empty box */
        RTBvhNode n;
        Box b = T ? B.bounds(0, T) : scene;
        for (int k = 0; k < 3; k++) {
            n.lo0[k] = b.lo[k] - B.margin; n.hi0[k] = b.hi[k] + B.margin;
            n.lo1[k] = std::numeric_limits<float>::infinity(); n.hi1[k] = -std::numeric_limits<float>::infinity();
        }
        n.ref0 = RT_BVH_LEAF_FLAG | ((T ? T - 1u : 0u) << 27); n.ref1 = RT_BVH_LEAF_FLAG; n.pad0 = n.pad1 = 0;
        if (!T) { for (int k = 0; k < 3; k++) { n.lo0[k] = n.lo1[k]; n.hi0[k] = n.hi1[k]; } }
        nodes.push_back(n);
        return 1;
    }
    B.build(0, T, rb, 1);       /* T > LEAF: the root is inner node 0 */
    return B.depth;
}
