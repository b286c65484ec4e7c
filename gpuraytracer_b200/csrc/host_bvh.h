/*
 * host_bvh.h -- the acceleration structure for scenes too large for the brute-force kernels
 * (reference: RTrace/computeShader.swift:45-97 setupAccelerationStructures, which hands the
 * triangles to Metal's builder).  Built on the HOST when the structure is first needed (init-time work, like the
 * reference's build): a binary BVH by median splits of the triangle centroids along the longest axis, at most
 * RT_BVH_LEAF triangles per leaf.  Every node holds the boxes of its two children, grown by `margin`, so that a
 * ray the exact triangle test could accept always passes the boxes on the way down (same argument as the object
 * groups of sorted.cuh).  The device side is csrc/bvh.cuh.
 */
#ifndef RT_HOST_BVH_H
#define RT_HOST_BVH_H
#include <stdint.h>
#include <vector>

#define RT_BVH_LEAF 12                     /* measured 2 / 4 / 8 / 12 / 16 (profiles/r03g_, r03h_bvh_forms.txt): each step up to 12 is faster
                                              (8 over 4: 8-12 %, 12 over 8: 0-20 %), 16 gives the same trees as 12 on the test scenes */
#define RT_BVH_LEAF_FLAG 0x80000000u       /* ref = FLAG | (count - 1) << 27 | first slot; else index of an inner node */

struct RTBvhNode {                         /* 64 bytes = four float4 loads */
    float lo0[3]; uint32_t ref0;           /* child 0: box, reference */
    float hi0[3]; uint32_t ref1;           /* child 1's reference rides here */
    float lo1[3]; uint32_t pad0;
    float hi1[3]; uint32_t pad1;
};

/* vertices: 3 per triangle, 4 floats each (simd_float3).  order[slot] = triangle id of leaf slot `slot`.
 * Returns the depth of the tree (the traversal stack must hold that many entries). */
int rtBuildBvh(const float *vertices, uint32_t numTriangles, std::vector<RTBvhNode> &nodes, std::vector<uint32_t> &order,
               float *marginOut);

#endif
