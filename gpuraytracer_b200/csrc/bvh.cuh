/*
 * bvh.cuh -- `pathTrace` for scenes too large for the brute-force kernels (included by pathtrace.cu).
 * SURVEY.md 8(f) N4; reference: the Metal acceleration structure of RTrace/computeShader.swift:45-97 and the
 * intersector queries of RTrace/raytrace.metal:26-30,48-49,79-85.
 *
 * The structure is host_bvh.h's binary BVH (children's boxes in the parent, grown by a margin); the triangles
 * are rt_prep_scene's own intersection records gathered into leaf order (rt_bvh_gather), so a leaf runs the
 * SAME triangle test on the SAME operands as the brute-force loops.  What differs is the order in which the
 * triangles are met, so the closest hit is chosen by the order-independent rule "smaller t, and for an exactly
 * equal cross-multiplied comparison the lower primitive id" -- which is what the ascending strict-less loop
 * computes (pairwise: both orders compare the same two fp32 products).  A box is skipped only when the ray
 * misses it (slab test; a NaN keeps) or enters it beyond the best hit so far.
 *
 * One thread per pixel, per-lane stack in local memory; geometry and shading rows are read from global memory
 * (L2-resident), only the header and the lights are staged in shared memory.
 */
#pragma once

#define RT_BVH_STACK 48
#define RT_BVH_LEAF_FLAG 0x80000000u

struct BvhRef {
    const float4 *nodes;     /* 4 float4 per node: lo0|ref0, hi0|ref1, lo1|-, hi1|- */
    const float4 *tris;      /* 4 float4 per leaf slot: A, B, C (isect record), id */
};

/* slab test against a grown box; returns the entry distance (in units of |d|) through tEnter.
 * Written so that any NaN (0 * inf when the origin lies in a slab plane of a zero direction component) keeps. */
__device__ __forceinline__ bool bvh_box(f3 lo, f3 hi, f3 o, f3 inv, float &tEnter) {
    const float x0 = (lo.x - o.x) * inv.x, x1 = (hi.x - o.x) * inv.x;
    const float y0 = (lo.y - o.y) * inv.y, y1 = (hi.y - o.y) * inv.y;
    const float z0 = (lo.z - o.z) * inv.z, z1 = (hi.z - o.z) * inv.z;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
    tEnter = tn;
    return !(tn > tf * 1.00001f + 1e-30f);
}

/* one leaf triangle against the best hit so far, order-independent form of tri_accept's last comparison */
__device__ __forceinline__ void bvh_leaf_closest(const BvhRef &bv, uint32_t slot, f3 o, f3 d, float tmin,
                                                 float &bestN, float &bestD, uint32_t &best) {
    const float4 A = bv.tris[4 * slot], B = bv.tris[4 * slot + 1], C = bv.tris[4 * slot + 2];
    const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
    const f3 w = mk(A.x, A.y, A.z) - o;
    const float D = dot(d, n), Nn = dot(w, n);
    const f3 c = cross(d, w);
    const float U = dot(e2, c), V = dot(e1n, c);
    const uint32_t db = __float_as_uint(D);
    const int t2 = (int)(((__float_as_uint(U) ^ db) | (__float_as_uint(V) ^ db)) | (__float_as_uint(Nn) ^ db));
    const float s = U + V, a = fabsf(D), N = fabsf(Nn);
    const float p1 = N * bestD, p2 = bestN * a;
    if ((t2 >= 0) & (fabsf(s) <= a) & (N >= tmin * a) & (p1 <= p2)) {       /* (the id is fetched for hits and ties only) */
        const uint32_t id = __float_as_uint(bv.tris[4 * slot + 3].x);
        if ((p1 < p2) | ((id < best) & (best != 0xffffffffu))) { bestN = N; bestD = a; best = id; }
    }
}

/* 1 / d for the slab test; a zero component becomes +-1e30 so that (plane - origin) * inv is never 0 * inf */
__device__ __forceinline__ f3 bvh_inv(f3 d) {
    return mk(fabsf(d.x) > 1e-30f ? 1.0f / d.x : copysignf(1e30f, d.x), fabsf(d.y) > 1e-30f ? 1.0f / d.y : copysignf(1e30f, d.y),
              fabsf(d.z) > 1e-30f ? 1.0f / d.z : copysignf(1e30f, d.z));
}

/* The traversal stack of one lane (local memory). */
struct BvhStack {
    uint32_t s[RT_BVH_STACK];
    int sp;
    __device__ __forceinline__ void reset() { sp = 0; }
    __device__ __forceinline__ void push(uint32_t v) { if (sp < RT_BVH_STACK) s[sp++] = v; }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
    __device__ __forceinline__ uint32_t pop() { return s[--sp]; }
};

/* One step down the tree: the node's two child boxes against the ray; `limit` = the parameter beyond which a box
 * cannot matter, as the pair (limN, limD), t = limN / limD.  Returns the next reference to visit (pushing the
 * farther child when both are hit) or pops; RT_BVH_DONE when nothing is left. */
#define RT_BVH_DONE 0xffffffffu
__device__ __forceinline__ uint32_t bvh_step(const BvhRef &bv, uint32_t ref, f3 o, f3 inv, float limN, float limD, BvhStack &st) {
    const float4 n0 = bv.nodes[4 * ref], n1 = bv.nodes[4 * ref + 1], n2 = bv.nodes[4 * ref + 2], n3 = bv.nodes[4 * ref + 3];
    float t0, t1;
    bool h0 = bvh_box(mk(n0.x, n0.y, n0.z), mk(n1.x, n1.y, n1.z), o, inv, t0);
    bool h1 = bvh_box(mk(n2.x, n2.y, n2.z), mk(n3.x, n3.y, n3.z), o, inv, t1);
    h0 = h0 && !(t0 * limD > limN * 1.0001f);            /* beyond the limit?  a NaN keeps */
    h1 = h1 && !(t1 * limD > limN * 1.0001f);
    const uint32_t r0 = __float_as_uint(n0.w), r1 = __float_as_uint(n1.w);
    if (h0 && h1) {
        const bool swap = t1 < t0;                       /* nearer child first */
        st.push(swap ? r0 : r1);
        return swap ? r1 : r0;
    }
    if (h0 || h1) return h0 ? r0 : r1;
    return st.empty() ? RT_BVH_DONE : st.pop();
}

/* closest hit (raytrace.metal:48-49).  "while-while": every lane walks down to its next leaf, then the warp does
 * its leaves together (one loop mixing both ran 8 of 32 lanes per instruction).  `budget` bounds the walk
 * (2 x nodes + 8 steps): a corrupt structure cannot hang. */
__device__ __forceinline__ bool bvh_closest(const BvhRef &bv, BvhStack &st, f3 o, f3 d, float tmin, float tmax, uint32_t budget,
                                            float &dist, uint32_t &id) {
    const f3 inv = bvh_inv(d);
    float bestN = tmax, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    st.reset();
    uint32_t ref = 0u;                                   /* the root is inner node 0 */
    uint32_t it = 0;
    while (ref != RT_BVH_DONE && it < budget) {
        while (!(ref & RT_BVH_LEAF_FLAG) && it < budget) { ref = bvh_step(bv, ref, o, inv, bestN, bestD, st); it++; }
        if (ref == RT_BVH_DONE || !(ref & RT_BVH_LEAF_FLAG)) break;
        const uint32_t first = ref & 0x07ffffffu, count = ((ref >> 27) & 15u) + 1u;
        for (uint32_t k = 0; k < count; k++) bvh_leaf_closest(bv, first + k, o, d, tmin, bestN, bestD, best);
        ref = st.empty() ? RT_BVH_DONE : st.pop();
        it++;
    }
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/* any hit (accept_any_intersection, raytrace.metal:79-85): the order of the triangles does not matter */
__device__ __forceinline__ bool bvh_any(const BvhRef &bv, BvhStack &st, f3 o, f3 d, float tmin, float tmax, uint32_t budget) {
    const f3 inv = bvh_inv(d);
    st.reset();
    uint32_t ref = 0u, it = 0;
    bool hit = false;
    while (ref != RT_BVH_DONE && it < budget) {
        while (!(ref & RT_BVH_LEAF_FLAG) && it < budget) { ref = bvh_step(bv, ref, o, inv, tmax, 1.0f, st); it++; }
        if (ref == RT_BVH_DONE || !(ref & RT_BVH_LEAF_FLAG)) break;
        const uint32_t first = ref & 0x07ffffffu, count = ((ref >> 27) & 15u) + 1u;
        for (uint32_t k = 0; k < count; k++) {
            const uint32_t slot = first + k;
            float N, a;
            hit |= tri_test(bv.tris[4 * slot], bv.tris[4 * slot + 1], bv.tris[4 * slot + 2], o, d, tmin, tmax, 1.0f, N, a);
        }
        ref = (hit || st.empty()) ? RT_BVH_DONE : st.pop();
        it++;
    }
    return hit;
}

/* leaf slot k <- the intersection record rt_prep_scene made for triangle order[k] (bit for bit) + its id */
extern "C" __global__ void rt_bvh_gather(const float *__restrict__ scene, RTSceneLayout lay, const uint32_t *__restrict__ order,
                                         float4 *__restrict__ tris) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= lay.T) return;
    const uint32_t id = order[k];
    const float4 *rec = (const float4 *)(scene + lay.isect) + 4 * (size_t)id;
    tris[4 * (size_t)k] = rec[0]; tris[4 * (size_t)k + 1] = rec[1]; tris[4 * (size_t)k + 2] = rec[2];
    tris[4 * (size_t)k + 3] = make_float4(__uint_as_float(id), 0.0f, 0.0f, 0.0f);
}

struct RTBvhArgs { const float4 *nodes; const float4 *tris; uint32_t numNodes, pad; };

/* raytrace.metal:11-111 statement by statement (as pathtrace_body), every ray query through the BVH.
 * Measured and not kept (profiles/r03f_bvh_forms.txt; all forms gave the same bits): the sample loop and the bounce
 * loop merged into one, so that a lane whose path ended starts its pixel's next sample at once (10-18 % SLOWER:
 * camera rays, which walk the tree together, get mixed with scattered bounce rays), and the traversal stacks in
 * shared memory instead of local memory (+-2 %).  What pays is residency: more warps hide more of the node-fetch
 * latency -- six blocks per SM (80 registers) against four +15 to +17 % at 36 to 36 864 triangles, eight (64
 * registers, 64 bytes of spills) another +3 % (profiles/r03g_bvh_forms.txt). */
template <bool MIS = false>
__device__ __forceinline__ void pathtrace_bvh_body(const float *__restrict__ scene, const uint32_t *__restrict__ offsets,
                                                   float4 *__restrict__ out, const RTLaunchParams &p, const RTSceneLayout &lay,
                                                   const RTBvhArgs &ba, float *sm, const RTHaltonTables &ht) {
    for (uint32_t k = threadIdx.x; k < RT_HDR_WORDS; k += RT_BLOCK_THREADS) sm[k] = scene[k];
    for (uint32_t k = threadIdx.x; k < RT_LIGHT_WORDS * lay.L; k += RT_BLOCK_THREADS) sm[RT_HDR_WORDS + k] = scene[lay.light + k];
    __syncthreads();
    const float *lights = sm + RT_HDR_WORDS;
    const float *shade = scene + lay.shade;                /* global: SoA rows of Tp words */
    BvhStack st;

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t x = blockIdx.x * RT_BLOCK_W + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = blockIdx.y * RT_BLOCK_H + (warp >> 1) * 4 + (lane >> 3);
    if (x >= (uint32_t)p.width || ly >= (uint32_t)p.rows) return;
    const uint32_t y = p.tileRows ? (uint32_t)p.row0 + (ly / p.tileRows) * p.tileStride + (ly % p.tileRows)
                                  : (uint32_t)p.row0 + ly;
    const size_t pix = (size_t)ly * (uint32_t)p.width + x;

    BvhRef bv = { ba.nodes, ba.tris };
    const uint32_t budget = 2u * ba.numNodes + 8u;
    const f3 cam_o = mk(sm[RT_HDR_CAM_O], sm[RT_HDR_CAM_O + 1], sm[RT_HDR_CAM_O + 2]);
    const f3 cam_u = mk(sm[RT_HDR_CAM_U], sm[RT_HDR_CAM_U + 1], sm[RT_HDR_CAM_U + 2]);
    const f3 cam_v = mk(sm[RT_HDR_CAM_V], sm[RT_HDR_CAM_V + 1], sm[RT_HDR_CAM_V + 2]);
    const f3 cam_w = mk(sm[RT_HDR_CAM_W], sm[RT_HDR_CAM_W + 1], sm[RT_HDR_CAM_W + 2]);
    const float halfW = sm[RT_HDR_HALF_W], halfH = sm[RT_HDR_HALF_H];
    const float resX = sm[RT_HDR_RES_X], resY = sm[RT_HDR_RES_Y];
    const float fx = (float)(int)x, fy = (float)(int)y;
    const uint32_t offset = offsets[pix];                               /* raytrace.metal:37, hoisted */
    const uint32_t nb = (uint32_t)p.bounces, numLights = p.numLights;
    f3 lum = mk(0.0f, 0.0f, 0.0f);
    if (p.accumulate) { float4 o = out[pix]; lum = mk(o.x, o.y, o.z); }

#pragma unroll 1
    for (uint32_t k = 0; k < p.spp; k++) {
        const uint32_t i = offset + (p.firstSample + k);
        float jx = halton<0>(i), jy = halton_cam_y(ht, i);              /* sampling.metal:141-147 */
        float s = ((fx + jx) / resX) * 2.0f - 1.0f;
        float t = -(((fy + jy) / resY) * 2.0f - 1.0f);
        f3 rd = normalize((cam_u * (s * halfW) + cam_v * (t * halfH)) - cam_w);
        f3 ro = cam_o;
        f3 acc = mk(0.0f, 0.0f, 0.0f), color = mk(1.0f, 1.0f, 1.0f);
        float dist; uint32_t id;
        bool hit = bvh_closest(bv, st, ro, rd, 0.001f, 1000.0f, budget, dist, id);
#pragma unroll 1
        for (uint32_t b = 0; hit; b++) {                                /* raytrace.metal:47-101 */
            const float *sh = shade + id;
            if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) {                      /* :57-60; MIS: only a camera ray takes it (mis.cuh) */
                if (!MIS || b == 0) acc = mk(sh[(RT_SH_ER + 0) * lay.Tp], sh[(RT_SH_ER + 1) * lay.Tp], sh[(RT_SH_ER + 2) * lay.Tp]);
                break;
            }
            const f3 normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
            const f3 diffuse = mk(sh[(RT_SH_DR + 0) * lay.Tp], sh[(RT_SH_DR + 1) * lay.Tp], sh[(RT_SH_DR + 2) * lay.Tp]);
            f3 P = (ro + rd * dist) + normal * 1e-3f;                   /* :67 */
            const float4 h4 = halton_bounce4(ht, i, b);                 /* the bounce's four Halton values (RTHaltonTables) */
            const float wx = h4.x, wy = h4.y;                           /* :72-73 */
            color = color * diffuse;                                    /* :76 */
            for (uint32_t l = 0; l < numLights; l++) {
                const float *L = lights + RT_LIGHT_WORDS * l;
                f3 ldir; float ldist;
                f3 lightColor = sample_area_light(L, wx, wy, P, ldir, ldist);
                lightColor = lightColor * saturate(dot(normal, ldir));  /* :75 */
                if (MIS) lightColor = lightColor * misA_light_weight(L, P, normal, ldir);
                if (!bvh_any(bv, st, P, ldir, 0.0f, ldist - 1e-3f, budget)) acc = acc + lightColor * color;   /* :79-88 */
            }
            if (!MIS && b + 1 >= nb) break;
            const f3 right = mk(sh[(RT_SH_RX + 0) * lay.Tp], sh[(RT_SH_RX + 1) * lay.Tp], sh[(RT_SH_RX + 2) * lay.Tp]);
            const f3 fwd = mk(sh[(RT_SH_FX + 0) * lay.Tp], sh[(RT_SH_FX + 1) * lay.Tp], sh[(RT_SH_FX + 2) * lay.Tp]);
            const float ux = h4.z, uy = h4.w;                           /* :93-94 */
            f3 sdir = sample_cosine_hemisphere(ux, uy);
            rd = (right * sdir.x + normal * sdir.y) + fwd * sdir.z;     /* sampling.metal:65 */
            ro = P;                                                     /* :99-100 */
            hit = bvh_closest(bv, st, ro, rd, 0.001f, 1000.0f, budget, dist, id);
            if (MIS) {
                for (uint32_t l = 0; l < numLights; l++)
                    acc = acc + misA_cosine_term(lights + RT_LIGHT_WORDS * l, P, normal, rd, hit, dist, color);
                if (b + 1 >= nb) break;
            }
        }
        lum = lum + acc;                                                /* raytrace.metal:103 */
    }
    float4 o;
    if (p.accumulate) o = make_float4(lum.x, lum.y, lum.z, (float)(p.firstSample + p.spp));
    else { float samples = (float)p.spp; o = make_float4(lum.x / samples, lum.y / samples, lum.z / samples, 1.0f); }
    out[pix] = o;
}

#define RT_BVH_ENTRY(name, MINB, MIS)                                                                                 \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, const RTBvhArgs ba, const RTHaltonTables ht) {         \
        extern __shared__ __align__(16) float sm[];          /* header, then the lights */                       \
        pathtrace_bvh_body<MIS>(scene, offsets, out, p, lay, ba, sm, ht);                                        \
    }
RT_BVH_ENTRY(pathTrace_bvh, 8, false)
RT_BVH_ENTRY(pathTrace_bvh_mis, 8, true)
#ifdef RT_TUNING_VARIANTS   /* RT_BVH_FORM=<suffix> (rtcuda.cpp), tools/time_bvh.py: blocks per SM */
RT_BVH_ENTRY(pathTrace_bvh_l4, 4, false)
RT_BVH_ENTRY(pathTrace_bvh_l6, 6, false)
RT_BVH_ENTRY(pathTrace_bvh_l10, 10, false)
#endif
