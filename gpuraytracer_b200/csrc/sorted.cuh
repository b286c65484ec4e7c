/*
 * sorted.cuh -- `pathTrace` with warp-level path compaction AND rays sorted by the object groups they can reach
 * (included by pathtrace.cu after wave.cuh).
 *
 * Same arithmetic, same per-pixel results (bit for bit) as pathtrace_body and pathtrace_wave_body: every
 * accepted hit is found by the same triangle test on the same operands, visited in ascending primitive id.
 * What changes is WHICH triangles a ray is tested against.  rt_prep_scene cuts the scene into object groups
 * (connected sets of quads: the room, each box, the light) and gives each cullable group a bounding box grown
 * by 1/128 of the scene's extent.  A ray whose line stays clear of that box cannot be accepted by any triangle
 * inside it -- the exact test's own rounding lets it accept rays a few 1e-6 (relative) outside a triangle, the
 * margin is four orders of magnitude more; the only rays for which the error could reach the margin run within
 * 1e-4 rad of a triangle's plane AND start within 1e-5 of it, and every ray here starts 1e-3 off its surface
 * (oracle-side check over every query of a frame: tests/test_group_cull_cpu.py).
 *
 * Skipping per lane would gain nothing: a warp runs the loop as long as any lane needs it.  So the paths of a
 * round are SORTED by the set of groups their rays of the bounce can reach (the key, <= 4 bits):
 *
 *   1. camera rays as in wave.cuh; survivors are listed (hitlist) and keep a slot = sample*32 + lane for the
 *      whole round: the path state lives in that slot, only one-byte slot numbers move;
 *   2. per bounce
 *      A. full warps walk the hit list: shade set-up, light sample, continuation direction; both rays of the
 *         bounce start at P, so one box test per group and ray gives the key; everything the trace needs is
 *         left in the slot;
 *      S. counting sort of the slot numbers by key (match_any + per-key counters), descending, stable;
 *      B. full warps walk the sorted list 32 entries at a time: the quads to test are the always-tested ones
 *         plus those of every group in the OR of the chunk's keys (warp-uniform), two quads per pass of the
 *         paired dual loop (pathtrace.cu: PAIRED, dual_hit); survivors are listed for the next bounce;
 *   3. radiance is kept in res[slot] from the start (emissive hit: assignment; unshadowed light: +=), and
 *      each lane adds its R results IN SAMPLE ORDER (raytrace.metal:103).
 *
 * Cornell frame: 67 % of the bounce steps need only the room and the light (6 of 18 quads), 31 % one box more
 * (12), 1 % everything; the last bounce's shadow rays: 78 % / 22 %.
 *
 * The kernel is picked by the host for launches whose scene CAN be grouped (T even and <= 64, no spheres, one
 * light); whether it IS (pairs share v0 bit for bit, at least one cullable group) is known on the device only,
 * and a block that finds the flag clear runs pathtrace_wave_body<R = 4> in the same shared memory instead.
 */
#pragma once

#ifndef RT_EXP_TRACE
#define RT_EXP_TRACE 1         /* A/B builds (tools/build_exp.sh): 0 separate closest / shadow loops, 1 one masked loop, 2 the r03 dual loop, no shell rule, 3 = 2 with the shell rule on the last bounce */
#endif
/* RT_SORT_FIELDS = 11 (X.xyz: hit point, then P; D.xyz: continuation ray; colour.xyz; light sample point x, z) and
 * RT_SORT_KEYS come from rt_kernel_smem.h */

template <int R>
struct SortSmem {
    static constexpr uint32_t CAP = 32u * R;                      /* slots per warp = paths per round; one byte names a slot */
    static_assert(CAP <= 256u, "slot numbers are bytes");
    /* fields + res.xyz + four byte lists (hit list, sorted list; surface id and key per slot) + the lanes' offsets + key counters */
    static constexpr uint32_t WORDS_PER_WARP = rtSortedWordsPerWarp(R);
};

/* line of the ray against a grown box, w = box centre - origin, H = half extent: the three cross-product axes
 * and "the whole box is behind the origin".  Conservative (a subset of the separating axes); a NaN keeps. */
__device__ __forceinline__ bool ray_clear_of_box(f3 w, f3 H, f3 d) {
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    bool clear = fabsf(fmaf(w.y, d.z, -(w.z * d.y))) > fmaf(H.y, az, H.z * ay);
    clear |= fabsf(fmaf(w.z, d.x, -(w.x * d.z))) > fmaf(H.z, ax, H.x * az);
    clear |= fabsf(fmaf(w.x, d.y, -(w.y * d.x))) > fmaf(H.x, ay, H.y * ax);
    clear |= dot(w, d) + fmaf(H.z, az, fmaf(H.y, ay, H.x * ax)) < 0.0f;
    return clear;
}

/* dual_hit<PAIRED> over a LIST of quads: bit k of `pairs` = triangles 2k, 2k+1 (warp-uniform word), two quads per
 * pass, ascending, so the strict-less tie rule picks the same primitive as the full loop; an odd last quad is
 * taken twice (a second test of a triangle never passes the strict-less comparison against its own distance). */
__device__ __forceinline__ bool dual_hit_pairs(const SceneRef<false> &sc, uint32_t pairs, f3 o,
                                               f3 dc, float tminC, float tmaxC, float &dist, uint32_t &id,
                                               f3 ds, float tminS, float tmaxS, bool &shadowed) {
    float bestN = tmaxC, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    bool sh = false;
#pragma unroll 1
    while (pairs) {
        const uint32_t pa = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u;
        uint32_t pb = pa;
        if (pairs) { pb = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u; }
#pragma unroll
        for (uint32_t h = 0; h < 2; h++) {
            if (h == 1 && pb == pa) break;                                         /* an odd last quad: alone (warp-uniform) */
            const uint32_t t0 = 2u * (h ? pb : pa);
            f3 w = mk(0.0f, 0.0f, 0.0f), cC = w, cS = w;
#pragma unroll
            for (uint32_t u = 0; u < 2; u++) {
                const uint32_t i = t0 + u;
                const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
                const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
                if (u == 0) { w = mk(A.x, A.y, A.z) - o; cC = cross(dc, w); cS = cross(ds, w); }
                const float Nn = dot(w, n);
                float N, a;
                {
                    float D = dot(dc, n);
                    float U = dot(e2, cC), V = dot(e1n, cC);
                    if (tri_accept(D, Nn, U, V, tminC, bestN, bestD, N, a)) { bestN = N; bestD = a; best = i; }
                }
                {
                    float D = dot(ds, n);
                    float U = dot(e2, cS), V = dot(e1n, cS);
                    sh |= tri_accept(D, Nn, U, V, tminS, tmaxS, 1.0f, N, a);
                }
            }
        }
    }
    shadowed = sh;
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/* dual_hit_pairs with the shadow ray tested against a SUBSET of the quads (pairsS, a subset of pairs: the shell rule,
 * rt_scene_layout.h): one ascending pass, so the closest hit's tie rule is untouched; whether a quad's shadow half
 * runs is warp-uniform. */
__device__ __forceinline__ bool dual_hit_pairs_masked(const SceneRef<false> &sc, uint32_t pairs, uint32_t pairsS, f3 o,
                                                      f3 dc, float tminC, float tmaxC, float &dist, uint32_t &id,
                                                      f3 ds, float tminS, float tmaxS, bool &shadowed) {
    float bestN = tmaxC, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    bool sh = false;
#pragma unroll 1
    while (pairs) {
        const uint32_t pa = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u;
        uint32_t pb = pa;
        if (pairs) { pb = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u; }
#pragma unroll
        for (uint32_t h = 0; h < 2; h++) {
            if (h == 1 && pb == pa) break;                                         /* an odd last quad: alone (warp-uniform) */
            const uint32_t q = h ? pb : pa, t0 = 2u * q;
            const bool doS = (pairsS >> q) & 1u;                                   /* warp-uniform */
            f3 w = mk(0.0f, 0.0f, 0.0f), cC = w, cS = w;
#pragma unroll
            for (uint32_t u = 0; u < 2; u++) {
                const uint32_t i = t0 + u;
                const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
                const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
                if (u == 0) { w = mk(A.x, A.y, A.z) - o; cC = cross(dc, w); if (doS) cS = cross(ds, w); }
                const float Nn = dot(w, n);
                float N, a;
                {
                    float D = dot(dc, n);
                    float U = dot(e2, cC), V = dot(e1n, cC);
                    if (tri_accept(D, Nn, U, V, tminC, bestN, bestD, N, a)) { bestN = N; bestD = a; best = i; }
                }
                if (doS) {
                    float D = dot(ds, n);
                    float U = dot(e2, cS), V = dot(e1n, cS);
                    sh |= tri_accept(D, Nn, U, V, tminS, tmaxS, 1.0f, N, a);
                }
            }
        }
    }
    shadowed = sh;
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/* the closest-hit half of dual_hit_pairs alone (the shadow ray of a bounce is traced against its own, shorter quad
 * list since the shell rule: rt_scene_layout.h) */
__device__ __forceinline__ bool closest_hit_pairs(const SceneRef<false> &sc, uint32_t pairs, f3 o, f3 d, float tmin, float tmax,
                                                  float &dist, uint32_t &id) {
    float bestN = tmax, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
#pragma unroll 1
    while (pairs) {
        const uint32_t pa = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u;
        uint32_t pb = pa;
        if (pairs) { pb = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u; }
#pragma unroll
        for (uint32_t h = 0; h < 2; h++) {
            if (h == 1 && pb == pa) break;                                         /* an odd last quad: alone (warp-uniform) */
            const uint32_t t0 = 2u * (h ? pb : pa);
            f3 w = mk(0.0f, 0.0f, 0.0f), c = w;
#pragma unroll
            for (uint32_t u = 0; u < 2; u++) {
                const uint32_t i = t0 + u;
                const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
                const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
                if (u == 0) { w = mk(A.x, A.y, A.z) - o; c = cross(d, w); }
                float N, a;
                if (tri_accept(dot(d, n), dot(w, n), dot(e2, c), dot(e1n, c), tmin, bestN, bestD, N, a)) { bestN = N; bestD = a; best = i; }
            }
        }
    }
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/* is the segment from P to the light sample point inside the shrunk shell box?  (rt_scene_layout.h: then no shell quad
 * can be accepted by the shadow ray; SB = centre xyz - | half extent xyz -; a NaN says no) */
__device__ __forceinline__ bool shadow_inside_shell(const float *SB, f3 P, f3 sp) {
    return (fabsf(P.x - SB[0]) < SB[4]) & (fabsf(P.y - SB[1]) < SB[5]) & (fabsf(P.z - SB[2]) < SB[6]) &
           (fabsf(sp.x - SB[0]) < SB[4]) & (fabsf(sp.y - SB[1]) < SB[5]) & (fabsf(sp.z - SB[2]) < SB[6]);
}
/* The exit rule (rt_scene_layout.h): through which face of the shell box can the closest-hit ray from P along d leave it?
 * 0..5 = face 2 * axis + (1 = high side): only that face's shell quads can accept the ray; 6 = none can (it leaves through a
 * face without quads, or runs in a face's plane); 7 = test them all (P is not inside the shrunk box, or the crossing points
 * of two faces lie within the margin of the box: an edge).  The crossing point comes from an approximate reciprocal -- it
 * is compared with a face grown by 1/128 of the scene, four orders above the exact test's own reach; a NaN keeps the face.
 * CPU side: oracle/rt_groups.c exit_rule, over every query of a frame. */
__device__ __forceinline__ uint32_t exit_face_code(const float *SB, f3 P, f3 d) {
    const bool inside = (fabsf(P.x - SB[0]) < SB[4]) & (fabsf(P.y - SB[1]) < SB[5]) & (fabsf(P.z - SB[2]) < SB[6]);
    const float s = SB[7], m = s + SB[3];
    const float hx = SB[4] + s, hy = SB[5] + s, hz = SB[6] + s;
    uint32_t n = 0u, code = 6u;
    {
        const float t = __fdividef((d.x > 0.0f ? SB[0] + hx : SB[0] - hx) - P.x, d.x);
        const float py = fmaf(t, d.y, P.y), pz = fmaf(t, d.z, P.z);
        if ((d.x != 0.0f) & !(fabsf(py - SB[1]) > SB[5] + m) & !(fabsf(pz - SB[2]) > SB[6] + m)) { n++; code = d.x > 0.0f ? 1u : 0u; }
    }
    {
        const float t = __fdividef((d.y > 0.0f ? SB[1] + hy : SB[1] - hy) - P.y, d.y);
        const float px = fmaf(t, d.x, P.x), pz = fmaf(t, d.z, P.z);
        if ((d.y != 0.0f) & !(fabsf(px - SB[0]) > SB[4] + m) & !(fabsf(pz - SB[2]) > SB[6] + m)) { n++; code = d.y > 0.0f ? 3u : 2u; }
    }
    {
        const float t = __fdividef((d.z > 0.0f ? SB[2] + hz : SB[2] - hz) - P.z, d.z);
        const float px = fmaf(t, d.x, P.x), py = fmaf(t, d.y, P.y);
        if ((d.z != 0.0f) & !(fabsf(px - SB[0]) > SB[4] + m) & !(fabsf(py - SB[1]) > SB[5] + m)) { n++; code = d.z > 0.0f ? 5u : 4u; }
    }
    return (!inside | (n > 1u)) ? 7u : code;
}
#define RT_KEY_SHELL 0x80u     /* bit of keyarr[]: the shadow ray must be tested against the shell quads too */

__device__ __forceinline__ bool any_hit_pairs(const SceneRef<false> &sc, uint32_t pairs, f3 o, f3 d, float tmin, float tmax) {
    bool hit = false;
#pragma unroll 1
    while (pairs) {
        const uint32_t pa = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u;
        uint32_t pb = pa;
        if (pairs) { pb = (uint32_t)__ffs((int)pairs) - 1u; pairs &= pairs - 1u; }
#pragma unroll
        for (uint32_t h = 0; h < 2; h++) {
            if (h == 1 && pb == pa) break;                                         /* an odd last quad: alone (warp-uniform) */
            const uint32_t t0 = 2u * (h ? pb : pa);
            f3 w = mk(0.0f, 0.0f, 0.0f), c = w;
#pragma unroll
            for (uint32_t u = 0; u < 2; u++) {
                const uint32_t i = t0 + u;
                const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
                const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
                if (u == 0) { w = mk(A.x, A.y, A.z) - o; c = cross(d, w); }
                float N, a;
                hit |= tri_accept(dot(d, n), dot(w, n), dot(e2, c), dot(e1n, c), tmin, tmax, 1.0f, N, a);
            }
        }
    }
    return hit;
}

template <int R, bool PERSIST>
__device__ __forceinline__ void pathtrace_sorted_body(const float *__restrict__ scene, const uint32_t *__restrict__ offsets,
                                                      float4 *__restrict__ out, const RTLaunchParams p,
                                                      const RTSceneLayout lay, float *sm, const RTTileSched ts,
                                                      const RTHaltonTables ht) {
    constexpr uint32_t CAP = SortSmem<R>::CAP;
    {
        const float4 *src = (const float4 *)scene;
        float4 *dst = (float4 *)sm;
        const uint32_t n4 = lay.words >> 2;
        for (uint32_t k = threadIdx.x; k < n4; k += RT_BLOCK_THREADS) dst[k] = src[k];
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;

    /* this warp's slots, lists and counters, after the scene */
    float *Q = sm + lay.words + warp * SortSmem<R>::WORDS_PER_WARP;
    float *res = Q + RT_SORT_FIELDS * CAP;
    uint8_t *hitlist = (uint8_t *)(res + 3 * CAP), *idarr = hitlist + CAP, *sorted = idarr + CAP, *keyarr = sorted + CAP;
    uint32_t *offs = (uint32_t *)(keyarr + CAP);
    uint32_t *cnts = offs + 32;
#define SF(field, s) Q[(field) * CAP + (s)]

    SceneRef<false> sc;
    sc.base = (const float4 *)(sm + lay.isect);
    sc.T = (lay.T + 3u) & ~3u; sc.S = 0u; sc.Treal = lay.T;
    sc.sph = 4 * lay.Tp;

    const f3 cam_o = mk(sm[RT_HDR_CAM_O], sm[RT_HDR_CAM_O + 1], sm[RT_HDR_CAM_O + 2]);
    const f3 cam_u = mk(sm[RT_HDR_CAM_U], sm[RT_HDR_CAM_U + 1], sm[RT_HDR_CAM_U + 2]);
    const f3 cam_v = mk(sm[RT_HDR_CAM_V], sm[RT_HDR_CAM_V + 1], sm[RT_HDR_CAM_V + 2]);
    const f3 cam_w = mk(sm[RT_HDR_CAM_W], sm[RT_HDR_CAM_W + 1], sm[RT_HDR_CAM_W + 2]);
    const float halfW = sm[RT_HDR_HALF_W], halfH = sm[RT_HDR_HALF_H];
    const float resX = sm[RT_HDR_RES_X], resY = sm[RT_HDR_RES_Y];
    const uint32_t nb = (uint32_t)p.bounces;
    const uint32_t *grp = (const uint32_t *)(sm + lay.group);
    const uint32_t K = grp[RT_GRP_K], shellQuads = grp[RT_GRP_SHELL];
    const float *shellBox = sm + lay.group + RT_GRP_SHELL_BOX;
    const uint32_t mwT = (lay.T + 31u) >> 5;
    uint32_t *tail = (uint32_t *)(sm + lay.words + (RT_BLOCK_THREADS / 32) * SortSmem<R>::WORDS_PER_WARP);
    /* the quads each key names: always-tested ones + those of the groups whose bit is set */
    uint32_t *keyPairs = tail + (RT_BLOCK_THREADS / 32) * mwT + 1;
    if (threadIdx.x < RT_SORT_KEYS) {
        uint32_t pm = grp[RT_GRP_ALWAYS];
        for (uint32_t g = 0; g < K; g++) if ((threadIdx.x >> g) & 1u) pm |= grp[4 + 8 * g + 3];
        keyPairs[threadIdx.x] = pm;
    }
    __syncthreads();

    for (;;) {
    uint32_t x0, ly0, tile = 0, chunk = 0, kBegin = 0, kEnd = p.spp;
    if (PERSIST) {
        /* the block's unit comes from a ticket (wave.cuh: RTTileSched) */
        uint32_t *s_unit = tail + (RT_BLOCK_THREADS / 32) * mwT;
        __syncthreads();
        if (threadIdx.x == 0) *s_unit = atomicAdd(ts.counter, 1u);
        __syncthreads();
        const uint32_t unit = *s_unit;
        const uint32_t blocksX = ((uint32_t)p.width + RT_BLOCK_W - 1u) / RT_BLOCK_W, blocksY = ((uint32_t)p.rows + RT_BLOCK_H - 1u) / RT_BLOCK_H;
        const uint32_t nBlocks = blocksX * blocksY;
        if (unit >= nBlocks * ts.chunks) break;
        chunk = unit / nBlocks;
        const uint32_t bt = unit - chunk * nBlocks;
        tile = bt * (RT_BLOCK_THREADS / 32) + warp;
        kBegin = chunk * ts.chunkSpp; kEnd = min(p.spp, kBegin + ts.chunkSpp);
        x0 = (bt % blocksX) * RT_BLOCK_W + (warp & 1) * 8; ly0 = (bt / blocksX) * RT_BLOCK_H + (warp >> 1) * 4;
    } else {
        x0 = blockIdx.x * RT_BLOCK_W + (warp & 1) * 8; ly0 = blockIdx.y * RT_BLOCK_H + (warp >> 1) * 4;
    }
    const uint32_t x = x0 + (lane & 7), ly = ly0 + (lane >> 3);
    const bool owns = x < (uint32_t)p.width && ly < (uint32_t)p.rows;
    const uint32_t y = p.tileRows ? (uint32_t)p.row0 + (ly / p.tileRows) * p.tileStride + (ly % p.tileRows)
                                  : (uint32_t)p.row0 + ly;
    const size_t pix = (size_t)ly * (uint32_t)p.width + x;
    const float fx = (float)(int)x, fy = (float)(int)y;

    /* ---- 0. camera-ray candidates of this warp's pixels (pathtrace.cu: camera_candidates) ---- */
    uint32_t *cullT = tail + warp * mwT;
    {
        CameraFrame cf = { cam_o, cam_u, cam_v, cam_w, halfW, halfH, resX, resY };
        const int big = 0x3fffffff;
        const int cx0 = __reduce_min_sync(0xffffffffu, owns ? (int)x : big), cx1 = __reduce_max_sync(0xffffffffu, owns ? (int)x : -big);
        const int cy0 = __reduce_min_sync(0xffffffffu, owns ? (int)y : big), cy1 = __reduce_max_sync(0xffffffffu, owns ? (int)y : -big);
        camera_candidates(sc, cf, lay.T, 0u, (float)cx0 - 1.0f, (float)cx1 + 2.0f, (float)cy0 - 1.0f, (float)cy1 + 2.0f,
                          cullT, mwT, cullT, 0u, lane);
    }

    const uint32_t offset = owns ? offsets[pix] : 0u;
    offs[lane] = offset;
    if (lane < RT_SORT_KEYS) cnts[lane] = 0u;
    f3 lum = mk(0.0f, 0.0f, 0.0f);
    if (PERSIST && chunk > 0) {
        const float4 o = tiles_wait_for_previous_chunk(ts.flags + tile, chunk, out + pix, owns, ts.status);
        lum = mk(o.x, o.y, o.z);
    } else if (p.accumulate && owns) { float4 o = out[pix]; lum = mk(o.x, o.y, o.z); }
    __syncwarp();

#pragma unroll 1
    for (uint32_t k0 = kBegin; k0 < kEnd; k0 += R) {
        const uint32_t nround = min((uint32_t)R, kEnd - k0);                  /* warp-uniform */
        const uint32_t ibase = p.firstSample + k0;
        uint32_t count = 0;                                                   /* hit-list length, warp-uniform */

        /* ---- 1. camera rays (as wave.cuh) ---------------------------------------------------- */
#pragma unroll 1
        for (uint32_t s = 0; s < nround; s += 2) {
            const bool two = s + 1 < nround;                                   /* warp-uniform */
            bool alive[2] = { false, false };
            uint32_t id[2] = { 0u, 0u };
            f3 hp[2], r0[2];
            hp[0] = hp[1] = r0[0] = r0[1] = mk(0.0f, 0.0f, 0.0f);
            if (owns) {
                f3 rd[2];
#pragma unroll
                for (int q = 0; q < 2; q++) {                                  /* sampling.metal:141-147 */
                    const uint32_t idx = offset + (ibase + s + q);
                    float jx = halton<0>(idx), jy = halton_cam_y(ht, idx);
                    float sx = ((fx + jx) / resX) * 2.0f - 1.0f;
                    float ty = -(((fy + jy) / resY) * 2.0f - 1.0f);
                    rd[q] = normalize((cam_u * (sx * halfW) + cam_v * (ty * halfH)) - cam_w);
                }
                bool hit[2]; float dist[2];
                closest_primary2_list(sc, cullT, mwT, cullT, 0u, rd[0], rd[1], 0.001f, 1000.0f, hit[0], dist[0], id[0], hit[1], dist[1], id[1]);
                hit[1] = hit[1] && two;
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    if (hit[q]) {
                        const float *sh = sm + lay.shade + id[q];
                        if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) r0[q] = mk(sh[(RT_SH_ER + 0) * lay.Tp], sh[(RT_SH_ER + 1) * lay.Tp], sh[(RT_SH_ER + 2) * lay.Tp]);
                        else { alive[q] = true; hp[q] = cam_o + rd[q] * dist[q]; }
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < 2; q++) {
                if (q == 1 && !two) break;
                const uint32_t mask = __ballot_sync(0xffffffffu, alive[q]);
                const uint32_t slot = (s + q) * 32u + lane;
                if (alive[q]) {
                    const uint32_t e = count + __popc(mask & lt);
                    SF(0, slot) = hp[q].x; SF(1, slot) = hp[q].y; SF(2, slot) = hp[q].z;
                    SF(6, slot) = 1.0f; SF(7, slot) = 1.0f; SF(8, slot) = 1.0f;
                    hitlist[e] = (uint8_t)slot; idarr[slot] = (uint8_t)id[q];
                }
                res[0 * CAP + slot] = r0[q].x; res[1 * CAP + slot] = r0[q].y; res[2 * CAP + slot] = r0[q].z;
                count += __popc(mask);
            }
        }
        __syncwarp();

        /* ---- 2. bounces ------------------------------------------------------------------------ */
#pragma unroll 1
        for (uint32_t b = 0; b < nb && count; b++) {
            const bool cont = b + 1 < nb;                                      /* warp-uniform */
            const bool sort = count > 32u;          /* one chunk: nothing to sort, the hit list is the order */

            /* -- A. shade set-up, the two rays of the bounce, the key -- */
#pragma unroll 1
            for (uint32_t c0 = 0; c0 < count; c0 += 32) {
                const uint32_t e = c0 + lane;
                const bool active = e < count;
                uint32_t key = 0xffu;
                if (active) {
                    const uint32_t slot = hitlist[e], id = idarr[slot];
                    const uint32_t i = offs[slot & 31u] + (ibase + (slot >> 5));
                    const float4 h4 = halton_bounce4(ht, i, b);                                 /* :72-73, :93-94 */
                    const f3 hp = mk(SF(0, slot), SF(1, slot), SF(2, slot));
                    f3 color = mk(SF(6, slot), SF(7, slot), SF(8, slot));
                    const float *sh = sm + lay.shade + id;
                    const f3 normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
                    const f3 diffuse = mk(sh[(RT_SH_DR + 0) * lay.Tp], sh[(RT_SH_DR + 1) * lay.Tp], sh[(RT_SH_DR + 2) * lay.Tp]);
                    const f3 P = hp + normal * 1e-3f;                                           /* raytrace.metal:67 */
                    const float wx = h4.x, wy = h4.y;                                           /* :72-73 */
                    color = color * diffuse;                                                    /* :76 */
                    f3 rd = mk(0.0f, 0.0f, 0.0f);
                    if (cont) {
                        const f3 right = mk(sh[(RT_SH_RX + 0) * lay.Tp], sh[(RT_SH_RX + 1) * lay.Tp], sh[(RT_SH_RX + 2) * lay.Tp]);
                        const f3 fwd = mk(sh[(RT_SH_FX + 0) * lay.Tp], sh[(RT_SH_FX + 1) * lay.Tp], sh[(RT_SH_FX + 2) * lay.Tp]);
                        const float ux = h4.z, uy = h4.w;                                       /* :93-94 */
                        f3 sdir = sample_cosine_hemisphere(ux, uy);
                        rd = (right * sdir.x + normal * sdir.y) + fwd * sdir.z;                 /* sampling.metal:65 */
                    }
                    const float *L = sm + lay.light;
                    const f3 sp = light_sample_point(mk(L[0], L[1], L[2]), L[3], L[7], wx, wy);             /* sampling.metal:207-212 */
                    /* which groups can the two rays reach?  (the box rule does not depend on the length of the
                     * direction: the shadow ray's is taken un-normalised, B makes the real one) */
                    const f3 ldir = sp - P;
                    key = 0u;
                    for (uint32_t g = 0; g < K; g++) {
                        const float *G = sm + lay.group + 4 + 8 * g;
                        const f3 w = mk(G[0], G[1], G[2]) - P, H = mk(G[4], G[5], G[6]);
                        bool clear = ray_clear_of_box(w, H, ldir);
                        if (cont) clear = clear && ray_clear_of_box(w, H, rd);
                        if (!clear) key |= 1u << g;
                    }
                    SF(0, slot) = P.x; SF(1, slot) = P.y; SF(2, slot) = P.z;
                    if (cont) { SF(3, slot) = rd.x; SF(4, slot) = rd.y; SF(5, slot) = rd.z; }
                    SF(6, slot) = color.x; SF(7, slot) = color.y; SF(8, slot) = color.z;
                    SF(9, slot) = sp.x; SF(10, slot) = sp.z;
                    keyarr[slot] = (uint8_t)(key | (shadow_inside_shell(shellBox, P, sp) ? 0u : RT_KEY_SHELL));
                }
                if (sort) {
                    const uint32_t same = __match_any_sync(0xffffffffu, key);
                    if (active && lane == (uint32_t)__ffs((int)same) - 1u) cnts[key] += __popc(same);
                }
                __syncwarp();
            }

            /* -- S. counting sort of the slot numbers, keys descending, stable -- */
            const uint8_t *order = hitlist;
            if (sort) {
                {
                    const uint32_t c = lane < RT_SORT_KEYS ? cnts[RT_SORT_KEYS - 1u - lane] : 0u;
                    uint32_t incl = c;
#pragma unroll
                    for (int d = 1; d < RT_SORT_KEYS; d <<= 1) {
                        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= (uint32_t)d) incl += up;
                    }
                    __syncwarp();
                    if (lane < RT_SORT_KEYS) cnts[RT_SORT_KEYS - 1u - lane] = incl - c;       /* start of each key's run */
                    __syncwarp();
                }
#pragma unroll 1
                for (uint32_t c0 = 0; c0 < count; c0 += 32) {
                    const uint32_t e = c0 + lane;
                    const bool active = e < count;
                    const uint32_t slot = active ? hitlist[e] : 0u;
                    const uint32_t key = active ? (keyarr[slot] & (RT_SORT_KEYS - 1u)) : 0xffu;
                    const uint32_t same = __match_any_sync(0xffffffffu, key);
                    const uint32_t leader = (uint32_t)__ffs((int)same) - 1u;
                    uint32_t start = 0u;
                    if (active && lane == leader) { start = cnts[key]; cnts[key] = start + __popc(same); }
                    start = __shfl_sync(0xffffffffu, start, leader);
                    if (active) sorted[start + __popc(same & lt)] = (uint8_t)slot;
                    __syncwarp();
                }
                if (lane < RT_SORT_KEYS) cnts[lane] = 0u;
                __syncwarp();
                order = sorted;
            }

            /* -- B. trace, 32 sorted entries at a time, against the quads their keys name -- */
            uint32_t m = 0;                                                    /* survivors listed so far */
#pragma unroll 1
            for (uint32_t c0 = 0; c0 < count; c0 += 32) {
                const uint32_t e = c0 + lane;
                const bool active = e < count;
                const uint32_t slot = active ? order[e] : 0u;
                const uint32_t key = active ? keyarr[slot] : 0u;
                __syncwarp();                                                  /* the chunk's entries are read: the hit list may be rewritten */
                const uint32_t keys = __reduce_or_sync(0xffffffffu, key);
                const uint32_t pairs = keyPairs[keys & (RT_SORT_KEYS - 1u)];
                /* the shadow rays skip the shell quads unless an entry of the chunk starts or ends outside the shell box */
                const uint32_t pairsS = ((keys & RT_KEY_SHELL) ? pairs : (pairs & ~shellQuads)) & ~grp[RT_GRP_LIGHT_PLANE];
                bool alive = false;
                uint32_t id = 0;
                if (active) {
                    const f3 P = mk(SF(0, slot), SF(1, slot), SF(2, slot));
                    const float *L = sm + lay.light;
                    f3 ldir; float ldist, inv;
                    light_direction(mk(SF(9, slot), L[1], SF(10, slot)), P, ldir, ldist, inv);  /* sampling.metal:214-220 */
                    const float ltmax = ldist - 1e-3f;                                          /* raytrace.metal:81 */
                    const uint32_t surf = idarr[slot];
                    bool hitC = false;
                    float dist = 0.0f;
                    f3 rd = mk(0.0f, 0.0f, 0.0f);
#if RT_EXP_TRACE == 0
                    if (cont) {
                        rd = mk(SF(3, slot), SF(4, slot), SF(5, slot));
                        hitC = closest_hit_pairs(sc, pairs, P, rd, 0.001f, 1000.0f, dist, id);   /* :49 */
                    }
                    const bool shadowed = any_hit_pairs(sc, pairsS, P, ldir, 0.0f, ltmax);       /* :79-88 */
#elif RT_EXP_TRACE == 1
                    bool shadowed;
                    if (cont) {
                        rd = mk(SF(3, slot), SF(4, slot), SF(5, slot));
                        hitC = dual_hit_pairs_masked(sc, pairs, pairsS, P, rd, 0.001f, 1000.0f, dist, id, ldir, 0.0f, ltmax, shadowed);
                    } else {
                        shadowed = any_hit_pairs(sc, pairsS, P, ldir, 0.0f, ltmax);
                    }
#else
                    bool shadowed;
                    if (cont) {
                        rd = mk(SF(3, slot), SF(4, slot), SF(5, slot));
                        hitC = dual_hit_pairs(sc, pairs, P, rd, 0.001f, 1000.0f, dist, id, ldir, 0.0f, ltmax, shadowed);
                    } else {
                        shadowed = any_hit_pairs(sc, RT_EXP_TRACE == 3 ? pairsS : pairs, P, ldir, 0.0f, ltmax);
                    }
#endif
                    if (!shadowed) {                                                            /* :74-75, :88 */
                        const float *sh = sm + lay.shade + surf;
                        const f3 normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
                        f3 lightColor = light_colour(mk(L[4], L[5], L[6]), inv, ldir);
                        lightColor = lightColor * saturate(dot(normal, ldir));
                        const f3 lc = lightColor * mk(SF(6, slot), SF(7, slot), SF(8, slot));
                        res[0 * CAP + slot] = res[0 * CAP + slot] + lc.x;
                        res[1 * CAP + slot] = res[1 * CAP + slot] + lc.y;
                        res[2 * CAP + slot] = res[2 * CAP + slot] + lc.z;
                    }
                    if (hitC) {
                        const float *sh = sm + lay.shade + id;
                        if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) {                                  /* :59 */
                            res[0 * CAP + slot] = sh[(RT_SH_ER + 0) * lay.Tp]; res[1 * CAP + slot] = sh[(RT_SH_ER + 1) * lay.Tp];
                            res[2 * CAP + slot] = sh[(RT_SH_ER + 2) * lay.Tp];
                        } else {
                            alive = true;
                            const f3 hp = P + rd * dist;
                            SF(0, slot) = hp.x; SF(1, slot) = hp.y; SF(2, slot) = hp.z;
                        }
                    }
                }
                const uint32_t mask = __ballot_sync(0xffffffffu, alive);
                if (alive) {
                    const uint32_t o = m + __popc(mask & lt);
                    hitlist[o] = (uint8_t)slot; idarr[slot] = (uint8_t)id;
                }
                m += __popc(mask);
                __syncwarp();
            }
            count = m;
        }

        /* ---- 3. in-order accumulation, raytrace.metal:103 ---------------------------------- */
        __syncwarp();
        for (uint32_t s = 0; s < nround; s++)
            lum = lum + mk(res[0 * CAP + s * 32u + lane], res[1 * CAP + s * 32u + lane], res[2 * CAP + s * 32u + lane]);
        __syncwarp();
    }

    if (PERSIST && kEnd < p.spp) {
        tiles_publish_chunk(ts.flags + tile, chunk, out + pix, make_float4(lum.x, lum.y, lum.z, 0.0f), owns);
    } else if (owns) {
        float4 o;
        if (p.accumulate) {
            o = make_float4(lum.x, lum.y, lum.z, (float)(p.firstSample + p.spp));
        } else {
            float samples = (float)p.spp;                                   /* raytrace.metal:106-109 */
            o = make_float4(lum.x / samples, lum.y / samples, lum.z / samples, 1.0f);
        }
        out[pix] = o;
    }
    break;
    }
#undef SF
}
