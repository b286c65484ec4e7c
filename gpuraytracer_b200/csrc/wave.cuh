/*
 * wave.cuh -- `pathTrace` with warp-level path compaction (included by pathtrace.cu).
 *
 * Same arithmetic, same per-pixel results (bit for bit) as pathtrace_body, different
 * scheduling.  In the per-pixel loop a lane whose path ended (miss / emissive hit) idles
 * until the slowest path of its warp has finished its bounces: ncu shows 26.8 of 32 lanes
 * active on average, 24.7 in the shadow loop (profiles/r01b_*).  Here a warp owns its 32
 * pixels but traces R samples of each per ROUND (32*R paths):
 *
 *   1. camera rays: R/2 passes over the 32 pixels, two samples per pass sharing the triangle loads
 *      (closest_primary2), camera-ray form of the triangle test;
 *      paths that land on a diffuse surface are appended to a queue in shared memory
 *      (ballot + popc give each survivor its slot), the others deliver their radiance;
 *   2. per bounce: the queue is consumed 32 entries at a time by FULL warps -- shade,
 *      shadow ray, cosine continuation, closest hit -- and the survivors are appended
 *      to the same queue in place (a chunk is read into registers before its survivors
 *      are written, and survivors never outnumber the entries consumed);
 *   3. every path delivers its radiance into res[sample][lane]; after the round each
 *      lane adds its R results IN SAMPLE ORDER, so the fp32 sum is the reference's
 *      (RTrace/raytrace.metal:103).
 *
 * Only the last chunk of a bounce runs with idle lanes.  No block-level barrier is
 * involved after the scene staging: warps run independently (__syncwarp only).
 * Queue entry (SoA, 12 words): hit point, colour (throughput), accumulated radiance,
 * primitive id, Halton index, tag = sample*32 + lane.
 */
#pragma once

#include "rt_kernel_smem.h"     /* RT_WAVE_FIELDS, RT_WAVE_RES and the sizes the host launches with */

/* Work distribution of pathTrace_tiles: a unit is one 16x8-pixel block tile x one CHUNK of chunkSpp consecutive
 * samples, one unit per block, the grid is tiles x chunks blocks.  A block takes a ticket from an atomic counter
 * and the ticket, not blockIdx, names its unit, chunk-major (every tile's chunk 0, then every tile's chunk 1, ...),
 * so units start in that order.  Chunk c of a warp's 8x4 pixels continues the running sum chunk c-1 left in the
 * output buffer -- the same in-order fp32 sum as one pass over all samples -- and waits for
 * flags[that warp tile] == c first; the unit it waits for started a whole image earlier and is running or done,
 * so the wait is short and cannot deadlock.  Small units are what keeps every SM busy to the end when a GPU has
 * only a few blocks per SM (1080p over 8 GPUs: 3.4 waves), where whole tiles leave the last wave half idle.
 * (Tried first: a persistent grid whose warps, then blocks, loop over units -- same loops, 5 % slower.) */
struct RTTileSched {
    unsigned int *counter;     /* next unit; zeroed by rt_prep_scene (prepared-scene header) */
    unsigned int *flags;       /* per warp tile (4 per block tile): chunks completed; zeroed by the host before the launch */
    unsigned int *status;      /* host-mapped word of the queue: a bit is set when a wait below gives up (rtcuda.cpp turns it
                                  into RT_ERR_DEVICE_TIMEOUT at rtWaitUntilCompleted); never written otherwise */
    uint32_t chunks, chunkSpp;
};
#define RT_STATUS_CHUNK_WAIT_TIMEOUT 1u
/* how long a warp waits for the chunk before its own, in ns of %globaltimer.  The producer holds a smaller ticket, so
 * it is running or done and the legitimate wait is at most one unit's run time (milliseconds; under
 * compute-sanitizer hundreds of them): the bound only exists so that a lost flag write or a corrupt flag array
 * ends as an error code instead of a hung GPU, like the bounded BVH walk (bvh.cuh). */
#define RT_CHUNK_WAIT_NS 8000000000ull

/* the two rare steps of a chunked unit, out of line so that they do not disturb the register allocation of the loops */
__device__ __forceinline__ unsigned long long rt_globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __noinline__ float4 tiles_wait_for_previous_chunk(const unsigned int *flag, uint32_t chunk, const float4 *src, bool owns,
                                                             unsigned int *status) {
    if ((threadIdx.x & 31) == 0) {
        unsigned int done;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(done) : "l"(flag) : "memory");
        if (done < chunk) {
            const unsigned long long t0 = rt_globaltimer();
            uint32_t spins = 0;
            do {
                __nanosleep(64);
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(done) : "l"(flag) : "memory");
                /* the clock is read every 256th poll only */
                if (done < chunk && (++spins & 255u) == 0u && rt_globaltimer() - t0 > RT_CHUNK_WAIT_NS) {
                    atomicOr_system(status, RT_STATUS_CHUNK_WAIT_TIMEOUT);
                    break;
                }
            } while (done < chunk);
        }
    }
    __syncwarp();
    return owns ? __ldcg(src) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}
__device__ __noinline__ void tiles_publish_chunk(unsigned int *flag, uint32_t chunk, float4 *dst, float4 sum, bool owns) {
    if (owns) __stcg(dst, sum);
    __threadfence();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(flag), "r"(chunk + 1u) : "memory");
}

/* triangles per iteration of the several-rays-a-pass loops (multi_hit, any_hit_multi): 1 -- five rays against one triangle
 * are 2.7 KB of straight-line code already, and the 256-sphere frame is bound by instruction fetch, not by the loop's overhead */
#ifndef RT_MULTI_UNROLL
#define RT_MULTI_UNROLL 1
#endif

template <int R>
struct WaveSmem {
    static constexpr uint32_t CAP = 32u * R;                                  /* paths per warp per round */
    static constexpr uint32_t WORDS_PER_WARP = rtWaveWordsPerWarp(R);
};

template <bool COUNT, int UNROLL, int R, bool PERSIST, bool SPATIAL, bool MIS = false>
__device__ __forceinline__ void pathtrace_wave_body(const float *__restrict__ scene, const uint32_t *__restrict__ offsets,
                                                    float4 *__restrict__ out, const RTLaunchParams p,
                                                    const RTSceneLayout lay, float *sm,
                                                    unsigned long long *__restrict__ counters, const RTTileSched ts,
                                                    const RTHaltonTables ht) {
    constexpr uint32_t CAP = WaveSmem<R>::CAP;
    {
        const float4 *src = (const float4 *)scene;
        float4 *dst = (float4 *)sm;
        const uint32_t n4 = lay.words >> 2;
        for (uint32_t k = threadIdx.x; k < n4; k += RT_BLOCK_THREADS) dst[k] = src[k];
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    /* this warp's queue and result slots, after the scene */
    float *Q = sm + lay.words + warp * WaveSmem<R>::WORDS_PER_WARP;
    float *res = Q + RT_WAVE_FIELDS * CAP;
    uint32_t *Qu = (uint32_t *)Q;
#define QF(field, e) Q[(field) * CAP + (e)]
#define QU(field, e) Qu[(field) * CAP + (e)]

    SceneRef<false> sc;
    sc.base = (const float4 *)(sm + lay.isect);
    sc.T = (lay.T + (UNROLL - 1)) / UNROLL * UNROLL; sc.S = (lay.S + 3u) & ~3u; sc.Treal = lay.T;
    sc.sph = 4 * lay.Tp;
    /* spheres: walk the grid rt_prep_scene built (sgrid.cuh) unless it is flagged unusable */
    if (lay.S && !(lay.flags & RT_LAYOUT_NO_SGRID) && ((const uint32_t *)scene)[lay.sgrid + 3] != 0u) sc.grid = scene + lay.sgrid;

    const f3 cam_o = mk(sm[RT_HDR_CAM_O], sm[RT_HDR_CAM_O + 1], sm[RT_HDR_CAM_O + 2]);
    const f3 cam_u = mk(sm[RT_HDR_CAM_U], sm[RT_HDR_CAM_U + 1], sm[RT_HDR_CAM_U + 2]);
    const f3 cam_v = mk(sm[RT_HDR_CAM_V], sm[RT_HDR_CAM_V + 1], sm[RT_HDR_CAM_V + 2]);
    const f3 cam_w = mk(sm[RT_HDR_CAM_W], sm[RT_HDR_CAM_W + 1], sm[RT_HDR_CAM_W + 2]);
    const float halfW = sm[RT_HDR_HALF_W], halfH = sm[RT_HDR_HALF_H];
    const float resX = sm[RT_HDR_RES_X], resY = sm[RT_HDR_RES_Y];
    const uint32_t nb = (uint32_t)p.bounces, numLights = p.numLights;
    const bool paired = ((const uint32_t *)sm)[RT_HDR_PAIRED] != 0u;           /* quads cut in two: pathtrace.cu, PAIRED */
    /* the shell quads the shadow rays of a several-lights pass may skip (0: none, or some light reaches out of the shell box) */
    const uint32_t lightPlane = ((const uint32_t *)sm)[lay.group + RT_GRP_LIGHT_PLANE];
    const uint32_t shellSkip = ((const uint32_t *)sm)[lay.group + RT_GRP_LIGHTS_IN] ? ((const uint32_t *)sm)[lay.group + RT_GRP_SHELL] : 0u;
    Counters cnt = { 0u, 0u, 0u };
    unsigned long long paths = 0ull;

    /* One 8x4-pixel tile per warp: its place in the block's 16x8 pixels; PERSIST: the block's unit comes from a ticket. */
    /* SPATIAL (launches with fewer samples per pixel than the R slots of a round, progressive rendering): the
     * warp takes G = R / spp tiles side by side (sub-tile t = pixels x + 8t) and a round's slot j is sample
     * j % spp of sub-tile j / spp, so the queue is as full as with R samples of one tile.  One round per warp;
     * each pixel's samples are consecutive slots, added in order in phase 3. */
    const uint32_t G = SPATIAL ? max(1u, (uint32_t)R / max(p.spp, 1u)) : 1u;
    for (;;) {
    uint32_t x0, ly0, tile = 0, chunk = 0, kBegin = 0, kEnd = p.spp;
    if (PERSIST) {
        /* the block draws its unit from the counter instead of reading blockIdx: units START in unit order whatever
         * order the hardware dispatches blocks in, which is what makes waiting for the previous chunk safe */
        uint32_t *s_unit = (uint32_t *)(sm + lay.words + (RT_BLOCK_THREADS / 32) * WaveSmem<R>::WORDS_PER_WARP) +
                           (RT_BLOCK_THREADS / 32) * (((lay.T + 31u) >> 5) + ((lay.S + 31u) >> 5));
        __syncthreads();
        if (threadIdx.x == 0) *s_unit = atomicAdd(ts.counter, 1u);
        __syncthreads();
        const uint32_t unit = *s_unit;
        const uint32_t blocksX = ((uint32_t)p.width + RT_BLOCK_W - 1u) / RT_BLOCK_W, blocksY = ((uint32_t)p.rows + RT_BLOCK_H - 1u) / RT_BLOCK_H;
        const uint32_t nBlocks = blocksX * blocksY;
        if (unit >= nBlocks * ts.chunks) break;
        chunk = unit / nBlocks;
        const uint32_t bt = unit - chunk * nBlocks;
        tile = bt * (RT_BLOCK_THREADS / 32) + warp;                    /* this warp's flag: same pixels in every chunk */
        kBegin = chunk * ts.chunkSpp; kEnd = min(p.spp, kBegin + ts.chunkSpp);
        x0 = (bt % blocksX) * RT_BLOCK_W + (warp & 1) * 8; ly0 = (bt / blocksX) * RT_BLOCK_H + (warp >> 1) * 4;
    } else {
        x0 = (blockIdx.x * RT_BLOCK_W + (warp & 1) * 8) * G; ly0 = blockIdx.y * RT_BLOCK_H + (warp >> 1) * 4;
    }
    if (SPATIAL) kEnd = G * p.spp;                 /* one round: G sub-tiles x spp samples <= R slots */
    const uint32_t x = x0 + (lane & 7), ly = ly0 + (lane >> 3);
    /* out-of-image lanes stay in the warp (they help consume the queue) but own no pixel */
    const bool owns = x < (uint32_t)p.width && ly < (uint32_t)p.rows;
    const uint32_t y = p.tileRows ? (uint32_t)p.row0 + (ly / p.tileRows) * p.tileStride + (ly % p.tileRows)
                                  : (uint32_t)p.row0 + ly;
    const size_t pix = (size_t)ly * (uint32_t)p.width + x;
    const float fx = (float)(int)x, fy = (float)(int)y;

    /* ---- 0. which primitives can a camera ray of this warp's pixels hit?  (camera_candidates) ---- */
    const uint32_t mwT = (lay.T + 31u) >> 5, mwS = (lay.S + 31u) >> 5;
    uint32_t *cullT = (uint32_t *)(sm + lay.words + (RT_BLOCK_THREADS / 32) * WaveSmem<R>::WORDS_PER_WARP) + warp * (mwT + mwS);
    uint32_t *cullS = cullT + mwT;
    {
        CameraFrame cf = { cam_o, cam_u, cam_v, cam_w, halfW, halfH, resX, resY };
        const int big = 0x3fffffff;
        const int xr = SPATIAL ? (int)min(x + 8u * (G - 1u), (uint32_t)p.width - 1u) : (int)x;      /* rightmost sub-tile */
        const int x0 = __reduce_min_sync(0xffffffffu, owns ? (int)x : big), x1 = __reduce_max_sync(0xffffffffu, owns ? xr : -big);
        const int y0 = __reduce_min_sync(0xffffffffu, owns ? (int)y : big), y1 = __reduce_max_sync(0xffffffffu, owns ? (int)y : -big);
        camera_candidates(sc, cf, lay.T, lay.S, (float)x0 - 1.0f, (float)x1 + 2.0f, (float)y0 - 1.0f, (float)y1 + 2.0f,
                          cullT, mwT, cullS, mwS, lane);
    }

    const uint32_t offset = owns ? offsets[pix] : 0u;
    f3 lum = mk(0.0f, 0.0f, 0.0f);
    if (PERSIST && chunk > 0) {
        /* the running sum of the samples before kBegin, once the block that made it has published it */
        const float4 o = tiles_wait_for_previous_chunk(ts.flags + tile, chunk, out + pix, owns, ts.status);
        lum = mk(o.x, o.y, o.z);
    } else if (!SPATIAL && p.accumulate && owns) { float4 o = out[pix]; lum = mk(o.x, o.y, o.z); }

#pragma unroll 1
    for (uint32_t k0 = kBegin; k0 < kEnd; k0 += R) {
        const uint32_t nround = min((uint32_t)R, kEnd - k0);                  /* warp-uniform */
        uint32_t count = 0;                                                   /* queue length, warp-uniform */

        /* ---- 1. camera rays -------------------------------------------------------------- */
#pragma unroll 1
        for (uint32_t s = 0; s < nround; s += 2) {
            /* two samples of the pixel per pass (closest_primary2); the second one is void when nround is odd */
            const bool two = s + 1 < nround;                                   /* warp-uniform */
            bool alive[2] = { false, false };
            uint32_t id[2] = { 0u, 0u }, idx[2];
            f3 hp[2], r0[2];
            hp[0] = hp[1] = r0[0] = r0[1] = mk(0.0f, 0.0f, 0.0f);
            bool own[2] = { owns, owns };
            float fxq[2] = { fx, fx };
            if (SPATIAL) {
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    const uint32_t j = s + q, t = j / p.spp, xt = x + 8u * t;
                    own[q] = owns && xt < (uint32_t)p.width && j < nround;
                    fxq[q] = (float)(int)xt;
                    idx[q] = (own[q] ? offsets[pix + 8u * t] : 0u) + (p.firstSample + (j - t * p.spp));
                }
            } else {
                idx[0] = offset + (p.firstSample + k0 + s); idx[1] = idx[0] + 1u;
            }
            if (own[0] || own[1]) {
                f3 rd[2];
#pragma unroll
                for (int q = 0; q < 2; q++) {                                  /* sampling.metal:141-147 */
                    float jx = halton<0>(idx[q]), jy = halton_cam_y(ht, idx[q]);
                    float sx = ((fxq[q] + jx) / resX) * 2.0f - 1.0f;
                    float ty = -(((fy + jy) / resY) * 2.0f - 1.0f);
                    rd[q] = normalize((cam_u * (sx * halfW) + cam_v * (ty * halfH)) - cam_w);
                }
                bool hit[2]; float dist[2];
                /* (an odd last pass traces its second ray for nothing: one code path, smaller kernel) */
                closest_primary2_list(sc, cullT, mwT, cullS, mwS, rd[0], rd[1], 0.001f, 1000.0f, hit[0], dist[0], id[0], hit[1], dist[1], id[1]);
                hit[0] = hit[0] && own[0];
                hit[1] = hit[1] && two && own[1];
                if (COUNT) cnt.closest += SPATIAL ? (own[0] ? 1u : 0u) + (two && own[1] ? 1u : 0u) : (two ? 2u : 1u);
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    if (hit[q]) {
                        const bool tri = id[q] < lay.T;
                        const float *sh = tri ? sm + lay.shade + id[q] : sm + lay.sshade + (id[q] - lay.T);
                        const uint32_t stride = tri ? lay.Tp : lay.Sp;
                        const uint32_t fl = tri ? RT_SH_FLAG : RT_SSH_FLAG, er = tri ? RT_SH_ER : RT_SSH_ER;
                        if (sh[fl * stride] != 0.0f) r0[q] = mk(sh[(er + 0) * stride], sh[(er + 1) * stride], sh[(er + 2) * stride]);
                        else { alive[q] = true; hp[q] = cam_o + rd[q] * dist[q]; }
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < 2; q++) {
                if (q == 1 && !two) break;
                const uint32_t mask = __ballot_sync(0xffffffffu, alive[q]);
                const uint32_t tag = (s + q) * 32u + lane;
                if (alive[q]) {
                    const uint32_t e = count + __popc(mask & ((1u << lane) - 1u));
                    QF(0, e) = hp[q].x; QF(1, e) = hp[q].y; QF(2, e) = hp[q].z;
                    QF(3, e) = 1.0f; QF(4, e) = 1.0f; QF(5, e) = 1.0f;
                    QF(6, e) = 0.0f; QF(7, e) = 0.0f; QF(8, e) = 0.0f;
                    QU(9, e) = id[q]; QU(10, e) = idx[q]; QU(11, e) = tag;
                } else {
                    res[0 * CAP + tag] = r0[q].x; res[1 * CAP + tag] = r0[q].y; res[2 * CAP + tag] = r0[q].z;
                }
                count += __popc(mask);
            }
        }
        __syncwarp();

        /* ---- 2. bounces: consume the queue with full warps ------------------------------- */
#pragma unroll 1
        for (uint32_t b = 0; b < nb && count; b++) {
            uint32_t m = 0;                                                    /* survivors written so far */
#pragma unroll 1
            for (uint32_t c0 = 0; c0 < count; c0 += 32) {
                const uint32_t e = c0 + lane;
                const bool active = e < count;
                bool alive = false;
                f3 hp, color, acc; uint32_t id = 0, i = 0, tag = 0;
                if (active) {
                    hp = mk(QF(0, e), QF(1, e), QF(2, e));
                    color = mk(QF(3, e), QF(4, e), QF(5, e));
                    acc = mk(QF(6, e), QF(7, e), QF(8, e));
                    id = QU(9, e); i = QU(10, e); tag = QU(11, e);
                }
                /* the bounce's four Halton values: one load from the table (rt_device_math.cuh: RTHaltonTables) */
                const float4 h4 = active ? halton_bounce4(ht, i, b) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                __syncwarp();                                                  /* chunk is in registers: slots may be reused */
                if (active) {
                    f3 normal, right, fwd, diffuse;
                    if (id < lay.T) {
                        const float *sh = sm + lay.shade + id;
                        normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
                        right = mk(sh[(RT_SH_RX + 0) * lay.Tp], sh[(RT_SH_RX + 1) * lay.Tp], sh[(RT_SH_RX + 2) * lay.Tp]);
                        fwd = mk(sh[(RT_SH_FX + 0) * lay.Tp], sh[(RT_SH_FX + 1) * lay.Tp], sh[(RT_SH_FX + 2) * lay.Tp]);
                        diffuse = mk(sh[(RT_SH_DR + 0) * lay.Tp], sh[(RT_SH_DR + 1) * lay.Tp], sh[(RT_SH_DR + 2) * lay.Tp]);
                    } else {
                        const uint32_t j = id - lay.T;
                        const float *sh = sm + lay.sshade + j;
                        const float *S = sm + lay.sphere + RT_SPHERE_WORDS * j;
                        normal = normalize(hp - mk(S[0], S[1], S[2]));
                        right = normalize(cross(normal, mk(0.0072f, 1.0f, 0.0034f)));
                        fwd = cross(right, normal);
                        diffuse = mk(sh[(RT_SSH_DR + 0) * lay.Sp], sh[(RT_SSH_DR + 1) * lay.Sp], sh[(RT_SSH_DR + 2) * lay.Sp]);
                    }
                    f3 P = hp + normal * 1e-3f;                                                 /* raytrace.metal:67 */
                    if (COUNT) { cnt.shaded++; cnt.shadow += numLights; }
                    const float wx = h4.x, wy = h4.y;                                           /* :72-73 */
                    color = color * diffuse;                                                    /* :76 */
                    /* the continuation ray starts at the same point as the shadow ray: it is sampled first
                     * and traced together with the first light's shadow ray (dual_hit) */
                    const bool cont = b + 1 < nb;                                               /* warp-uniform */
                    /* MIS (mis.cuh): the continuation ray is traced after the last bounce too, for its cosine-strategy term */
                    const bool traceC = MIS || cont;
                    f3 rd = mk(0.0f, 0.0f, 0.0f);
                    if (traceC) {
                        const float ux = h4.z, uy = h4.w;                                       /* :93-94 */
                        f3 sdir = sample_cosine_hemisphere(ux, uy);
                        rd = (right * sdir.x + normal * sdir.y) + fwd * sdir.z;                 /* sampling.metal:65 */
                        if (COUNT) cnt.closest++;
                    }
                    bool hitC = false; float dist = 0.0f;
                    /* shadow rays, one per light (:79-88; the reference has one light).  The first one rides with
                     * the continuation ray (dual_hit; with three or four lights all of them in one multi_hit pass): every
                     * ray of a bounce starts at P, so the origin-dependent part of each test is shared. */
                    if (numLights == 1) {            /* the reference's case: its own straight-line code */
                        const float *L = sm + lay.light;
                        f3 ldir; float ldist;
                        f3 lightColor = sample_area_light(L, wx, wy, P, ldir, ldist);
                        lightColor = lightColor * saturate(dot(normal, ldir));                  /* :75 */
                        if (MIS) lightColor = lightColor * misA_light_weight(L, P, normal, ldir);
                        bool shadowed;
                        if (paired) {                                                           /* warp-uniform, per launch */
                            if (traceC) hitC = dual_hit<false, UNROLL, true>(sc, P, rd, 0.001f, 1000.0f, dist, id, ldir, 0.0f, ldist - 1e-3f, shadowed);
                            else shadowed = any_hit<false, UNROLL, true>(sc, P, ldir, 0.0f, ldist - 1e-3f);
                        } else {
                            if (traceC) hitC = dual_hit<false, UNROLL>(sc, P, rd, 0.001f, 1000.0f, dist, id, ldir, 0.0f, ldist - 1e-3f, shadowed);
                            else shadowed = any_hit<false, UNROLL>(sc, P, ldir, 0.0f, ldist - 1e-3f);
                        }
                        if (!shadowed) acc = acc + lightColor * color;                          /* :88 */
                    } else {
                        /* up to four lights a pass: their shadow rays, and on the first pass the continuation ray too, share
                         * one walk over the primitives (pathtrace.cu: multi_hit / any_hit_multi) */
                        uint32_t l = 0;
                        while (l < numLights) {
                            const uint32_t left = numLights - l;
                            const bool withC = traceC && l == 0;
                            const uint32_t n = left >= 4u ? 4u : left;
                            f3 ldir[4], lcol[4]; float ltmax[4]; bool shd[4] = { false, false, false, false };
#pragma unroll
                            for (uint32_t r = 0; r < 4; r++) {
                                /* (a ray that is not wanted: tmax = 0, nothing can be nearer) */
                                ldir[r] = mk(0.0f, 1.0f, 0.0f); lcol[r] = mk(0.0f, 0.0f, 0.0f); ltmax[r] = 0.0f;
                                if (r < n) {
                                    const float *L = sm + lay.light + RT_LIGHT_WORDS * (l + r);
                                    float4 dt, cl;
                                    sample_area_light_shaded(L, wx, wy, P, normal, dt, cl);         /* :74-75 */
                                    ldir[r] = mk(dt.x, dt.y, dt.z); ltmax[r] = dt.w; lcol[r] = mk(cl.x, cl.y, cl.z);
                                    if (MIS) lcol[r] = lcol[r] * misA_light_weight(L, P, normal, ldir[r]);
                                }
                            }
                            /* the shell rule for a pass of shadow rays (rt_scene_layout.h: RT_GRP_LIGHTS_IN): every light's
                             * sample rectangle lies inside the shrunk shell box, so a shadow ray from a point inside it cannot
                             * be accepted by a shell triangle; skipped when that holds for every lane here */
                            uint32_t skipS = 0u;
                            if (shellSkip) {
                                const float *SB = sm + lay.group + RT_GRP_SHELL_BOX;
                                const bool in = (fabsf(P.x - SB[0]) < SB[4]) & (fabsf(P.y - SB[1]) < SB[5]) & (fabsf(P.z - SB[2]) < SB[6]);
                                skipS = __reduce_and_sync(__activemask(), in ? shellSkip : 0u);
                            }
                            skipS |= lightPlane;            /* the lights' own quads: no shadow ray reaches them before its end (RT_GRP_LIGHT_PLANE) */
                            if (withC && n >= 3u) {
                                hitC = multi_hit<RT_MULTI_UNROLL, 4>(sc, P, rd, 0.001f, 1000.0f, dist, id, ldir, ltmax, shd, skipS);                   /* :49 + :79-88 */
                            } else if (withC) {
                                hitC = dual_hit<false, UNROLL>(sc, P, rd, 0.001f, 1000.0f, dist, id, ldir[0], 0.0f, ltmax[0], shd[0]);
                                if (n == 2u) shd[1] = any_hit<false, UNROLL>(sc, P, ldir[1], 0.0f, ltmax[1]);
                            } else if (n == 4u) {
                                any_hit_multi<false, RT_MULTI_UNROLL, 4>(sc, P, ldir, ltmax, shd, skipS);
                            } else if (n == 3u) {
                                f3 d3[3] = { ldir[0], ldir[1], ldir[2] }; float t3[3] = { ltmax[0], ltmax[1], ltmax[2] }; bool s3[3];
                                any_hit_multi<false, RT_MULTI_UNROLL, 3>(sc, P, d3, t3, s3, skipS);
                                shd[0] = s3[0]; shd[1] = s3[1]; shd[2] = s3[2];
                            } else if (n == 2u) {
                                f3 d2[2] = { ldir[0], ldir[1] }; float t2[2] = { ltmax[0], ltmax[1] }; bool s2[2];
                                any_hit_multi<false, RT_MULTI_UNROLL, 2>(sc, P, d2, t2, s2, skipS);
                                shd[0] = s2[0]; shd[1] = s2[1];
                            } else {
                                shd[0] = any_hit<false, UNROLL>(sc, P, ldir[0], 0.0f, ltmax[0]);
                            }
#pragma unroll
                            for (uint32_t r = 0; r < 4; r++)
                                if (r < n && !shd[r]) acc = acc + lcol[r] * color;                  /* :88, in light order */
                            l += n;
                        }
                    }
                    if (MIS) {      /* the cosine strategy's term of every light for this continuation ray, in light order */
                        for (uint32_t l = 0; l < numLights; l++)
                            acc = acc + misA_cosine_term(sm + lay.light + RT_LIGHT_WORDS * l, P, normal, rd, hitC, dist, color);
                    }
                    if (hitC) {
                        const bool tri = id < lay.T;
                        const float *sh = tri ? sm + lay.shade + id : sm + lay.sshade + (id - lay.T);
                        const uint32_t stride = tri ? lay.Tp : lay.Sp;
                        const uint32_t fl = tri ? RT_SH_FLAG : RT_SSH_FLAG, er = tri ? RT_SH_ER : RT_SSH_ER;
                        if (sh[fl * stride] != 0.0f) {                                          /* :59; MIS: the path ends, nothing is assigned */
                            if (!MIS) acc = mk(sh[(er + 0) * stride], sh[(er + 1) * stride], sh[(er + 2) * stride]);
                        } else if (cont) { alive = true; hp = P + rd * dist; }
                    }
                }
                const uint32_t mask = __ballot_sync(0xffffffffu, alive);
                if (alive) {
                    const uint32_t o = m + __popc(mask & ((1u << lane) - 1u));
                    QF(0, o) = hp.x; QF(1, o) = hp.y; QF(2, o) = hp.z;
                    QF(3, o) = color.x; QF(4, o) = color.y; QF(5, o) = color.z;
                    QF(6, o) = acc.x; QF(7, o) = acc.y; QF(8, o) = acc.z;
                    QU(9, o) = id; QU(10, o) = i; QU(11, o) = tag;
                } else if (active) {
                    res[0 * CAP + tag] = acc.x; res[1 * CAP + tag] = acc.y; res[2 * CAP + tag] = acc.z;
                }
                m += __popc(mask);
                __syncwarp();
            }
            count = m;
        }
        /* paths still queued after the last bounce cannot exist: b+1 == nb never enqueues */

        /* ---- 3. in-order accumulation, raytrace.metal:103 ---------------------------------- */
        __syncwarp();
        if (SPATIAL) {
            for (uint32_t t = 0; t < G; t++) {
                const uint32_t xt = x + 8u * t;
                if (!(owns && xt < (uint32_t)p.width)) continue;
                const size_t pt = pix + 8u * t;
                f3 l = mk(0.0f, 0.0f, 0.0f);
                if (p.accumulate) { float4 o = out[pt]; l = mk(o.x, o.y, o.z); }
                for (uint32_t s = t * p.spp; s < (t + 1u) * p.spp; s++)
                    l = l + mk(res[0 * CAP + s * 32u + lane], res[1 * CAP + s * 32u + lane], res[2 * CAP + s * 32u + lane]);
                float4 o;
                if (p.accumulate) o = make_float4(l.x, l.y, l.z, (float)(p.firstSample + p.spp));
                else { float samples = (float)p.spp; o = make_float4(l.x / samples, l.y / samples, l.z / samples, 1.0f); }
                out[pt] = o;
                if (COUNT) paths += p.spp;
            }
        } else {
            for (uint32_t s = 0; s < nround; s++)
                lum = lum + mk(res[0 * CAP + s * 32u + lane], res[1 * CAP + s * 32u + lane], res[2 * CAP + s * 32u + lane]);
        }
        __syncwarp();
    }

    if (SPATIAL) {
        /* stored per sub-tile above */
    } else if (PERSIST && kEnd < p.spp) {
        /* not the last chunk: leave the running sum for the next one and publish it */
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
    if (COUNT && owns && !SPATIAL) paths += kEnd - kBegin;
    break;           /* one unit per block: the grid is blocks x chunks (a persistent loop here was measured 5 % slower) */
    }
#undef QF
#undef QU
    if (COUNT) {
        atomicAdd(&counters[0], paths);
        atomicAdd(&counters[1], (unsigned long long)cnt.closest);
        atomicAdd(&counters[2], (unsigned long long)cnt.shadow);
        atomicAdd(&counters[3], (unsigned long long)cnt.shaded);
    }
}
