/*
 * sortedblk.cuh -- sorted.cuh with the sort taken over the whole BLOCK (included by pathtrace.cu after sorted.cuh).
 *
 * In sorted.cuh every warp sorts the <= 32 R paths of its own round.  With four key classes and ~170 entries per
 * bounce, half of the 32-entry chunks mix classes and are traced against the union of their quads: 9.9 quads
 * per bounce step where the keys alone would need 8.1 (oracle-side count over a frame).  Here the four
 * warps of a block pool their rounds: 128 R slots, one hit list, one counting sort per bounce (8.65 quads per
 * step at R = 6), chunks of the sorted list handed to whichever warp is free (a ticket per chunk: chunks cost
 * 3 to 9 passes of the quad loop depending on their key).  The slots still belong to (warp, sample, lane), so the
 * camera phase and the in-order accumulation (raytrace.metal:103) are each warp's own; in between, any warp may
 * trace any path.  Same arithmetic, same bits.
 *
 * Price: block barriers (camera | per bounce: A | sort, twice | B) instead of __syncwarp, shared-memory atomics for the
 * key counters and list appends.
 *
 * Since round 2 the key also says through which face of the room the continuation ray can leave it (the exit rule,
 * rt_scene_layout.h: three more bits, 128 keys; not in the wide form), the shadow rays skip the walls (shell rule) and
 * the light's own quads (light-plane rule), and an entry's rank within its key is the value pass A's counter returns, so
 * that the sort only turns ranks into places: 5.0 quads per step for the continuation ray, 80 warp instructions a path.
 */
#pragma once

#ifndef RT_EXP_SORT_UNROLL
#define RT_EXP_SORT_UNROLL 2   /* A/B (profiles/r04x_ab.log: 1: 55.54 ms, 2: 55.09, 3: 55.35 at 256 spp): iterations of the rank loop in flight (a chain of dependent shared-memory round trips each) */
#endif
#ifndef RT_EXP_RANK_IN_A
#define RT_EXP_RANK_IN_A 1     /* A/B: 1 = an entry's rank within its key is taken in pass A (the sort loses its match and its atomic, gains a barrier) */
#endif
#ifndef RT_EXP_TICKET_AHEAD
#define RT_EXP_TICKET_AHEAD 1  /* A/B (tools/build_exp.sh): 1 = pass B draws the next chunk's ticket before tracing the current chunk */
#endif

/* The slots of a block's round, in either layout of rt_kernel_smem.h (RT_SORTBLK_F4).  What a slot holds over a bounce:
 * camera phase / pass B leave the hit point X and the surface id; pass A turns X into P and adds the continuation
 * direction D, the throughput colour, x and z of the light sample point and the sort key; pass B reads all of it. */
template <uint32_t NSLOT, bool F4>
struct SlotStore {
    float *Q;            /* F4: three float4 arrays A, B, C of NSLOT; else RT_SORT_FIELDS float arrays of NSLOT */
    uint8_t *idarr, *keyarr;     /* !F4 only */
    __device__ __forceinline__ float4 *A() const { return (float4 *)Q; }
    __device__ __forceinline__ float4 *B() const { return (float4 *)Q + NSLOT; }
    __device__ __forceinline__ float4 *C() const { return (float4 *)Q + 2 * NSLOT; }
    __device__ __forceinline__ float &F(uint32_t field, uint32_t s) const { return Q[field * NSLOT + s]; }
    /* a path that landed on a diffuse surface: hit point, colour (1,1,1), surface id */
    __device__ __forceinline__ void put_hit_first(uint32_t s, f3 hp, uint32_t id) const {
        if (F4) { A()[s] = make_float4(hp.x, hp.y, hp.z, 0.0f); C()[s] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(id)); }
        else { F(0, s) = hp.x; F(1, s) = hp.y; F(2, s) = hp.z; F(6, s) = 1.0f; F(7, s) = 1.0f; F(8, s) = 1.0f; idarr[s] = (uint8_t)id; }
    }
    __device__ __forceinline__ void get_shade(uint32_t s, f3 &hp, f3 &color, uint32_t &id) const {
        if (F4) { const float4 a = A()[s], c = C()[s]; hp = mk(a.x, a.y, a.z); color = mk(c.x, c.y, c.z); id = __float_as_uint(c.w) & 0xffu; }
        else { id = idarr[s]; hp = mk(F(0, s), F(1, s), F(2, s)); color = mk(F(6, s), F(7, s), F(8, s)); }
    }
    __device__ __forceinline__ void put_rays(uint32_t s, f3 P, f3 rd, bool haveD, f3 color, float spx, float spz, uint32_t id, uint32_t key) const {
        if (F4) {
            A()[s] = make_float4(P.x, P.y, P.z, spx);
            B()[s] = make_float4(rd.x, rd.y, rd.z, spz);
            C()[s] = make_float4(color.x, color.y, color.z, __uint_as_float(id | (key << 8)));
        } else {
            F(0, s) = P.x; F(1, s) = P.y; F(2, s) = P.z;
            if (haveD) { F(3, s) = rd.x; F(4, s) = rd.y; F(5, s) = rd.z; }
            F(6, s) = color.x; F(7, s) = color.y; F(8, s) = color.z;
            F(9, s) = spx; F(10, s) = spz;
            keyarr[s] = (uint8_t)key;
        }
    }
    __device__ __forceinline__ uint32_t key(uint32_t s) const {
        return F4 ? (((const uint32_t *)C())[4u * s + 3u] >> 8) & 0xffu : (uint32_t)keyarr[s];
    }
    /* pass B: everything of the slot; colour and id come with the key */
    __device__ __forceinline__ void get_trace_head(uint32_t s, f3 &color, uint32_t &surf, uint32_t &key) const {
        if (F4) { const float4 c = C()[s]; color = mk(c.x, c.y, c.z); const uint32_t w = __float_as_uint(c.w); surf = w & 0xffu; key = (w >> 8) & 0xffu; }
        else { key = keyarr[s]; surf = idarr[s]; color = mk(0.0f, 0.0f, 0.0f); }
    }
    __device__ __forceinline__ f3 get_color(uint32_t s, f3 fromHead) const { return F4 ? fromHead : mk(F(6, s), F(7, s), F(8, s)); }
    __device__ __forceinline__ void get_trace(uint32_t s, bool haveD, f3 &P, float &spx, float &spz, f3 &rd) const {
        if (F4) {
            const float4 a = A()[s], b = B()[s];
            P = mk(a.x, a.y, a.z); spx = a.w; rd = mk(b.x, b.y, b.z); spz = b.w;
        } else {
            P = mk(F(0, s), F(1, s), F(2, s)); spx = F(9, s); spz = F(10, s);
            rd = haveD ? mk(F(3, s), F(4, s), F(5, s)) : mk(0.0f, 0.0f, 0.0f);
        }
    }
    /* the path goes on: next hit point and surface (the key is spent) */
    __device__ __forceinline__ void put_hit_next(uint32_t s, f3 hp, uint32_t id) const {
        if (F4) { A()[s] = make_float4(hp.x, hp.y, hp.z, 0.0f); ((uint32_t *)C())[4u * s + 3u] = id; }
        else { F(0, s) = hp.x; F(1, s) = hp.y; F(2, s) = hp.z; idarr[s] = (uint8_t)id; }
    }
};

template <int R, bool SPATIAL = false>
struct SortBlkSmem {
    static constexpr uint32_t CAP = 32u * R;                                   /* slots per warp */
    static constexpr uint32_t NSLOT = rtSortedBlkSlots(R);                     /* slots per block */
    static_assert(RT_BLOCK_THREADS == 32 * RT_BLOCK_WARPS, "rt_kernel_smem.h sizes the arrays for this block");
    static_assert(NSLOT <= 65536u, "slot numbers are 16 bits");
    /* fields + res.xyz, two 16-bit lists (hit list, sorted list), two byte arrays (surface id, key per slot),
     * the lanes' offsets, key counters (count + running), three block counters */
    /* SPATIAL (fewer samples per launch than R: G = R / spp pixel tiles side by side per warp, as pathTrace_wide):
     * the Halton index of every slot instead of one offset per lane */
    static constexpr uint32_t OFFS = SPATIAL ? NSLOT : RT_BLOCK_THREADS;
    static constexpr uint32_t WORDS = rtSortedBlkCoreWords(R, SPATIAL);        /* rt_kernel_smem.h: what the host launches with */
    static_assert(WORDS == (RT_SORTBLK_F4 ? 15u * NSLOT + NSLOT : (RT_SORT_FIELDS + 3u) * NSLOT + (6u * NSLOT) / 4u) + OFFS + 3u * rtSortedBlkKeys(SPATIAL) + 4u, "layout below");
};

template <int R, bool PERSIST, bool SPATIAL = false, bool MIS = false>
__device__ __forceinline__ void pathtrace_sortedblk_body(const float *__restrict__ scene, const uint32_t *__restrict__ offsets,
                                                         float4 *__restrict__ out, const RTLaunchParams p,
                                                         const RTSceneLayout lay, float *sm, const RTTileSched ts,
                                                         const RTHaltonTables ht) {
    static_assert(!(PERSIST && SPATIAL), "one or the other");
    constexpr uint32_t CAP = SortBlkSmem<R>::CAP, NSLOT = SortBlkSmem<R>::NSLOT;
    {
        const float4 *src = (const float4 *)scene;
        float4 *dst = (float4 *)sm;
        const uint32_t n4 = lay.words >> 2;
        for (uint32_t k = threadIdx.x; k < n4; k += RT_BLOCK_THREADS) dst[k] = src[k];
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;

    constexpr bool F4 = RT_SORTBLK_F4 != 0;
    float *Q = sm + lay.words;
    float *res = Q + (F4 ? 12u : (uint32_t)RT_SORT_FIELDS) * NSLOT;
    uint16_t *hitlist = (uint16_t *)(res + 3 * NSLOT), *sorted = hitlist + NSLOT;
    uint8_t *idarr = (uint8_t *)(sorted + NSLOT), *keyarr = idarr + NSLOT;              /* (not there with F4: id and key live in the slot) */
    uint32_t *offs = F4 ? (uint32_t *)(sorted + NSLOT) : (uint32_t *)(keyarr + NSLOT);
    const SlotStore<NSLOT, F4> SL = { Q, idarr, keyarr };
    constexpr int SORT_UNROLL = RT_EXP_SORT_UNROLL;
    /* sort keys: exit face (CB = 3 bits; none in the wide form, rt_kernel_smem.h) under the groups' bits; bit 7 of a key byte is RT_KEY_SHELL */
    constexpr bool EXIT = !SPATIAL;
    constexpr uint32_t NK = rtSortedBlkKeys(SPATIAL), CB = EXIT ? 3u : 0u;
    uint32_t *cnts = offs + SortBlkSmem<R, SPATIAL>::OFFS, *run = cnts + NK, *startOf = run + NK;
    uint32_t *ctr = startOf + NK;                      /* [0], [1] camera survivors of even / odd rounds  [2] survivors of a bounce  [3] chunk ticket */
    for (uint32_t k = threadIdx.x; k < 3 * NK + 4; k += RT_BLOCK_THREADS) cnts[k] = 0u;
    __syncthreads();

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
    uint32_t *tail = (uint32_t *)(sm + lay.words + SortBlkSmem<R, SPATIAL>::WORDS);
    /* the quads each key names.  Continuation ray (7 bits): the always-tested quads that are not shell quads, those of
     * the groups whose bit is set, and the shell quads of the key's exit face (0..5 one face, 6 none, 7 all of them);
     * shadow ray (the 4 group bits): the first two -- the shell quads are added when RT_KEY_SHELL asks for them */
    uint32_t *keyPairs = tail + (RT_BLOCK_THREADS / 32) * mwT + 1, *keyPairsS = keyPairs + NK;
    if (threadIdx.x < NK) {
        const uint32_t code = EXIT ? threadIdx.x & 7u : 7u, gb = threadIdx.x >> CB;
        uint32_t pm = grp[RT_GRP_ALWAYS] & ~shellQuads;
        for (uint32_t g = 0; g < K; g++) if ((gb >> g) & 1u) pm |= grp[4 + 8 * g + 3];
        if (!EXIT || code == 0u) keyPairsS[gb] = pm & ~grp[RT_GRP_LIGHT_PLANE];     /* (the light's own quads: rt_scene_layout.h) */
        keyPairs[threadIdx.x] = pm | (code < 6u ? grp[RT_GRP_FACES + code] : code == 6u ? 0u : shellQuads);
    }
    __syncthreads();

    /* SPATIAL: the warp takes G tiles side by side (sub-tile t = pixels x + 8t); slot j of its one round is sample
     * j % spp of sub-tile j / spp (wave.cuh) */
    const uint32_t G = SPATIAL ? max(1u, (uint32_t)R / max(p.spp, 1u)) : 1u;
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
        x0 = (blockIdx.x * RT_BLOCK_W + (warp & 1) * 8) * G; ly0 = blockIdx.y * RT_BLOCK_H + (warp >> 1) * 4;
    }
    if (SPATIAL) kEnd = G * p.spp;                 /* one round: G sub-tiles x spp samples <= R slots */
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
        const int xr = SPATIAL ? (int)min(x + 8u * (G - 1u), (uint32_t)p.width - 1u) : (int)x;      /* rightmost sub-tile */
        const int cx0 = __reduce_min_sync(0xffffffffu, owns ? (int)x : big), cx1 = __reduce_max_sync(0xffffffffu, owns ? xr : -big);
        const int cy0 = __reduce_min_sync(0xffffffffu, owns ? (int)y : big), cy1 = __reduce_max_sync(0xffffffffu, owns ? (int)y : -big);
        camera_candidates(sc, cf, lay.T, 0u, (float)cx0 - 1.0f, (float)cx1 + 2.0f, (float)cy0 - 1.0f, (float)cy1 + 2.0f,
                          cullT, mwT, cullT, 0u, lane);
    }

    const uint32_t offset = owns ? offsets[pix] : 0u;
    if (!SPATIAL) offs[threadIdx.x] = offset;
    f3 lum = mk(0.0f, 0.0f, 0.0f);
    if (SPATIAL) {
    } else if (PERSIST && chunk > 0) {
        const float4 o = tiles_wait_for_previous_chunk(ts.flags + tile, chunk, out + pix, owns, ts.status);
        lum = mk(o.x, o.y, o.z);
    } else if (p.accumulate && owns) { float4 o = out[pix]; lum = mk(o.x, o.y, o.z); }

#pragma unroll 1
    for (uint32_t k0 = kBegin; k0 < kEnd; k0 += R) {
        const uint32_t nround = min((uint32_t)R, kEnd - k0);                  /* block-uniform */
        const uint32_t ibase = p.firstSample + k0;
        /* a warp may start the next round's camera rays while another still reads this round's count: two counters */
        uint32_t *hits = ctr + (((k0 - kBegin) / R) & 1u);

        /* ---- 1. camera rays: each warp its own pixels, survivors onto the block's hit list ------------- */
        /* halton<1> of the two samples of a pass comes from the table (rt_device_math.cuh: RTHaltonTables), asked for
         * one pass ahead so that the load's latency is behind the previous pass's triangle tests */
        float jyNext[2] = { 0.0f, 0.0f };
        if (!SPATIAL) { jyNext[0] = halton_cam_y(ht, offset + ibase); jyNext[1] = halton_cam_y(ht, offset + ibase + 1u); }
#pragma unroll 1
        for (uint32_t s = 0; s < nround; s += 2) {
            const bool two = s + 1 < nround;
            bool alive[2] = { false, false };
            uint32_t id[2] = { 0u, 0u };
            f3 hp[2], r0[2];
            hp[0] = hp[1] = r0[0] = r0[1] = mk(0.0f, 0.0f, 0.0f);
            bool own[2] = { owns, owns };
            float fxq[2] = { fx, fx };
            uint32_t idx[2] = { offset + (ibase + s), offset + (ibase + s + 1u) };
            if (SPATIAL) {
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    const uint32_t j = s + q, t = j / p.spp, xt = x + 8u * t;
                    own[q] = owns && xt < (uint32_t)p.width && j < nround;
                    fxq[q] = (float)(int)xt;
                    idx[q] = (own[q] ? offsets[pix + 8u * t] : 0u) + (p.firstSample + (j - t * p.spp));
                    if (j < nround) offs[warp * CAP + j * 32u + lane] = idx[q];
                }
            }
            float jyq[2] = { jyNext[0], jyNext[1] };
            if (SPATIAL) { jyq[0] = halton_cam_y(ht, idx[0]); jyq[1] = halton_cam_y(ht, idx[1]); }
            else if (s + 2 < nround) { jyNext[0] = halton_cam_y(ht, idx[0] + 2u); jyNext[1] = halton_cam_y(ht, idx[1] + 2u); }
            if (own[0] || own[1]) {
                f3 rd[2];
#pragma unroll
                for (int q = 0; q < 2; q++) {                                  /* sampling.metal:141-147 */
                    float jx = halton<0>(idx[q]), jy = jyq[q];
                    float sx = ((fxq[q] + jx) / resX) * 2.0f - 1.0f;
                    float ty = -(((fy + jy) / resY) * 2.0f - 1.0f);
                    rd[q] = normalize((cam_u * (sx * halfW) + cam_v * (ty * halfH)) - cam_w);
                }
                bool hit[2]; float dist[2];
                closest_primary2_list(sc, cullT, mwT, cullT, 0u, rd[0], rd[1], 0.001f, 1000.0f, hit[0], dist[0], id[0], hit[1], dist[1], id[1]);
                hit[0] = hit[0] && own[0];
                hit[1] = hit[1] && two && own[1];
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    if (hit[q]) {
                        const float *sh = sm + lay.shade + id[q];
                        if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) r0[q] = mk(sh[(RT_SH_ER + 0) * lay.Tp], sh[(RT_SH_ER + 1) * lay.Tp], sh[(RT_SH_ER + 2) * lay.Tp]);
                        else { alive[q] = true; hp[q] = cam_o + rd[q] * dist[q]; }
                    }
                }
            }
            const uint32_t m0 = __ballot_sync(0xffffffffu, alive[0]), m1 = __ballot_sync(0xffffffffu, alive[1]);
            uint32_t base = 0u;
            if (lane == 0 && (m0 | m1)) base = atomicAdd(hits, (uint32_t)(__popc(m0) + __popc(m1)));
            base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
            for (int q = 0; q < 2; q++) {
                if (q == 1 && !two) break;
                const uint32_t slot = warp * CAP + (s + q) * 32u + lane;
                if (alive[q]) {
                    const uint32_t e = base + (q ? __popc(m0) : 0) + __popc((q ? m1 : m0) & lt);
                    SL.put_hit_first(slot, hp[q], id[q]);
                    hitlist[e] = (uint16_t)slot;
                }
                res[0 * NSLOT + slot] = r0[q].x; res[1 * NSLOT + slot] = r0[q].y; res[2 * NSLOT + slot] = r0[q].z;
            }
        }
        __syncthreads();
        uint32_t count = *hits;                                                /* hit-list length, block-uniform */

        /* ---- 2. bounces -------------------------------------------------------------------------------- */
#pragma unroll 1
        for (uint32_t b = 0; b < nb && count; b++) {
            const bool cont = b + 1 < nb;
            const bool traceC = MIS || cont;        /* MIS (mis.cuh): the continuation ray is traced after the last bounce too */
            const bool sort = count > 32u;          /* one chunk: nothing to sort, the hit list is the order */

            /* -- A. shade set-up, the two rays of the bounce, the key: chunk c of the hit list by warp c % 4 -- */
            /* (the entry's slot, Halton index and the bounce's four Halton values -- one 16-byte load from the table,
             * rt_device_math.cuh -- are fetched one chunk ahead) */
            uint32_t slotN = 0u, iN = 0u;
            float4 h4N = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (warp * 32u + lane < count) {
                slotN = hitlist[warp * 32u + lane];
                const uint32_t ow = slotN / CAP, os = slotN - ow * CAP;        /* owner warp, its slot */
                iN = SPATIAL ? offs[slotN] : offs[ow * 32u + (os & 31u)] + (ibase + (os >> 5));
                h4N = halton_bounce4(ht, iN, b);                               /* raytrace.metal:72-73, :93-94 */
            }
#pragma unroll 1
            for (uint32_t c0 = warp * 32u; c0 < count; c0 += RT_BLOCK_THREADS) {
                const uint32_t e = c0 + lane;
                const bool active = e < count;
                const uint32_t slot = slotN;
                const float4 h4 = h4N;
                if (e + RT_BLOCK_THREADS < count) {
                    slotN = hitlist[e + RT_BLOCK_THREADS];
                    const uint32_t ow = slotN / CAP, os = slotN - ow * CAP;
                    iN = SPATIAL ? offs[slotN] : offs[ow * 32u + (os & 31u)] + (ibase + (os >> 5));
                    h4N = halton_bounce4(ht, iN, b);
                }
                uint32_t key = 0xffu;
                if (active) {
                    uint32_t id; f3 hp, color;
                    SL.get_shade(slot, hp, color, id);
                    const float *sh = sm + lay.shade + id;
                    const f3 normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
                    const f3 diffuse = mk(sh[(RT_SH_DR + 0) * lay.Tp], sh[(RT_SH_DR + 1) * lay.Tp], sh[(RT_SH_DR + 2) * lay.Tp]);
                    const f3 P = hp + normal * 1e-3f;                                           /* raytrace.metal:67 */
                    const float wx = h4.x, wy = h4.y;                                           /* :72-73 */
                    color = color * diffuse;                                                    /* :76 */
                    f3 rd = mk(0.0f, 0.0f, 0.0f);
                    if (traceC) {
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
                    const f3 ld = sp - P;
                    key = 0u;
                    for (uint32_t g = 0; g < K; g++) {
                        const float *G = sm + lay.group + 4 + 8 * g;
                        const f3 w = mk(G[0], G[1], G[2]) - P, H = mk(G[4], G[5], G[6]);
                        bool clear = ray_clear_of_box(w, H, ld);
                        if (traceC) clear = clear && ray_clear_of_box(w, H, rd);
                        if (!clear) key |= 1u << g;
                    }
                    /* which face can the continuation ray leave the shell box through (the exit rule), and can the shadow ray
                     * skip the shell quads?  (rt_scene_layout.h; sorted.cuh: exit_face_code, shadow_inside_shell) */
                    if (EXIT) key = (key << 3) | (traceC ? exit_face_code(shellBox, P, rd) : 6u);
                    SL.put_rays(slot, P, rd, traceC, color, sp.x, sp.z, id, key | (shadow_inside_shell(shellBox, P, sp) ? 0u : RT_KEY_SHELL));
                }
                if (sort) {
                    const uint32_t same = __match_any_sync(0xffffffffu, key);
                    if (RT_EXP_RANK_IN_A) {
                        /* the entry's rank within its key comes back from the counter it is counted in (kept in the sorted
                         * list's own storage until the sort turns rank into place) */
                        const uint32_t leader = (uint32_t)__ffs((int)same) - 1u;
                        uint32_t at = 0u;
                        if (active && lane == leader) at = atomicAdd(&cnts[key], (uint32_t)__popc(same));
                        at = __shfl_sync(0xffffffffu, at, leader);
                        if (active) sorted[e] = (uint16_t)(at + __popc(same & lt));
                    } else {
                        if (active && lane == (uint32_t)__ffs((int)same) - 1u) atomicAdd(&cnts[key], (uint32_t)__popc(same));
                    }
                }
            }
            __syncthreads();
            if (threadIdx.x == 0) { *hits = 0u; ctr[2] = 0u; ctr[3] = 0u; }   /* every thread has read `count` */

            /* -- S. counting sort of the slot numbers, keys descending -- */
            const uint16_t *order = hitlist;
            if (sort) {
                /* where each key starts, keys descending: every warp works the same table out and writes it (same values) */
                {
                    constexpr uint32_t PER = (NK + 31u) / 32u;
                    uint32_t c[PER], sum = 0u;
#pragma unroll
                    for (uint32_t q = 0; q < PER; q++) { c[q] = lane * PER + q < NK ? cnts[NK - 1u - (lane * PER + q)] : 0u; sum += c[q]; }
                    uint32_t incl = sum;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= (uint32_t)d) incl += up;
                    }
                    uint32_t at = incl - sum;
#pragma unroll
                    for (uint32_t q = 0; q < PER; q++) { if (lane * PER + q < NK) startOf[NK - 1u - (lane * PER + q)] = at; at += c[q]; }
                    __syncwarp();
                }
                if (RT_EXP_RANK_IN_A) {
                    /* every entry's place = start of its key + its rank: all read (a thread has at most NSLOT / 128 entries),
                     * then, behind a barrier, all written -- ranks and places share the sorted list's storage */
                    constexpr uint32_t PT = NSLOT / RT_BLOCK_THREADS;
                    uint32_t place[PT];
#pragma unroll
                    for (uint32_t q = 0; q < PT; q++) {
                        const uint32_t e = warp * 32u + lane + q * RT_BLOCK_THREADS;
                        place[q] = 0xffffffffu;
                        if (e < count) {
                            const uint32_t slot = hitlist[e];
                            place[q] = ((startOf[SL.key(slot) & (NK - 1u)] + sorted[e]) << 16) | slot;
                        }
                    }
                    __syncthreads();
#pragma unroll
                    for (uint32_t q = 0; q < PT; q++)
                        if (place[q] != 0xffffffffu) sorted[place[q] >> 16] = (uint16_t)(place[q] & 0xffffu);
                } else
#pragma unroll SORT_UNROLL
                for (uint32_t c0 = warp * 32u; c0 < count; c0 += RT_BLOCK_THREADS) {
                    const uint32_t e = c0 + lane;
                    const bool active = e < count;
                    const uint32_t slot = active ? hitlist[e] : 0u;
                    const uint32_t key = active ? (SL.key(slot) & (NK - 1u)) : 0xffffu;
                    const uint32_t same = __match_any_sync(0xffffffffu, key);
                    const uint32_t leader = (uint32_t)__ffs((int)same) - 1u;
                    uint32_t at = 0u;
                    if (active && lane == leader) at = atomicAdd(&run[key], (uint32_t)__popc(same));
                    at = __shfl_sync(0xffffffffu, at, leader) + (active ? startOf[key] : 0u);
                    if (active) sorted[at + __popc(same & lt)] = (uint16_t)slot;
                }
                order = sorted;
            }
            __syncthreads();
            if (sort) for (uint32_t k = threadIdx.x; k < 2 * NK; k += RT_BLOCK_THREADS) cnts[k] = 0u;   /* cnts and run, for the next bounce */

            /* -- B. trace: chunks of the sorted list, a ticket each -- */
            /* (the next chunk's ticket is drawn before this chunk is traced: RT_EXP_TICKET_AHEAD) */
            uint32_t cNext = 0u;
            if (RT_EXP_TICKET_AHEAD) {
                if (lane == 0) cNext = atomicAdd(&ctr[3], 1u) * 32u;
                cNext = __shfl_sync(0xffffffffu, cNext, 0);
            }
#pragma unroll 1
            for (;;) {
                uint32_t c0 = 0u;
                if (RT_EXP_TICKET_AHEAD) {
                    c0 = cNext;
                    if (c0 >= count) break;
                    if (lane == 0) cNext = atomicAdd(&ctr[3], 1u) * 32u;
                } else {
                    if (lane == 0) c0 = atomicAdd(&ctr[3], 1u) * 32u;
                    c0 = __shfl_sync(0xffffffffu, c0, 0);
                    if (c0 >= count) break;
                }
                const uint32_t e = c0 + lane;
                const bool active = e < count;
                const uint32_t slot = active ? order[e] : 0u;
                uint32_t key = 0u, surf = 0u;
                f3 colorHead = mk(0.0f, 0.0f, 0.0f);
                if (active) SL.get_trace_head(slot, colorHead, surf, key);
                __syncwarp();                                                  /* (unsorted single chunk: its entries are read before the list is rewritten) */
                const uint32_t keys = __reduce_or_sync(0xffffffffu, key);
                /* the shadow rays skip the shell quads unless an entry of the chunk starts or ends outside the shell box;
                 * the continuation rays take the union of the entries' lists (exit faces do not OR as the groups' bits do) */
                const uint32_t pairsS = keyPairsS[(keys >> CB) & (RT_SORT_KEYS - 1u)] | ((keys & RT_KEY_SHELL) ? shellQuads : 0u);
                const uint32_t pairs = __reduce_or_sync(0xffffffffu, active ? keyPairs[key & (NK - 1u)] : 0u) | pairsS;
                bool alive = false;
                uint32_t id = 0;
                if (active) {
                    f3 P, rd; float spx, spz;
                    SL.get_trace(slot, traceC, P, spx, spz, rd);
                    const float *L = sm + lay.light;
                    f3 ldir; float ldist, inv;
                    light_direction(mk(spx, L[1], spz), P, ldir, ldist, inv);                   /* sampling.metal:214-220 */
                    const float ltmax = ldist - 1e-3f;                                          /* raytrace.metal:81 */
                    bool hitC = false;
                    float dist = 0.0f;
#if RT_EXP_TRACE == 0
                    if (traceC) {
                        hitC = closest_hit_pairs(sc, pairs, P, rd, 0.001f, 1000.0f, dist, id);   /* :49 */
                    }
                    const bool shadowed = any_hit_pairs(sc, pairsS, P, ldir, 0.0f, ltmax);       /* :79-88 */
#elif RT_EXP_TRACE == 1
                    bool shadowed;
                    if (traceC) {
                        hitC = dual_hit_pairs_masked(sc, pairs, pairsS, P, rd, 0.001f, 1000.0f, dist, id, ldir, 0.0f, ltmax, shadowed);
                    } else {
                        shadowed = any_hit_pairs(sc, pairsS, P, ldir, 0.0f, ltmax);
                    }
#else
                    bool shadowed;
                    if (traceC) {
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
                        if (MIS) lightColor = lightColor * misA_light_weight(L, P, normal, ldir);
                        const f3 lc = lightColor * SL.get_color(slot, colorHead);
                        res[0 * NSLOT + slot] = res[0 * NSLOT + slot] + lc.x;
                        res[1 * NSLOT + slot] = res[1 * NSLOT + slot] + lc.y;
                        res[2 * NSLOT + slot] = res[2 * NSLOT + slot] + lc.z;
                    }
                    if (MIS) {      /* the cosine strategy's term for this continuation ray (one light on this path) */
                        const float *sh = sm + lay.shade + surf;
                        const f3 normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
                        const f3 term = misA_cosine_term(L, P, normal, rd, hitC, dist, SL.get_color(slot, colorHead));
                        res[0 * NSLOT + slot] = res[0 * NSLOT + slot] + term.x;
                        res[1 * NSLOT + slot] = res[1 * NSLOT + slot] + term.y;
                        res[2 * NSLOT + slot] = res[2 * NSLOT + slot] + term.z;
                    }
                    if (hitC) {
                        const float *sh = sm + lay.shade + id;
                        if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) {                                  /* :59; MIS: the path ends, nothing is assigned */
                            if (!MIS) {
                                res[0 * NSLOT + slot] = sh[(RT_SH_ER + 0) * lay.Tp]; res[1 * NSLOT + slot] = sh[(RT_SH_ER + 1) * lay.Tp];
                                res[2 * NSLOT + slot] = sh[(RT_SH_ER + 2) * lay.Tp];
                            }
                        } else if (cont) {
                            alive = true;
                            SL.put_hit_next(slot, P + rd * dist, id);
                        }
                    }
                }
                const uint32_t mask = __ballot_sync(0xffffffffu, alive);
                uint32_t base = 0u;
                if (lane == 0 && mask) base = atomicAdd(&ctr[2], (uint32_t)__popc(mask));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (alive) {
                    hitlist[base + __popc(mask & lt)] = (uint16_t)slot;
                }
                if (RT_EXP_TICKET_AHEAD) cNext = __shfl_sync(0xffffffffu, cNext, 0);
            }
            __syncthreads();
            count = ctr[2];
        }
        /* (count == 0 when the loop ends -- the last bounce lists no survivors -- so ctr[2] is 0 for the next round) */

        /* ---- 3. in-order accumulation, raytrace.metal:103 ---------------------------------- */
        if (SPATIAL) {
            for (uint32_t t = 0; t < G; t++) {
                const uint32_t xt = x + 8u * t;
                if (!(owns && xt < (uint32_t)p.width)) continue;
                const size_t pt = pix + 8u * t;
                f3 l = mk(0.0f, 0.0f, 0.0f);
                if (p.accumulate) { float4 o = out[pt]; l = mk(o.x, o.y, o.z); }
                for (uint32_t s = t * p.spp; s < (t + 1u) * p.spp; s++) {
                    const uint32_t slot = warp * CAP + s * 32u + lane;
                    l = l + mk(res[0 * NSLOT + slot], res[1 * NSLOT + slot], res[2 * NSLOT + slot]);
                }
                float4 o;
                if (p.accumulate) o = make_float4(l.x, l.y, l.z, (float)(p.firstSample + p.spp));
                else { float samples = (float)p.spp; o = make_float4(l.x / samples, l.y / samples, l.z / samples, 1.0f); }
                out[pt] = o;
            }
        } else {
            for (uint32_t s = 0; s < nround; s++) {
                const uint32_t slot = warp * CAP + s * 32u + lane;
                lum = lum + mk(res[0 * NSLOT + slot], res[1 * NSLOT + slot], res[2 * NSLOT + slot]);
            }
        }
    }

    if (SPATIAL) {
        /* stored per sub-tile above */
    } else if (PERSIST && kEnd < p.spp) {
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
}
