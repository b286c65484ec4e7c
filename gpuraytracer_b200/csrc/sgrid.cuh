/*
 * sgrid.cuh -- a uniform grid over the SPHERES of a scene (config 4 of BASELINE.json: 256 spheres + 4 area lights;
 * included by pathtrace.cu before the intersector).
 *
 * The reference has no sphere primitives in its live kernels (the semantics come from the dead
 * shaders_old.metal:108-136, 172-200 -- a loop over all spheres); the brute-force loops of pathtrace.cu restate that and
 * stay the definition.  With 256 spheres they are two thirds of the kernel's time: every ray tests every sphere.  Here a
 * ray walks the cells of an 8 x 8 x 8 grid it passes through (3D-DDA, one lane = one ray) and tests the spheres listed
 * there: a handful instead of 256.
 *
 * Why the image cannot change.  A sphere is listed in every cell its bounding cube, grown by 1/64 of a cell, overlaps; the
 * grid's bounds are the spheres' bounds grown by the same margin.  A hit point lies on its sphere, hence inside a cell
 * that lists the sphere, and the walk visits the cells in the order the ray passes through them; rounding can make it
 * cut a corner by a few ulps, never by 1/64 of a cell.  So every sphere the brute-force loop could accept is tested,
 * with the SAME arithmetic on the SAME operands (m, C, B, disc, the nearer root first).  The spheres are met in another
 * order, so the closest hit is chosen by the order-independent form of the loop's rule, as in bvh.cuh: smaller
 * cross-multiplied distance, and on exact equality the lower primitive id (a sphere met twice changes nothing).
 * The walk stops once the next cell starts beyond the best hit (closest) or beyond the ray's end (shadow).
 * A scene whose cells overflow (more than RT_SGRID_CELL spheres in one) is flagged and traced by brute force.
 * The per-pixel kernel (pathTrace_smem) never uses the grid: it stays the independent second implementation the tests
 * compare every other form with.
 */
#pragma once

/* ---- build: block 0 of rt_prep_scene, all its threads ---- */
__device__ void prep_sphere_grid(const unsigned char *__restrict__ spheres, float *__restrict__ scene, const RTSceneLayout lay,
                                 int tid, int nth) {
    __shared__ float s_red[6][8];
    __shared__ uint32_t s_cnt[RT_SGRID_N * RT_SGRID_N * RT_SGRID_N];
    __shared__ uint32_t s_over;
    float *hdr = scene + lay.sgrid;
    uint32_t *cells = (uint32_t *)hdr + RT_SGRID_HDR;
    const float INF = __int_as_float(0x7f800000);
    float lo[3] = { INF, INF, INF }, hi[3] = { -INF, -INF, -INF };
    for (uint32_t j = tid; j < lay.S; j += nth) {
        const float *s = (const float *)(spheres + 80 * (size_t)j);
        const float r = fabsf(s[16]);
#pragma unroll
        for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], s[k] - r); hi[k] = fmaxf(hi[k], s[k] + r); }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int sft = 16; sft; sft >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], sft));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], sft));
        }
        if ((tid & 31) == 0) { s_red[k][tid >> 5] = lo[k]; s_red[3 + k][tid >> 5] = hi[k]; }
    }
    for (int c = tid; c < RT_SGRID_N * RT_SGRID_N * RT_SGRID_N; c += nth) s_cnt[c] = 0u;
    for (int c = tid; c < RT_SGRID_N * RT_SGRID_N * RT_SGRID_N * (RT_SGRID_CELL / 2); c += nth) cells[c] = 0xffffffffu;
    if (tid == 0) s_over = 0u;
    __syncthreads();
    float cell[3], inv[3];
    bool ok = lay.S > 0u;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float l = INF, h = -INF;
        for (int w = 0; w < (nth >> 5); w++) { l = fminf(l, s_red[k][w]); h = fmaxf(h, s_red[3 + k][w]); }
        /* grown by 1/64 of a cell on every side: cell = ext / N, margin = cell / 64  =>  ext' = ext (1 + 2 / (64 N)) */
        const float margin = (h - l) * (1.0f / (64.0f * RT_SGRID_N));
        lo[k] = l - margin; hi[k] = h + margin;
        cell[k] = (hi[k] - lo[k]) * (1.0f / RT_SGRID_N);
        ok = ok && cell[k] > 0.0f && cell[k] < INF;
        inv[k] = ok ? 1.0f / cell[k] : 0.0f;
    }
    if (ok) {
        for (uint32_t j = tid; j < lay.S; j += nth) {
            const float *s = (const float *)(spheres + 80 * (size_t)j);
            const float r = fabsf(s[16]);
            int a[3], b[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const float grow = r + cell[k] * (1.0f / 64.0f);
                a[k] = min(max((int)floorf((s[k] - grow - lo[k]) * inv[k]), 0), RT_SGRID_N - 1);
                b[k] = min(max((int)floorf((s[k] + grow - lo[k]) * inv[k]), 0), RT_SGRID_N - 1);
            }
            for (int z = a[2]; z <= b[2]; z++)
                for (int y = a[1]; y <= b[1]; y++)
                    for (int x = a[0]; x <= b[0]; x++) {
                        const int c = (z * RT_SGRID_N + y) * RT_SGRID_N + x;
                        const uint32_t slot = atomicAdd(&s_cnt[c], 1u);
                        if (slot < RT_SGRID_CELL) ((unsigned short *)cells)[c * RT_SGRID_CELL + slot] = (unsigned short)j;
                        else s_over = 1u;
                    }
        }
    }
    __syncthreads();
    if (tid == 0) {
        hdr[0] = lo[0]; hdr[1] = lo[1]; hdr[2] = lo[2]; ((uint32_t *)hdr)[3] = (ok && !s_over) ? 1u : 0u;
        hdr[4] = cell[0]; hdr[5] = cell[1]; hdr[6] = cell[2]; hdr[7] = 0.0f;
        hdr[8] = inv[0]; hdr[9] = inv[1]; hdr[10] = inv[2]; hdr[11] = 0.0f;
        hdr[12] = hdr[13] = hdr[14] = hdr[15] = 0.0f;
    }
}

/* ---- walk ---- */
struct SGridWalk {
    int ix, iy, iz, sx, sy, sz;
    float tx, ty, tz, dx, dy, dz, tExit;
};

/* sets up the walk of the ray (o, d); false when the ray misses the grid's box.  hdr: the grid's header in global memory
 * (read through the read-only path: every lane reads the same words) */
__device__ __forceinline__ bool sgrid_begin(const float *__restrict__ hdr, f3 o, f3 d, SGridWalk &w) {
    const f3 lo = mk(__ldg(hdr), __ldg(hdr + 1), __ldg(hdr + 2));
    const f3 cell = mk(__ldg(hdr + 4), __ldg(hdr + 5), __ldg(hdr + 6));
    const f3 inv = mk(__ldg(hdr + 8), __ldg(hdr + 9), __ldg(hdr + 10));
    const float BIG = 1e30f;
    const f3 id = mk(fabsf(d.x) > 1e-30f ? 1.0f / d.x : copysignf(BIG, d.x), fabsf(d.y) > 1e-30f ? 1.0f / d.y : copysignf(BIG, d.y),
                     fabsf(d.z) > 1e-30f ? 1.0f / d.z : copysignf(BIG, d.z));
    const f3 hi = mk(lo.x + cell.x * RT_SGRID_N, lo.y + cell.y * RT_SGRID_N, lo.z + cell.z * RT_SGRID_N);
    const float x0 = (lo.x - o.x) * id.x, x1 = (hi.x - o.x) * id.x;
    const float y0 = (lo.y - o.y) * id.y, y1 = (hi.y - o.y) * id.y;
    const float z0 = (lo.z - o.z) * id.z, z1 = (hi.z - o.z) * id.z;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
    if (tn > tf) return false;                                           /* (a NaN goes on: the index clamps below keep it inside) */
    const f3 p = o + d * tn;
    w.ix = min(max((int)floorf((p.x - lo.x) * inv.x), 0), RT_SGRID_N - 1);
    w.iy = min(max((int)floorf((p.y - lo.y) * inv.y), 0), RT_SGRID_N - 1);
    w.iz = min(max((int)floorf((p.z - lo.z) * inv.z), 0), RT_SGRID_N - 1);
    w.sx = d.x >= 0.0f ? 1 : -1; w.sy = d.y >= 0.0f ? 1 : -1; w.sz = d.z >= 0.0f ? 1 : -1;
    /* parameter at which the ray leaves the current cell along each axis; an axis the ray does not move along never steps */
    w.tx = fabsf(d.x) > 1e-30f ? (lo.x + cell.x * (float)(w.ix + (w.sx > 0 ? 1 : 0)) - o.x) * id.x : BIG;
    w.ty = fabsf(d.y) > 1e-30f ? (lo.y + cell.y * (float)(w.iy + (w.sy > 0 ? 1 : 0)) - o.y) * id.y : BIG;
    w.tz = fabsf(d.z) > 1e-30f ? (lo.z + cell.z * (float)(w.iz + (w.sz > 0 ? 1 : 0)) - o.z) * id.z : BIG;
    w.dx = cell.x * fabsf(id.x); w.dy = cell.y * fabsf(id.y); w.dz = cell.z * fabsf(id.z);
    w.tExit = tf;
    return true;
}
/* the cell's list (eight 16-bit sphere numbers) */
__device__ __forceinline__ uint4 sgrid_cell(const float *__restrict__ hdr, const SGridWalk &w) {
    const uint4 *cells = (const uint4 *)(hdr + RT_SGRID_HDR);
    return __ldg(cells + ((w.iz * RT_SGRID_N + w.iy) * RT_SGRID_N + w.ix));
}
/* steps to the next cell; returns the parameter at which the ray enters it, or -1 when it left the grid.  A ray that starts
 * ON a cell boundary can get a first parameter a few ulps below zero (the index came from a multiplication by 1/cell, the
 * boundary from a multiplication by cell): it is clamped, -1 stays the only negative value. */
__device__ __forceinline__ float sgrid_step(SGridWalk &w) {
    /* written with selects, not branches: the lanes of a warp step along different axes */
    const bool ax = (w.tx <= w.ty) & (w.tx <= w.tz), ay = !ax & (w.ty <= w.tz), az = !ax & !ay;
    const float t = ax ? w.tx : ay ? w.ty : w.tz;
    w.ix += ax ? w.sx : 0; w.iy += ay ? w.sy : 0; w.iz += az ? w.sz : 0;
    w.tx += ax ? w.dx : 0.0f; w.ty += ay ? w.dy : 0.0f; w.tz += az ? w.dz : 0.0f;
    /* every index is in -1 .. N: one of them outside 0 .. N-1 <=> a bit above the low three is set in their union */
    const bool outside = (unsigned)(w.ix | w.iy | w.iz) >= (unsigned)RT_SGRID_N;
    return (outside | (t > w.tExit)) ? -1.0f : fmaxf(t, 0.0f);
}
/* the cell's list as a queue: head() is its first sphere (0xffff: the list is over), pop() moves on */
struct SGridList {
    uint32_t a, b, c, d;
    __device__ __forceinline__ SGridList(uint4 v) : a(v.x), b(v.y), c(v.z), d(v.w) {}
    __device__ __forceinline__ uint32_t head() const { return a & 0xffffu; }
    __device__ __forceinline__ void pop() {
        a = __funnelshift_r(a, b, 16); b = __funnelshift_r(b, c, 16); c = __funnelshift_r(c, d, 16); d = (d >> 16) | 0xffff0000u;
    }
};

/* sph: the sphere records in shared memory (float4 pairs: centre.xyz radius | prepared camera part); Treal: the number of
 * triangles (sphere j has primitive id Treal + j).  Continues the closest-hit state (bestN, bestD, best) of the triangles. */
__device__ __noinline__ void sgrid_closest(const float *__restrict__ hdr, const float4 *sph, uint32_t Treal, f3 o, f3 d, float tmin,
                                           float &bestN_io, float &bestD_io, uint32_t &best_io) {
    SGridWalk w;
    if (!sgrid_begin(hdr, o, d, w)) return;
    float bestN = bestN_io, bestD = bestD_io;
    uint32_t best = best_io;
    const float A = dot(d, d), lim = tmin * A;
    /* the NEXT cell's list is asked for before this cell's spheres are tested, so that its way from the cache overlaps them */
    SGridList ids(sgrid_cell(hdr, w));
#pragma unroll 1
    for (int it = 0; it < 3 * RT_SGRID_N; it++) {
        const float t = sgrid_step(w);
        uint4 nxt = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
        if (t >= 0.0f) nxt = sgrid_cell(hdr, w);
#pragma unroll 1
        for (uint32_t j = ids.head(); j != 0xffffu; j = ids.head()) {     /* (rolled: the unrolled eight were 5 KB a walk) */
            ids.pop();
            const float4 R = sph[2 * j];
            const f3 m = o - mk(R.x, R.y, R.z);
            const float C = fmaf(-R.w, R.w, dot(m, m));
            const float B = dot(m, d);
            const float disc = fmaf(B, B, -(A * C));
            if (disc >= 0.0f) {
                const float sq = sqrtf(disc);
                const float N0 = -B - sq, N1 = -B + sq;
                const float N = N0 >= lim ? N0 : N1;                     /* the nearer root first (rt_detmath.h: rt_sphere_test) */
                if (N >= lim) {
                    const float p1 = N * bestD, p2 = bestN * A;
                    const uint32_t id = Treal + j;
                    if ((p1 < p2) | ((p1 == p2) & (id < best) & (best != 0xffffffffu))) { bestN = N; bestD = A; best = id; }
                }
            }
        }
        if (t < 0.0f) break;
        if (t * bestD > bestN * 1.0001f) break;                          /* the rest of the walk lies behind the best hit */
        ids = SGridList(nxt);
    }
    bestN_io = bestN; bestD_io = bestD; best_io = best;
}

/* any hit in [0, tmax] (the shadow ray's query; tmin = 0) */
__device__ __noinline__ bool sgrid_any(const float *__restrict__ hdr, const float4 *sph, f3 o, f3 d, float tmax) {
    SGridWalk w;
    if (!sgrid_begin(hdr, o, d, w)) return false;
    const float A = dot(d, d), far = tmax * A;
    SGridList ids(sgrid_cell(hdr, w));
#pragma unroll 1
    for (int it = 0; it < 3 * RT_SGRID_N; it++) {
        const float t = sgrid_step(w);
        const bool more = (t >= 0.0f) & (t <= tmax);
        uint4 nxt = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
        if (more) nxt = sgrid_cell(hdr, w);
#pragma unroll 1
        for (uint32_t j = ids.head(); j != 0xffffu; j = ids.head()) {
            ids.pop();
            const float4 R = sph[2 * j];
            const f3 m = o - mk(R.x, R.y, R.z);
            const float C = fmaf(-R.w, R.w, dot(m, m));
            const float B = dot(m, d);
            const float disc = fmaf(B, B, -(A * C));
            if (disc >= 0.0f) {
                const float sq = sqrtf(disc);
                const float N0 = -B - sq, N1 = -B + sq;
                /* sphere_roots with tmin = 0, bestN = tmax, bestD = 1 */
                if (((N0 >= 0.0f * A) & (N0 * 1.0f < far)) | ((N1 >= 0.0f * A) & (N1 * 1.0f < far))) return true;
            }
        }
        if (!more) break;
        ids = SGridList(nxt);
    }
    return false;
}
