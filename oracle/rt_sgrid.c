/*
 * rt_sgrid.c -- TEST INFRASTRUCTURE.  CPU restatement of the sphere grid of gpuraytracer_b200/csrc/sgrid.cuh (build and
 * 3D-DDA walk, the same fp32 operations in the same order) and the check that it cannot change a result:
 *
 *   rt_oracle_sgridcheck   traces a frame with the oracle and, for EVERY ray query of it, repeats the sphere part of the
 *                          query by walking the grid: hit / miss, primitive id and distance bits must be the full
 *                          loop's (rt_detmath.h: rt_intersect).  Also counts cells visited and sphere tests made.
 *
 * The reference has no counterpart: its live kernels have no sphere primitives, and nothing here is part
 * of the arithmetic contract -- the grid only says which spheres a ray need not be tested against.
 */
#include "rt_oracle.h"
#include "rt_detmath.h"
#include <string.h>
#include <stdlib.h>

#define GN 8
#define GCELL 8

extern void (*rt_oracle_query_hook)(const float o[3], const float d[3], float tmin, float tmax, int any,
                                    int hit, float distance, uint32_t primitive_id);

static struct {
    float lo[3], cell[3], inv[3];
    int ok;
    uint16_t cells[GN * GN * GN][GCELL];
    const float *vertices; uint32_t T;
    const float *spheres; uint32_t S;
    uint64_t queries, mismatches, cellsVisited, sphereTests, firstBad[8];
} g;

static void build(void) {
    float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (uint32_t j = 0; j < g.S; j++) {
        const float *s = g.spheres + (size_t)j * 20;
        const float r = fabsf(s[16]);
        for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], s[k] - r); hi[k] = fmaxf(hi[k], s[k] + r); }
    }
    g.ok = g.S > 0;
    for (int k = 0; k < 3; k++) {
        const float margin = (hi[k] - lo[k]) * (1.0f / (64.0f * GN));
        g.lo[k] = lo[k] - margin;
        const float h = hi[k] + margin;
        g.cell[k] = (h - g.lo[k]) * (1.0f / GN);
        g.ok = g.ok && g.cell[k] > 0.0f && g.cell[k] < INFINITY;
        g.inv[k] = g.ok ? 1.0f / g.cell[k] : 0.0f;
    }
    memset(g.cells, 0xff, sizeof g.cells);
    static int cnt[GN * GN * GN];
    memset(cnt, 0, sizeof cnt);
    if (!g.ok) return;
    for (uint32_t j = 0; j < g.S; j++) {
        const float *s = g.spheres + (size_t)j * 20;
        const float r = fabsf(s[16]);
        int a[3], b[3];
        for (int k = 0; k < 3; k++) {
            const float grow = r + g.cell[k] * (1.0f / 64.0f);
            int x = (int)floorf((s[k] - grow - g.lo[k]) * g.inv[k]), y = (int)floorf((s[k] + grow - g.lo[k]) * g.inv[k]);
            a[k] = x < 0 ? 0 : x > GN - 1 ? GN - 1 : x;
            b[k] = y < 0 ? 0 : y > GN - 1 ? GN - 1 : y;
        }
        for (int z = a[2]; z <= b[2]; z++)
            for (int y = a[1]; y <= b[1]; y++)
                for (int x = a[0]; x <= b[0]; x++) {
                    const int c = (z * GN + y) * GN + x;
                    if (cnt[c] < GCELL) g.cells[c][cnt[c]++] = (uint16_t)j; else g.ok = 0;
                }
    }
}

typedef struct { int i[3], s[3]; float t[3], dt[3], tExit; } walk_t;

static int walk_begin(const float o[3], const float d[3], walk_t *w) {
    const float BIG = 1e30f;
    float id[3], t0[3], t1[3];
    for (int k = 0; k < 3; k++) {
        id[k] = fabsf(d[k]) > 1e-30f ? 1.0f / d[k] : copysignf(BIG, d[k]);
        const float hi = g.lo[k] + g.cell[k] * GN;
        t0[k] = (g.lo[k] - o[k]) * id[k]; t1[k] = (hi - o[k]) * id[k];
    }
    const float tn = fmaxf(fmaxf(fminf(t0[0], t1[0]), fminf(t0[1], t1[1])), fmaxf(fminf(t0[2], t1[2]), 0.0f));
    const float tf = fminf(fminf(fmaxf(t0[0], t1[0]), fmaxf(t0[1], t1[1])), fmaxf(t0[2], t1[2]));
    if (tn > tf) return 0;
    for (int k = 0; k < 3; k++) {
        const float p = o[k] + d[k] * tn;
        int x = (int)floorf((p - g.lo[k]) * g.inv[k]);
        w->i[k] = x < 0 ? 0 : x > GN - 1 ? GN - 1 : x;
        w->s[k] = d[k] >= 0.0f ? 1 : -1;
        w->t[k] = fabsf(d[k]) > 1e-30f ? (g.lo[k] + g.cell[k] * (float)(w->i[k] + (w->s[k] > 0 ? 1 : 0)) - o[k]) * id[k] : BIG;
        w->dt[k] = g.cell[k] * fabsf(id[k]);
    }
    w->tExit = tf;
    return 1;
}
static float walk_step(walk_t *w) {
    int k = (w->t[0] <= w->t[1] && w->t[0] <= w->t[2]) ? 0 : (w->t[1] <= w->t[2]) ? 1 : 2;
    const float t = w->t[k];
    w->i[k] += w->s[k]; w->t[k] += w->dt[k];
    if ((unsigned)w->i[k] >= GN) return -1.0f;
    return t > w->tExit ? -1.0f : fmaxf(t, 0.0f);
}

/* the sphere part of a closest-hit query, continuing (bestN, bestD, best) of the triangles */
static void grid_closest(const float o[3], const float d[3], float tmin, float *bestN, float *bestD, uint32_t *best,
                         uint64_t *cells, uint64_t *tests) {
    walk_t w;
    if (!walk_begin(o, d, &w)) return;
    const float A = rt_dot3(d[0], d[1], d[2], d[0], d[1], d[2]), lim = tmin * A;
    for (int it = 0; it < 3 * GN; it++) {
        const uint16_t *ids = g.cells[(w.i[2] * GN + w.i[1]) * GN + w.i[0]];
        (*cells)++;
        for (int e = 0; e < GCELL && ids[e] != 0xffff; e++) {
            const float *s = g.spheres + (size_t)ids[e] * 20;
            (*tests)++;
            const float mx = o[0] - s[0], my = o[1] - s[1], mz = o[2] - s[2];
            const float C = fmaf(-s[16], s[16], rt_dot3(mx, my, mz, mx, my, mz));
            const float B = rt_dot3(mx, my, mz, d[0], d[1], d[2]);
            const float disc = fmaf(B, B, -(A * C));
            if (disc >= 0.0f) {
                const float sq = sqrtf(disc), N0 = -B - sq, N1 = -B + sq;
                const float N = N0 >= lim ? N0 : N1;
                if (N >= lim) {
                    const float p1 = N * *bestD, p2 = *bestN * A;
                    const uint32_t id = g.T + ids[e];
                    if (p1 < p2 || (p1 == p2 && id < *best && *best != 0xffffffffu)) { *bestN = N; *bestD = A; *best = id; }
                }
            }
        }
        const float t = walk_step(&w);
        if (t < 0.0f) break;
        if (t * *bestD > *bestN * 1.0001f) break;
    }
}
static int grid_any(const float o[3], const float d[3], float tmax, uint64_t *cells, uint64_t *tests) {
    walk_t w;
    if (!walk_begin(o, d, &w)) return 0;
    const float A = rt_dot3(d[0], d[1], d[2], d[0], d[1], d[2]), far = tmax * A;
    for (int it = 0; it < 3 * GN; it++) {
        const uint16_t *ids = g.cells[(w.i[2] * GN + w.i[1]) * GN + w.i[0]];
        (*cells)++;
        for (int e = 0; e < GCELL && ids[e] != 0xffff; e++) {
            const float *s = g.spheres + (size_t)ids[e] * 20;
            (*tests)++;
            const float mx = o[0] - s[0], my = o[1] - s[1], mz = o[2] - s[2];
            const float C = fmaf(-s[16], s[16], rt_dot3(mx, my, mz, mx, my, mz));
            const float B = rt_dot3(mx, my, mz, d[0], d[1], d[2]);
            const float disc = fmaf(B, B, -(A * C));
            if (disc >= 0.0f) {
                const float sq = sqrtf(disc), N0 = -B - sq, N1 = -B + sq;
                if ((N0 >= 0.0f * A && N0 * 1.0f < far) || (N1 >= 0.0f * A && N1 * 1.0f < far)) return 1;
            }
        }
        const float t = walk_step(&w);
        if (t < 0.0f || t > tmax) break;
    }
    return 0;
}

static void hook(const float o[3], const float d[3], float tmin, float tmax, int any,
                 int hit, float distance, uint32_t primitive_id) {
    /* the triangles as the full loop does them, then the spheres through the grid */
    float bestN = tmax, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    int h = 0;
    uint64_t cells = 0, tests = 0;
    for (uint32_t i = 0; i < g.T; i++) {
        const float *v = g.vertices + (size_t)i * 12;
        if (rt_triangle_test(v, v + 4, v + 8, o, d, tmin, &bestN, &bestD)) { h = 1; best = i; if (any) break; }
    }
    int bad;
    if (any) {
        if (!h) h = grid_any(o, d, tmax, &cells, &tests);
        bad = h != hit;
    } else {
        grid_closest(o, d, tmin, &bestN, &bestD, &best, &cells, &tests);
        h = best != 0xffffffffu;
        const float dist = h ? bestN / bestD : 0.0f;
        bad = h != hit || (h && (best != primitive_id || memcmp(&dist, &distance, 4)));
    }
#pragma omp critical
    {
        g.queries++; g.cellsVisited += cells; g.sphereTests += tests;
        if (bad) {
            if (!g.mismatches) {
                float rec[8] = { o[0], o[1], o[2], d[0], d[1], d[2], tmax, (float)any };
                memcpy(g.firstBad, rec, sizeof rec);
            }
            g.mismatches++;
        }
    }
}

/* stats: queries, mismatches, cells visited, sphere tests made, grid usable; then the first failing query (8 floats: o, d, tmax, any) */
int rt_oracle_sgridcheck(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                         const simd_float3 *vertices, const SphereGPU *spheres, const uint32_t *offsets, float *out,
                         const RTLaunchParams *params, uint64_t stats[5], float firstBad[8], int nthreads) {
    memset(&g, 0, sizeof g);
    g.vertices = (const float *)vertices; g.T = params->numTriangles;
    g.spheres = (const float *)spheres; g.S = params->numSpheres;
    build();
    rt_oracle_query_hook = g.ok ? hook : 0;
    uint64_t cnt[RT_CNT_COUNT];
    int rc = rt_oracle_pathtrace(cameras, materials, lights, vertices, spheres, offsets, out, params, cnt, nthreads);
    rt_oracle_query_hook = 0;
    stats[0] = g.queries; stats[1] = g.mismatches; stats[2] = g.cellsVisited; stats[3] = g.sphereTests; stats[4] = (uint64_t)g.ok;
    memcpy(firstBad, g.firstBad, 32);
    return rc;
}
