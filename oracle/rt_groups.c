/*
 * rt_groups.c -- TEST INFRASTRUCTURE.  CPU restatement of what the sorted kernel adds to the path
 * (gpuraytracer_b200/csrc/pathtrace.cu: prep_groups; csrc/sorted.cuh: ray_clear_of_box), and the check that it
 * cannot change a result:
 *
 *   rt_oracle_groups      the object groups of a scene, word for word the "group" section of the prepared scene
 *                         (rt_scene_layout.h): quads (triangles 2k, 2k+1 on one v0) are connected when they have a
 *                         bit-equal vertex; a group = a connected component; a group that spans 90 % of the scene
 *                         on every axis or holds a light's centre is always tested; up to 4 others get a box
 *                         (bounds grown by 1/128 of the scene's largest extent) and a key bit
 *   rt_oracle_cullcheck   traces a frame with the oracle and, for EVERY ray query of it, repeats the query against
 *                         only the quads of the groups whose box the ray does not stay clear of -- and, for a shadow
 *                         ray whose two ends (shading point, light sample point) lie inside the shrunk shell box,
 *                         without the shell quads (sortedblk.cuh pass A / B) --: hit / miss, primitive id and
 *                         distance must be the full query's, bit for bit.  Also counts the keys.
 *
 * The reference has no counterpart (its queries go to Metal's acceleration structure); nothing here is part of
 * the arithmetic contract -- a group only says which quads a ray may skip.
 */
#include "rt_oracle.h"
#include "rt_detmath.h"
#include <string.h>

#define RT_MAX_GROUPS 4
#define RT_GRP_SHELL_BOX (4 + 8 * RT_MAX_GROUPS)
#define RT_GRP_FACES (RT_GRP_SHELL_BOX + 8)
#define RT_GRP_LIGHT_PLANE (RT_GRP_FACES + 6)
#define RT_GRP_LIGHTS_IN (RT_GRP_FACES + 7)
#define RT_GROUP_WORDS (RT_GRP_FACES + 8)

extern void (*rt_oracle_query_hook)(const float o[3], const float d[3], float tmin, float tmax, int any,
                                    int hit, float distance, uint32_t primitive_id);
extern _Thread_local float rt_oracle_light_sample[3];     /* rt_oracle.c: the sample point the shadow ray aims at */

static int g_scene_has_spheres = 0;              /* rt_oracle_groups_spheres: the triangles of a scene with spheres are grouped for the shell's sake only */
static const int g_light_normals_bound = 0;      /* (the library builds no groups' mask when a light-normal table is bound; the checker never binds one here) */
int rt_oracle_groups(const simd_float3 *vertices, uint32_t T, const SquareLightGPU *lights, uint32_t L,
                     uint32_t out[RT_GROUP_WORDS]) {
    const float *v = (const float *)vertices;
    memset(out, 0, 4 * RT_GROUP_WORDS);
    int paired = T >= 2 && T % 2 == 0;
    for (uint32_t i = 1; i < T && paired; i += 2)
        if (memcmp(v + (size_t)(i - 1) * 12, v + (size_t)i * 12, 12)) paired = 0;
    if (!paired || T > 64) return 0;
    const uint32_t P = T / 2;
    uint32_t label[32];
    float lo[32][3], hi[32][3], slo[3] = { INFINITY, INFINITY, INFINITY }, shi[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (uint32_t p = 0; p < P; p++) {
        label[p] = p;
        for (int k = 0; k < 3; k++) { lo[p][k] = INFINITY; hi[p][k] = -INFINITY; }
        for (int a = 0; a < 6; a++)
            for (int k = 0; k < 3; k++) {
                float x = v[((size_t)p * 6 + a) * 4 + k];
                lo[p][k] = fminf(lo[p][k], x); hi[p][k] = fmaxf(hi[p][k], x);
            }
        for (int k = 0; k < 3; k++) { slo[k] = fminf(slo[k], lo[p][k]); shi[k] = fmaxf(shi[k], hi[p][k]); }
    }
    for (int changed = 1; changed;) {                       /* smallest quad index of the component */
        changed = 0;
        for (uint32_t p = 0; p < P; p++)
            for (uint32_t q = 0; q < P; q++) {
                if (label[q] >= label[p]) continue;
                int share = 0;
                for (int a = 0; a < 6 && !share; a++)
                    for (int b = 0; b < 6 && !share; b++)
                        share = !memcmp(v + ((size_t)p * 6 + a) * 4, v + ((size_t)q * 6 + b) * 4, 12);
                if (share) { label[p] = label[q]; changed = 1; }
            }
    }
    float ext = 0.0f;
    for (int k = 0; k < 3; k++) ext = fmaxf(ext, shi[k] - slo[k]);
    const float margin = ext * (1.0f / 128.0f);
    uint32_t K = 0, always = 0;
    for (uint32_t c = 0; c < P; c++) {                      /* components in the order of their first quad */
        if (label[c] != c) continue;
        uint32_t pairs = 0;
        float glo[3] = { INFINITY, INFINITY, INFINITY }, ghi[3] = { -INFINITY, -INFINITY, -INFINITY };
        for (uint32_t p = 0; p < P; p++)
            if (label[p] == c) {
                pairs |= 1u << p;
                for (int k = 0; k < 3; k++) { glo[k] = fminf(glo[k], lo[p][k]); ghi[k] = fmaxf(ghi[k], hi[p][k]); }
            }
        int spans = 1, lit = 0;
        for (int k = 0; k < 3; k++) spans = spans && (ghi[k] - glo[k] >= 0.9f * (shi[k] - slo[k]));
        for (uint32_t l = 0; l < L; l++) {
            const float *ctr = (const float *)&lights[l].center;
            int in = 1;
            for (int k = 0; k < 3; k++) in = in && ctr[k] >= glo[k] - margin && ctr[k] <= ghi[k] + margin;
            lit = lit || in;
        }
        if (!spans && !lit && K < RT_MAX_GROUPS) {
            float *r = (float *)out + 4 + 8 * K;
            for (int k = 0; k < 3; k++) { r[k] = (glo[k] + ghi[k]) * 0.5f; r[4 + k] = (ghi[k] - glo[k]) * 0.5f + margin; }
            out[4 + 8 * K + 3] = pairs;
            K++;
        } else {
            always |= pairs;
        }
    }
    /* the shell (rt_scene_layout.h): always-tested quads lying in a face plane of the always-tested quads' own bounds;
     * then those bounds as centre | half extent shrunk by 1/16384 of the scene's largest extent */
    float alo[3] = { INFINITY, INFINITY, INFINITY }, ahi[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (uint32_t p = 0; p < P; p++)
        if ((always >> p) & 1u)
            for (int k = 0; k < 3; k++) { alo[k] = fminf(alo[k], lo[p][k]); ahi[k] = fmaxf(ahi[k], hi[p][k]); }
    uint32_t shell = 0;
    for (uint32_t p = 0; p < P; p++)
        if ((always >> p) & 1u)
            for (int k = 0; k < 3; k++)
                if (lo[p][k] == hi[p][k] && (lo[p][k] == alo[k] || lo[p][k] == ahi[k])) shell |= 1u << p;
    if (shell) {
        float *r = (float *)out + RT_GRP_SHELL_BOX;
        const float shrink = ext * (1.0f / 16384.0f);
        for (int k = 0; k < 3; k++) { r[k] = (alo[k] + ahi[k]) * 0.5f; r[4 + k] = (ahi[k] - alo[k]) * 0.5f - shrink; }
        r[7] = shrink; r[3] = margin;
        /* the shell quads face by face (the exit rule, rt_scene_layout.h) */
        for (uint32_t p = 0; p < P; p++)
            if ((shell >> p) & 1u)
                for (int k = 0; k < 3; k++)
                    if (lo[p][k] == hi[p][k]) {
                        if (lo[p][k] == alo[k]) out[RT_GRP_FACES + 2 * k] |= 1u << p;
                        if (lo[p][k] == ahi[k]) out[RT_GRP_FACES + 2 * k + 1] |= 1u << p;
                    }
    }
    /* the quads in the plane of the one light: no shadow ray can be accepted by them (rt_scene_layout.h) */
    int onePlane = L >= 1 && ext < 64.0f && !g_light_normals_bound;
    for (uint32_t l = 1; l < L; l++) onePlane = onePlane && ((const float *)&lights[l].center)[1] == ((const float *)&lights[0].center)[1];
    if (onePlane) {
        const float cy = ((const float *)&lights[0].center)[1];
        for (uint32_t p = 0; p < P; p++)
            if (lo[p][1] == cy && hi[p][1] == cy) out[RT_GRP_LIGHT_PLANE] |= 1u << p;
    }
    /* does every light's sample rectangle lie inside the shrunk shell box?  (rt_scene_layout.h: RT_GRP_LIGHTS_IN) */
    {
        const float *bx = (const float *)out + RT_GRP_SHELL_BOX;
        int in = shell != 0 && !g_light_normals_bound;
        for (uint32_t l = 0; l < L; l++) {
            const float *c = (const float *)&lights[l];
            const float hx = fmaxf(0.25f, 0.5f * c[12]), hz = fmaxf(0.25f, 0.5f * c[13]);
            in = in && fabsf(c[0] - bx[0]) + hx < bx[4] && fabsf(c[1] - bx[1]) < bx[5] && fabsf(c[2] - bx[2]) + hz < bx[6];
        }
        out[RT_GRP_LIGHTS_IN] = in ? 1u : 0u;
    }
    out[0] = K; out[1] = always; out[2] = (K >= 1 && !g_scene_has_spheres) ? 1u : 0u; out[3] = shell;
    return (int)K;
}

/* csrc/sorted.cuh: ray_clear_of_box, w = box centre - origin */
static int ray_clear_of_box(const float w[3], const float H[3], const float d[3]) {
    const float ax = fabsf(d[0]), ay = fabsf(d[1]), az = fabsf(d[2]);
    int clear = fabsf(fmaf(w[1], d[2], -(w[2] * d[1]))) > fmaf(H[1], az, H[2] * ay);
    clear |= fabsf(fmaf(w[2], d[0], -(w[0] * d[2]))) > fmaf(H[2], ax, H[0] * az);
    clear |= fabsf(fmaf(w[0], d[1], -(w[1] * d[0]))) > fmaf(H[0], ay, H[1] * ax);
    clear |= rt_dot3(w[0], w[1], w[2], d[0], d[1], d[2]) + fmaf(H[2], az, fmaf(H[1], ay, H[0] * ax)) < 0.0f;
    return clear;
}

static struct {
    uint32_t words[RT_GROUP_WORDS];
    const float *vertices; uint32_t T;
    uint64_t queries, mismatches, culledTests, fullTests;
    uint64_t keysDual[16], keysSingle[16];
    uint64_t shadowQueries, shadowInShell;
} g_chk;
static _Thread_local int t_pending; static _Thread_local float t_origin[3]; static _Thread_local uint32_t t_key;

/* the exit rule (rt_scene_layout.h; csrc/sortedblk.cuh: exit_face_code): the shell quads a closest-hit ray from o along d
 * must still be tested against.  (The kernel multiplies by an approximate reciprocal where this divides: its crossing points
 * differ in the seventh digit, the margin is 1/128 of the scene.) */
static uint32_t exit_rule(const float o[3], const float d[3]) {
    const float *bx = (const float *)g_chk.words + RT_GRP_SHELL_BOX;
    const uint32_t shell = g_chk.words[3];
    for (int k = 0; k < 3; k++)
        if (!(fabsf(o[k] - bx[k]) < bx[4 + k])) return shell;               /* starts outside the shrunk box: all of them */
    uint32_t out = 0; int n = 0;
    for (int k = 0; k < 3; k++) {
        if (d[k] == 0.0f) continue;                                          /* parallel to both faces of the axis */
        const int hiSide = d[k] > 0.0f;
        const float hu = bx[4 + k] + bx[7];
        const float b = hiSide ? bx[k] + hu : bx[k] - hu;
        const float t = (b - o[k]) / d[k];
        int in = 1;
        for (int j = 0; j < 3; j++) {
            if (j == k) continue;
            const float pj = fmaf(t, d[j], o[j]);
            if (fabsf(pj - bx[j]) > bx[4 + j] + bx[7] + bx[3]) in = 0;       /* (a NaN keeps the face) */
        }
        if (in) { out |= g_chk.words[RT_GRP_FACES + 2 * k + hiSide]; n++; }
    }
    return n > 1 ? shell : out;                                              /* the kernel's key has room for one face, none or all */
}

static void cull_hook(const float o[3], const float d[3], float tmin, float tmax, int any,
                      int hit, float distance, uint32_t primitive_id) {
    const uint32_t K = g_chk.words[0];
    uint32_t key = 0, pairs = g_chk.words[1];
    for (uint32_t g = 0; g < K; g++) {
        const float *r = (const float *)g_chk.words + 4 + 8 * g;
        const float w[3] = { r[0] - o[0], r[1] - o[1], r[2] - o[2] };
        if (!ray_clear_of_box(w, r + 4, d)) { key |= 1u << g; pairs |= g_chk.words[4 + 8 * g + 3]; }
    }
    if (!any) pairs = (pairs & ~g_chk.words[3]) | exit_rule(o, d);
    if (any) {
        /* the shell rule: both ends of the shadow ray inside the shrunk box of the always-tested quads' bounds */
        const float *bx = (const float *)g_chk.words + RT_GRP_SHELL_BOX, *sp = rt_oracle_light_sample;
        int in = 1;
        for (int k = 0; k < 3; k++) in = in && fabsf(o[k] - bx[k]) < bx[4 + k] && fabsf(sp[k] - bx[k]) < bx[4 + k];
        if (in) pairs &= ~g_chk.words[3];
        pairs &= ~g_chk.words[RT_GRP_LIGHT_PLANE];
#pragma omp atomic
        g_chk.shadowQueries++;
        if (in) {
#pragma omp atomic
            g_chk.shadowInShell++;
        }
    }
    int h = 0; uint32_t id = 0, tests = 0;
    float bestN = tmax, bestD = 1.0f;
    for (uint32_t i = 0; i < g_chk.T; i++) {
        if (!((pairs >> (i / 2)) & 1u)) continue;
        const float *v = g_chk.vertices + (size_t)i * 12;
        tests++;
        if (rt_triangle_test(v, v + 4, v + 8, o, d, tmin, &bestN, &bestD)) { h = 1; id = i; if (any) break; }
    }
    const float dist = h ? bestN / bestD : 0.0f;
    const int bad = any ? (h != hit) : (h != hit || (h && (id != primitive_id || memcmp(&dist, &distance, 4))));
#pragma omp atomic
    g_chk.queries++;
    if (bad) {
#pragma omp atomic
        g_chk.mismatches++;
    }
#pragma omp atomic
    g_chk.culledTests += tests;
#pragma omp atomic
    g_chk.fullTests += g_chk.T;
    /* the keys the kernel sorts by: a bounce's shadow ray and continuation ray start at one point and share a key;
     * the last bounce has a shadow ray only */
    if (any) {
        if (t_pending) {
#pragma omp atomic
            g_chk.keysSingle[t_key]++;
        }
        t_pending = 1; memcpy(t_origin, o, 12); t_key = key;
    } else {
        if (t_pending && !memcmp(t_origin, o, 12)) {
#pragma omp atomic
            g_chk.keysDual[t_key | key]++;
        } else if (t_pending) {
#pragma omp atomic
            g_chk.keysSingle[t_key]++;
        }
        t_pending = 0;
    }
}

/* stats: queries, mismatches, triangle tests with culling, without, then keysDual[16], keysSingle[16];
 * rt_oracle_cullcheck_shell: shadow queries, and how many of them had both ends inside the shell box */
void rt_oracle_cullcheck_shell(uint64_t out[2]) { out[0] = g_chk.shadowQueries; out[1] = g_chk.shadowInShell; }
int rt_oracle_cullcheck(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                        const simd_float3 *vertices, const uint32_t *offsets, float *out,
                        const RTLaunchParams *params, uint64_t stats[36], int nthreads) {
    memset(&g_chk, 0, sizeof g_chk);
    rt_oracle_groups(vertices, params->numTriangles, lights, params->numLights, g_chk.words);
    g_chk.vertices = (const float *)vertices; g_chk.T = params->numTriangles;
    rt_oracle_query_hook = cull_hook;
    uint64_t cnt[RT_CNT_COUNT];
    int rc = rt_oracle_pathtrace(cameras, materials, lights, vertices, 0, offsets, out, params, cnt, nthreads);
    rt_oracle_query_hook = 0;
    stats[0] = g_chk.queries; stats[1] = g_chk.mismatches; stats[2] = g_chk.culledTests; stats[3] = g_chk.fullTests;
    for (int k = 0; k < 16; k++) { stats[4 + k] = g_chk.keysDual[k]; stats[20 + k] = g_chk.keysSingle[k]; }
    return rc;
}

/* ---- mode B (csrc/modeb.cuh: closest_b): the same box rule per ray, and the shell rule for the rays that are segments towards
 * a light sample point.  Mode B's rays start 1e-4 off their surface (shaders.metal:527), not 1e-3, so the origin is asked
 * to lie inside the box shrunk by a QUARTER of the shrink only (1/65536 of the extent; the end as in mode A), and the
 * ray's tmin = 1e-3 does the rest: a shell quad's plane is crossed behind the origin or beyond the end.
 * Per ray here; the kernel tests the union of a warp's lists, a superset in the same ascending order. */
extern void (*rt_oracle_modeb_query_hook)(const float o[3], const float d[3], float tmin, float tmax, int seg, const float end[3],
                                          int hit, float distance, uint32_t primitive_id);
static struct { uint64_t queries, mismatches, culledTests, fullTests, segQueries, segInShell; } g_b;

static void cull_hook_b(const float o[3], const float d[3], float tmin, float tmax, int seg, const float end[3],
                        int hit, float distance, uint32_t primitive_id) {
    const uint32_t K = g_chk.words[0];
    uint32_t pairs = g_chk.words[1];
    for (uint32_t g = 0; g < K; g++) {
        const float *r = (const float *)g_chk.words + 4 + 8 * g;
        const float w[3] = { r[0] - o[0], r[1] - o[1], r[2] - o[2] };
        if (!ray_clear_of_box(w, r + 4, d)) pairs |= g_chk.words[4 + 8 * g + 3];
    }
    int in = 0;
    if (seg) {
        const float *bx = (const float *)g_chk.words + RT_GRP_SHELL_BOX;
        in = 1;
        for (int k = 0; k < 3; k++)
            in = in && fabsf(o[k] - bx[k]) < bx[4 + k] + 0.75f * bx[7] && fabsf(end[k] - bx[k]) < bx[4 + k];
        if (in) pairs &= ~g_chk.words[3];
    }
    int h = 0; uint32_t id = 0, tests = 0;
    float bestN = tmax, bestD = 1.0f;
    for (uint32_t i = 0; i < g_chk.T; i++) {
        if (!((pairs >> (i / 2)) & 1u)) continue;
        const float *v = g_chk.vertices + (size_t)i * 12;
        tests++;
        if (rt_triangle_test(v, v + 4, v + 8, o, d, tmin, &bestN, &bestD)) { h = 1; id = i; }
    }
    const float dist = h ? bestN / bestD : 0.0f;
    const int bad = h != hit || (h && (id != primitive_id || memcmp(&dist, &distance, 4)));
#pragma omp critical
    {
        g_b.queries++; g_b.mismatches += bad; g_b.culledTests += tests; g_b.fullTests += g_chk.T;
        g_b.segQueries += seg != 0; g_b.segInShell += in;
    }
}

/* stats: queries, mismatches, triangle tests with culling, without, segment queries, of those inside the shell box */
int rt_oracle_cullcheck_modeb(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                              const simd_float3 *vertices, uint8_t *pixels, float *text,
                              const RTLaunchParams *params, uint64_t stats[6], int nthreads) {
    memset(&g_chk, 0, sizeof g_chk); memset(&g_b, 0, sizeof g_b);
    rt_oracle_groups(vertices, params->numTriangles, lights, params->numLights, g_chk.words);
    g_chk.vertices = (const float *)vertices; g_chk.T = params->numTriangles;
    rt_oracle_modeb_query_hook = cull_hook_b;
    uint64_t cnt[RT_CNT_COUNT];
    int rc = rt_oracle_drawtriangle(cameras, materials, lights, vertices, pixels, text, params, cnt, nthreads);
    rt_oracle_modeb_query_hook = 0;
    stats[0] = g_b.queries; stats[1] = g_b.mismatches; stats[2] = g_b.culledTests; stats[3] = g_b.fullTests;
    stats[4] = g_b.segQueries; stats[5] = g_b.segInShell;
    return rc;
}


/* ---- the shell rule of the several-lights passes (csrc/wave.cuh: skipS), scenes with spheres included ----
 * Every shadow query of a frame whose origin lies inside the shrunk shell box is repeated WITHOUT the shell triangles
 * (all other triangles, then all spheres, the full loop's order): same answer required. */
int rt_oracle_groups_spheres(const simd_float3 *vertices, uint32_t T, const SquareLightGPU *lights, uint32_t L, uint32_t out[RT_GROUP_WORDS]) {
    g_scene_has_spheres = 1;
    int K = rt_oracle_groups(vertices, T, lights, L, out);
    g_scene_has_spheres = 0;
    return K;
}
static struct { const float *spheres; uint32_t S; uint64_t shadow, skipped, mismatches; } g_sh;
static void shell_hook(const float o[3], const float d[3], float tmin, float tmax, int any,
                       int hit, float distance, uint32_t primitive_id) {
    (void)distance; (void)primitive_id;
    if (!any) return;
    const float *bx = (const float *)g_chk.words + RT_GRP_SHELL_BOX;
    int in = g_chk.words[RT_GRP_LIGHTS_IN] != 0;
    for (int k = 0; k < 3; k++) in = in && fabsf(o[k] - bx[k]) < bx[4 + k];
    int h = 0;
    const uint32_t skip = (in ? g_chk.words[3] : 0u) | g_chk.words[RT_GRP_LIGHT_PLANE];     /* the quads in the lights' plane: from anywhere */
    in = skip != 0;
    if (in) {
        float bestN = tmax, bestD = 1.0f;
        for (uint32_t i = 0; i < g_chk.T && !h; i++) {
            if ((skip >> (i / 2)) & 1u) continue;
            const float *v = g_chk.vertices + (size_t)i * 12;
            h = rt_triangle_test(v, v + 4, v + 8, o, d, tmin, &bestN, &bestD);
        }
        for (uint32_t j = 0; j < g_sh.S && !h; j++) {
            const float *sp = g_sh.spheres + (size_t)j * 20;
            h = rt_sphere_test(sp, sp[16], o, d, tmin, &bestN, &bestD);
        }
    }
#pragma omp critical
    { g_sh.shadow++; g_sh.skipped += in; g_sh.mismatches += in && h != hit; }
}
/* stats: shadow queries, of those traced without the shell triangles, mismatches (must be 0) */
int rt_oracle_shellcheck(const CameraGPU *cameras, const MaterialGPU *materials, const SquareLightGPU *lights,
                         const simd_float3 *vertices, const SphereGPU *spheres, const uint32_t *offsets, float *out,
                         const RTLaunchParams *params, uint64_t stats[3], int nthreads) {
    memset(&g_chk, 0, sizeof g_chk); memset(&g_sh, 0, sizeof g_sh);
    if (params->numSpheres) rt_oracle_groups_spheres(vertices, params->numTriangles, lights, params->numLights, g_chk.words);
    else rt_oracle_groups(vertices, params->numTriangles, lights, params->numLights, g_chk.words);
    g_chk.vertices = (const float *)vertices; g_chk.T = params->numTriangles;
    g_sh.spheres = (const float *)spheres; g_sh.S = params->numSpheres;
    rt_oracle_query_hook = shell_hook;
    uint64_t cnt[RT_CNT_COUNT];
    int rc = rt_oracle_pathtrace(cameras, materials, lights, vertices, spheres, offsets, out, params, cnt, nthreads);
    rt_oracle_query_hook = 0;
    stats[0] = g_sh.shadow; stats[1] = g_sh.skipped; stats[2] = g_sh.mismatches;
    return rc;
}
