/*
 * rt_oracle.c -- TEST INFRASTRUCTURE.  CPU restatement of the reference kernel
 * `pathTrace` (mode A), one pixel at a time, OpenMP over rows.
 *
 * Follows, statement by statement:
 *     RTrace/raytrace.metal:22-109        kernel body
 *     RTrace/sampling.metal:16-26         getTriangleNormal
 *     RTrace/sampling.metal:39-49         sampleCosineWeightedHemisphere
 *     RTrace/sampling.metal:51-66         alignHemisphereWithNormal
 *     RTrace/sampling.metal:68-75         hash
 *     RTrace/sampling.metal:97-122        primes[], halton
 *     RTrace/sampling.metal:125-157       generateCameraRay
 *     RTrace/sampling.metal:198-236       sampleAreaLight
 *     RTrace/image.swift:40-65            tone map (resolve)
 * Arithmetic rules: rt_detmath.h (R1-R3).  Shader-level expressions below are
 * written operator by operator in the order MSL parses them; do not "simplify".
 * Must be compiled with -ffp-contract=off and without -ffast-math.
 *
 * Extensions that the reference kernel does not have, all inert at the
 * reference's settings: row tiles (row0/rows), spp / firstSample / bounces as
 * parameters, progressive accumulation, sphere primitives and more than one
 * light (config 4; see DESIGN.md "oracle-defined extensions").
 */
#include "rt_oracle.h"
#include "rt_detmath.h"

#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#if defined(__FAST_MATH__)
#error "the oracle must not be built with -ffast-math"
#endif

typedef struct { float x, y, z; } f3;

static inline f3 f3_make(float x, float y, float z) { f3 r = { x, y, z }; return r; }
static inline f3 f3_ld(const simd_float3 *p) { return f3_make(p->x, p->y, p->z); }
static inline f3 f3_add(f3 a, f3 b) { return f3_make(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline f3 f3_sub(f3 a, f3 b) { return f3_make(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline f3 f3_mul(f3 a, f3 b) { return f3_make(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline f3 f3_scale(f3 a, float s) { return f3_make(a.x * s, a.y * s, a.z * s); }
static inline f3 f3_neg(f3 a) { return f3_make(-a.x, -a.y, -a.z); }
static inline float f3_dot(f3 a, f3 b) { return rt_dot3(a.x, a.y, a.z, b.x, b.y, b.z); }
static inline f3 f3_cross(f3 a, f3 b) { f3 r; rt_cross3(a.x, a.y, a.z, b.x, b.y, b.z, &r.x, &r.y, &r.z); return r; }
static inline f3 f3_normalize(f3 a) { rt_normalize3(&a.x, &a.y, &a.z); return a; }
static inline float f3_length(f3 a) { return rt_length3(a.x, a.y, a.z); }

/* sampling.metal:68-75 */
uint32_t rt_oracle_hash(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

/* sampling.metal:97-104 */
static const uint32_t rt_primes[24] = {
    2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37,
    41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89
};

/* sampling.metal:107-122 */
float rt_oracle_halton(uint32_t i, uint32_t d) {
    uint32_t b = rt_primes[d];
    float f = 1.0f;
    float invB = 1.0f / (float)b;
    float r = 0.0f;
    while (i > 0) {
        f = f * invB;
        r = r + f * (float)(i % b);
        i = i / b;
    }
    return r;
}

/* sampling.metal:16-26 */
static inline f3 triangle_normal(const simd_float3 *vertices, uint32_t primitiveId) {
    f3 v0 = f3_ld(&vertices[primitiveId * 3 + 0]);
    f3 v1 = f3_ld(&vertices[primitiveId * 3 + 1]);
    f3 v2 = f3_ld(&vertices[primitiveId * 3 + 2]);
    f3 edge1 = f3_sub(v1, v0);
    f3 edge2 = f3_sub(v2, v0);
    return f3_normalize(f3_cross(edge1, edge2));
}

/* sampling.metal:39-49 */
static inline f3 sample_cosine_hemisphere(float ux, float uy) {
    float phi = (2.0f * 3.14159265358979323846264338327950288f) * ux;
    float sin_phi, cos_phi;
    rt_sincosf(phi, &sin_phi, &cos_phi);
    float cos_theta = sqrtf(uy);
    float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
    return f3_make(sin_theta * cos_phi, cos_theta, sin_theta * sin_phi);
}

/* sampling.metal:51-66 */
static inline f3 align_hemisphere(f3 sample, f3 normal) {
    f3 up = normal;
    f3 right = f3_normalize(f3_cross(normal, f3_make(0.0072f, 1.0f, 0.0034f)));
    f3 forward = f3_cross(right, up);
    return f3_add(f3_add(f3_scale(right, sample.x), f3_scale(up, sample.y)), f3_scale(forward, sample.z));
}

/* sampling.metal:125-157.  Note the INTEGER division in aspectRatio (:132). */
static inline void camera_ray(const CameraGPU *camera, uint32_t ix, uint32_t iy, float jx, float jy,
                              f3 *origin, f3 *direction) {
    int x = (int)ix;
    int y = (int)iy;
    float aspectRatio = (float)(camera->resolution.x / camera->resolution.y);
    float halfWidth = rt_tanf(camera->horizontalFov / 2.0f);
    float halfHeight = halfWidth / aspectRatio;

    f3 w = f3_neg(f3_normalize(f3_ld(&camera->direction)));
    f3 u = f3_normalize(f3_cross(f3_ld(&camera->up), w));
    f3 v = f3_normalize(f3_cross(w, u));

    float s = (((float)x + jx) / (float)camera->resolution.x) * 2.0f - 1.0f;
    float t = -((((float)y + jy) / (float)camera->resolution.y) * 2.0f - 1.0f);

    f3 dir = f3_normalize(f3_sub(f3_add(f3_scale(u, s * halfWidth), f3_scale(v, t * halfHeight)), w));
    *origin = f3_ld(&camera->position);
    *direction = dir;
}

/* ---- light geometry -------------------------------------------------------------------------------
 * The reference hard-codes the light's normal (0,-1,0) and spans its sample rectangle with right = (0.25,0,0),
 * up = (0,0,0.25) (sampling.metal:205-209; the TODO of :175 asks for the generalisation).  Extensions, inert unless asked for:
 *   RT_VARIANT_LIGHT_EXTENT   right / up span the light's own width x depth (half extents width/2, depth/2)
 *   a per-light NORMAL table  (rt_oracle_set_light_normals; the library: a buffer bound at RT_INDEX_LIGHT_NORMALS, float4
 *                             per light): right = tangent * hx, up = bitangent * hz with buildOrthonormalBasis(normal)
 *                             (sampling.metal:158-171), as directSquareLightRay (:173-196) builds its rectangle.  For
 *                             (0,-1,0) that basis is (1,0,0), (0,0,1): the reference's literals.
 * Without a table the literals themselves are used, so the reference's bits cannot move. */
static const float *g_light_normals = 0;
void rt_oracle_set_light_normals(const float *normals_float4_per_light) { g_light_normals = normals_float4_per_light; }

typedef struct { f3 center, normal, tangent, bitangent, right, up; float hx, hz; } light_geom;

/* sampling.metal:158-171 */
static inline void build_orthonormal_basis(f3 normal, f3 *tangent, f3 *bitangent) {
    f3 t;
    if (fabsf(normal.x) > 0.9f) t = f3_normalize(f3_sub(f3_make(0, 1, 0), f3_scale(normal, f3_dot(f3_make(0, 1, 0), normal))));
    else                        t = f3_normalize(f3_sub(f3_make(1, 0, 0), f3_scale(normal, f3_dot(f3_make(1, 0, 0), normal))));
    *tangent = t;
    *bitangent = f3_cross(normal, t);
}

static inline light_geom light_geometry(const SquareLightGPU *light, uint32_t l, int light_extent) {
    light_geom g;
    g.center = f3_ld(&light->center);
    g.hx = light_extent ? 0.5f * light->width : 0.25f;
    g.hz = light_extent ? 0.5f * light->depth : 0.25f;
    if (g_light_normals) {
        g.normal = f3_normalize(f3_make(g_light_normals[4 * l], g_light_normals[4 * l + 1], g_light_normals[4 * l + 2]));
        build_orthonormal_basis(g.normal, &g.tangent, &g.bitangent);
        g.right = f3_scale(g.tangent, g.hx);
        g.up = f3_scale(g.bitangent, g.hz);
    } else {
        g.normal = f3_make(0.0f, -1.0f, 0.0f);
        g.tangent = f3_make(1.0f, 0.0f, 0.0f);
        g.bitangent = f3_make(0.0f, 0.0f, 1.0f);
        g.right = f3_make(g.hx, 0.0f, 0.0f);
        g.up = f3_make(0.0f, 0.0f, g.hz);
    }
    return g;
}

/* Test hook: the light sample point of the calling thread's last sample_area_light (rt_groups.c restates the sorted
 * kernel's culling rules, one of which looks at the shadow ray's far end). */
_Thread_local float rt_oracle_light_sample[3];

/* sampling.metal:198-236 over the geometry above */
static inline f3 sample_area_light(const SquareLightGPU *light, const light_geom *g, float ux, float uy, f3 position,
                                   f3 *lightDirection, float *lightDistance) {
    ux = ux * 2.0f - 1.0f;
    uy = uy * 2.0f - 1.0f;

    f3 samplePosition = f3_add(f3_add(g->center, f3_scale(g->right, ux)), f3_scale(g->up, uy));

    rt_oracle_light_sample[0] = samplePosition.x; rt_oracle_light_sample[1] = samplePosition.y; rt_oracle_light_sample[2] = samplePosition.z;
    f3 ld = f3_sub(samplePosition, position);
    float dist = f3_length(ld);
    float inverseLightDistance = 1.0f / fmaxf(dist, 1e-3f);
    ld = f3_scale(ld, inverseLightDistance);

    f3 lightColor = f3_make(light->color.x, light->color.y, light->color.z);
    lightColor = f3_scale(lightColor, inverseLightDistance * inverseLightDistance);
    lightColor = f3_scale(lightColor, rt_saturate(f3_dot(f3_neg(ld), g->normal)));

    *lightDirection = ld;
    *lightDistance = dist;
    return lightColor;
}

/* ---- multiple importance sampling in mode A (RT_VARIANT_MIS) ---------------------------------------------
 * The reference's pathTrace has no MIS weights (its own TODO: the commented-out block of sampling.metal:238-260, which
 * mode B carries out).  With the flag, the direct light of every shaded bounce is estimated by TWO strategies
 * combined with the balance heuristic -- powerHeuristic with beta = 1 (shaders.metal:132-137), over
 * calculateSquareLightPdf (:315-326, with the light's normal) and calculateCosineWeightedPdf (:376-380):
 *   light strategy   the reference's own sample (sampleAreaLight + shadow ray), its term g = lightColor * color weighted
 *                    by w_l = heuristic(p_light, p_cosine)
 *   cosine strategy  the continuation ray the kernel samples anyway (cosine-weighted): when it passes through the light's
 *                    sample rectangle, front side, with nothing nearer than the rectangle minus 1e-3 (the shadow ray's
 *                    own visibility rule), it adds  light.color * (pi / A_s) * color * w_c,  A_s = the rectangle's area,
 *                    w_c = heuristic(p_cosine, p_light)
 * Both estimate  (light.color * color / A_s) * integral over the rectangle's solid angle of cos(theta_surface) V dw:
 * the light strategy because dA cos(theta_light) / d^2 = dw, the cosine strategy because its pdf is cos(theta_surface) / pi.
 * The weights add up to one (up to the heuristic's 1e-6), so the expectation is the flag-off light term's.
 * Two things change with the flag, both needed for that: the continuation ray is traced after the LAST bounce too (for
 * its cosine term only), and an emissive hit after a shaded bounce ends the path WITHOUT `accumulated = emissive`
 * (raytrace.metal:59 discards the light gathered so far; a camera ray that sees the light still takes its emissive).
 * Checker-only selectors (the library has no such forms): RT_ORACLE_MIS_LIGHT_ONLY w_l = 1, w_c = 0;
 * RT_ORACLE_MIS_COSINE_ONLY w_l = 0, w_c = 1 -- same expectation, which tests/test_mis_cpu.py checks statistically. */
static inline float mis_heuristic(float pdf1, float pdf2, float pdf3) {
    const float n = 1.0f;                       /* samplesPerStrategy = 1; pow(x, 1.0) = x */
    float p1 = n * pdf1;
    float sum = p1 + n * pdf2 + n * pdf3;
    return p1 / (sum + 1e-6f);
}
static inline float mis_square_light_pdf(f3 origin, const SquareLightGPU *light, const light_geom *g, f3 direction) {
    f3 toLight = f3_sub(g->center, origin);
    float distance = f3_length(toLight);
    float cosTheta = fmaxf(0.0f, f3_dot(f3_neg(direction), g->normal));
    float area = light->width * light->depth;
    return (distance * distance) / (area * cosTheta + 1e-6f);
}
static inline float mis_cosine_pdf(f3 normal, f3 direction) {
    return fmaxf(0.0f, f3_dot(normal, direction)) / 3.14159265358979323846264338327950288f;
}
/* the cosine strategy's term of one light for the continuation ray (P, rd) whose closest hit is (hit, hitDistance) */
static inline f3 mis_cosine_term(const SquareLightGPU *light, const light_geom *g, f3 P, f3 normal, f3 rd,
                                 int hit, float hitDistance, f3 color, int mode) {
    const f3 zero = f3_make(0.0f, 0.0f, 0.0f);
    if (mode == 2) return zero;                                        /* light strategy alone */
    float denom = f3_dot(rd, g->normal);
    if (!(denom < 0.0f)) return zero;                                  /* the rectangle's front side faces -normal */
    float t = f3_dot(f3_sub(g->center, P), g->normal) / denom;
    if (!(t > 0.0f)) return zero;
    if (hit && hitDistance < t - 1e-3f) return zero;                   /* something nearer: raytrace.metal:81's rule */
    f3 v = f3_sub(f3_add(P, f3_scale(rd, t)), g->center);
    if (!(fabsf(f3_dot(v, g->tangent)) <= g->hx && fabsf(f3_dot(v, g->bitangent)) <= g->hz)) return zero;
    float weight = 1.0f;
    if (mode == 1) weight = mis_heuristic(mis_cosine_pdf(normal, rd), mis_square_light_pdf(P, light, g, rd), 0.0f);
    float area = (2.0f * g->hx) * (2.0f * g->hz);
    f3 term = f3_make(light->color.x, light->color.y, light->color.z);
    term = f3_scale(term, 3.14159265358979323846264338327950288f / area);
    term = f3_scale(term, weight);
    return f3_mul(term, color);
}

typedef struct {
    const CameraGPU *camera;
    const MaterialGPU *materials;
    const SquareLightGPU *lights;
    const simd_float3 *vertices;
    const SphereGPU *spheres;
    uint32_t T, S, L;
    int bounces;
    int light_extent;
    int mis;          /* 0 off, 1 RT_VARIANT_MIS, 2 / 3 the checker-only single-strategy selectors */
} scene_view;

/* Test hook (rt_groups.c: the group-culling checker sees every ray query of a frame and its answer); NULL otherwise. */
void (*rt_oracle_query_hook)(const float o[3], const float d[3], float tmin, float tmax, int any,
                             int hit, float distance, uint32_t primitive_id) = 0;

static inline rt_hit scene_intersect(const scene_view *sc, f3 o, f3 d, float tmin, float tmax, int any) {
    float oo[3] = { o.x, o.y, o.z }, dd[3] = { d.x, d.y, d.z };
    rt_hit h = rt_intersect((const float *)sc->vertices, sc->T, (const float *)sc->spheres, sc->S,
                            oo, dd, tmin, tmax, any);
    if (rt_oracle_query_hook) rt_oracle_query_hook(oo, dd, tmin, tmax, any, h.hit, h.distance, h.primitive_id);
    return h;
}

/* One camera sample: raytrace.metal:37-102.  Returns accumulatedColor. */
static inline f3 trace_path(const scene_view *sc, uint32_t ix, uint32_t iy, uint32_t offset, uint32_t n,
                            uint64_t *cnt) {
    float uvx = rt_oracle_halton(offset + n, 0);
    float uvy = rt_oracle_halton(offset + n, 1);

    f3 ro, rd;
    camera_ray(&sc->camera[0], ix, iy, uvx, uvy, &ro, &rd);
    const float rmin = 0.001f, rmax = 1000.0f;     /* sampling.metal:154-155; kept across bounces */

    f3 accumulatedColor = f3_make(0.0f, 0.0f, 0.0f);
    f3 color = f3_make(1.0f, 1.0f, 1.0f);

    cnt[RT_CNT_CLOSEST_RAYS]++;
    rt_hit hit = scene_intersect(sc, ro, rd, rmin, rmax, 0);
    for (int bounce = 0; bounce < sc->bounces; bounce++) {
        if (!hit.hit) break;

        /* materials[primitive_id]; sphere primitives (extension) carry their own */
        const MaterialGPU *m = hit.primitive_id < sc->T ? &sc->materials[hit.primitive_id]
                                                        : &sc->spheres[hit.primitive_id - sc->T].material;
        if (f3_length(f3_ld(&m->emissive)) > 0.0f) {
            /* assignment, raytrace.metal:59.  MIS: only a camera ray takes the emissive (see above) */
            if (!sc->mis || bounce == 0) accumulatedColor = f3_ld(&m->emissive);
            break;
        }
        cnt[RT_CNT_SHADED_BOUNCES]++;

        f3 normal;
        if (hit.primitive_id < sc->T) {
            normal = triangle_normal(sc->vertices, hit.primitive_id);
        } else {  /* extension: outward sphere normal at the un-offset hit point */
            f3 hp = f3_add(ro, f3_scale(rd, hit.distance));
            normal = f3_normalize(f3_sub(hp, f3_ld(&sc->spheres[hit.primitive_id - sc->T].center)));
        }
        f3 P = f3_add(f3_add(ro, f3_scale(rd, hit.distance)), f3_scale(normal, 1e-3f));

        float wx = rt_oracle_halton(offset + n, 2 + bounce * 5 + 0);
        float wy = rt_oracle_halton(offset + n, 2 + bounce * 5 + 1);
        color = f3_mul(color, f3_make(m->diffuse.x, m->diffuse.y, m->diffuse.z));

        /* reference: lights[0] only (raytrace.metal:22); extension: every light,
         * same w, one shadow ray each, summed in index order */
        for (uint32_t l = 0; l < sc->L; l++) {
            const light_geom g = light_geometry(&sc->lights[l], l, sc->light_extent);
            f3 ldir; float ldist;
            f3 lightColor = sample_area_light(&sc->lights[l], &g, wx, wy, P, &ldir, &ldist);
            lightColor = f3_scale(lightColor, rt_saturate(f3_dot(normal, ldir)));
            if (sc->mis) {
                float weight = sc->mis == 1 ? mis_heuristic(mis_square_light_pdf(P, &sc->lights[l], &g, ldir), mis_cosine_pdf(normal, ldir), 0.0f)
                                            : (sc->mis == 2 ? 1.0f : 0.0f);
                lightColor = f3_scale(lightColor, weight);
            }
            /* default-constructed ray: min_distance = 0 (raytrace.metal:79-82) */
            cnt[RT_CNT_SHADOW_RAYS]++;
            rt_hit sh = scene_intersect(sc, P, ldir, 0.0f, ldist - 1e-3f, 1);
            if (!sh.hit) accumulatedColor = f3_add(accumulatedColor, f3_mul(lightColor, color));
        }

        float ux = rt_oracle_halton(offset + n, 2 + bounce * 5 + 2);
        float uy = rt_oracle_halton(offset + n, 2 + bounce * 5 + 3);
        f3 sd = sample_cosine_hemisphere(ux, uy);
        sd = align_hemisphere(sd, normal);
        ro = P;
        rd = sd;
        /* the next bounce's closest hit (raytrace.metal:49 at the top of the reference's loop); MIS: after the last
         * bounce too, for the cosine strategy's term */
        if (bounce + 1 < sc->bounces || sc->mis) {
            cnt[RT_CNT_CLOSEST_RAYS]++;
            hit = scene_intersect(sc, ro, rd, rmin, rmax, 0);
            if (sc->mis)
                for (uint32_t l = 0; l < sc->L; l++) {
                    const light_geom g = light_geometry(&sc->lights[l], l, sc->light_extent);
                    accumulatedColor = f3_add(accumulatedColor, mis_cosine_term(&sc->lights[l], &g, P, normal, rd, hit.hit,
                                                                                 hit.distance, color, sc->mis));
                }
        }
    }
    cnt[RT_CNT_PATHS]++;
    return accumulatedColor;
}

static int check_params(const RTLaunchParams *p) {
    if (!p) return RT_ERR_INVALID_ARGUMENT;
    if (p->width <= 0 || p->height <= 0 || p->row0 < 0 || p->rows < 0) return RT_ERR_INVALID_ARGUMENT;
    if (!p->tileRows && p->row0 + p->rows > p->height) return RT_ERR_INVALID_ARGUMENT;
    return RT_OK;
}

int rt_oracle_pathtrace(const CameraGPU *cameras, const MaterialGPU *materials,
                        const SquareLightGPU *lights, const simd_float3 *vertices,
                        const SphereGPU *spheres, const uint32_t *offsets,
                        float *out, const RTLaunchParams *p,
                        uint64_t counters[RT_CNT_COUNT], int nthreads) {
    int rc = check_params(p);
    if (rc) return rc;
    if (!cameras || !materials || !lights || !vertices || !offsets || !out) return RT_ERR_INVALID_ARGUMENT;
    if (p->spp == 0 || p->numLights == 0) return RT_ERR_INVALID_ARGUMENT;
    if (p->bounces < 1 || p->bounces > 4) return RT_ERR_UNSUPPORTED;   /* primes[2+5b+3] must exist */
    if (p->numSpheres && !spheres) return RT_ERR_INVALID_ARGUMENT;

    scene_view sc = { cameras, materials, lights, vertices, spheres,
                      p->numTriangles, p->numSpheres, p->numLights, p->bounces,
                      (p->variant & RT_VARIANT_LIGHT_EXTENT) != 0,
                      (p->variant & RT_ORACLE_MIS_LIGHT_ONLY) ? 2 : (p->variant & RT_ORACLE_MIS_COSINE_ONLY) ? 3 : (p->variant & RT_VARIANT_MIS) ? 1 : 0 };
    uint64_t total[RT_CNT_COUNT] = { 0 };
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
    const int W = p->width;

#pragma omp parallel num_threads(nthreads)
    {
        uint64_t cnt[RT_CNT_COUNT] = { 0 };
        /* jobs of 32 pixels of one row, so that a few rows still keep every core busy */
        const int chunks = (W + 31) / 32;
#pragma omp for schedule(dynamic, 1)
        for (int job = 0; job < p->rows * chunks; job++) {
            const int ly = job / chunks, xb = (job % chunks) * 32, xe = xb + 32 < W ? xb + 32 : W;
            const uint32_t iy = p->tileRows
                ? (uint32_t)p->row0 + ((uint32_t)ly / p->tileRows) * p->tileStride + ((uint32_t)ly % p->tileRows)
                : (uint32_t)(p->row0 + ly);
            for (int x = xb; x < xe; x++) {
                const size_t px = (size_t)ly * W + x;
                const uint32_t offset = offsets[px];          /* raytrace.metal:37 */
                float *o = out + px * 4;
                f3 luminance = p->accumulate ? f3_make(o[0], o[1], o[2]) : f3_make(0.0f, 0.0f, 0.0f);
                for (uint32_t k = 0; k < p->spp; k++) {
                    uint32_t n = p->firstSample + k;
                    f3 a = trace_path(&sc, (uint32_t)x, iy, offset, n, cnt);
                    luminance = f3_add(luminance, a);          /* raytrace.metal:103 */
                }
                if (p->accumulate) {
                    o[0] = luminance.x; o[1] = luminance.y; o[2] = luminance.z;
                    o[3] = (float)(p->firstSample + p->spp);
                } else {
                    float samples = (float)p->spp;             /* raytrace.metal:106 */
                    o[0] = luminance.x / samples; o[1] = luminance.y / samples; o[2] = luminance.z / samples;
                    o[3] = 1.0f;                               /* raytrace.metal:109 */
                }
            }
        }
#pragma omp critical
        for (int c = 0; c < RT_CNT_COUNT; c++) total[c] += cnt[c];
    }
    if (counters) for (int c = 0; c < RT_CNT_COUNT; c++) counters[c] = total[c];
    return RT_OK;
}

/* ---- exported pieces for known-answer tests --------------------------------- */
void rt_oracle_camera_ray(const CameraGPU *cam, uint32_t x, uint32_t y, float jx, float jy,
                          float origin[3], float dir[3]) {
    f3 o, d;
    camera_ray(cam, x, y, jx, jy, &o, &d);
    origin[0] = o.x; origin[1] = o.y; origin[2] = o.z;
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}

void rt_oracle_sample_area_light(const SquareLightGPU *light, float ux, float uy, const float pos[3],
                                 float color[3], float ldir[3], float *ldist) {
    f3 d; float dist;
    const light_geom g = light_geometry(light, 0, 0);
    f3 c = sample_area_light(light, &g, ux, uy, f3_make(pos[0], pos[1], pos[2]), &d, &dist);
    color[0] = c.x; color[1] = c.y; color[2] = c.z;
    ldir[0] = d.x; ldir[1] = d.y; ldir[2] = d.z;
    *ldist = dist;
}

void rt_oracle_sample_area_light_ex(const SquareLightGPU *light, float ux, float uy, const float pos[3],
                                    float color[3], float ldir[3], float *ldist, int light_extent) {
    f3 d; float dist;
    const light_geom g = light_geometry(light, 0, light_extent);      /* (light 0's row of the normal table, if one is set) */
    f3 c = sample_area_light(light, &g, ux, uy, f3_make(pos[0], pos[1], pos[2]), &d, &dist);
    color[0] = c.x; color[1] = c.y; color[2] = c.z;
    ldir[0] = d.x; ldir[1] = d.y; ldir[2] = d.z;
    *ldist = dist;
}

void rt_oracle_cosine_hemisphere(float ux, float uy, const float normal[3], float out[3]) {
    f3 s = sample_cosine_hemisphere(ux, uy);
    s = align_hemisphere(s, f3_make(normal[0], normal[1], normal[2]));
    out[0] = s.x; out[1] = s.y; out[2] = s.z;
}

void rt_oracle_sincos(float x, float *s, float *c) { rt_sincosf(x, s, c); }

int rt_oracle_intersect(const simd_float3 *vertices, uint32_t numTriangles,
                        const SphereGPU *spheres, uint32_t numSpheres,
                        const float o[3], const float d[3], float tmin, float tmax,
                        int acceptAny, float *distance, uint32_t *primitive_id) {
    rt_hit h = rt_intersect((const float *)vertices, numTriangles, (const float *)spheres, numSpheres,
                            o, d, tmin, tmax, acceptAny);
    if (distance) *distance = h.distance;
    if (primitive_id) *primitive_id = h.primitive_id;
    return h.hit;
}

/* The closest hit over the triangles ids[0..n) met IN THAT ORDER, chosen by the order-independent form of the
 * loop's rule that the BVH kernel uses (gpuraytracer_b200/csrc/bvh.cuh, bvh_leaf_closest): smaller cross-multiplied
 * distance, and on exact equality of the two products the lower primitive id.  Test hook: tests/test_bvh_cpu.py
 * checks that every order gives rt_oracle_intersect's answer (the ascending strict-less loop). */
int rt_oracle_intersect_ordered(const simd_float3 *vertices, const uint32_t *ids, uint32_t n,
                                const float o[3], const float d[3], float tmin, float tmax,
                                float *distance, uint32_t *primitive_id) {
    const float *vs = (const float *)vertices;
    float bestN = tmax, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    for (uint32_t k = 0; k < n; k++) {
        const uint32_t id = ids[k];
        const float *v = vs + (size_t)id * 12;
        float N = tmax, a = 1.0f;
        if (!rt_triangle_test(v, v + 4, v + 8, o, d, tmin, &N, &a)) continue;      /* valid and nearer than tmax: (N, a) */
        const float p1 = N * bestD, p2 = bestN * a;
        if (p1 < p2 || (p1 == p2 && id < best && best != 0xffffffffu)) { bestN = N; bestD = a; best = id; }
    }
    if (distance) *distance = bestN / bestD;
    if (primitive_id) *primitive_id = best;
    return best != 0xffffffffu;
}

/* RTrace/image.swift:40-65.  pow() is the one place the host image writer uses
 * libm; the GPU resolve kernel is checked against this with +-1 LSB. */
static uint16_t f32_to_f16_rne(float f);
static float f16_to_f32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1fu, m = h & 0x3ffu;
    uint32_t x;
    if (e == 0x1fu) x = sign | 0x7f800000u | (m << 13);
    else if (e) x = sign | ((e + 112u) << 23) | (m << 13);
    else if (!m) x = sign;
    else { int k = 0; uint32_t mm = m; while (!(mm & 0x400u)) { mm <<= 1; k++; } x = sign | ((uint32_t)(113 - k) << 23) | ((mm & 0x3ffu) << 13); }
    float f;
    memcpy(&f, &x, 4);
    return f;
}

/* image.swift:25-65: the render target is rgba16Float (renderer.swift:74-82), so what saveTextureToImage tone-maps
 * is the shader's value rounded to half and widened again (Float(Float16)); then x2 exposure, Reinhard, gamma 2.2 */
void rt_oracle_tonemap_rgba8(const float *rgba, uint8_t *out, uint64_t npixels) {
    const float exposure = 2.0f, gamma = 2.2f;
    for (uint64_t i = 0; i < npixels; i++) {
        for (int c = 0; c < 3; c++) {
            float value = f16_to_f32(f32_to_f16_rne(rgba[i * 4 + c]));
            value *= exposure;
            value = value / (value + 1.0f);
            value = powf(value, 1.0f / gamma);
            value = fmaxf(0.0f, fminf(1.0f, value));
            out[i * 4 + c] = (uint8_t)(value * 255.0f);
        }
        out[i * 4 + 3] = 255;
    }
}

/* fp32 -> fp16, round to nearest even: what a shader write to the reference's rgba16Float render target does
 * (RTrace/renderer.swift:74-82) and what image.swift:25-38 reads back.  IEEE 754 binary16 by bit arithmetic. */
static uint16_t f32_to_f16_rne(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u, e = (x >> 23) & 0xffu;
    uint32_t m = x & 0x7fffffu;
    if (e == 0xffu) return (uint16_t)(sign | 0x7c00u | (m ? 0x200u | (m >> 13) : 0u));     /* inf, NaN (quiet) */
    const int32_t E = (int32_t)e - 127 + 15;
    if (E >= 0x1f) return (uint16_t)(sign | 0x7c00u);                                       /* overflow -> inf */
    if (E <= 0) {                                                                           /* subnormal half or zero */
        if (E < -10) return (uint16_t)sign;                                                 /* below 2^-25 */
        m |= 0x800000u;
        const uint32_t shift = (uint32_t)(14 - E);                                          /* 14 .. 24 */
        uint32_t h = m >> shift;
        const uint32_t rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1u);
        if (rem > half || (rem == half && (h & 1u))) h++;
        return (uint16_t)(sign | h);
    }
    uint32_t h = ((uint32_t)E << 10) | (m >> 13);
    const uint32_t rem = m & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) h++;                                 /* a carry may reach inf: correct */
    return (uint16_t)(sign | h);
}

/* checker of rtResolveToRGBA16F: float4 -> half4; divide_by_alpha: xyz / w first, alpha 1 */
void rt_oracle_resolve_rgba16f(const float *rgba, uint16_t *out, uint64_t npixels, int divide_by_alpha) {
    for (uint64_t i = 0; i < npixels; i++) {
        float v[4] = { rgba[i * 4], rgba[i * 4 + 1], rgba[i * 4 + 2], rgba[i * 4 + 3] };
        if (divide_by_alpha) { v[0] = v[0] / v[3]; v[1] = v[1] / v[3]; v[2] = v[2] / v[3]; v[3] = 1.0f; }
        for (int c = 0; c < 4; c++) out[i * 4 + c] = f32_to_f16_rne(v[c]);
    }
}

int rt_oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
