/*
 * rt_oracle_modeb.c -- TEST INFRASTRUCTURE.  CPU restatement of the reference's MIS
 * kernel `drawTriangle` (mode B), one pixel at a time, OpenMP over rows.
 *
 * Follows, statement by statement, Sources/gpuRaytracer/shaders.metal:
 *     :22-56     primes, halton, haltonRandom
 *     :58-85     hash, randomFloat, hashRandom (800 / 600 are the shader's literals)
 *     :132-137   powerHeuristic (called with beta = 1.0)
 *     :145-157   cameraExposure, reinhartToneMapping
 *     :159-172   buildOrthonormalBasis
 *     :186-212   smithG1_GGX, D_GGX, F_Schlick, V_SmithGGXCorrelated, Fd_Lambert
 *     :214-246   generateCameraRay
 *     :248-257   writeToPixelBuffer
 *     :259-289   calculateBRDFContribution
 *     :291-326   directSquareLightRay, calculateSquareLightPdf
 *     :355-445   cosineWeightedRay, calculateCosineWeightedPdf, vndfRay, calculateVNDFPdf
 *     :447-509   getTriangleNormal, getClosestIntersection
 *     :519-541   calculateDirectLightSamplingContribution
 *     :543-625   recursiveMultiImportanceSampling
 *     :627-633   writeToTextBuffer
 *     :635-707   drawTriangle
 * Arithmetic rules: rt_detmath.h (R1-R3): every shader-level operator is one rounding,
 * written here in the order MSL parses it.  Pinned bit for bit against the unmodified
 * shader compiled as C++ (oracle/ref_driver_b.cpp) by tests/test_oracle_vs_ref.py.
 *
 * Parameters the shader hard-codes and RTLaunchParams exposes: cameraRays (6, :644) and
 * misSamples (300, :648).  Row tiles as in mode A.
 */
#include "rt_oracle.h"
#include "rt_detmath.h"

#ifdef _OPENMP
#include <omp.h>
#endif

#define RT_PI_F 3.14159265358979323846264338327950288f

typedef struct { float x, y, z; } v3;

static inline v3 mk(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 ld3(const simd_float3 *p) { return mk(p->x, p->y, p->z); }
static inline v3 add(v3 a, v3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, v3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 divv(v3 a, v3 b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline v3 scale(v3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
static inline v3 divs(v3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
static inline v3 neg(v3 a) { return mk(-a.x, -a.y, -a.z); }
static inline float dot(v3 a, v3 b) { return rt_dot3(a.x, a.y, a.z, b.x, b.y, b.z); }
static inline v3 cross(v3 a, v3 b) { v3 r; rt_cross3(a.x, a.y, a.z, b.x, b.y, b.z, &r.x, &r.y, &r.z); return r; }
static inline v3 normalize(v3 a) { rt_normalize3(&a.x, &a.y, &a.z); return a; }
static inline float length(v3 a) { return rt_length3(a.x, a.y, a.z); }

typedef struct { v3 origin, direction; float min_distance, max_distance; } ray_t;
typedef struct { v3 tangent, bitangent; } basis_t;
enum { HIT = 0, HITLIGHT = 1, MISS = 2 };          /* IntersectionTypeGPU, shaderTypes.h:71-75 */
typedef struct {
    int type;
    v3 point;
    ray_t ray;
    v3 normal;
    v3 diffuse, emissive;                          /* MaterialGPU */
    float metallic, roughness;
} isect_t;

typedef struct {
    const CameraGPU *camera;
    const MaterialGPU *materials;
    const SquareLightGPU *light;
    const simd_float3 *vertices;
    uint32_t T;
    uint64_t *cnt;
} ctx_t;

/* :58-69 */
static inline float randomFloat(uint32_t seed) {
    return (float)rt_oracle_hash(seed) / ((float)0xffffffffU + 1.0f);
}

/* :71-85 */
static inline void hashRandom(uint32_t x, uint32_t y, uint32_t i, float *u1, float *u2) {
    uint32_t sampleId = (y * 800u + x) * i;
    uint32_t seed1 = rt_oracle_hash(x + y * 800u + sampleId);
    uint32_t seed2 = rt_oracle_hash(y + x * 600u + sampleId + 12345u);
    *u1 = randomFloat(seed1);
    *u2 = randomFloat(seed2);
}

/* :49-56 */
static inline void haltonRandom(uint32_t i, uint32_t d, float *u1, float *u2) {
    *u1 = rt_oracle_halton(i, 0 + d);
    *u2 = rt_oracle_halton(i, 1 + d);
}

/* :132-137 */
static inline float powerHeuristic(float pdf1, float pdf2, float pdf3, uint32_t samplesPerStrategy, float beta) {
    float n = (float)samplesPerStrategy;
    float p1 = rt_powf(n * pdf1, beta);
    float sum = p1 + rt_powf(n * pdf2, beta) + rt_powf(n * pdf3, beta);
    return p1 / (sum + 1e-6f);
}

/* :145-150 */
static inline float cameraExposure(const CameraGPU *camera) {
    float maxLuminance = 1.2f * rt_powf(2.0f, camera->ev100);
    return 1.0f / maxLuminance;
}

static inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

/* :152-157 */
static inline v3 reinhartToneMapping(v3 color) {
    v3 finalColor = divv(color, add(color, mk(1.0f, 1.0f, 1.0f)));
    return mk(clampf(finalColor.x, 0.0f, 1.0f), clampf(finalColor.y, 0.0f, 1.0f), clampf(finalColor.z, 0.0f, 1.0f));
}

/* :159-172 */
static inline basis_t buildOrthonormalBasis(v3 normal) {
    v3 tangent;
    if (fabsf(normal.x) > 0.9f)
        tangent = normalize(sub(mk(0, 1, 0), scale(normal, dot(mk(0, 1, 0), normal))));
    else
        tangent = normalize(sub(mk(1, 0, 0), scale(normal, dot(mk(1, 0, 0), normal))));
    basis_t b;
    b.tangent = tangent;
    b.bitangent = cross(normal, tangent);
    return b;
}

/* :186-191 */
static inline float smithG1_GGX(float NoV, float roughness) {
    float a = roughness * roughness;
    float a2 = a * a;
    float NoV2 = NoV * NoV;
    return 2.0f / (1.0f + sqrtf(1.0f + a2 * (1.0f - NoV2) / NoV2));
}

/* :193-197 */
static inline float D_GGX(float NoH, float a) {
    float a2 = a * a;
    float f = (NoH * a2 - NoH) * NoH + 1.0f;
    return a2 / (RT_PI_F * f * f);
}

/* :199-201 */
static inline v3 F_Schlick(float LoH, v3 f0) {
    return add(f0, scale(sub(mk(1.0f, 1.0f, 1.0f), f0), rt_powf(1.0f - LoH, 5.0f)));
}

/* :203-208 */
static inline float V_SmithGGXCorrelated(float NoV, float NoL, float a) {
    float a2 = a * a;
    float GGXL = NoV * sqrtf((-NoL * a2 + NoL) * NoL + a2);
    float GGXV = NoL * sqrtf((-NoV * a2 + NoV) * NoV + a2);
    return 0.5f / (GGXV + GGXL);
}

/* :214-246 (same as mode A's, index is uint2) */
static inline ray_t generateCameraRay(const CameraGPU *camera, uint32_t ix, uint32_t iy, float jx, float jy) {
    int x = (int)ix, y = (int)iy;
    float aspectRatio = (float)(camera->resolution.x / camera->resolution.y);
    float halfWidth = rt_tanf(camera->horizontalFov / 2.0f);
    float halfHeight = halfWidth / aspectRatio;
    v3 w = neg(normalize(ld3(&camera->direction)));
    v3 u = normalize(cross(ld3(&camera->up), w));
    v3 v = normalize(cross(w, u));
    float s = (((float)x + jx) / (float)camera->resolution.x) * 2.0f - 1.0f;
    float t = -((((float)y + jy) / (float)camera->resolution.y) * 2.0f - 1.0f);
    v3 dir = normalize(sub(add(scale(u, s * halfWidth), scale(v, t * halfHeight)), w));
    ray_t r;
    r.origin = ld3(&camera->position);
    r.direction = dir;
    r.min_distance = 0.001f;
    r.max_distance = 1000.0f;
    return r;
}

/* :259-289.  `rayDirection` is ray.direction of the reference's first argument. */
static inline v3 calculateBRDFContribution(v3 rayDirection, v3 normal, const isect_t *m, v3 lightDir) {
    v3 v = neg(normalize(rayDirection));
    v3 n = normal;
    v3 l = lightDir;
    v3 h = normalize(add(v, l));

    float NoV = fabsf(dot(n, v)) + 1e-5f;
    float NoL = fminf(1.0f, fmaxf(0.0f, dot(n, l)));
    float NoH = fminf(1.0f, fmaxf(0.0f, dot(n, h)));
    float LoH = fminf(1.0f, fmaxf(0.0f, dot(l, h)));

    v3 dielectricF0 = mk(0.04f, 0.04f, 0.04f);
    v3 f0 = add(dielectricF0, scale(sub(m->diffuse, dielectricF0), m->metallic));   /* mix(x,y,a) = x + (y-x)*a */

    float D = D_GGX(NoH, m->roughness);
    v3 F = F_Schlick(LoH, f0);
    float G = V_SmithGGXCorrelated(NoV, NoL, m->roughness);

    v3 Fr = divs(scale(F, D * G), 4.0f * NoV * NoL + 1e-7f);
    v3 Fd = scale(m->diffuse, 1.0f / RT_PI_F);

    v3 energyCompensation = sub(mk(1.0f, 1.0f, 1.0f), F);
    v3 kD = scale(energyCompensation, 1.0f - m->metallic);
    v3 BRDF = mul(kD, add(Fd, Fr));
    return scale(BRDF, NoL);
}

/* test hook (rt_groups.c: rt_oracle_cullcheck_modeb): every closest-hit query of the frame with its answer; seg = the ray is
 * a segment towards a light sample point (directSquareLightRay), whose end is then in rt_oracle_modeb_segment_end */
void (*rt_oracle_modeb_query_hook)(const float o[3], const float d[3], float tmin, float tmax, int seg, const float end[3],
                                   int hit, float distance, uint32_t primitive_id);
static _Thread_local float t_segment_end[3];
static _Thread_local int t_segment;

/* :291-313 */
static inline ray_t directSquareLightRay(v3 origin, const SquareLightGPU *light, float rx, float ry) {
    v3 lightNormal = mk(0, -1, 0);
    basis_t basis = buildOrthonormalBasis(lightNormal);
    v3 center = ld3(&light->center);
    float x = (rx - 0.5f) * light->width;
    float y = (ry - 0.5f) * light->depth;
    v3 samplePos = add(add(center, scale(basis.tangent, x)), scale(basis.bitangent, y));
    v3 toLight = sub(samplePos, origin);
    float distance = length(toLight);
    v3 lightDir = divs(toLight, distance);
    ray_t r;
    r.origin = origin;
    r.direction = lightDir;
    r.min_distance = 0.001f;
    r.max_distance = distance;
    t_segment_end[0] = samplePos.x; t_segment_end[1] = samplePos.y; t_segment_end[2] = samplePos.z; t_segment = 1;
    return r;
}

/* :315-326 */
static inline float calculateSquareLightPdf(v3 origin, const SquareLightGPU *light, v3 direction) {
    v3 lightNormal = mk(0, -1, 0);
    v3 toLight = sub(ld3(&light->center), origin);
    float distance = length(toLight);
    float cosTheta = fmaxf(0.0f, dot(neg(direction), lightNormal));
    float area = light->width * light->depth;
    return (distance * distance) / (area * cosTheta + 1e-6f);
}

/* :355-374 */
static inline ray_t cosineWeightedRay(v3 origin, v3 normal, float rx, float ry) {
    float phi = 2.0f * RT_PI_F * rx;
    float cosTheta = sqrtf(ry);
    float sinTheta = sqrtf(1.0f - ry);
    basis_t basis = buildOrthonormalBasis(normal);
    v3 direction = normalize(add(add(scale(basis.tangent, rt_cosf(phi) * sinTheta),
                                     scale(basis.bitangent, rt_sinf(phi) * sinTheta)),
                                 scale(normal, cosTheta)));
    ray_t r;
    r.origin = origin;
    r.direction = direction;
    r.min_distance = 0.001f;
    r.max_distance = 1000.0f;
    return r;
}

/* :376-380 */
static inline float calculateCosineWeightedPdf(v3 normal, v3 direction) {
    float cosTheta = fmaxf(0.0f, dot(normal, direction));
    return cosTheta / RT_PI_F;
}

/* :382-434 */
static inline ray_t vndfRay(v3 origin, v3 viewDir, v3 normal, float roughness, float rx, float ry) {
    float alpha = roughness * roughness;
    basis_t basis = buildOrthonormalBasis(normal);
    v3 Ve = normalize(mk(alpha * dot(viewDir, basis.tangent), alpha * dot(viewDir, basis.bitangent), dot(viewDir, normal)));
    v3 T1 = normalize(mk(Ve.z, 0, -Ve.x));
    v3 T2 = cross(Ve, T1);
    float phi = 2.0f * RT_PI_F * rx;
    float lenVe = length(Ve);
    float cosThetaMax = lenVe / sqrtf(1.0f + lenVe * lenVe);
    float cosTheta = cosThetaMax + (1.0f - cosThetaMax) * ry;
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    v3 h = normalize(add(add(scale(T1, rt_cosf(phi) * sinTheta), scale(T2, rt_sinf(phi) * sinTheta)), scale(Ve, cosTheta)));
    v3 Nh = normalize(mk(alpha * h.x, alpha * h.y, fmaxf(0.0f, h.z)));
    v3 worldH = normalize(add(add(scale(basis.tangent, Nh.x), scale(basis.bitangent, Nh.y)), scale(normal, Nh.z)));
    /* reflect(i, n) = i - 2 * dot(n, i) * n */
    v3 i = neg(viewDir);
    v3 direction = sub(i, scale(worldH, 2.0f * dot(worldH, i)));
    ray_t r;
    r.origin = origin;
    r.direction = direction;
    r.min_distance = 0.001f;
    r.max_distance = 1000.0f;
    return r;
}

/* :436-445 */
static inline float calculateVNDFPdf(v3 viewDir, v3 normal, v3 lightDir, float roughness) {
    v3 h = normalize(add(viewDir, lightDir));
    float NoH = fabsf(dot(normal, h));
    float VoH = fabsf(dot(viewDir, h));
    float NoV = fabsf(dot(normal, viewDir));
    float D = D_GGX(NoH, roughness);
    float G1 = smithG1_GGX(NoV, roughness);
    return (D * G1 * VoH) / (4.0f * NoV);
}

/* :447-457 */
static inline v3 getTriangleNormal(const simd_float3 *vertices, uint32_t primitiveId) {
    v3 v0 = ld3(&vertices[primitiveId * 3 + 0]);
    v3 v1 = ld3(&vertices[primitiveId * 3 + 1]);
    v3 v2 = ld3(&vertices[primitiveId * 3 + 2]);
    return normalize(cross(sub(v1, v0), sub(v2, v0)));
}

/* :459-509 */
static inline isect_t getClosestIntersection(const ctx_t *c, ray_t in) {
    float o[3] = { in.origin.x, in.origin.y, in.origin.z }, d[3] = { in.direction.x, in.direction.y, in.direction.z };
    c->cnt[RT_CNT_CLOSEST_RAYS]++;
    rt_hit h = rt_intersect((const float *)c->vertices, c->T, NULL, 0, o, d, in.min_distance, in.max_distance, 0);
    if (rt_oracle_modeb_query_hook)
        rt_oracle_modeb_query_hook(o, d, in.min_distance, in.max_distance, t_segment, t_segment_end, h.hit, h.distance, h.primitive_id);
    t_segment = 0;
    isect_t r;
    memset(&r, 0, sizeof r);
    if (!h.hit) { r.type = MISS; return r; }
    const MaterialGPU *m = &c->materials[h.primitive_id];
    r.diffuse = mk(m->diffuse.x, m->diffuse.y, m->diffuse.z);
    r.emissive = ld3(&m->emissive);
    r.metallic = m->metallic;
    r.roughness = m->roughness;
    r.point = add(in.origin, scale(in.direction, h.distance));
    if (length(r.emissive) > 0.0f) { r.type = HITLIGHT; return r; }
    r.type = HIT;
    r.ray = in;
    r.normal = getTriangleNormal(c->vertices, h.primitive_id);
    return r;
}

/* :519-541 */
static inline v3 calculateDirectLightSamplingContribution(const ctx_t *c, const isect_t *in, float rx, float ry,
                                                          uint32_t samplesPerStrategy, int usePowerHeuristic) {
    const SquareLightGPU *light = c->light;
    v3 directLight = mk(0.0f, 0.0f, 0.0f);
    ray_t lightRay = directSquareLightRay(add(in->point, scale(in->normal, 1e-4f)), light, rx, ry);
    float directLightPDF = calculateSquareLightPdf(in->point, light, lightRay.direction);
    float cosinePDF = calculateCosineWeightedPdf(in->normal, lightRay.direction);
    float vndfPDF = calculateVNDFPdf(neg(in->ray.direction), in->normal, lightRay.direction, in->roughness);

    isect_t li = getClosestIntersection(c, lightRay);
    if (li.type == HITLIGHT) {
        v3 brdf = calculateBRDFContribution(in->ray.direction, in->normal, in, lightRay.direction);
        v3 er = ld3(&light->emittedRadiance);
        if (usePowerHeuristic) {
            float weight = powerHeuristic(directLightPDF, cosinePDF, vndfPDF, samplesPerStrategy, 1.0f);
            directLight = add(directLight, divs(mul(scale(brdf, weight), er), directLightPDF));
        } else {
            directLight = add(directLight, divs(mul(brdf, er), directLightPDF));
        }
    }
    return directLight;
}

/* :543-625 */
static inline v3 recursiveMultiImportanceSampling(const ctx_t *c, const isect_t *in, uint32_t samples) {
    const SquareLightGPU *light = c->light;
    uint32_t sps = samples / 3;
    v3 directLight = mk(0, 0, 0), cosine = mk(0, 0, 0), vndf = mk(0, 0, 0);
    v3 er = ld3(&light->emittedRadiance);
    v3 origin = add(in->point, scale(in->normal, 1e-4f));
    v3 view = neg(in->ray.direction);

    for (uint32_t i = 0; i < sps; i++) {
        float ux, uy;
        haltonRandom(i, 0, &ux, &uy);
        directLight = add(directLight, calculateDirectLightSamplingContribution(c, in, ux, uy, sps, 1));
    }
    for (uint32_t i = 0; i < sps; i++) {
        float ux, uy;
        haltonRandom(i + sps, 2, &ux, &uy);
        ray_t lightRay = cosineWeightedRay(origin, in->normal, ux, uy);
        float cosinePDF = calculateCosineWeightedPdf(in->normal, lightRay.direction);
        float directLightPDF = calculateSquareLightPdf(in->point, light, lightRay.direction);
        float vndfPDF = calculateVNDFPdf(view, in->normal, lightRay.direction, in->roughness);
        float weight = powerHeuristic(cosinePDF, directLightPDF, vndfPDF, sps, 1.0f);
        isect_t ci = getClosestIntersection(c, lightRay);
        v3 brdf = calculateBRDFContribution(in->ray.direction, in->normal, in, lightRay.direction);
        if (ci.type == HITLIGHT) {
            cosine = add(cosine, divs(mul(scale(brdf, weight), er), cosinePDF));
        } else if (ci.type == HIT) {
            float vx, vy;
            haltonRandom(i, 6, &vx, &vy);
            cosine = add(cosine, mul(divs(brdf, cosinePDF), calculateDirectLightSamplingContribution(c, &ci, vx, vy, 1, 0)));
        }
    }
    for (uint32_t i = 0; i < sps; i++) {
        float ux, uy;
        haltonRandom(i + 2 * sps, 4, &ux, &uy);
        ray_t lightRay = vndfRay(origin, view, in->normal, in->roughness, ux, uy);
        float vndfPDF = calculateVNDFPdf(view, in->normal, lightRay.direction, in->roughness);
        float cosinePDF = calculateCosineWeightedPdf(in->normal, lightRay.direction);
        float directLightPDF = calculateSquareLightPdf(in->point, light, lightRay.direction);
        float weight = powerHeuristic(vndfPDF, directLightPDF, cosinePDF, sps, 1.0f);
        isect_t vi = getClosestIntersection(c, lightRay);
        v3 brdf = calculateBRDFContribution(in->ray.direction, in->normal, in, lightRay.direction);
        if (vi.type == HITLIGHT) {
            vndf = add(vndf, divs(mul(scale(brdf, weight), er), vndfPDF));
        } else if (vi.type == HIT) {
            float vx, vy;
            haltonRandom(i + sps, 6, &vx, &vy);
            vndf = add(vndf, mul(divs(brdf, vndfPDF), calculateDirectLightSamplingContribution(c, &vi, vx, vy, 1, 0)));
        }
    }
    return divs(add(add(directLight, cosine), vndf), (float)sps);
}

/* :635-707, one grid point */
static inline void drawTrianglePixel(const ctx_t *c, uint32_t ix, uint32_t iy, uint32_t cameraRaysPerPixel,
                                     uint32_t misSamples, uint8_t *pixel4, float *text3) {
    const CameraGPU *camera = c->camera;
    v3 accumulatedColor = mk(0.0f, 0.0f, 0.0f);
    for (uint32_t i = 0; i < cameraRaysPerPixel; i++) {
        float jx, jy;
        hashRandom(ix, iy, i, &jx, &jy);
        ray_t r = generateCameraRay(camera, ix, iy, jx, jy);
        isect_t in = getClosestIntersection(c, r);
        c->cnt[RT_CNT_PATHS]++;
        if (in.type == MISS) {
        } else if (in.type == HITLIGHT) {
            accumulatedColor = add(accumulatedColor, ld3(&c->light->emittedRadiance));
        } else {
            c->cnt[RT_CNT_SHADED_BOUNCES]++;
            accumulatedColor = add(accumulatedColor, recursiveMultiImportanceSampling(c, &in, misSamples));
        }
    }
    v3 exposedColor = scale(divs(accumulatedColor, (float)cameraRaysPerPixel), cameraExposure(camera));
    v3 toneMappedColor = reinhartToneMapping(exposedColor);
    float gamma = 2.2f;
    float e = 1.0f / gamma;
    v3 g = mk(rt_powf(toneMappedColor.x, e), rt_powf(toneMappedColor.y, e), rt_powf(toneMappedColor.z, e));
    text3[0] = accumulatedColor.x; text3[1] = accumulatedColor.y; text3[2] = accumulatedColor.z;   /* :627-633 */
    pixel4[0] = (uint8_t)(g.x * 255);                                                               /* :248-257 */
    pixel4[1] = (uint8_t)(g.y * 255);
    pixel4[2] = (uint8_t)(g.z * 255);
    pixel4[3] = 255;
}

int rt_oracle_drawtriangle(const CameraGPU *cameras, const MaterialGPU *materials,
                           const SquareLightGPU *lights, const simd_float3 *vertices,
                           uint8_t *pixels, float *text, const RTLaunchParams *p,
                           uint64_t counters[RT_CNT_COUNT], int nthreads) {
    if (!p || !cameras || !materials || !lights || !vertices || !pixels || !text) return RT_ERR_INVALID_ARGUMENT;
    if (p->width <= 0 || p->height <= 0 || p->row0 < 0 || p->rows < 0) return RT_ERR_INVALID_ARGUMENT;
    if (!p->tileRows && p->row0 + p->rows > p->height) return RT_ERR_INVALID_ARGUMENT;
    if (p->cameraRays == 0 || p->misSamples < 3) return RT_ERR_INVALID_ARGUMENT;
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
        ctx_t c = { cameras, materials, lights, vertices, p->numTriangles, cnt };
        const int chunks = (W + 15) / 16;        /* jobs of 16 pixels of one row */
#pragma omp for schedule(dynamic, 1)
        for (int job = 0; job < p->rows * chunks; job++) {
            const int ly = job / chunks, xb = (job % chunks) * 16, xe = xb + 16 < W ? xb + 16 : W;
            const uint32_t iy = p->tileRows
                ? (uint32_t)p->row0 + ((uint32_t)ly / p->tileRows) * p->tileStride + ((uint32_t)ly % p->tileRows)
                : (uint32_t)(p->row0 + ly);
            for (int x = xb; x < xe; x++) {
                const size_t px = (size_t)ly * W + x;
                /* shaders.metal:642: threads outside the camera resolution do nothing */
                if (x >= cameras[0].resolution.x || (int)iy >= cameras[0].resolution.y) continue;
                drawTrianglePixel(&c, (uint32_t)x, iy, p->cameraRays, p->misSamples, pixels + px * 4, text + px * 3);
            }
        }
#pragma omp critical
        for (int k = 0; k < RT_CNT_COUNT; k++) total[k] += cnt[k];
    }
    if (counters) for (int k = 0; k < RT_CNT_COUNT; k++) counters[k] = total[k];
    return RT_OK;
}
