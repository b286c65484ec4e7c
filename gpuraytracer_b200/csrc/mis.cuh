/*
 * mis.cuh -- multiple importance sampling in mode A (RT_VARIANT_MIS; included by pathtrace.cu).
 *
 * The reference's pathTrace weights nothing: its own TODO is the commented-out block of RTrace/sampling.metal:238-260,
 * which mode B carries out (Sources/gpuRaytracer/shaders.metal:519-541).  With the flag, the direct light of every
 * shaded bounce is estimated by the reference's light sample AND by the cosine-sampled continuation ray, combined with
 * the balance heuristic: powerHeuristic with beta = 1 (shaders.metal:132-137) over calculateSquareLightPdf (:315-326,
 * with the light's own normal) and calculateCosineWeightedPdf (:376-380).  oracle/rt_oracle.c ("multiple importance
 * sampling in mode A") states the estimator and why both strategies estimate the same integral; these are the same
 * operations in the same order, so the kernels stay bit-equal to it.
 */
#pragma once

#define RT_MIS_PI 3.14159265358979323846264338327950288f

__device__ __forceinline__ float misA_heuristic(float pdf1, float pdf2, float pdf3) {
    const float n = 1.0f;                        /* samplesPerStrategy = 1; pow(x, 1.0) = x */
    const float p1 = n * pdf1;
    const float sum = p1 + n * pdf2 + n * pdf3;
    return p1 / (sum + 1e-6f);
}
/* L: the prepared light (rt_scene_layout.h) */
__device__ __forceinline__ float misA_square_light_pdf(const float *L, f3 origin, f3 direction) {
    const f3 toLight = mk(L[0], L[1], L[2]) - origin;
    const float distance = length(toLight);
    const float cosTheta = fmaxf(0.0f, dot(-direction, mk(L[RT_LIGHT_NORMAL], L[RT_LIGHT_NORMAL + 1], L[RT_LIGHT_NORMAL + 2])));
    const float area = L[12] * L[13];
    return (distance * distance) / (area * cosTheta + 1e-6f);
}
__device__ __forceinline__ float misA_cosine_pdf(f3 normal, f3 direction) {
    return fmaxf(0.0f, dot(normal, direction)) / RT_MIS_PI;
}
/* the light strategy's weight for the sampled direction ldir */
__device__ __forceinline__ float misA_light_weight(const float *L, f3 P, f3 normal, f3 ldir) {
    return misA_heuristic(misA_square_light_pdf(L, P, ldir), misA_cosine_pdf(normal, ldir), 0.0f);
}
/* the cosine strategy's term of one light for the continuation ray (P, rd) whose closest hit is (hit, hitDistance) */
__device__ __forceinline__ f3 misA_cosine_term(const float *L, f3 P, f3 normal, f3 rd, bool hit, float hitDistance, f3 color) {
    const f3 zero = mk(0.0f, 0.0f, 0.0f);
    const f3 nL = mk(L[RT_LIGHT_NORMAL], L[RT_LIGHT_NORMAL + 1], L[RT_LIGHT_NORMAL + 2]);
    const f3 center = mk(L[0], L[1], L[2]);
    const float denom = dot(rd, nL);
    if (!(denom < 0.0f)) return zero;                                   /* the rectangle's front side faces -normal */
    const float t = dot(center - P, nL) / denom;
    if (!(t > 0.0f)) return zero;
    if (hit && hitDistance < t - 1e-3f) return zero;                    /* something nearer: raytrace.metal:81's rule */
    const f3 v = (P + rd * t) - center;
    const float hx = L[3], hz = L[7];
    if (!(fabsf(dot(v, mk(L[RT_LIGHT_TANGENT], L[RT_LIGHT_TANGENT + 1], L[RT_LIGHT_TANGENT + 2]))) <= hx &&
          fabsf(dot(v, mk(L[RT_LIGHT_BITANGENT], L[RT_LIGHT_BITANGENT + 1], L[RT_LIGHT_BITANGENT + 2]))) <= hz)) return zero;
    const float weight = misA_heuristic(misA_cosine_pdf(normal, rd), misA_square_light_pdf(L, P, rd), 0.0f);
    const float area = (2.0f * hx) * (2.0f * hz);
    f3 term = mk(L[4], L[5], L[6]);
    term = term * (RT_MIS_PI / area);
    term = term * weight;
    return term * color;
}
