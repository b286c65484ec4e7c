/*
 * rt_device_math.cuh -- device-side arithmetic of the path-tracing kernels.
 *
 * The translation unit is compiled with --fmad=false, IEEE division and square
 * root and no flush-to-zero, so that
 *   - a * b + c written in the kernels is two roundings (the reference's shader
 *     expressions, RTrace/raytrace.metal and RTrace/sampling.metal, evaluated
 *     operator by operator), and
 *   - a fused multiply-add happens exactly where fmaf() is written: inside the
 *     functions that stand in for the Metal standard library and the Metal
 *     ray/triangle query (both closed source; definitions in DESIGN.md section
 *     "Arithmetic contract").
 * With that, per-pixel results do not depend on how the kernel schedules the
 * work, and discrete outcomes (hit / miss, shadowed / lit) cannot flip.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rt {

struct f3 { float x, y, z; };

__device__ __forceinline__ f3 mk(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ f3 operator-(f3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ f3 operator/(f3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ f3 operator/(f3 a, f3 b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }

/* ---- Metal stdlib stand-ins (fused where written) ---------------------------- */
__device__ __forceinline__ float dot(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
    return mk(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
__device__ __forceinline__ float length(f3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ f3 normalize(f3 a) { float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
__device__ __forceinline__ float saturate(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }

/* sincos: two-step Cody-Waite reduction by pi/2, Cephes single-precision polynomials */
__device__ __forceinline__ void sincos_det(float x, float &s_out, float &c_out) {
    const float TWO_OVER_PI = 0.6366197466850281f;
    const float PIO2_HI = 1.5707963705062866f;
    const float PIO2_LO = -4.371138828673793e-08f;
    float kf = floorf(fmaf(x, TWO_OVER_PI, 0.5f));
    int k = (int)kf;
    float r = fmaf(-kf, PIO2_HI, x);
    r = fmaf(-kf, PIO2_LO, r);
    float r2 = r * r;
    float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = fmaf(r2, ps, -1.6666654611e-1f);
    float s = fmaf(r * r2, ps, r);
    float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = fmaf(r2, pc, 4.166664568298827e-2f);
    float c = fmaf(r2 * r2, pc, fmaf(r2, -0.5f, 1.0f));
    /* quadrant rotation without a switch: swap on odd k, negate by bit tests */
    bool sw = k & 1;
    float ss = sw ? c : s;
    float cc = sw ? s : c;
    s_out = (k & 2) ? -ss : ss;
    c_out = ((k + 1) & 2) ? -cc : cc;
}
__device__ __forceinline__ float tan_det(float x) { float s, c; sincos_det(x, s, c); return s / c; }

/* ---- pow = exp2(y * log2(x)): same split, polynomials and special cases on the CPU side
 * of the arithmetic contract.  Call sites are mode B's:
 * Sources/gpuRaytracer/shaders.metal:132-137,145-150,199-201,702-703. */
__device__ __forceinline__ float log2_det(float x) {
    int e = 0;
    if (x < 1.17549435e-38f) { x = x * 16777216.0f; e = -24; }
    uint32_t bits = __float_as_uint(x);
    e += (int)((bits >> 23) & 0xffu) - 126;
    float m = __uint_as_float((bits & 0x007fffffu) | 0x3f000000u);
    if (m < 0.70710678118654752440f) { e -= 1; m = (m + m) - 1.0f; } else { m = m - 1.0f; }
    float z = m * m;
    float p = fmaf(7.0376836292e-2f, m, -1.1514610310e-1f);
    p = fmaf(p, m, 1.1676998740e-1f);
    p = fmaf(p, m, -1.2420140846e-1f);
    p = fmaf(p, m, 1.4249322787e-1f);
    p = fmaf(p, m, -1.6668057665e-1f);
    p = fmaf(p, m, 2.0000714765e-1f);
    p = fmaf(p, m, -2.4999993993e-1f);
    p = fmaf(p, m, 3.3333331174e-1f);
    float y = (p * m) * z;
    y = fmaf(-0.5f, z, y);
    float ln = m + y;
    return fmaf(ln, 1.44269504088896341f, (float)e);
}
__device__ __forceinline__ float exp2_det(float t) {
    if (!(t > -126.0f)) return 0.0f;
    if (t >= 128.0f) return __int_as_float(0x7f800000);
    float nf = floorf(t + 0.5f);
    float f = t - nf;
    float p = fmaf(1.535336188319500e-4f, f, 1.339887440266574e-3f);
    p = fmaf(p, f, 9.618437357674640e-3f);
    p = fmaf(p, f, 5.550332471162809e-2f);
    p = fmaf(p, f, 2.402264791363012e-1f);
    p = fmaf(p, f, 6.931472028550421e-1f);
    p = fmaf(p, f, 1.0f);
    int n = (int)nf;
    int n1 = n / 2, n2 = n - n1;
    float s1 = __uint_as_float((uint32_t)(n1 + 127) << 23), s2 = __uint_as_float((uint32_t)(n2 + 127) << 23);
    return (p * s1) * s2;
}
__device__ __forceinline__ float pow_det(float x, float y) {
    if (y == 1.0f) return x;
    if (y == 0.0f || x == 1.0f) return 1.0f;
    if (x == 0.0f) return y > 0.0f ? 0.0f : __int_as_float(0x7f800000);
    if (!(x > 0.0f)) return __int_as_float(0x7fc00000);
    if (isinf(x)) return y > 0.0f ? __int_as_float(0x7f800000) : 0.0f;
    return exp2_det(y * log2_det(x));
}

/* buildOrthonormalBasis, Sources/gpuRaytracer/shaders.metal:159-172 */
__device__ __forceinline__ void orthonormal_basis(f3 normal, f3 &tangent, f3 &bitangent) {
    if (fabsf(normal.x) > 0.9f) tangent = normalize(mk(0.0f, 1.0f, 0.0f) - normal * dot(mk(0.0f, 1.0f, 0.0f), normal));
    else                        tangent = normalize(mk(1.0f, 0.0f, 0.0f) - normal * dot(mk(1.0f, 0.0f, 0.0f), normal));
    bitangent = cross(normal, tangent);
}

/* ---- Halton, RTrace/sampling.metal:97-122 ------------------------------------
 * halton(i, d): f = 1; invB = 1.0f / b; while (i) { f *= invB; r += f * (i % b); i /= b; }
 * The products f_k = f_{k-1} * invB are the same rounded sequence whatever i is,
 * so they are compile-time constants here; the digits come out of divisions by a
 * compile-time base (multiply-high + shift).  DIGITS covers i < 2^22 (B^DIGITS >=
 * 2^22); larger i fall through to the generic loop, so the function is exact for
 * every 32-bit index.  A zero digit adds f*0 = +0 and leaves r unchanged, which
 * is why running all DIGITS steps equals the reference's early-exit loop.
 */
__host__ __device__ constexpr uint32_t rt_prime(int d) {
    constexpr uint32_t P[24] = { 2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37,
                                 41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89 };
    return P[d];
}
__host__ __device__ constexpr int rt_digits_for(uint32_t b, uint64_t limit) {
    int k = 0; uint64_t p = 1;
    while (p < limit) { p *= b; k++; }
    return k;
}
__host__ __device__ constexpr float rt_halton_f(uint32_t b, int k) {   /* f after k+1 products */
    float invB = 1.0f / (float)b;
    float f = 1.0f;
    for (int j = 0; j <= k; j++) f = f * invB;
    return f;
}

template <int DIM>
__device__ __forceinline__ float halton(uint32_t i) {
    constexpr uint32_t B = rt_prime(DIM);
    if (DIM == 0 && i < (1u << 24)) {
        /* base 2: every partial sum is exact while i < 2^24 */
        return (float)__brev(i) * 2.3283064365386963e-10f;
    }
    constexpr int DIGITS = rt_digits_for(B, 1ull << 22);
    float r = 0.0f;
#pragma unroll
    for (int k = 0; k < DIGITS; k++) {
        uint32_t q = i / B;
        uint32_t dgt = i - q * B;
        r = r + rt_halton_f(B, k) * (float)dgt;
        i = q;
    }
    if (i > 0) {                                   /* never taken for i < 2^22 */
        float invB = 1.0f / (float)B;
        float f = rt_halton_f(B, DIGITS - 1);
        while (i > 0) {
            f = f * invB;
            r = r + f * (float)(i % B);
            i = i / B;
        }
    }
    return r;
}

/* ---- Halton with a warp-uniform run-time dimension ------------------------------
 * The path tracer's bounce loop is rolled, so the dimension 2 + 5*bounce + SLOT is a
 * run-time value, but the same for every active lane of the warp.  Base, magic
 * multiplier and the f_k products come out of a constant-bank table through the
 * uniform datapath; the digit count is fixed per SLOT at compile time (the largest
 * any bounce needs: extra digits are zero and add +0).  Digits: q = umulhi(i, M)
 * with M = ceil(2^32 / B) is floor(i / B) exactly for i < 2^32 / B (error term
 * i * (M*B - 2^32) < 2^32 needs i < 2^32 / 89), far above the 2^22 fast-path bound.
 */
#define RT_HAL_COLS 16
struct HaltonTable {
    float    f[24][RT_HAL_COLS];     /* f[d][k] = invB^(k+1), rounded after every product */
    uint32_t B[24], M[24];
};
__host__ __device__ constexpr HaltonTable rt_make_halton_table() {
    HaltonTable t{};
    for (int d = 0; d < 24; d++) {
        uint32_t b = rt_prime(d);
        t.B[d] = b;
        t.M[d] = (uint32_t)(((1ull << 32) + b - 1) / b);
        for (int k = 0; k < RT_HAL_COLS; k++) t.f[d][k] = rt_halton_f(b, k);
    }
    return t;
}
__constant__ HaltonTable c_halton = rt_make_halton_table();

__device__ __noinline__ float halton_rt(uint32_t i, uint32_t d);

/* SLOT 0..3 of bounce b: dimension 2 + 5*b + SLOT (raytrace.metal:72-73,93-94) */
template <int DIGITS>
__device__ __forceinline__ float halton_digits(uint32_t i, uint32_t dim) {
    static_assert(DIGITS <= RT_HAL_COLS, "table too narrow");
    const uint32_t B = c_halton.B[dim], M = c_halton.M[dim];
    float r = 0.0f;
#pragma unroll
    for (int k = 0; k < DIGITS; k++) {
        uint32_t q = __umulhi(i, M);
        uint32_t dgt = i - q * B;
        r = r + c_halton.f[dim][k] * (float)dgt;
        i = q;
    }
    return r;
}
template <int SLOT>
__device__ __forceinline__ float halton_slot(uint32_t i, uint32_t bounce) {
    /* digits needed for i < 2^22: in the slot's base of bounce 0 (5, 7, 11, 13), and in the smallest base any later
     * bounce gives it (19, 23, 29, 31) -- 31 digit steps per bounce against 21.  The bounce is warp-uniform. */
    constexpr int DIGITS0 = rt_digits_for(rt_prime(2 + SLOT), 1ull << 22);
    constexpr int DIGITS1 = rt_digits_for(rt_prime(7 + SLOT), 1ull << 22);
    const uint32_t dim = 2u + 5u * bounce + (uint32_t)SLOT;
    if (i >= (1u << 22)) return halton_rt(i, dim);
    return bounce == 0u ? halton_digits<DIGITS0>(i, dim) : halton_digits<DIGITS1>(i, dim);
}

/* run-time dimension (self-test and mode B) */
__device__ __noinline__ float halton_rt(uint32_t i, uint32_t d) {
    uint32_t b = c_halton.B[d];
    float f = 1.0f, invB = 1.0f / (float)b, r = 0.0f;
    while (i > 0) { f = f * invB; r = r + f * (float)(i % b); i = i / b; }
    return r;
}

/* ---- Halton values out of a table in global memory ---------------------------------
 * halton(i, d) depends on the index and the dimension only, and the indices of a frame all lie in
 * [0, 2^20 + firstSample + spp) (offsets are arc4random() % 2^20, renderer.swift:100).  The four values of a
 * bounce (31 digit steps, ~170 instructions, a fifth of the sorted kernel) are therefore looked up: one 16-byte
 * load per path and bounce from a table of n x bounces float4 (55 MB for three bounces, 3 % of B200's HBM and
 * within its L2) that rt_build_halton filled with halton_slot<0..3>(i, b) -- the SAME device function, so the
 * values are the computed ones bit for bit; camY holds halton<1>(i) (14 base-3 digits) for the camera jitter.
 * An index beyond the table (offsets a caller chose larger) falls back to the computation, out of line.
 * rtcuda.cpp owns the tables (per device, built on first use, grown on demand). */
struct RTHaltonTables {
    const float4 *bounce;      /* [bounces][n] */
    const float  *camY;        /* [n] */
    uint32_t n, bounces;
};
__device__ __noinline__ float4 halton_bounce4_compute(uint32_t i, uint32_t b) {
    return make_float4(halton_slot<0>(i, b), halton_slot<1>(i, b), halton_slot<2>(i, b), halton_slot<3>(i, b));
}
__device__ __noinline__ float halton_cam_y_compute(uint32_t i) { return halton<1>(i); }
__device__ __forceinline__ float4 halton_bounce4(const RTHaltonTables &ht, uint32_t i, uint32_t b) {
    if (i < ht.n && b < ht.bounces) return __ldg(ht.bounce + (size_t)b * ht.n + i);
    return halton_bounce4_compute(i, b);
}
__device__ __forceinline__ float halton_cam_y(const RTHaltonTables &ht, uint32_t i) {
    if (i < ht.n) return __ldg(ht.camY + i);
    return halton_cam_y_compute(i);
}

/* sampling.metal:68-75 */
__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

}  // namespace rt
