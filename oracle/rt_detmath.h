/*
 * rt_detmath.h -- TEST INFRASTRUCTURE (oracle).  Not part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build or call anything under oracle/.
 *
 * What this file is: the fp32 definition of everything the reference shaders
 * get from OUTSIDE their own source, i.e. from the Metal standard library and
 * from the Metal ray-tracing runtime, neither of which has source in
 * /root/reference ("third-party arithmetic"):
 *
 *   dot / cross / length / normalize / saturate / sqrt     <metal_stdlib>
 *   sincos / sin / cos / tan                               <metal_stdlib>
 *   intersector<triangle_data>::intersect (closest / any)  <metal_raytracing>
 *       call sites: RTrace/raytrace.metal:26-30,48-49,83-85
 *                   Sources/gpuRaytracer/shaders.metal:459-483
 *
 * Metal's own versions are closed source and the reference builds with
 * MTL_FAST_MATH=YES (RTrace.xcodeproj/project.pbxproj:198,256), so there is no
 * bit-level ground truth to copy.  The rules fixed here are:
 *
 *   R1  all arithmetic is IEEE-754 binary32, round to nearest even, denormals
 *       kept; `/` and sqrtf are correctly rounded.
 *   R2  the LIBRARY functions below use fused multiply-add exactly where an
 *       fmaf() is written, nowhere else.
 *   R3  SHADER-level expressions (everything written in the .metal files) are
 *       evaluated as C would evaluate them in float: one rounding per
 *       operator, left to right, NO contraction.  That is what makes the
 *       unmodified .metal source compiled by g++ (-ffp-contract=off
 *       -fsingle-precision-constant, oracle/_ref) bit-identical to the C
 *       restatement in rt_oracle.c and to the CUDA kernel.
 *
 * Build flags that these rules need: gcc -O2 -ffp-contract=off -fno-fast-math
 * (-mfma only makes fmaf() an instruction instead of a libm call).
 */
#ifndef RT_DETMATH_H
#define RT_DETMATH_H

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __cplusplus
#define RT_INLINE static inline
#else
#define RT_INLINE static inline
#endif

/* ---- <metal_stdlib> geometric functions ---------------------------------- */

/* dot(a,b) = fma(az,bz, fma(ay,by, ax*bx)) */
RT_INLINE float rt_dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return fmaf(az, bz, fmaf(ay, by, ax * bx));
}

/* cross(a,b): each component is  fma(p, q, -(r*s))  */
RT_INLINE void rt_cross3(float ax, float ay, float az, float bx, float by, float bz,
                         float *cx, float *cy, float *cz) {
    *cx = fmaf(ay, bz, -(az * by));
    *cy = fmaf(az, bx, -(ax * bz));
    *cz = fmaf(ax, by, -(ay * bx));
}

/* length(v) = sqrt(dot(v,v)) */
RT_INLINE float rt_length3(float x, float y, float z) {
    return sqrtf(rt_dot3(x, y, z, x, y, z));
}

/* normalize(v) = v * (1 / sqrt(dot(v,v)))   (one sqrt, one division, three products) */
RT_INLINE void rt_normalize3(float *x, float *y, float *z) {
    float inv = 1.0f / sqrtf(rt_dot3(*x, *y, *z, *x, *y, *z));
    *x = *x * inv; *y = *y * inv; *z = *z * inv;
}

/* saturate(x) = min(max(x,0),1); a NaN input gives 0 (fmaxf drops the NaN) */
RT_INLINE float rt_saturate(float x) {
    return fminf(fmaxf(x, 0.0f), 1.0f);
}

/* ---- <metal_stdlib> trigonometry ----------------------------------------- */
/*
 * sincos(x): Cody-Waite reduction by pi/2 in two steps, then the classic
 * minimax polynomials on [-pi/4, pi/4] (single-precision Cephes coefficients).
 * Accurate to about 1 ulp for |x| < 1e3, which covers phi = 2*pi*u, u in [0,1).
 */
#define RT_TWO_OVER_PI 0.6366197466850281f        /* (float)(2/pi), 0x3f22f983              */
#define RT_PIO2_HI     1.5707963705062866f        /* (float)(pi/2), 0x3fc90fdb              */
#define RT_PIO2_LO     (-4.371138828673793e-08f)  /* (float)(pi/2 - RT_PIO2_HI), 0xb33bbd2e */

RT_INLINE void rt_sincosf(float x, float *s_out, float *c_out) {
    float kf = floorf(fmaf(x, RT_TWO_OVER_PI, 0.5f));
    int   k  = (int)kf;
    float r  = fmaf(-kf, RT_PIO2_HI, x);
    r        = fmaf(-kf, RT_PIO2_LO, r);
    float r2 = r * r;
    /* sin(r) = r + r*r2*(S1 + r2*(S2 + r2*S3)) */
    float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps       = fmaf(r2, ps, -1.6666654611e-1f);
    float s  = fmaf(r * r2, ps, r);
    /* cos(r) = 1 - r2/2 + r2*r2*(C1 + r2*(C2 + r2*C3)) */
    float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc       = fmaf(r2, pc, 4.166664568298827e-2f);
    float c  = fmaf(r2 * r2, pc, fmaf(r2, -0.5f, 1.0f));
    switch (k & 3) {
        case 0:  *s_out =  s; *c_out =  c; break;
        case 1:  *s_out =  c; *c_out = -s; break;
        case 2:  *s_out = -s; *c_out = -c; break;
        default: *s_out = -c; *c_out =  s; break;
    }
}
RT_INLINE float rt_sinf(float x) { float s, c; rt_sincosf(x, &s, &c); return s; }
RT_INLINE float rt_cosf(float x) { float s, c; rt_sincosf(x, &s, &c); return c; }
/* tan(x) = sin(x) / cos(x) */
RT_INLINE float rt_tanf(float x) { float s, c; rt_sincosf(x, &s, &c); return s / c; }

/* ---- <metal_stdlib> pow ----------------------------------------------------
 * pow(x, y) = exp2(y * log2(x)), which is how a fast-math GPU library evaluates
 * it (Sources/gpuRaytracer/shaders.metal:132-137,145-150,199-201,702-703 are the
 * call sites).  log2: exponent/mantissa split with the mantissa in
 * [sqrt(1/2), sqrt(2)), Cephes logf polynomial, one fma by log2(e).  exp2: split
 * into integer and fraction in [-1/2, 1/2], Cephes exp2f polynomial, exponent
 * added by bit arithmetic.  Exact special cases, as IEEE pow has them:
 * pow(x, 1) = x, pow(x, 0) = 1, pow(1, y) = 1, pow(0, y > 0) = 0; x < 0 is NaN.
 */
RT_INLINE float rt_log2f(float x) {          /* x > 0, finite */
    uint32_t bits; int e = 0;
    if (x < 1.17549435e-38f) { x = x * 16777216.0f; e = -24; }
    memcpy(&bits, &x, 4);
    e += (int)((bits >> 23) & 0xffu) - 126;
    bits = (bits & 0x007fffffu) | 0x3f000000u;           /* m in [0.5, 1) */
    float m; memcpy(&m, &bits, 4);
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

RT_INLINE float rt_exp2f(float t) {
    if (!(t > -126.0f)) return 0.0f;                     /* underflow (and NaN) -> 0 */
    if (t >= 128.0f) return INFINITY;
    float nf = floorf(t + 0.5f);
    float f = t - nf;                                     /* [-0.5, 0.5], exact */
    float p = fmaf(1.535336188319500e-4f, f, 1.339887440266574e-3f);
    p = fmaf(p, f, 9.618437357674640e-3f);
    p = fmaf(p, f, 5.550332471162809e-2f);
    p = fmaf(p, f, 2.402264791363012e-1f);
    p = fmaf(p, f, 6.931472028550421e-1f);
    p = fmaf(p, f, 1.0f);                                 /* 2^f in [0.707, 1.415] */
    int n = (int)nf;
    /* scale by 2^n in two exact steps so that n = 128 / -126 stay in range */
    int n1 = n / 2, n2 = n - n1;
    uint32_t b1 = (uint32_t)(n1 + 127) << 23, b2 = (uint32_t)(n2 + 127) << 23;
    float s1, s2; memcpy(&s1, &b1, 4); memcpy(&s2, &b2, 4);
    return (p * s1) * s2;
}

RT_INLINE float rt_powf(float x, float y) {
    if (y == 1.0f) return x;
    if (y == 0.0f || x == 1.0f) return 1.0f;
    if (x == 0.0f) return y > 0.0f ? 0.0f : INFINITY;
    if (!(x > 0.0f)) return NAN;
    if (isinf(x)) return y > 0.0f ? INFINITY : 0.0f;
    return rt_exp2f(y * rt_log2f(x));
}

/* ---- <metal_raytracing> intersector ---------------------------------------
 *
 * Replacement for the hardware BVH query: a brute-force loop over the T
 * triangles of the vertex buffer (stride 16 bytes, three vertices per
 * triangle, primitive_id = triangle index; reference
 * RTrace/computeShader.swift:47-66, RTrace/renderer.swift:51-57) with a
 * division-free ray/triangle test in the "precomputed normal" form of
 * Moller-Trumbore.  Per triangle (v0,v1,v2), independent of the ray:
 *
 *     e1 = v1-v0   e2 = v2-v0   n = cross(e1, e2)   e1n = -e1
 *
 * and per ray (o,d):
 *
 *     w  = v0 - o              D  = dot(d, n)         Nn = dot(w, n)         (n = e2 x e1, e1n = v0 - v1)
 *     c  = cross(d, w)         U  = dot(e2, c)        V  = dot(e1n, c)
 *     a = |D|
 *
 * which are the classic quantities scaled by the determinant: (u,v,t) =
 * (U,V,Nn)/D, never formed inside the loop.  The hit is accepted when
 *     signbit(U) == signbit(V) == signbit(Nn) == signbit(D)     u >= 0, v >= 0, t >= 0, no culling
 *     |U+V| <= a                                   u + v <= 1, edge inclusive
 *     |Nn| >= tmin*a                               t >= tmin   (tmin == 0: the sign rule alone, no product)
 *     |Nn|*bestD < bestN*a                         t <  best so far (strict)
 * where (bestN,bestD) starts as (tmax,1), so the same comparison enforces
 * t < tmax, and a tie between two triangles keeps the LOWER primitive id.
 * (The sign rule is the integer form of "U, V, Nn flipped by sign(D) are >= 0": on the GPU it is two
 * three-input logic operations and one integer compare for all three quantities, where flipping and
 * comparing each one costs three more issue slots per test.  The two forms differ only for an exact
 * zero whose sign bit is not D's.  An exact zero of U or V -- a ray through an edge -- comes out of a
 * cancellation as +0, and with n = e2 x e1, c = d x w D is POSITIVE where the ray meets the side the
 * winding normal e1 x e2 points to: there both triangles of a shared edge accept and the lower id wins,
 * as with the inclusive comparison; on a back face such a ray is accepted by neither.  Because the sign
 * of such a zero now decides, NO quantity on the way to U, V, D, Nn is formed by a unary minus: a compiler
 * may move a negation through a product or into the logic operation that tests the sign bit, which is
 * exact for every value except the sign of an exact zero -- and exact zeros do occur, about twice per
 * 1e8 paths of the Cornell frame.)
 * D == 0 can never pass (|Nn|*bestD < bestN*0 is false).  The reported
 * distance is the single division bestN/bestD after the loop.
 * Cost: 21 multiply/add instructions per test (w 3, D 3, Nn 3, c 6, U 3, V 3),
 * 32 flop, + 4 flop for the comparisons' products and sum.
 *
 * accept_any_intersection(true) (shadow rays, raytrace.metal:83) answers only
 * hit / no hit; it is the same test and may stop at the first accepted
 * triangle.  Sphere primitives (config 4, an extension the live reference
 * kernels do not have; arithmetic after shaders_old.metal:108-136) are tested
 * after the triangles and carry primitive ids T..T+S-1; see rt_sphere_test.
 */
typedef struct {
    int      hit;           /* 0 = intersection_type::none */
    float    distance;
    uint32_t primitive_id;
} rt_hit;

/* One triangle test in the scaled form above.  Returns 1 and updates
 * (bestN,bestD) when the triangle is the new closest hit. */
RT_INLINE int rt_triangle_test(const float *v0, const float *v1, const float *v2,
                               const float o[3], const float d[3], float tmin,
                               float *bestN, float *bestD) {
    float e1x = v1[0] - v0[0], e1y = v1[1] - v0[1], e1z = v1[2] - v0[2];
    float e2x = v2[0] - v0[0], e2y = v2[1] - v0[1], e2z = v2[2] - v0[2];
    float nx, ny, nz, cx, cy, cz;
    float e1nx = v0[0] - v1[0], e1ny = v0[1] - v1[1], e1nz = v0[2] - v1[2];
    rt_cross3(e2x, e2y, e2z, e1x, e1y, e1z, &nx, &ny, &nz);             /* n = e2 x e1 */
    float wx = v0[0] - o[0], wy = v0[1] - o[1], wz = v0[2] - o[2];
    float D  = rt_dot3(d[0], d[1], d[2], nx, ny, nz);
    float Nn = rt_dot3(wx, wy, wz, nx, ny, nz);
    rt_cross3(d[0], d[1], d[2], wx, wy, wz, &cx, &cy, &cz);             /* c = d x w */
    float U = rt_dot3(e2x, e2y, e2z, cx, cy, cz);
    float V = rt_dot3(e1nx, e1ny, e1nz, cx, cy, cz);
    /* every scaled quantity must carry the SIGN BIT of D (an exact zero with the other sign bit does not) */
    int sD = signbit(D) != 0;
    if ((signbit(U) != 0) != sD || (signbit(V) != 0) != sD || (signbit(Nn) != 0) != sD) return 0;
    float a = fabsf(D), N = fabsf(Nn);
    /* t >= tmin: for tmin == 0 (shadow rays) the sign rule is the whole test and the product 0*a is not formed */
    int farEnough = tmin == 0.0f ? 1 : (N >= tmin * a);
    if (fabsf(U + V) <= a && farEnough && N * *bestD < *bestN * a) {
        *bestN = N; *bestD = a;
        return 1;
    }
    return 0;
}

/* Sphere test, same rational bookkeeping.  With m = o - c, A = dot(d,d),
 * B = dot(m,d), C = dot(m,m) - r*r the roots are t = (-B -+ sqrt(B*B - A*C))/A.
 * The nearer root is tried first, then the farther one (so a ray that starts
 * inside a sphere hits its far wall; shaders_old.metal:121-122 picks
 * min(t1,t2) whenever either is positive, which is wrong from inside). */
RT_INLINE int rt_sphere_test(const float c[3], float radius,
                             const float o[3], const float d[3], float tmin,
                             float *bestN, float *bestD) {
    float mx = o[0] - c[0], my = o[1] - c[1], mz = o[2] - c[2];
    float A = rt_dot3(d[0], d[1], d[2], d[0], d[1], d[2]);
    float B = rt_dot3(mx, my, mz, d[0], d[1], d[2]);
    float C = fmaf(-radius, radius, rt_dot3(mx, my, mz, mx, my, mz));
    float disc = fmaf(B, B, -(A * C));
    if (!(disc >= 0.0f)) return 0;
    float sq = sqrtf(disc);
    float N0 = -B - sq, N1 = -B + sq;   /* scaled by A > 0 */
    if (N0 >= tmin * A && N0 * *bestD < *bestN * A) { *bestN = N0; *bestD = A; return 1; }
    if (N1 >= tmin * A && N1 * *bestD < *bestN * A) { *bestN = N1; *bestD = A; return 1; }
    return 0;
}

/* vertices: 3*T entries of 4 floats (simd_float3, 16-byte stride);
 * spheres: S entries of 20 floats (SphereGPU, 80 bytes: center@0, radius@64). */
RT_INLINE rt_hit rt_intersect(const float *vertices, uint32_t numTriangles,
                              const float *spheres, uint32_t numSpheres,
                              const float o[3], const float d[3],
                              float tmin, float tmax, int acceptAny) {
    rt_hit h; h.hit = 0; h.distance = 0.0f; h.primitive_id = 0;
    float bestN = tmax, bestD = 1.0f;
    for (uint32_t i = 0; i < numTriangles; i++) {
        const float *v = vertices + (size_t)i * 12;
        if (rt_triangle_test(v, v + 4, v + 8, o, d, tmin, &bestN, &bestD)) {
            h.hit = 1; h.primitive_id = i;
            if (acceptAny) break;
        }
    }
    if (!(acceptAny && h.hit)) {
        for (uint32_t j = 0; j < numSpheres; j++) {
            const float *s = spheres + (size_t)j * 20;
            if (rt_sphere_test(s, s[16], o, d, tmin, &bestN, &bestD)) {
                h.hit = 1; h.primitive_id = numTriangles + j;
                if (acceptAny) break;
            }
        }
    }
    if (h.hit) h.distance = bestN / bestD;
    return h;
}

#endif /* RT_DETMATH_H */
