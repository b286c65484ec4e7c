/*
 * modeb.cuh -- sm_100a kernel `drawTriangle_mis`: the reference's MIS kernel
 * `drawTriangle`, Sources/gpuRaytracer/shaders.metal:635-707, with
 * recursiveMultiImportanceSampling (:543-625): per pixel 6 camera rays, and for each
 * one that lands on geometry 100 light + 100 cosine + 100 GGX-VNDF samples combined
 * with a three-strategy balance heuristic, one indirect bounce with next-event
 * estimation.  Included by pathtrace.cu (same module, same prepared scene, same
 * brute-force intersector).
 *
 * GPU shape: one thread per pixel, 8x4 pixel warps.  Every Halton point the shader
 * uses is pixel-independent (:556,564,584,595,617), so each block computes the five
 * tables once into shared memory and all lanes read them by broadcast.  The three
 * strategy loops are one rolled loop whose strategy index is warp-uniform; the two
 * divergent branches inside it (the sampled ray hit the light / hit geometry -> a
 * second ray towards the light) share ONE out-of-line copy of the primitive loop, so
 * the SIMT stack reconverges right after it and the kernel stays inside the
 * instruction cache.  Arithmetic follows oracle/rt_oracle_modeb.c operator by operator
 * (DESIGN.md section 3).
 */
#pragma once

namespace rt {

#define RT_PI_F 3.14159265358979323846264338327950288f

struct SurfB {            /* IntersectionGPU of a `Hit`, shaders.metal:14-20 */
    f3 point, normal, rdir;        /* rdir = ray.direction of the ray that found it */
    f3 diffuse;
    float metallic, roughness;
    f3 tangent, bitangent;         /* buildOrthonormalBasis(normal), prepared per triangle */
};

struct LightB { f3 center, er; float width, depth; f3 tangent, bitangent; };

/* :186-191 */
__device__ __forceinline__ float smithG1_GGX(float NoV, float roughness) {
    float a = roughness * roughness;
    float a2 = a * a;
    float NoV2 = NoV * NoV;
    return 2.0f / (1.0f + sqrtf(1.0f + a2 * (1.0f - NoV2) / NoV2));
}
/* :193-197 */
__device__ __forceinline__ float D_GGX(float NoH, float a) {
    float a2 = a * a;
    float f = (NoH * a2 - NoH) * NoH + 1.0f;
    return a2 / (RT_PI_F * f * f);
}
/* :203-208 */
__device__ __forceinline__ float V_SmithGGXCorrelated(float NoV, float NoL, float a) {
    float a2 = a * a;
    float GGXL = NoV * sqrtf((-NoL * a2 + NoL) * NoL + a2);
    float GGXV = NoL * sqrtf((-NoV * a2 + NoV) * NoV + a2);
    return 0.5f / (GGXV + GGXL);
}
/* :132-137 with beta = 1.0 (pow(x, 1) = x exactly) */
__device__ __forceinline__ float power_heuristic(float pdf1, float pdf2, float pdf3, float n) {
    float p1 = n * pdf1;
    float sum = p1 + n * pdf2 + n * pdf3;
    return p1 / (sum + 1e-6f);
}

/* :259-289.  ONE copy for its three callers (RT_MODEB_OUTLINE): with the primitive loop cut down by the group rule the
 * kernel's time went to fetching the shading code, 32 KB of it executed against a 32 KB instruction cache
 * (profiles/r04t_modeb_ncu_full_summary.txt: 22 % of the stall samples "no instruction") */
#ifndef RT_MODEB_OUTLINE
#define RT_MODEB_OUTLINE 1
#endif
#if RT_MODEB_OUTLINE
#define RT_MODEB_FN __noinline__
#else
#define RT_MODEB_FN __forceinline__
#endif
/* What the BRDF, the VNDF pdf and the light pdf need of the SURFACE alone: the three strategy loops evaluate them 300 times
 * for one camera-ray surface (the compiler cannot hoist this out of an out-of-line function), the next-event estimate at the
 * second vertex once for its own.  Same operations on the same operands as the one-piece forms of the shader. */
struct SurfPre {
    f3 v; float NoV;                    /* :260, :263  -normalize(ray.direction), |n.v| + 1e-5 */
    f3 f0, omf0, Fd; float omMetal;     /* :268, :271 (1 - f0), :277, :280 (1 - metallic) */
    float sV;                           /* :206  sqrt((-NoV a2 + NoV) NoV + a2) of V_SmithGGXCorrelated */
    float pdfNoV, G1;                   /* :440, :442  |n.viewDir| with viewDir = -ray.direction as it is, smithG1 of it */
    float dist2;                        /* :317-318, :324  distance * distance to the light's centre */
};
__device__ __forceinline__ void surf_pre_brdf(const SurfB &s, SurfPre &q) {
    q.v = -normalize(s.rdir);
    q.NoV = fabsf(dot(s.normal, q.v)) + 1e-5f;
    f3 d0 = mk(0.04f, 0.04f, 0.04f);
    q.f0 = d0 + (s.diffuse - d0) * s.metallic;
    q.omf0 = mk(1.0f, 1.0f, 1.0f) - q.f0;
    q.Fd = s.diffuse * (1.0f / RT_PI_F);
    q.omMetal = 1.0f - s.metallic;
    float a2 = s.roughness * s.roughness;
    q.sV = sqrtf((-q.NoV * a2 + q.NoV) * q.NoV + a2);
}
__device__ __forceinline__ void surf_pre_pdfs(const SurfB &s, const LightB &L, SurfPre &q) {
    q.pdfNoV = fabsf(dot(s.normal, -s.rdir));
    q.G1 = smithG1_GGX(q.pdfNoV, s.roughness);
    float distance = length(L.center - s.point);
    q.dist2 = distance * distance;
}
__device__ RT_MODEB_FN f3 brdf_contribution(const SurfB &s, const SurfPre &q, f3 l) {
    f3 n = s.normal;
    f3 h = normalize(q.v + l);
    float NoV = q.NoV;
    float NoL = fminf(1.0f, fmaxf(0.0f, dot(n, l)));
    float NoH = fminf(1.0f, fmaxf(0.0f, dot(n, h)));
    float LoH = fminf(1.0f, fmaxf(0.0f, dot(l, h)));
    float D = D_GGX(NoH, s.roughness);
    f3 F = q.f0 + q.omf0 * pow_det(1.0f - LoH, 5.0f);                             /* F_Schlick :199-201 */
    float a2 = s.roughness * s.roughness;                                         /* V_SmithGGXCorrelated :203-208 */
    float GGXL = NoV * sqrtf((-NoL * a2 + NoL) * NoL + a2);
    float GGXV = NoL * q.sV;
    float G = 0.5f / (GGXV + GGXL);
    f3 Fr = (F * (D * G)) / (4.0f * NoV * NoL + 1e-7f);
    f3 kD = (mk(1.0f, 1.0f, 1.0f) - F) * q.omMetal;
    f3 BRDF = kD * (q.Fd + Fr);
    return BRDF * NoL;
}

/* :315-326 (dist2: SurfPre) */
__device__ __forceinline__ float square_light_pdf(float dist2, const LightB &L, f3 direction) {
    float cosTheta = fmaxf(0.0f, dot(-direction, mk(0.0f, -1.0f, 0.0f)));
    float area = L.width * L.depth;
    return dist2 / (area * cosTheta + 1e-6f);
}
/* :376-380 */
__device__ __forceinline__ float cosine_pdf(f3 normal, f3 direction) {
    return fmaxf(0.0f, dot(normal, direction)) / RT_PI_F;
}
/* :436-445 (NoV and G1: SurfPre) */
__device__ __forceinline__ float vndf_pdf(f3 viewDir, f3 normal, f3 lightDir, float roughness, const SurfPre &q) {
    f3 h = normalize(viewDir + lightDir);
    float NoH = fabsf(dot(normal, h));
    float VoH = fabsf(dot(viewDir, h));
    float D = D_GGX(NoH, roughness);
    return (D * q.G1 * VoH) / (4.0f * q.pdfNoV);
}

/* :355-374: direction only (origin, min 0.001, max 1000 are fixed) */
__device__ __forceinline__ f3 cosine_weighted_dir(const SurfB &s, float rx, float ry) {
    float phi = 6.2831854820251465f * rx;             /* (2.0 * M_PI_F) * randomPoint.x */
    float cosTheta = sqrtf(ry);
    float sinTheta = sqrtf(1.0f - ry);
    float sp, cp;
    sincos_det(phi, sp, cp);
    return normalize((s.tangent * (cp * sinTheta) + s.bitangent * (sp * sinTheta)) + s.normal * cosTheta);
}

/* :382-434 */
__device__ __forceinline__ f3 vndf_dir(const SurfB &s, f3 viewDir, float rx, float ry) {
    float alpha = s.roughness * s.roughness;
    f3 Ve = normalize(mk(alpha * dot(viewDir, s.tangent), alpha * dot(viewDir, s.bitangent), dot(viewDir, s.normal)));
    f3 T1 = normalize(mk(Ve.z, 0.0f, -Ve.x));
    f3 T2 = cross(Ve, T1);
    float phi = 6.2831854820251465f * rx;
    float lenVe = length(Ve);
    float cosThetaMax = lenVe / sqrtf(1.0f + lenVe * lenVe);
    float cosTheta = cosThetaMax + (1.0f - cosThetaMax) * ry;
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    float sp, cp;
    sincos_det(phi, sp, cp);
    f3 h = normalize((T1 * (cp * sinTheta) + T2 * (sp * sinTheta)) + Ve * cosTheta);
    f3 Nh = normalize(mk(alpha * h.x, alpha * h.y, fmaxf(0.0f, h.z)));
    f3 worldH = normalize((s.tangent * Nh.x + s.bitangent * Nh.y) + s.normal * Nh.z);
    f3 i = -viewDir;
    return i - worldH * (2.0f * dot(worldH, i));      /* reflect(-viewDir, worldH) */
}

}  // namespace rt

/* ------------------------------------------------------------------ kernel ---- */
using namespace rt;

/* One out-of-line copy of the closest-hit primitive loop for every secondary ray.
 *
 * On a scene rt_prep_scene could cut into object groups (sorted.cuh; the Cornell box: room, two boxes, light) a ray is
 * tested against the always-tested quads plus those of every group whose grown box its line does not stay clear of;
 * a ray that is a SEGMENT towards a light sample point (directSquareLightRay: tmax = the distance to `end`) and has both
 * ends inside the shell box is not tested against the shell quads (the room's walls) either.  Mode B's rays start 1e-4 off
 * their surface, not 1e-3 as mode A's, so the origin is asked for a quarter of the shrink only; tmin = 1e-3 does the
 * rest (a wall's plane is crossed behind the origin or beyond the end).  The lanes of a warp walk ONE list, the union of
 * theirs (a skip per lane gains nothing), in ascending id as the full loop: same hit, same distance bits.
 * CPU side: oracle/rt_groups.c rt_oracle_cullcheck_modeb repeats every query of a frame under the per-ray rule
 * (tests/test_group_cull_cpu.py).  grp: the prepared scene's group section in shared memory. */
__device__ __noinline__ bool closest_b(const float4 *base, const float *grp, uint32_t T4, uint32_t Treal, f3 o, f3 d, float tmin,
                                       float tmax, bool segment, f3 end, float &dist, uint32_t &id) {
    SceneRef<false> sc;
    sc.base = base; sc.S = 0; sc.Treal = Treal; sc.sph = 0;
    /* T4's top bit: the scene's triangles come in pairs on one v0 (rt_prep_scene: RT_HDR_PAIRED) */
    sc.T = T4 & 0x7fffffffu;
    const uint32_t *gu = (const uint32_t *)grp;
    if ((T4 & 0x80000000u) && gu[RT_GRP_OK]) {                                       /* per launch */
        uint32_t pairs = gu[RT_GRP_ALWAYS];
        const uint32_t K = gu[RT_GRP_K];
        for (uint32_t g = 0; g < K; g++) {
            const float *G = grp + 4 + 8 * g;
            if (!ray_clear_of_box(mk(G[0], G[1], G[2]) - o, mk(G[4], G[5], G[6]), d)) pairs |= gu[4 + 8 * g + 3];
        }
        if (segment) {
            const float *SB = grp + RT_GRP_SHELL_BOX;
            const float q = 0.75f * SB[7];
            const bool in = (fabsf(o.x - SB[0]) < SB[4] + q) & (fabsf(o.y - SB[1]) < SB[5] + q) & (fabsf(o.z - SB[2]) < SB[6] + q) &
                            (fabsf(end.x - SB[0]) < SB[4]) & (fabsf(end.y - SB[1]) < SB[5]) & (fabsf(end.z - SB[2]) < SB[6]);
            if (in) pairs &= ~gu[RT_GRP_SHELL];
        }
        pairs = __reduce_or_sync(__activemask(), pairs);
        return closest_hit_pairs(sc, pairs, o, d, tmin, tmax, dist, id);
    }
    if (T4 & 0x80000000u) return closest_hit<false, false, 4, true>(sc, o, d, tmin, tmax, dist, id);
    return closest_hit<false, false, 4>(sc, o, d, tmin, tmax, dist, id);
}

/* getClosestIntersection's classification (:484-508) from the shade rows: 0 Hit, 1 HitLight */
__device__ __forceinline__ bool load_surface(const float *sm, const RTSceneLayout &lay, uint32_t id, f3 ro, f3 rd, float dist,
                                             SurfB &s) {
    const float *sh = sm + lay.shade + id;
    if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) return false;                 /* length(emissive) > 0: HitLight */
    s.point = ro + rd * dist;
    s.rdir = rd;
    s.normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
    s.tangent = mk(sh[(RT_SH_TX + 0) * lay.Tp], sh[(RT_SH_TX + 1) * lay.Tp], sh[(RT_SH_TX + 2) * lay.Tp]);
    s.bitangent = mk(sh[(RT_SH_BX + 0) * lay.Tp], sh[(RT_SH_BX + 1) * lay.Tp], sh[(RT_SH_BX + 2) * lay.Tp]);
    s.diffuse = mk(sh[(RT_SH_DR + 0) * lay.Tp], sh[(RT_SH_DR + 1) * lay.Tp], sh[(RT_SH_DR + 2) * lay.Tp]);
    s.metallic = sh[RT_SH_METAL * lay.Tp];
    s.roughness = sh[RT_SH_ROUGH * lay.Tp];
    return true;
}

/* calculateDirectLightSamplingContribution, :519-541 (directSquareLightRay :291-313 inlined) */
__device__ RT_MODEB_FN f3 direct_light(const float *sm, const RTSceneLayout &lay, const float4 *base, uint32_t T4,
                                           const LightB &L, const SurfB &s, const SurfPre &q, float rx, float ry, float n,
                                           bool usePower, uint32_t &rays) {
    f3 origin = s.point + s.normal * 1e-4f;
    float x = (rx - 0.5f) * L.width;
    float y = (ry - 0.5f) * L.depth;
    f3 samplePos = (L.center + L.tangent * x) + L.bitangent * y;
    f3 toLight = samplePos - origin;
    float distance = length(toLight);
    f3 ldir = toLight / distance;
    float dist; uint32_t id;
    rays++;
    if (!closest_b(base, sm + lay.group, T4, lay.T, origin, ldir, 0.001f, distance, true, samplePos, dist, id)) return mk(0.0f, 0.0f, 0.0f);
    if (sm[lay.shade + id + RT_SH_FLAG * lay.Tp] == 0.0f) return mk(0.0f, 0.0f, 0.0f);       /* not HitLight */
    f3 brdf = brdf_contribution(s, q, ldir);
    float directLightPDF = square_light_pdf(q.dist2, L, ldir);
    if (usePower) {
        float cosinePDF = cosine_pdf(s.normal, ldir);
        float vndfPDF = vndf_pdf(-s.rdir, s.normal, ldir, s.roughness, q);
        float weight = power_heuristic(directLightPDF, cosinePDF, vndfPDF, n);
        return ((brdf * weight) * L.er) / directLightPDF;
    }
    return (brdf * L.er) / directLightPDF;
}

#define RT_MODEB_THREADS 128
#ifndef RT_MODEB_MINBLOCKS
#define RT_MODEB_MINBLOCKS 4      /* 128 registers, no spills: 45.3 M camera rays/s; 3 (149 registers): 42.7; 5 (96): 44.5 */
#endif

extern "C" __global__ void __launch_bounds__(RT_MODEB_THREADS, RT_MODEB_MINBLOCKS)
drawTriangle_mis(const float *__restrict__ scene, uchar4 *__restrict__ pixels, float *__restrict__ text,
                 const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters) {
    extern __shared__ __align__(16) float sm[];
    {
        const float4 *src = (const float4 *)scene;
        float4 *dst = (float4 *)sm;
        const uint32_t n4 = lay.words >> 2;
        for (uint32_t k = threadIdx.x; k < n4; k += RT_MODEB_THREADS) dst[k] = src[k];
    }
    /* the shader's five pixel-independent Halton streams, :556,564,584,595,617 */
    const uint32_t sps = p.misSamples / 3u;
    float2 *tabL = (float2 *)(sm + lay.words);        /* haltonRandom(i, 0)            i < sps   */
    float2 *tabC = tabL + sps;                        /* haltonRandom(i + sps, 2)                */
    float2 *tabV = tabC + sps;                        /* haltonRandom(i + 2 sps, 4)              */
    float2 *tab6 = tabV + sps;                        /* haltonRandom(i, 6)            i < 2 sps */
    for (uint32_t k = threadIdx.x; k < 5u * sps; k += RT_MODEB_THREADS) {
        uint32_t which = k / sps, i = k - which * sps;
        uint32_t idx = which == 0 ? i : which == 1 ? i + sps : which == 2 ? i + 2u * sps : which == 3 ? i : i + sps;
        uint32_t dim = which == 0 ? 0u : which == 1 ? 2u : which == 2 ? 4u : 6u;
        tabL[k] = make_float2(halton_rt(idx, dim), halton_rt(idx, dim + 1u));
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    if (x >= (uint32_t)p.width || ly >= (uint32_t)p.rows) return;
    const uint32_t y = p.tileRows ? (uint32_t)p.row0 + (ly / p.tileRows) * p.tileStride + (ly % p.tileRows)
                                  : (uint32_t)p.row0 + ly;
    const float resX = sm[RT_HDR_RES_X], resY = sm[RT_HDR_RES_Y];
    if ((int)x >= (int)resX || (int)y >= (int)resY) return;                                /* :642 */
    const size_t pix = (size_t)ly * (uint32_t)p.width + x;

    const float4 *base = (const float4 *)(sm + lay.isect);
    const uint32_t Tpad = (lay.T + 3u) & ~3u;
    const uint32_t T4 = Tpad | (((const uint32_t *)sm)[RT_HDR_PAIRED] ? 0x80000000u : 0u);   /* flag for closest_b */
    SceneRef<false> sc;
    sc.base = base; sc.T = Tpad; sc.S = 0; sc.Treal = lay.T; sc.sph = 0;

    const f3 cam_o = mk(sm[RT_HDR_CAM_O], sm[RT_HDR_CAM_O + 1], sm[RT_HDR_CAM_O + 2]);
    const f3 cam_u = mk(sm[RT_HDR_CAM_U], sm[RT_HDR_CAM_U + 1], sm[RT_HDR_CAM_U + 2]);
    const f3 cam_v = mk(sm[RT_HDR_CAM_V], sm[RT_HDR_CAM_V + 1], sm[RT_HDR_CAM_V + 2]);
    const f3 cam_w = mk(sm[RT_HDR_CAM_W], sm[RT_HDR_CAM_W + 1], sm[RT_HDR_CAM_W + 2]);
    const float halfW = sm[RT_HDR_HALF_W], halfH = sm[RT_HDR_HALF_H];
    const float fx = (float)(int)x, fy = (float)(int)y;

    LightB L;
    {
        const float *l = sm + lay.light;                                                    /* squareLights[0], :638 */
        L.center = mk(l[0], l[1], l[2]);
        L.er = mk(l[8], l[9], l[10]);
        L.width = l[12]; L.depth = l[13];
        orthonormal_basis(mk(0.0f, -1.0f, 0.0f), L.tangent, L.bitangent);                   /* :292-293 */
    }
    const float nf = (float)sps;
    uint32_t rays = 0, shaded = 0;
    f3 accumulated = mk(0.0f, 0.0f, 0.0f);

#pragma unroll 1
    for (uint32_t i = 0; i < p.cameraRays; i++) {
        /* hashRandom, :71-85 (800 and 600 are the shader's literals) */
        uint32_t sampleId = (y * 800u + x) * i;
        uint32_t seed1 = hash_u32(x + y * 800u + sampleId);
        uint32_t seed2 = hash_u32(y + x * 600u + sampleId + 12345u);
        float jx = (float)hash_u32(seed1) / 4294967296.0f;
        float jy = (float)hash_u32(seed2) / 4294967296.0f;
        /* generateCameraRay, :214-246 */
        float s = ((fx + jx) / resX) * 2.0f - 1.0f;
        float t = -(((fy + jy) / resY) * 2.0f - 1.0f);
        f3 rd = normalize((cam_u * (s * halfW) + cam_v * (t * halfH)) - cam_w);
        float dist; uint32_t id;
        rays++;
        if (!closest_hit<false, true, 4>(sc, cam_o, rd, 0.001f, 1000.0f, dist, id)) continue;          /* Miss */
        SurfB in;
        if (!load_surface(sm, lay, id, cam_o, rd, dist, in)) { accumulated = accumulated + L.er; continue; }   /* HitLight :669-670 */
        shaded++;

        /* recursiveMultiImportanceSampling, :543-625: three strategies, one rolled loop */
        const f3 origin = in.point + in.normal * 1e-4f;
        const f3 view = -in.rdir;
        SurfPre pre;
        surf_pre_brdf(in, pre);
        surf_pre_pdfs(in, L, pre);
        f3 sumL = mk(0.0f, 0.0f, 0.0f), sumC = sumL, sumV = sumL;
#pragma unroll 1
        for (uint32_t k = 0; k < sps; k++) {                                                   /* :553-560 */
            float2 u = tabL[k];
            sumL = sumL + direct_light(sm, lay, base, T4, L, in, pre, u.x, u.y, nf, true, rays);
        }
#pragma unroll 1
        for (uint32_t k = 0; k < 2u * sps; k++) {                                              /* :562-591 and :593-623 */
            const bool vndf = k >= sps;                                                        /* warp-uniform */
            const uint32_t j = vndf ? k - sps : k;
            float2 u = vndf ? tabV[j] : tabC[j];
            f3 dir = vndf ? vndf_dir(in, view, u.x, u.y) : cosine_weighted_dir(in, u.x, u.y);
            float cosinePDF = cosine_pdf(in.normal, dir);
            float directLightPDF = square_light_pdf(pre.dist2, L, dir);
            float vndfPDF = vndf_pdf(view, in.normal, dir, in.roughness, pre);
            float ownPDF = vndf ? vndfPDF : cosinePDF;
            float weight = vndf ? power_heuristic(vndfPDF, directLightPDF, cosinePDF, nf)
                                : power_heuristic(cosinePDF, directLightPDF, vndfPDF, nf);
            float d2; uint32_t id2;
            rays++;
            bool hit2 = closest_b(base, sm + lay.group, T4, lay.T, origin, dir, 0.001f, 1000.0f, false, origin, d2, id2);
            f3 brdf = brdf_contribution(in, pre, dir);
            f3 add = mk(0.0f, 0.0f, 0.0f);
            bool contributes = false;
            if (hit2) {
                SurfB s2;
                if (!load_surface(sm, lay, id2, origin, dir, d2, s2)) {                        /* HitLight */
                    add = ((brdf * weight) * L.er) / ownPDF;
                } else {                                                                        /* Hit: NEE at the second vertex */
                    float2 u2 = tab6[vndf ? j + sps : j];                                       /* haltonRandom(i, 6) / (i + sps, 6) */
                    SurfPre pre2;                                                               /* (the pdfs' part is not used without the power heuristic) */
                    surf_pre_brdf(s2, pre2);
                    pre2.pdfNoV = 0.0f; pre2.G1 = 0.0f;
                    { float distance = length(L.center - s2.point); pre2.dist2 = distance * distance; }
                    add = (brdf / ownPDF) * direct_light(sm, lay, base, T4, L, s2, pre2, u2.x, u2.y, 1.0f, false, rays);
                }
                contributes = true;
            }
            if (contributes) { if (vndf) sumV = sumV + add; else sumC = sumC + add; }
        }
        accumulated = accumulated + ((sumL + sumC) + sumV) / nf;                               /* :624 */
    }

    /* :688-706 */
    const float exposure = sm[RT_HDR_EXPOSURE];
    f3 exposed = (accumulated / (float)p.cameraRays) * exposure;
    f3 tone = exposed / (exposed + mk(1.0f, 1.0f, 1.0f));
    tone = mk(fminf(fmaxf(tone.x, 0.0f), 1.0f), fminf(fmaxf(tone.y, 0.0f), 1.0f), fminf(fmaxf(tone.z, 0.0f), 1.0f));
    const float e = 1.0f / 2.2f;
    f3 g = mk(pow_det(tone.x, e), pow_det(tone.y, e), pow_det(tone.z, e));
    text[pix * 3 + 0] = accumulated.x; text[pix * 3 + 1] = accumulated.y; text[pix * 3 + 2] = accumulated.z;
    pixels[pix] = make_uchar4((unsigned char)(g.x * 255.0f), (unsigned char)(g.y * 255.0f), (unsigned char)(g.z * 255.0f), 255);
    if (counters) {
        atomicAdd(&counters[0], (unsigned long long)p.cameraRays);
        atomicAdd(&counters[1], (unsigned long long)rays);
        atomicAdd(&counters[3], (unsigned long long)shaded);
    }
}
