/*
 * pathtrace.cu -- sm_100a kernels behind librtcuda.so.
 *
 *   rt_prep_scene     one block: turns the buffers the host uploaded (shaderTypes.h
 *                     layout, AoS) into the prepared scene of rt_scene_layout.h
 *   pathTrace_*       the per-pixel path tracer; replaces the Metal kernel
 *                     `pathTrace`, reference RTrace/raytrace.metal:11-111 with
 *                     RTrace/sampling.metal:16-26,39-66,97-157,198-236
 *   rt_resolve_rgba8  RTrace/image.swift:40-65 tone map as a device pass
 *   rt_resolve_rgba16f  float4 -> half4, the reference's render-target format (renderer.swift:74-82)
 *   rt_debug_eval     evaluates single device functions for bit-compare tests
 *
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -prec-div=true
 *        -prec-sqrt=true -ftz=false -lineinfo  (see rt_device_math.cuh for why).
 *
 * Kernel shape (DESIGN.md has the numbers): one thread per pixel, a warp covers an
 * 8x4 pixel tile, a block 16x8 pixels.  Each block first stages the prepared scene
 * from HBM into shared memory; per sample the thread walks the (rolled) bounce loop,
 * and every ray query is a loop over ALL primitives: there is no acceleration
 * structure.
 */
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "../../include/rtcuda.h"
#include "rt_scene_layout.h"
#include "rt_device_math.cuh"

using namespace rt;

/* constant-bank mirror of the intersection part of the prepared scene (variant
 * "const"): every lane reads the same primitive, so the operands come through the
 * uniform datapath and cost neither LSU issue slots nor vector registers. */
__constant__ float4 c_scene[RT_CONST_SLOTS];

#include "sgrid.cuh"      /* the uniform grid over a scene's spheres: built by rt_prep_scene, walked by the intersector */

/* ------------------------------------------------------------------ prep ---- */
struct SimdF3 { float x, y, z, pad; };

/*
 * Object groups for the sorted kernel (sorted.cuh; layout: rt_scene_layout.h "group").  One warp, lane = quad
 * (triangles 2*lane, 2*lane+1).  Two quads are connected when some vertex of one is bit-equal to some vertex of
 * the other; a group is a connected component (the Cornell scene: the room, the tall box, the short box, the
 * light).  A group whose bounds span 90 % of the scene on every axis, or hold a light's centre (every shadow
 * ray goes there), is "always tested"; up to RT_MAX_GROUPS
 * others get a box (bounds of the vertices grown by 1/128 of the scene's largest extent) and a sort-key bit;
 * what is left over is always tested too.  ANY partition gives the same image -- a group only says which quads
 * a ray that misses its box may skip -- so nothing here is part of the arithmetic contract.
 */
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int s = 16; s; s >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, s));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int s = 16; s; s >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, s));
    return v;
}
/* which quads share a vertex with quad p?  Every warp of the block takes its share of the partners q (lane = p);
 * the bits are OR-ed into s_adj[p].  (One warp doing all of it took 15 us: three times the rest of the prep kernel,
 * and a progressive 1-spp frame is only 0.45 ms.) */
__device__ void prep_adjacency(const unsigned char *__restrict__ vertices, uint32_t P, uint32_t lane, uint32_t warp,
                               uint32_t nwarps, uint32_t *s_adj) {
    if (lane >= P) return;
    const uint32_t *v = (const uint32_t *)vertices;
    uint32_t vb[18];
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
        for (int k = 0; k < 3; k++) vb[3 * a + k] = v[(6u * lane + a) * 4u + k];
    uint32_t adj = 0u;
    for (uint32_t q = warp; q < P; q += nwarps) {
        bool share = false;
#pragma unroll
        for (int b = 0; b < 6; b++) {
            const uint32_t x = v[(6u * q + b) * 4u], y = v[(6u * q + b) * 4u + 1u], z = v[(6u * q + b) * 4u + 2u];
#pragma unroll
            for (int a = 0; a < 6; a++) share |= vb[3 * a] == x && vb[3 * a + 1] == y && vb[3 * a + 2] == z;
        }
        if (share) adj |= 1u << q;
    }
    if (adj) atomicOr(&s_adj[lane], adj);
}

__device__ void prep_groups(const unsigned char *__restrict__ vertices, const unsigned char *__restrict__ lights,
                            float *__restrict__ scene, const RTSceneLayout lay, bool paired, uint32_t lane,
                            const uint32_t *s_adj) {
    uint32_t *gu = (uint32_t *)scene + lay.group;
    float *gf = scene + lay.group;
    for (uint32_t k = lane; k < RT_GROUP_WORDS; k += 32u) gu[k] = 0u;
    __syncwarp();
    const uint32_t P = lay.T / 2u;
    if (!(paired && lay.T >= 2u && lay.T <= 64u)) return;                         /* warp-uniform */
    const bool has = lane < P;
    const float INF = __int_as_float(0x7f800000);
    float lo[3] = { INF, INF, INF }, hi[3] = { -INF, -INF, -INF };
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (has) { const float x = ((const float *)vertices)[(6u * lane + a) * 4u + k]; lo[k] = fminf(lo[k], x); hi[k] = fmaxf(hi[k], x); }
    const uint32_t adj = has ? s_adj[lane] : 0u;
    uint32_t label = lane;
    for (uint32_t it = 0; it < P; it++) {
        uint32_t best = label;
        for (uint32_t q = 0; q < P; q++) {
            const uint32_t l = __shfl_sync(0xffffffffu, label, q);
            if ((adj >> q) & 1u) best = min(best, l);
        }
        const bool changed = best != label;
        label = best;
        if (!__any_sync(0xffffffffu, changed)) break;
    }
    const uint32_t leaders = __ballot_sync(0xffffffffu, has && label == lane);
    const uint32_t comp = __popc(leaders & ((1u << label) - 1u)), ncomp = __popc(leaders);
    float slo[3], shi[3], ext = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; k++) { slo[k] = warp_min(lo[k]); shi[k] = warp_max(hi[k]); ext = fmaxf(ext, shi[k] - slo[k]); }
    const float margin = ext * (1.0f / 128.0f);
    uint32_t K = 0u, always = 0u;
    for (uint32_t c = 0; c < ncomp; c++) {
        const bool member = has && comp == c;
        const uint32_t pairs = __ballot_sync(0xffffffffu, member);
        float glo[3], ghi[3];
        bool spans = true;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            glo[k] = warp_min(member ? lo[k] : INF); ghi[k] = warp_max(member ? hi[k] : -INF);
            spans = spans && (ghi[k] - glo[k] >= 0.9f * (shi[k] - slo[k]));
        }
        /* a group that holds a light's centre is reached by every shadow ray: a key bit would always be set */
        bool lit = false;
        for (uint32_t l = 0; l < lay.L; l++) {
            const float *c = (const float *)(lights + 64 * (size_t)l);
            bool in = true;
#pragma unroll
            for (int k = 0; k < 3; k++) in = in && c[k] >= glo[k] - margin && c[k] <= ghi[k] + margin;
            lit = lit || in;
        }
        if (!spans && !lit && K < RT_MAX_GROUPS) {
            if (lane == 0) {
                float *r = gf + 4 + 8 * K;
#pragma unroll
                for (int k = 0; k < 3; k++) { r[k] = (glo[k] + ghi[k]) * 0.5f; r[4 + k] = (ghi[k] - glo[k]) * 0.5f + margin; }
                ((uint32_t *)r)[3] = pairs; r[7] = 0.0f;
            }
            K++;
        } else {
            always |= pairs;
        }
    }
    /* the shell: always-tested quads in a face plane of the always-tested quads' own bounds (rt_scene_layout.h) */
    const bool mine = has && ((always >> lane) & 1u);
    float alo[3], ahi[3];
    bool face = false;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        alo[k] = warp_min(mine ? lo[k] : INF); ahi[k] = warp_max(mine ? hi[k] : -INF);
        face = face || (mine && lo[k] == hi[k] && (lo[k] == alo[k] || lo[k] == ahi[k]));
    }
    const uint32_t shell = __ballot_sync(0xffffffffu, face);
    /* the shell quads face by face (the exit rule, rt_scene_layout.h) */
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const uint32_t loFace = __ballot_sync(0xffffffffu, face && lo[k] == hi[k] && lo[k] == alo[k]);
        const uint32_t hiFace = __ballot_sync(0xffffffffu, face && lo[k] == hi[k] && lo[k] == ahi[k]);
        if (lane == 0) { gu[RT_GRP_FACES + 2 * k] = loFace; gu[RT_GRP_FACES + 2 * k + 1] = hiFace; }
    }
    /* the quads in the plane of the light(s) (rt_scene_layout.h: RT_GRP_LIGHT_PLANE) */
    {
        const float cy = ((const float *)lights)[1];
        bool ok1 = lay.L >= 1u && !(lay.flags & RT_LAYOUT_LIGHT_NORMALS) && ext < 64.0f;
        for (uint32_t l = 1; l < lay.L; l++) ok1 = ok1 && ((const float *)(lights + 64 * (size_t)l))[1] == cy;     /* several lights: in ONE plane */
        const uint32_t inPlane = __ballot_sync(0xffffffffu, has && ok1 && lo[1] == cy && hi[1] == cy);
        if (lane == 0) gu[RT_GRP_LIGHT_PLANE] = inPlane;
    }
    if (lane == 0) {
        float *r = gf + RT_GRP_SHELL_BOX;
        const float shrink = ext * (1.0f / 16384.0f);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            r[k] = shell ? (alo[k] + ahi[k]) * 0.5f : 0.0f;
            r[4 + k] = shell ? (ahi[k] - alo[k]) * 0.5f - shrink : 0.0f;
        }
        r[3] = shell ? margin : 0.0f; r[7] = shell ? shrink : 0.0f;      /* (modeb.cuh asks the origin for a quarter of the shrink) */
        /* (a scene with spheres has its triangles grouped for the shell's sake only: the sorted kernels are not for it) */
        gu[RT_GRP_K] = K; gu[RT_GRP_ALWAYS] = always; gu[RT_GRP_OK] = (K >= 1u && lay.S == 0u) ? 1u : 0u; gu[RT_GRP_SHELL] = shell;
        /* does every light's sample rectangle lie inside the shrunk shell box?  (rt_scene_layout.h: RT_GRP_LIGHTS_IN) */
        bool in = shell != 0u && !(lay.flags & RT_LAYOUT_LIGHT_NORMALS);
        for (uint32_t l = 0; l < lay.L; l++) {
            const float *c = (const float *)(lights + 64 * (size_t)l);
            const float hx = fmaxf(0.25f, 0.5f * c[12]), hz = fmaxf(0.25f, 0.5f * c[13]);
            in = in && fabsf(c[0] - r[0]) + hx < r[4] && fabsf(c[1] - r[1]) < r[5] && fabsf(c[2] - r[2]) + hz < r[6];
        }
        gu[RT_GRP_LIGHTS_IN] = in ? 1u : 0u;
    }
}

extern "C" __global__ void rt_prep_scene(const unsigned char *__restrict__ camera,
                                         const unsigned char *__restrict__ materials,
                                         const unsigned char *__restrict__ lights,
                                         const unsigned char *__restrict__ vertices,
                                         const unsigned char *__restrict__ spheres,
                                         float *__restrict__ scene, RTSceneLayout lay,
                                         const float4 *__restrict__ lightNormals) {
    /* One block for the scenes the brute-force kernels take; for the BVH kernel's scenes (up to 4 M triangles) the host
     * launches a grid: every block takes its share of the per-primitive records, block 0 alone writes the header, the
     * lights and the object groups. */
    const int tid = threadIdx.x, nth = blockDim.x;
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gnth = gridDim.x * blockDim.x;
    const bool lead = blockIdx.x == 0;
    const float *cam = (const float *)camera;
    const f3 position = mk(cam[0], cam[1], cam[2]);
    if (lead && tid == 0) {
        /* generateCameraRay's per-launch part, sampling.metal:131-139 */
        const int *cami = (const int *)camera;
        f3 direction = mk(cam[4], cam[5], cam[6]);
        f3 up = mk(cam[8], cam[9], cam[10]);
        int resx = cami[12], resy = cami[13];
        float fov = cam[14], ev100 = cam[15];
        float aspectRatio = (float)(resy != 0 ? resx / resy : 0);   /* integer division, :132 */
        float halfWidth = tan_det(fov / 2.0f);
        float halfHeight = halfWidth / aspectRatio;
        f3 w = -normalize(direction);
        f3 u = normalize(cross(up, w));
        f3 v = normalize(cross(w, u));
        scene[RT_HDR_CAM_O + 0] = position.x; scene[RT_HDR_CAM_O + 1] = position.y; scene[RT_HDR_CAM_O + 2] = position.z;
        scene[RT_HDR_CAM_U + 0] = u.x; scene[RT_HDR_CAM_U + 1] = u.y; scene[RT_HDR_CAM_U + 2] = u.z;
        scene[RT_HDR_CAM_V + 0] = v.x; scene[RT_HDR_CAM_V + 1] = v.y; scene[RT_HDR_CAM_V + 2] = v.z;
        scene[RT_HDR_CAM_W + 0] = w.x; scene[RT_HDR_CAM_W + 1] = w.y; scene[RT_HDR_CAM_W + 2] = w.z;
        scene[RT_HDR_HALF_W] = halfWidth;
        scene[RT_HDR_HALF_H] = halfHeight;
        scene[RT_HDR_RES_X] = (float)resx;
        scene[RT_HDR_RES_Y] = (float)resy;
        scene[RT_HDR_EXPOSURE] = 1.0f / (1.2f * pow_det(2.0f, ev100));   /* cameraExposure, shaders.metal:145-150 */
        for (int i = RT_HDR_EXPOSURE + 1; i < RT_HDR_WORDS; i++) if (i != RT_HDR_PAIRED) scene[i] = 0.0f;
    }
    for (uint32_t i = gtid; i < lay.Tp; i += gnth) {
        float *is = scene + lay.isect + RT_ISECT_WORDS * i;
        float *ps = is + 12;
        float *sh = scene + lay.shade + i;
        if (i < lay.T) {
            const SimdF3 *vv = (const SimdF3 *)vertices + 3 * i;
            f3 v0 = mk(vv[0].x, vv[0].y, vv[0].z);
            f3 v1 = mk(vv[1].x, vv[1].y, vv[1].z);
            f3 v2 = mk(vv[2].x, vv[2].y, vv[2].z);
            f3 e1 = v1 - v0, e2 = v2 - v0;
            f3 nn = cross(e1, e2);                 /* un-normalised winding normal (shading)                  */
            f3 ni = cross(e2, e1);                 /* the intersector's n, and e1n = v0 - v1: no unary minus  */
            f3 e1n = v0 - v1;                      /* anywhere on the way to U, V, D, Nn (rt_detmath.h)       */
            f3 w = v0 - position;                  /* camera rays: w and Nn do not depend on the ray */
            is[0] = v0.x; is[1] = v0.y; is[2] = v0.z; is[3] = ni.x;
            is[4] = ni.y; is[5] = ni.z; is[6] = e2.x; is[7] = e2.y;
            is[8] = e2.z; is[9] = e1n.x; is[10] = e1n.y; is[11] = e1n.z;
            ps[0] = w.x; ps[1] = w.y; ps[2] = w.z; ps[3] = dot(w, ni);
            f3 n = normalize(nn);                                            /* sampling.metal:16-26 */
            f3 right = normalize(cross(n, mk(0.0072f, 1.0f, 0.0034f)));      /* sampling.metal:57   */
            f3 fwd = cross(right, n);                                        /* sampling.metal:61   */
            const float *m = (const float *)(materials + 48 * (size_t)i);
            f3 em = mk(m[8], m[9], m[10]);
            sh[(RT_SH_NX + 0) * lay.Tp] = n.x; sh[(RT_SH_NX + 1) * lay.Tp] = n.y; sh[(RT_SH_NX + 2) * lay.Tp] = n.z;
            sh[(RT_SH_RX + 0) * lay.Tp] = right.x; sh[(RT_SH_RX + 1) * lay.Tp] = right.y; sh[(RT_SH_RX + 2) * lay.Tp] = right.z;
            sh[(RT_SH_FX + 0) * lay.Tp] = fwd.x; sh[(RT_SH_FX + 1) * lay.Tp] = fwd.y; sh[(RT_SH_FX + 2) * lay.Tp] = fwd.z;
            sh[(RT_SH_DR + 0) * lay.Tp] = m[0]; sh[(RT_SH_DR + 1) * lay.Tp] = m[1]; sh[(RT_SH_DR + 2) * lay.Tp] = m[2];
            sh[(RT_SH_ER + 0) * lay.Tp] = em.x; sh[(RT_SH_ER + 1) * lay.Tp] = em.y; sh[(RT_SH_ER + 2) * lay.Tp] = em.z;
            sh[RT_SH_FLAG * lay.Tp] = (length(em) > 0.0f) ? 1.0f : 0.0f;     /* raytrace.metal:57   */
            sh[RT_SH_METAL * lay.Tp] = m[4];
            sh[RT_SH_ROUGH * lay.Tp] = m[5];
            f3 tg, bt;
            orthonormal_basis(n, tg, bt);                                    /* shaders.metal:159-172 */
            sh[(RT_SH_TX + 0) * lay.Tp] = tg.x; sh[(RT_SH_TX + 1) * lay.Tp] = tg.y; sh[(RT_SH_TX + 2) * lay.Tp] = tg.z;
            sh[(RT_SH_BX + 0) * lay.Tp] = bt.x; sh[(RT_SH_BX + 1) * lay.Tp] = bt.y; sh[(RT_SH_BX + 2) * lay.Tp] = bt.z;
        } else {
            /* padding triangles are degenerate (all zero): D == 0, never accepted */
            for (int k = 0; k < RT_ISECT_WORDS; k++) is[k] = 0.0f;
            for (int k = 0; k < RT_SHADE_ROWS; k++) sh[k * lay.Tp] = 0.0f;
        }
    }
    for (uint32_t j = gtid; j < lay.Sp; j += gnth) {
        float *sp = scene + lay.sphere + RT_SPHERE_WORDS * j;
        float *pp = sp + 4;
        float *sh = scene + lay.sshade + j;
        if (j < lay.S) {
            const float *s = (const float *)(spheres + 80 * (size_t)j);   /* SphereGPU: center@0 material@16 radius@64 */
            sp[0] = s[0]; sp[1] = s[1]; sp[2] = s[2]; sp[3] = s[16];
            f3 m = position - mk(s[0], s[1], s[2]);
            pp[0] = m.x; pp[1] = m.y; pp[2] = m.z; pp[3] = fmaf(-s[16], s[16], dot(m, m));
            f3 em = mk(s[12], s[13], s[14]);
            sh[(RT_SSH_DR + 0) * lay.Sp] = s[4]; sh[(RT_SSH_DR + 1) * lay.Sp] = s[5]; sh[(RT_SSH_DR + 2) * lay.Sp] = s[6];
            sh[(RT_SSH_ER + 0) * lay.Sp] = em.x; sh[(RT_SSH_ER + 1) * lay.Sp] = em.y; sh[(RT_SSH_ER + 2) * lay.Sp] = em.z;
            sh[RT_SSH_FLAG * lay.Sp] = (length(em) > 0.0f) ? 1.0f : 0.0f;
        } else {
            /* padding sphere: radius 0 far away; C > 0 and disc < 0 for any finite ray */
            sp[0] = 0.0f; sp[1] = 3.0e18f; sp[2] = 0.0f; sp[3] = 0.0f;
            pp[0] = 0.0f; pp[1] = -3.0e18f; pp[2] = 0.0f; pp[3] = 9.0e36f;
            for (int k = 0; k < RT_SSHADE_ROWS; k++) sh[k * lay.Sp] = 0.0f;
        }
    }
    if (!lead) return;                                    /* (block-uniform: the barriers below are block 0's own) */
    if (lay.S && !(lay.flags & RT_LAYOUT_NO_SGRID)) prep_sphere_grid(spheres, scene, lay, tid, nth);     /* sgrid.cuh */
    {
        /* is every triangle 2k+1 on the v0 of triangle 2k, bit for bit?  (PAIRED loops; padding pairs are all zero.)
         * Only the brute-force kernels ask: above their limit the flag stays clear and the scan is skipped. */
        int ok = (lay.T % 2u) == 0u && lay.T <= RT_MAX_TRIANGLES;
        for (uint32_t i = 2u * tid + 1u; i < lay.T && lay.T <= RT_MAX_TRIANGLES; i += 2u * nth) {
            const uint32_t *a = (const uint32_t *)((const SimdF3 *)vertices + 3 * (i - 1u));
            const uint32_t *b = (const uint32_t *)((const SimdF3 *)vertices + 3 * i);
            ok = ok && a[0] == b[0] && a[1] == b[1] && a[2] == b[2];
        }
        __shared__ uint32_t s_adj[32];
        if (tid < 32) s_adj[tid] = 0u;
        ok = __syncthreads_and(ok);
        if (tid == 0) ((uint32_t *)scene)[RT_HDR_PAIRED] = ok ? 1u : 0u;
        const bool groupable = ok && lay.T >= 2u && lay.T <= 64u;                             /* block-uniform */
        if (groupable) prep_adjacency(vertices, lay.T / 2u, (uint32_t)tid & 31u, (uint32_t)tid >> 5, (uint32_t)nth >> 5, s_adj);
        __syncthreads();
        if (tid < 32) prep_groups(vertices, lights, scene, lay, ok != 0, (uint32_t)tid, s_adj);
    }
    for (uint32_t l = tid; l < lay.L; l += nth) {
        const float *s = (const float *)(lights + 64 * (size_t)l);        /* SquareLightGPU: center@0 color@16 */
        float *d = scene + lay.light + RT_LIGHT_WORDS * l;
        const bool extent = (lay.flags & RT_LAYOUT_LIGHT_EXTENT) != 0u;
        d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = extent ? 0.5f * s[12] : 0.25f;   /* right.x, sampling.metal:208 */
        d[4] = s[4]; d[5] = s[5]; d[6] = s[6]; d[7] = extent ? 0.5f * s[13] : 0.25f;   /* up.z,    sampling.metal:209 */
        d[8] = s[8]; d[9] = s[9]; d[10] = s[10];                          /* emittedRadiance@32 */
        d[12] = s[12]; d[13] = s[13]; d[14] = 0.0f; d[15] = 0.0f;         /* width@48 depth@52  */
        /* the light's normal and the basis of its sample rectangle: the reference's literals, or -- with a
         * normal table bound -- the table's normal and buildOrthonormalBasis of it (sampling.metal:158-171) */
        const bool general = (lay.flags & RT_LAYOUT_LIGHT_NORMALS) != 0u && lightNormals != nullptr;
        f3 nL = mk(0.0f, -1.0f, 0.0f), tg = mk(1.0f, 0.0f, 0.0f), bt = mk(0.0f, 0.0f, 1.0f);
        if (general) {
            const float4 t = lightNormals[l];
            nL = normalize(mk(t.x, t.y, t.z));
            orthonormal_basis(nL, tg, bt);
        }
        ((uint32_t *)d)[RT_LIGHT_GENERAL] = general ? 1u : 0u;
        d[RT_LIGHT_NORMAL] = nL.x; d[RT_LIGHT_NORMAL + 1] = nL.y; d[RT_LIGHT_NORMAL + 2] = nL.z; d[RT_LIGHT_NORMAL + 3] = 0.0f;
        d[RT_LIGHT_TANGENT] = tg.x; d[RT_LIGHT_TANGENT + 1] = tg.y; d[RT_LIGHT_TANGENT + 2] = tg.z; d[RT_LIGHT_TANGENT + 3] = 0.0f;
        d[RT_LIGHT_BITANGENT] = bt.x; d[RT_LIGHT_BITANGENT + 1] = bt.y; d[RT_LIGHT_BITANGENT + 2] = bt.z; d[RT_LIGHT_BITANGENT + 3] = 0.0f;
    }
}

/* ------------------------------------------------------------- intersector ---- */
/*
 * Division-free ray/triangle test in the form fixed by the arithmetic contract
 * (oracle/rt_detmath.h states the same thing for the CPU): with the
 * per-triangle constants n = cross(e2,e1), e2, e1n = v0 - v1 and, per ray,
 *     w = v0 - o, D = dot(d,n), Nn = dot(w,n), c = cross(d,w), U = dot(e2,c), V = dot(e1n,c),
 * and a = |D|, a hit needs U, V and Nn to carry the sign bit of D and
 *     |U + V| <= a, |Nn| >= tmin*a, |Nn|*bestD < bestN*a.
 * (bestN,bestD) starts at (tmax,1).  One division, bestN/bestD, after the loop.
 * 21 FP32 instructions + the comparisons; for camera rays w and Nn come from the
 * prepared scene (they depend on the camera position only), leaving 15.
 */
__device__ __forceinline__ bool tri_accept(float D, float Nn, float U, float V, float tmin,
                                           float bestN, float bestD, float &N_out, float &a_out) {
    /* sign rule: U, V and Nn must carry D's sign bit -- two LOP3 and one ISETP for all three.
     * (History, all measured on B200 with the results unchanged: flipping each quantity by XOR and comparing
     * it with 0 costs three more issue slots per test, 4 % of the frame; flipping by integer add, or by
     * multiplication with +-1 to move work to the FMA pipe, was slower still: the loop is bound by issue
     * slots, and a half-rate ALU instruction costs 1.2-1.5 of them, tools/gen_issue_bench.py.) */
    const uint32_t db = __float_as_uint(D);
    const uint32_t t1 = (__float_as_uint(U) ^ db) | (__float_as_uint(V) ^ db);
    const int t2 = (int)(t1 | (__float_as_uint(Nn) ^ db));
    const float s = U + V;
    const float a = fabsf(D), N = fabsf(Nn);
    N_out = N; a_out = a;
    /* tmin == 0 (shadow rays; a compile-time constant after inlining): the sign rule is the whole test */
    if (tmin == 0.0f) return (t2 >= 0) & (fabsf(s) <= a) & (N * bestD < bestN * a);
    return (t2 >= 0) & (fabsf(s) <= a) & (N >= tmin * a) & (N * bestD < bestN * a);
}

/* A = v0.xyz n.x | B = n.yz e2.xy | C = e2.z e1n.xyz */
__device__ __forceinline__ bool tri_test(float4 A, float4 B, float4 C, f3 o, f3 d, float tmin,
                                         float bestN, float bestD, float &N_out, float &a_out) {
    f3 v0 = mk(A.x, A.y, A.z), n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
    f3 w = v0 - o;
    float D = dot(d, n);
    float Nn = dot(w, n);
    f3 c = cross(d, w);
    float U = dot(e2, c);
    float V = dot(e1n, c);
    return tri_accept(D, Nn, U, V, tmin, bestN, bestD, N_out, a_out);
}

/* camera ray: P = w.xyz Nn (prepared) */
__device__ __forceinline__ bool tri_test_primary(float4 P, float4 A, float4 B, float4 C, f3 d, float tmin,
                                                 float bestN, float bestD, float &N_out, float &a_out) {
    f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
    f3 w = mk(P.x, P.y, P.z);
    float D = dot(d, n);
    f3 c = cross(d, w);
    float U = dot(e2, c);
    float V = dot(e1n, c);
    return tri_accept(D, P.w, U, V, tmin, bestN, bestD, N_out, a_out);
}

/* sphere: m = o - c, A = d.d, B = m.d, C = m.m - r*r, disc = B*B - A*C, roots
 * (-B -+ sqrt(disc)) / A, nearer root first (rt_detmath.h: rt_sphere_test). */
/* stage 1: B and the discriminant.  R = center.xyz radius, or for camera rays the prepared m.xyz C */
template <bool PRIMARY>
__device__ __forceinline__ void sphere_disc(float4 R, f3 o, f3 d, float A, float &B, float &disc) {
    f3 m; float C;
    if (PRIMARY) { m = mk(R.x, R.y, R.z); C = R.w; }
    else { m = o - mk(R.x, R.y, R.z); C = fmaf(-R.w, R.w, dot(m, m)); }
    B = dot(m, d);
    disc = fmaf(B, B, -(A * C));
}
/* stage 2, only where some lane has a real root: the square root and the two root tests */
__device__ __forceinline__ bool sphere_roots(float B, float disc, float A, float tmin, float bestN, float bestD, float &N_out) {
    if (!(disc >= 0.0f)) return false;
    float sq = sqrtf(disc);
    float N0 = -B - sq, N1 = -B + sq;
    bool ok0 = (N0 >= tmin * A) & (N0 * bestD < bestN * A);
    bool ok1 = (N1 >= tmin * A) & (N1 * bestD < bestN * A);
    N_out = ok0 ? N0 : N1;
    return ok0 | ok1;
}

/* The intersection part of the prepared scene, as float4 slots counted from the
 * start of the isect section: triangle i = slots 4i..4i+3, sphere j = slots
 * sph+2j, sph+2j+1.  CONST_BANK reads the constant-bank mirror, otherwise shared memory. */
template <bool CONST_BANK>
struct SceneRef {
    const float4 *base;
    uint32_t T, S;               /* PADDED loop bounds: T rounded up to 4, S to 2 */
    uint32_t Treal, sph;
    const float *grid = nullptr;   /* the sphere grid's header in global memory (sgrid.cuh), or null: brute force over the spheres */
    __device__ __forceinline__ float4 ld(uint32_t slot) const { return CONST_BANK ? c_scene[slot] : base[slot]; }
};

/* Both loops run over the PADDED counts T4 = T rounded up to 4 and S4 = S rounded up
 * to 4 (the prepared scene pads with primitives that can never be accepted), so the
 * unrolled bodies need no remainder code: less SASS, same results. */
template <bool CONST_BANK, bool PRIMARY, int UNROLL, bool PAIRED = false>
__device__ __forceinline__ bool closest_hit(const SceneRef<CONST_BANK> &sc, f3 o, f3 d, float tmin, float tmax,
                                            float &dist, uint32_t &id) {
    static_assert(!PAIRED || (UNROLL % 2 == 0 && !PRIMARY), "PAIRED: see any_hit");
    float bestN = tmax, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < sc.T; i0 += UNROLL) {
        f3 w = mk(0.0f, 0.0f, 0.0f), c = w;
#pragma unroll
        for (uint32_t u = 0; u < UNROLL; u++) {
            const uint32_t i = i0 + u;
            float N, a;
            bool hit;
            if (PAIRED) {                              /* second triangle of a pair: same v0, so same w and c */
                const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
                const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
                if (!(u & 1)) { w = mk(A.x, A.y, A.z) - o; c = cross(d, w); }
                hit = tri_accept(dot(d, n), dot(w, n), dot(e2, c), dot(e1n, c), tmin, bestN, bestD, N, a);
            } else {
                hit = PRIMARY ? tri_test_primary(sc.ld(4 * i + 3), sc.ld(4 * i), sc.ld(4 * i + 1), sc.ld(4 * i + 2), d, tmin, bestN, bestD, N, a)
                              : tri_test(sc.ld(4 * i), sc.ld(4 * i + 1), sc.ld(4 * i + 2), o, d, tmin, bestN, bestD, N, a);
            }
            if (hit) { bestN = N; bestD = a; best = i; }
        }
    }
    if (sc.S) {
        const float A = dot(d, d);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < sc.S; j0 += 4) {
            /* four discriminants first (the common case: all negative, nothing else to do) */
            float Bq[4], disc[4];
#pragma unroll
            for (uint32_t u = 0; u < 4; u++) sphere_disc<PRIMARY>(sc.ld(sc.sph + 2 * (j0 + u) + (PRIMARY ? 1 : 0)), o, d, A, Bq[u], disc[u]);
            if (fmaxf(fmaxf(disc[0], disc[1]), fmaxf(disc[2], disc[3])) >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    float N;
                    if (sphere_roots(Bq[u], disc[u], A, tmin, bestN, bestD, N)) { bestN = N; bestD = A; best = sc.Treal + j0 + u; }
                }
            }
        }
    }
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/*
 * PAIRED (all loops below): rt_prep_scene found that every triangle 2k+1 has bit for bit the v0 of triangle 2k
 * (a quad cut in two, the way scene.swift builds every wall, box face and the light).  Then w = v0 - o and
 * c = cross(d, w) of the second triangle ARE the first one's -- same operands, same operations -- and are not
 * computed again: 9 of the 21 FP32 instructions of every second test.  Nothing is approximated; the results are
 * the unpaired loop's bit for bit (tests run both).
 */
template <bool CONST_BANK, int UNROLL, bool PAIRED = false>
__device__ __forceinline__ bool any_hit(const SceneRef<CONST_BANK> &sc, f3 o, f3 d, float tmin, float tmax) {
    static_assert(!PAIRED || UNROLL % 2 == 0, "pairs must not straddle an unrolled block");
    bool hit = false;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < sc.T; i0 += UNROLL) {
        f3 w = mk(0.0f, 0.0f, 0.0f), c = w;
#pragma unroll
        for (uint32_t u = 0; u < UNROLL; u++) {
            const uint32_t i = i0 + u;
            float N, a;
            if (!PAIRED) {
                hit |= tri_test(sc.ld(4 * i), sc.ld(4 * i + 1), sc.ld(4 * i + 2), o, d, tmin, tmax, 1.0f, N, a);
            } else {
                const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
                const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
                if (!(u & 1)) { w = mk(A.x, A.y, A.z) - o; c = cross(d, w); }
                hit |= tri_accept(dot(d, n), dot(w, n), dot(e2, c), dot(e1n, c), tmin, tmax, 1.0f, N, a);
            }
        }
    }
    if (sc.S && !CONST_BANK && sc.grid) {
        if (!hit) hit = sgrid_any(sc.grid, sc.base + sc.sph, o, d, tmax);      /* (shadow rays: tmin = 0) */
    } else if (sc.S) {
        const float A = dot(d, d);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < sc.S; j0 += 4) {
            float Bq[4], disc[4];
#pragma unroll
            for (uint32_t u = 0; u < 4; u++) sphere_disc<false>(sc.ld(sc.sph + 2 * (j0 + u)), o, d, A, Bq[u], disc[u]);
            if (fmaxf(fmaxf(disc[0], disc[1]), fmaxf(disc[2], disc[3])) >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    float N;
                    hit |= sphere_roots(Bq[u], disc[u], A, tmin, tmax, 1.0f, N);
                }
            }
        }
    }
    return hit;
}

/*
 * Two rays from ONE origin in one pass over the primitives: the closest-hit query of the
 * continuation ray (direction dc) and the any-hit query of the shadow ray (direction ds) of a
 * bounce both start at the shading point (raytrace.metal:79-100).  w = v0 - o and Nn = dot(w, n)
 * do not depend on the direction, so they -- and the three LDS.128 of the triangle record -- are
 * shared: 36 instead of 42 FP32 instructions per triangle for the pair, and two independent
 * dependency chains in flight.  Same operations on the same operands as tri_test, so the
 * results are bit-identical to closest_hit + any_hit.  (Spheres: m = o - c and C are shared.)
 */
template <bool CONST_BANK, int UNROLL, bool PAIRED = false>
__device__ __forceinline__ bool dual_hit(const SceneRef<CONST_BANK> &sc, f3 o,
                                         f3 dc, float tminC, float tmaxC, float &dist, uint32_t &id,
                                         f3 ds, float tminS, float tmaxS, bool &shadowed) {
    static_assert(!PAIRED || UNROLL % 2 == 0, "pairs must not straddle an unrolled block");
    float bestN = tmaxC, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    bool sh = false;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < sc.T; i0 += UNROLL) {
        f3 w = mk(0.0f, 0.0f, 0.0f), cC = w, cS = w;
#pragma unroll
        for (uint32_t u = 0; u < UNROLL; u++) {
            const uint32_t i = i0 + u;
            const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
            const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
            if (!PAIRED || !(u & 1)) {                 /* PAIRED: the second triangle of a pair has the same v0 */
                w = mk(A.x, A.y, A.z) - o;
                cC = cross(dc, w);
                cS = cross(ds, w);
            }
            const float Nn = dot(w, n);
            float N, a;
            {
                float D = dot(dc, n);
                float U = dot(e2, cC), V = dot(e1n, cC);
                if (tri_accept(D, Nn, U, V, tminC, bestN, bestD, N, a)) { bestN = N; bestD = a; best = i; }
            }
            {
                float D = dot(ds, n);
                float U = dot(e2, cS), V = dot(e1n, cS);
                sh |= tri_accept(D, Nn, U, V, tminS, tmaxS, 1.0f, N, a);
            }
        }
    }
    if (sc.S && !CONST_BANK && sc.grid) {
        sgrid_closest(sc.grid, sc.base + sc.sph, sc.Treal, o, dc, tminC, bestN, bestD, best);
        if (!sh) sh = sgrid_any(sc.grid, sc.base + sc.sph, o, ds, tmaxS);
    } else if (sc.S) {
        const float Ac = dot(dc, dc), As = dot(ds, ds);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < sc.S; j0 += 4) {
            float Bc[4], discC[4], Bs[4], discS[4];
#pragma unroll
            for (uint32_t u = 0; u < 4; u++) {
                const float4 R = sc.ld(sc.sph + 2 * (j0 + u));
                const f3 m = o - mk(R.x, R.y, R.z);
                const float Cq = fmaf(-R.w, R.w, dot(m, m));
                Bc[u] = dot(m, dc); discC[u] = fmaf(Bc[u], Bc[u], -(Ac * Cq));
                Bs[u] = dot(m, ds); discS[u] = fmaf(Bs[u], Bs[u], -(As * Cq));
            }
            if (fmaxf(fmaxf(discC[0], discC[1]), fmaxf(discC[2], discC[3])) >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    float N;
                    if (sphere_roots(Bc[u], discC[u], Ac, tminC, bestN, bestD, N)) { bestN = N; bestD = Ac; best = sc.Treal + j0 + u; }
                }
            }
            if (fmaxf(fmaxf(discS[0], discS[1]), fmaxf(discS[2], discS[3])) >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    float N;
                    sh |= sphere_roots(Bs[u], discS[u], As, tminS, tmaxS, 1.0f, N);
                }
            }
        }
    }
    shadowed = sh;
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/*
 * N shadow rays from ONE origin (scenes with several lights: one shadow ray per light per bounce, all
 * from the shading point): w, Nn (triangles) and m, C (spheres) are computed once per primitive.
 */
template <bool CONST_BANK, int UNROLL, int N>
__device__ __forceinline__ void any_hit_multi(const SceneRef<CONST_BANK> &sc, f3 o, const f3 (&d)[N], const float (&tmax)[N],
                                              bool (&shadowed)[N], uint32_t skipPairs = 0u) {
    bool sh[N];
#pragma unroll
    for (int r = 0; r < N; r++) sh[r] = false;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < sc.T; i0 += UNROLL) {
#pragma unroll
        for (uint32_t u = 0; u < UNROLL; u++) {
            const uint32_t i = i0 + u;
            if (skipPairs && ((skipPairs >> ((i >> 1) & 31u)) & 1u)) continue;   /* shell triangles no shadow ray of this pass can be accepted by (warp-uniform; wave.cuh) */
            const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
            const f3 v0 = mk(A.x, A.y, A.z), n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
            const f3 w = v0 - o;
            const float Nn = dot(w, n);
#pragma unroll
            for (int r = 0; r < N; r++) {
                float Nq, a;
                float D = dot(d[r], n);
                f3 c = cross(d[r], w);
                float U = dot(e2, c), V = dot(e1n, c);
                sh[r] |= tri_accept(D, Nn, U, V, 0.0f, tmax[r], 1.0f, Nq, a);
            }
        }
    }
    if (sc.S && !CONST_BANK && sc.grid) {
#pragma unroll
        for (int r = 0; r < N; r++)
            if (!sh[r] && tmax[r] > 0.0f) sh[r] = sgrid_any(sc.grid, sc.base + sc.sph, o, d[r], tmax[r]);
    } else if (sc.S) {
        float Aq[N];
#pragma unroll
        for (int r = 0; r < N; r++) Aq[r] = dot(d[r], d[r]);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < sc.S; j0 += 2) {
            float Bq[2][N], disc[2][N];
            float mx = -1.0f;
#pragma unroll
            for (uint32_t u = 0; u < 2; u++) {
                const float4 R = sc.ld(sc.sph + 2 * (j0 + u));
                const f3 m = o - mk(R.x, R.y, R.z);
                const float Cq = fmaf(-R.w, R.w, dot(m, m));
#pragma unroll
                for (int r = 0; r < N; r++) {
                    Bq[u][r] = dot(m, d[r]);
                    disc[u][r] = fmaf(Bq[u][r], Bq[u][r], -(Aq[r] * Cq));
                    mx = fmaxf(mx, disc[u][r]);
                }
            }
            if (mx >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 2; u++)
#pragma unroll
                    for (int r = 0; r < N; r++) {
                        float Nq;
                        sh[r] |= sphere_roots(Bq[u][r], disc[u][r], Aq[r], 0.0f, tmax[r], 1.0f, Nq);
                    }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < N; r++) shadowed[r] = sh[r];
}

/*
 * ALL rays of a bounce in ONE pass over the primitives: the continuation ray's closest-hit query and up to NS shadow rays
 * (scenes with several lights), every one of them starting at the shading point.  dual_hit followed by any_hit_multi
 * walked the scene twice and computed w, Nn (triangles) and m, C (spheres) twice; here they are computed once per
 * primitive and bounce (256 spheres: 39 -> 32 instructions per sphere and bounce).  Same operations on the same
 * operands in the same order per ray as tri_test / the sphere quadratic, the closest hit's candidates still met in
 * ascending id: bit-identical results.  A shadow ray that is not wanted is given tmax = 0: nothing can be nearer.
 */
template <int UNROLL, int NS>
__device__ __forceinline__ bool multi_hit(const SceneRef<false> &sc, f3 o, f3 dc, float tminC, float tmaxC, float &dist, uint32_t &id,
                                          const f3 (&ds)[NS], const float (&tmaxS)[NS], bool (&shadowed)[NS], uint32_t skipPairs = 0u) {
    float bestN = tmaxC, bestD = 1.0f;
    uint32_t best = 0xffffffffu;
    bool sh[NS];
#pragma unroll
    for (int r = 0; r < NS; r++) sh[r] = false;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < sc.T; i0 += UNROLL) {
#pragma unroll
        for (uint32_t u = 0; u < UNROLL; u++) {
            const uint32_t i = i0 + u;
            const float4 A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
            const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
            const f3 w = mk(A.x, A.y, A.z) - o;
            const float Nn = dot(w, n);
            float N, a;
            {
                const float D = dot(dc, n);
                const f3 c = cross(dc, w);
                if (tri_accept(D, Nn, dot(e2, c), dot(e1n, c), tminC, bestN, bestD, N, a)) { bestN = N; bestD = a; best = i; }
            }
            if (skipPairs && ((skipPairs >> ((i >> 1) & 31u)) & 1u)) continue;   /* shell triangles: the continuation ray only (warp-uniform; wave.cuh) */
#pragma unroll
            for (int r = 0; r < NS; r++) {
                const float D = dot(ds[r], n);
                const f3 c = cross(ds[r], w);
                sh[r] |= tri_accept(D, Nn, dot(e2, c), dot(e1n, c), 0.0f, tmaxS[r], 1.0f, N, a);
            }
        }
    }
    if (sc.S && sc.grid) {
        sgrid_closest(sc.grid, sc.base + sc.sph, sc.Treal, o, dc, tminC, bestN, bestD, best);
#pragma unroll
        for (int r = 0; r < NS; r++)
            if (!sh[r] && tmaxS[r] > 0.0f) sh[r] = sgrid_any(sc.grid, sc.base + sc.sph, o, ds[r], tmaxS[r]);
    } else if (sc.S) {
        const float Ac = dot(dc, dc);
        float As[NS];
#pragma unroll
        for (int r = 0; r < NS; r++) As[r] = dot(ds[r], ds[r]);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < sc.S; j0 += 2) {
            float Bc[2], discC[2], Bs[2][NS], discS[2][NS];
            float mx = -1.0f;
#pragma unroll
            for (uint32_t u = 0; u < 2; u++) {
                const float4 R = sc.ld(sc.sph + 2 * (j0 + u));
                const f3 m = o - mk(R.x, R.y, R.z);
                const float Cq = fmaf(-R.w, R.w, dot(m, m));
                Bc[u] = dot(m, dc); discC[u] = fmaf(Bc[u], Bc[u], -(Ac * Cq));
                mx = fmaxf(mx, discC[u]);
#pragma unroll
                for (int r = 0; r < NS; r++) {
                    Bs[u][r] = dot(m, ds[r]);
                    discS[u][r] = fmaf(Bs[u][r], Bs[u][r], -(As[r] * Cq));
                    mx = fmaxf(mx, discS[u][r]);
                }
            }
            if (mx >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 2; u++) {
                    float N;
                    if (sphere_roots(Bc[u], discC[u], Ac, tminC, bestN, bestD, N)) { bestN = N; bestD = Ac; best = sc.Treal + j0 + u; }
#pragma unroll
                    for (int r = 0; r < NS; r++) sh[r] |= sphere_roots(Bs[u][r], discS[u][r], As[r], 0.0f, tmaxS[r], 1.0f, N);
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < NS; r++) shadowed[r] = sh[r];
    id = best;
    dist = bestN / bestD;
    return best != 0xffffffffu;
}

/*
 * Two CAMERA rays (two samples of one pixel) in one pass: the origin is the camera for both, so the
 * four LDS.128 of a triangle record and the loop bookkeeping are shared and two independent chains
 * are in flight.  Same operations as closest_hit<PRIMARY>: bit-identical results.
 */
template <bool CONST_BANK, int UNROLL>
__device__ __forceinline__ void closest_primary2(const SceneRef<CONST_BANK> &sc, f3 d0, f3 d1, float tmin, float tmax,
                                                 bool &hit0, float &dist0, uint32_t &id0,
                                                 bool &hit1, float &dist1, uint32_t &id1) {
    float bN0 = tmax, bD0 = 1.0f, bN1 = tmax, bD1 = 1.0f;
    uint32_t b0 = 0xffffffffu, b1 = 0xffffffffu;
#pragma unroll 1
    for (uint32_t i0 = 0; i0 < sc.T; i0 += UNROLL) {
#pragma unroll
        for (uint32_t u = 0; u < UNROLL; u++) {
            const uint32_t i = i0 + u;
            const float4 P = sc.ld(4 * i + 3), A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
            float N, a;
            if (tri_test_primary(P, A, B, C, d0, tmin, bN0, bD0, N, a)) { bN0 = N; bD0 = a; b0 = i; }
            if (tri_test_primary(P, A, B, C, d1, tmin, bN1, bD1, N, a)) { bN1 = N; bD1 = a; b1 = i; }
        }
    }
    if (sc.S) {
        const float A0 = dot(d0, d0), A1 = dot(d1, d1);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < sc.S; j0 += 4) {
            float B0[4], disc0[4], B1[4], disc1[4];
#pragma unroll
            for (uint32_t u = 0; u < 4; u++) {
                const float4 R = sc.ld(sc.sph + 2 * (j0 + u) + 1);
                sphere_disc<true>(R, d0, d0, A0, B0[u], disc0[u]);
                sphere_disc<true>(R, d1, d1, A1, B1[u], disc1[u]);
            }
            if (fmaxf(fmaxf(disc0[0], disc0[1]), fmaxf(disc0[2], disc0[3])) >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    float N;
                    if (sphere_roots(B0[u], disc0[u], A0, tmin, bN0, bD0, N)) { bN0 = N; bD0 = A0; b0 = sc.Treal + j0 + u; }
                }
            }
            if (fmaxf(fmaxf(disc1[0], disc1[1]), fmaxf(disc1[2], disc1[3])) >= 0.0f) {
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    float N;
                    if (sphere_roots(B1[u], disc1[u], A1, tmin, bN1, bD1, N)) { bN1 = N; bD1 = A1; b1 = sc.Treal + j0 + u; }
                }
            }
        }
    }
    id0 = b0; dist0 = bN0 / bD0; hit0 = b0 != 0xffffffffu;
    id1 = b1; dist1 = bN1 / bD1; hit1 = b1 != 0xffffffffu;
}

/*
 * closest_primary2 over a CANDIDATE LIST: bit k of cullT[wd] / cullS[wd] says primitive 32*wd + k may be
 * hit by a camera ray through this warp's pixels (camera_candidates below; warp-uniform words in shared
 * memory).  Candidates are visited in ascending index, like the full loop, so the strict-less tie rule
 * picks the same primitive; a primitive that is left out could not have been accepted.  Two candidates per
 * pass; an odd one is paired with itself (a second test of the same primitive never passes the
 * strict-less comparison against its own distance).
 */
__device__ __forceinline__ void closest_primary2_list(const SceneRef<false> &sc, const uint32_t *cullT, uint32_t mwT,
                                                      const uint32_t *cullS, uint32_t mwS,
                                                      f3 d0, f3 d1, float tmin, float tmax,
                                                      bool &hit0, float &dist0, uint32_t &id0,
                                                      bool &hit1, float &dist1, uint32_t &id1) {
    float bN0 = tmax, bD0 = 1.0f, bN1 = tmax, bD1 = 1.0f;
    uint32_t b0 = 0xffffffffu, b1 = 0xffffffffu;
#pragma unroll 1
    for (uint32_t wd = 0; wd < mwT; wd++) {
        uint32_t m = cullT[wd];
#pragma unroll 1
        while (m) {
            const uint32_t i = (wd << 5) + (uint32_t)__ffs((int)m) - 1u; m &= m - 1u;
            uint32_t j = i;
            if (m) { j = (wd << 5) + (uint32_t)__ffs((int)m) - 1u; m &= m - 1u; }
            const float4 Pi = sc.ld(4 * i + 3), Ai = sc.ld(4 * i), Bi = sc.ld(4 * i + 1), Ci = sc.ld(4 * i + 2);
            const float4 Pj = sc.ld(4 * j + 3), Aj = sc.ld(4 * j), Bj = sc.ld(4 * j + 1), Cj = sc.ld(4 * j + 2);
            float N, a;
            if (tri_test_primary(Pi, Ai, Bi, Ci, d0, tmin, bN0, bD0, N, a)) { bN0 = N; bD0 = a; b0 = i; }
            if (tri_test_primary(Pi, Ai, Bi, Ci, d1, tmin, bN1, bD1, N, a)) { bN1 = N; bD1 = a; b1 = i; }
            if (tri_test_primary(Pj, Aj, Bj, Cj, d0, tmin, bN0, bD0, N, a)) { bN0 = N; bD0 = a; b0 = j; }
            if (tri_test_primary(Pj, Aj, Bj, Cj, d1, tmin, bN1, bD1, N, a)) { bN1 = N; bD1 = a; b1 = j; }
        }
    }
    if (mwS) {
        const float A0 = dot(d0, d0), A1 = dot(d1, d1);
#pragma unroll 1
        for (uint32_t wd = 0; wd < mwS; wd++) {
            uint32_t m = cullS[wd];
#pragma unroll 1
            while (m) {
                const uint32_t i = (wd << 5) + (uint32_t)__ffs((int)m) - 1u; m &= m - 1u;
                uint32_t j = i;
                if (m) { j = (wd << 5) + (uint32_t)__ffs((int)m) - 1u; m &= m - 1u; }
                const float4 Ri = sc.ld(sc.sph + 2 * i + 1), Rj = sc.ld(sc.sph + 2 * j + 1);
                float B, disc, N;
                sphere_disc<true>(Ri, d0, d0, A0, B, disc);
                if (sphere_roots(B, disc, A0, tmin, bN0, bD0, N)) { bN0 = N; bD0 = A0; b0 = sc.Treal + i; }
                sphere_disc<true>(Ri, d1, d1, A1, B, disc);
                if (sphere_roots(B, disc, A1, tmin, bN1, bD1, N)) { bN1 = N; bD1 = A1; b1 = sc.Treal + i; }
                sphere_disc<true>(Rj, d0, d0, A0, B, disc);
                if (sphere_roots(B, disc, A0, tmin, bN0, bD0, N)) { bN0 = N; bD0 = A0; b0 = sc.Treal + j; }
                sphere_disc<true>(Rj, d1, d1, A1, B, disc);
                if (sphere_roots(B, disc, A1, tmin, bN1, bD1, N)) { bN1 = N; bD1 = A1; b1 = sc.Treal + j; }
            }
        }
    }
    id0 = b0; dist0 = bN0 / bD0; hit0 = b0 != 0xffffffffu;
    id1 = b1; dist1 = bN1 / bD1; hit1 = b1 != 0xffffffffu;
}

/*
 * Camera-ray candidates of one warp.  Every camera ray of the warp passes through the pixel rectangle
 * [x0, x1+1) x [y0, y1+1) of the image plane (sampling.metal:141-147: pixel + jitter in [0,1)), so a
 * primitive whose perspective projection stays clear of that rectangle by more than one pixel cannot be
 * hit: lane l projects primitive 32*wd + l (its three vertices; for a sphere the box bound of x/z and
 * y/z over its bounding cube), and the ballot of "not provably outside" is the candidate word.
 * The exact test's own rounding lets it accept rays a few 1e-4 pixel outside the true silhouette
 * (error of U, V ~ 1e-5 against a = |d.n| ~ |n| cos(theta)); the one-pixel margin covers that unless
 * the primitive is seen within 0.6 degrees of edge-on (or reaches behind the camera, or is a speck),
 * and those are always kept.  Every comparison is written so that a NaN keeps the primitive.
 */
struct CameraFrame { f3 o, u, v, w; float halfW, halfH, resX, resY; };

__device__ __forceinline__ void project_px(const CameraFrame &cf, f3 q, float &X, float &Y, float &zc) {
    zc = -dot(q, cf.w);
    const float xc = dot(q, cf.u), yc = dot(q, cf.v);
    X = (xc / (zc * cf.halfW) + 1.0f) * (0.5f * cf.resX);
    Y = (1.0f - yc / (zc * cf.halfH)) * (0.5f * cf.resY);
}

__device__ __forceinline__ void camera_candidates(const SceneRef<false> &sc, const CameraFrame &cf, uint32_t nT, uint32_t nS,
                                                  float rx0, float rx1, float ry0, float ry1,
                                                  uint32_t *cullT, uint32_t mwT, uint32_t *cullS, uint32_t mwS, uint32_t lane) {
    for (uint32_t wd = 0; wd < mwT; wd++) {
        const uint32_t i = (wd << 5) + lane;
        bool keep = false;
        if (i < nT) {
            const float4 P = sc.ld(4 * i + 3), A = sc.ld(4 * i), B = sc.ld(4 * i + 1), C = sc.ld(4 * i + 2);
            const f3 n = mk(A.w, B.x, B.y), e2 = mk(B.z, B.w, C.x), e1n = mk(C.y, C.z, C.w);
            const f3 q0 = mk(P.x, P.y, P.z), q1 = q0 - e1n, q2 = q0 + e2;
            float X0, Y0, z0, X1, Y1, z1, X2, Y2, z2;
            project_px(cf, q0, X0, Y0, z0); project_px(cf, q1, X1, Y1, z1); project_px(cf, q2, X2, Y2, z2);
            const float far2 = fmaxf(fmaxf(dot(q0, q0), dot(q1, q1)), dot(q2, q2));
            const float far = sqrtf(far2);
            const bool safe = (fminf(fminf(z0, z1), z2) > 0.01f * far) && (fabsf(P.w) > 0.01f * far * sqrtf(dot(n, n)));
            const bool outside = (fmaxf(fmaxf(X0, X1), X2) < rx0) || (fminf(fminf(X0, X1), X2) > rx1) ||
                                 (fmaxf(fmaxf(Y0, Y1), Y2) < ry0) || (fminf(fminf(Y0, Y1), Y2) > ry1);
            keep = !(safe && outside);
        }
        const uint32_t word = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) cullT[wd] = word;
    }
    for (uint32_t wd = 0; wd < mwS; wd++) {
        const uint32_t j = (wd << 5) + lane;
        bool keep = false;
        if (j < nS) {
            const float4 Sg = sc.ld(sc.sph + 2 * j);
            const f3 q = mk(Sg.x, Sg.y, Sg.z) - cf.o;
            const float r = fabsf(Sg.w);
            const float zc = -dot(q, cf.w), xc = dot(q, cf.u), yc = dot(q, cf.v);
            const float far = sqrtf(dot(q, q));
            const bool safe = (zc - r > 0.01f * (far + r)) && (r > 1e-3f * far);
            const float zn = zc - r, zf = zc + r;
            const float txh = (xc + r) / ((xc + r >= 0.0f) ? zn : zf), txl = (xc - r) / ((xc - r <= 0.0f) ? zn : zf);
            const float tyh = (yc + r) / ((yc + r >= 0.0f) ? zn : zf), tyl = (yc - r) / ((yc - r <= 0.0f) ? zn : zf);
            const float Xh = (txh / cf.halfW + 1.0f) * (0.5f * cf.resX), Xl = (txl / cf.halfW + 1.0f) * (0.5f * cf.resX);
            const float Yl = (1.0f - tyh / cf.halfH) * (0.5f * cf.resY), Yh = (1.0f - tyl / cf.halfH) * (0.5f * cf.resY);
            const bool outside = (Xh < rx0) || (Xl > rx1) || (Yh < ry0) || (Yl > ry1);
            keep = !(safe && outside);
        }
        const uint32_t word = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) cullS[wd] = word;
    }
    __syncwarp();
}

/* ---------------------------------------------------------------- shading ---- */
/* sampling.metal:198-236 with the zeros of :207-209 folded:
 * samplePosition = center + (hx,0,0)*ux + (0,0,hz)*uy.  The products with the
 * zero components are +-0 and drop out exactly, so only x and z move.  (hx,hz) come
 * from the prepared light (rt_prep_scene): the reference's 0.25, 0.25 unless the
 * launch asks for the light's own extent (RT_VARIANT_LIGHT_EXTENT). */
/* the three steps of sampleAreaLight, separately callable (sorted.cuh keeps the sample point and repeats steps
 * 2 and 3 when it traces): the same operations in the same order as the one-piece form below */
__device__ __forceinline__ f3 light_sample_point(f3 center, float hx, float hz, float ux, float uy) {
    ux = ux * 2.0f - 1.0f;
    uy = uy * 2.0f - 1.0f;
    return mk(center.x + hx * ux, center.y, center.z + hz * uy);
}
__device__ __forceinline__ void light_direction(f3 samplePosition, f3 P, f3 &ldir, float &ldist, float &inv) {
    f3 ld = samplePosition - P;
    float dist = length(ld);
    inv = 1.0f / fmaxf(dist, 1e-3f);
    ldir = ld * inv; ldist = dist;
}
__device__ __forceinline__ f3 light_colour(f3 lcolor, float inv, f3 ldir) {
    f3 c = lcolor * (inv * inv);
    /* dot(-ld, (0,-1,0)) = fma(-z,0, fma(-y,-1, -x*0)) = ld.y (the zero terms are exact) */
    return c * saturate(ldir.y);
}
__device__ __forceinline__ f3 sample_area_light(const float *L, float ux, float uy, f3 P, f3 &ldir, float &ldist) {
    const f3 lcolor = mk(L[4], L[5], L[6]);
    float inv;
    if (((const uint32_t *)L)[RT_LIGHT_GENERAL] != 0u) {
        /* a light with a normal of its own (prepared by rt_prep_scene): sampling.metal:198-236 written out over
         * right = tangent * hx, up = bitangent * hz and that normal -- nothing folds to zero here */
        ux = ux * 2.0f - 1.0f;
        uy = uy * 2.0f - 1.0f;
        const f3 right = mk(L[RT_LIGHT_TANGENT], L[RT_LIGHT_TANGENT + 1], L[RT_LIGHT_TANGENT + 2]) * L[3];
        const f3 up = mk(L[RT_LIGHT_BITANGENT], L[RT_LIGHT_BITANGENT + 1], L[RT_LIGHT_BITANGENT + 2]) * L[7];
        const f3 sp = (mk(L[0], L[1], L[2]) + right * ux) + up * uy;
        light_direction(sp, P, ldir, ldist, inv);
        f3 c = lcolor * (inv * inv);
        return c * saturate(dot(-ldir, mk(L[RT_LIGHT_NORMAL], L[RT_LIGHT_NORMAL + 1], L[RT_LIGHT_NORMAL + 2])));
    }
    light_direction(light_sample_point(mk(L[0], L[1], L[2]), L[3], L[7], ux, uy), P, ldir, ldist, inv);
    return light_colour(lcolor, inv, ldir);
}
/* The form of the passes over several lights (wave.cuh): ONE copy of the code whichever light it is called for, with the
 * cosine factor of raytrace.metal:75 applied.  Inlined into the unrolled four-light setup it came to 14 KB of the 42 KB the
 * 256-sphere frame keeps executing -- against a 32 KB instruction cache, and the kernel waited for instructions 43 % of the
 * time (profiles/r04r_c_ncu_full_summary.txt).  out: (ldir.xyz, ldist - 1e-3), (colour.xyz, -). */
__device__ __noinline__ void sample_area_light_shaded(const float *L, float ux, float uy, f3 P, f3 normal, float4 &dirTmax, float4 &col) {
    f3 ldir; float ldist;
    f3 c = sample_area_light(L, ux, uy, P, ldir, ldist);
    c = c * saturate(dot(normal, ldir));
    dirTmax = make_float4(ldir.x, ldir.y, ldir.z, ldist - 1e-3f);
    col = make_float4(c.x, c.y, c.z, 0.0f);
}
#include "mis.cuh"

/* sampling.metal:39-49 */
__device__ __forceinline__ f3 sample_cosine_hemisphere(float ux, float uy) {
    float phi = 6.2831854820251465f * ux;          /* (2.0f * M_PI_F) * u.x */
    float sp, cp;
    sincos_det(phi, sp, cp);
    float ct = sqrtf(uy);
    float st = sqrtf(1.0f - ct * ct);
    return mk(st * cp, ct, st * sp);
}

/* -------------------------------------------------------------- the kernel ---- */
#define RT_BLOCK_W 16
#define RT_BLOCK_H 8
#define RT_BLOCK_THREADS 128

struct Counters { uint32_t closest, shadow, shaded; };

/*
 * One thread per pixel; the sample loop and the bounce loop are both ROLLED: the
 * whole kernel is ~25 KB of SASS and its two primitive loops fit the 6 KB L0
 * instruction cache (profiles/: the first version unrolled the bounces to get
 * compile-time Halton bases, was 112 KB, and stalled on instruction fetch).
 * The bounce index is warp-uniform (lanes whose path ended are merely inactive),
 * so the Halton base of bounce b comes out of a constant table through the
 * uniform datapath (halton_slot).  Bounce 0 uses the camera-ray form of the
 * triangle test.
 */
template <bool CONST_BANK, bool COUNT, int UNROLL, bool MIS = false>
__device__ __forceinline__ void pathtrace_body(const float *__restrict__ scene, const uint32_t *__restrict__ offsets,
                                               float4 *__restrict__ out, const RTLaunchParams p,
                                               const RTSceneLayout lay, float *sm,
                                               unsigned long long *__restrict__ counters) {
    /* stage the prepared scene HBM -> shared memory (coalesced 128-bit copies) */
    {
        const float4 *src = (const float4 *)scene;
        float4 *dst = (float4 *)sm;
        const uint32_t n4 = lay.words >> 2;
        for (uint32_t k = threadIdx.x; k < n4; k += RT_BLOCK_THREADS) dst[k] = src[k];
    }
    __syncthreads();

    /* warp = 8x4 pixels, block = 2x2 warps = 16x8 pixels */
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t x = blockIdx.x * RT_BLOCK_W + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = blockIdx.y * RT_BLOCK_H + (warp >> 1) * 4 + (lane >> 3);
    if (x >= (uint32_t)p.width || ly >= (uint32_t)p.rows) return;
    /* contiguous rows, or interleaved row tiles (multi-GPU load balance) */
    const uint32_t y = p.tileRows ? (uint32_t)p.row0 + (ly / p.tileRows) * p.tileStride + (ly % p.tileRows)
                                  : (uint32_t)p.row0 + ly;
    const size_t pix = (size_t)ly * (uint32_t)p.width + x;

    SceneRef<CONST_BANK> sc;
    sc.base = (const float4 *)(sm + lay.isect);
    sc.T = (lay.T + (UNROLL - 1)) / UNROLL * UNROLL; sc.S = (lay.S + 3u) & ~3u; sc.Treal = lay.T;
    sc.sph = 4 * lay.Tp;

    const f3 cam_o = mk(sm[RT_HDR_CAM_O], sm[RT_HDR_CAM_O + 1], sm[RT_HDR_CAM_O + 2]);
    const f3 cam_u = mk(sm[RT_HDR_CAM_U], sm[RT_HDR_CAM_U + 1], sm[RT_HDR_CAM_U + 2]);
    const f3 cam_v = mk(sm[RT_HDR_CAM_V], sm[RT_HDR_CAM_V + 1], sm[RT_HDR_CAM_V + 2]);
    const f3 cam_w = mk(sm[RT_HDR_CAM_W], sm[RT_HDR_CAM_W + 1], sm[RT_HDR_CAM_W + 2]);
    const float halfW = sm[RT_HDR_HALF_W], halfH = sm[RT_HDR_HALF_H];
    const float resX = sm[RT_HDR_RES_X], resY = sm[RT_HDR_RES_Y];
    const float fx = (float)(int)x, fy = (float)(int)y;

    const uint32_t offset = offsets[pix];                               /* raytrace.metal:37, hoisted */
    const uint32_t nb = (uint32_t)p.bounces, numLights = p.numLights;
    Counters cnt = { 0u, 0u, 0u };
    f3 lum = mk(0.0f, 0.0f, 0.0f);
    if (p.accumulate) { float4 o = out[pix]; lum = mk(o.x, o.y, o.z); }

#pragma unroll 1
    for (uint32_t k = 0; k < p.spp; k++) {
        const uint32_t i = offset + (p.firstSample + k);
        /* camera ray, sampling.metal:141-147 */
        float jx = halton<0>(i), jy = halton<1>(i);
        float s = ((fx + jx) / resX) * 2.0f - 1.0f;
        float t = -(((fy + jy) / resY) * 2.0f - 1.0f);
        f3 rd = normalize((cam_u * (s * halfW) + cam_v * (t * halfH)) - cam_w);
        f3 ro = cam_o;

        f3 acc = mk(0.0f, 0.0f, 0.0f);
        f3 color = mk(1.0f, 1.0f, 1.0f);
        float dist; uint32_t id;
        if (COUNT) cnt.closest++;
        bool hit = closest_hit<CONST_BANK, true, UNROLL>(sc, ro, rd, 0.001f, 1000.0f, dist, id);

#pragma unroll 1
        for (uint32_t b = 0; hit; b++) {                                                        /* raytrace.metal:47-101 */
            f3 normal, right, fwd, diffuse;
            if (id < lay.T) {
                const float *sh = sm + lay.shade + id;
                if (sh[RT_SH_FLAG * lay.Tp] != 0.0f) {                                          /* :57-60 */
                    /* (MIS: only a camera ray takes the emissive, mis.cuh) */
                    if (!MIS || b == 0) acc = mk(sh[(RT_SH_ER + 0) * lay.Tp], sh[(RT_SH_ER + 1) * lay.Tp], sh[(RT_SH_ER + 2) * lay.Tp]);
                    break;
                }
                normal = mk(sh[(RT_SH_NX + 0) * lay.Tp], sh[(RT_SH_NX + 1) * lay.Tp], sh[(RT_SH_NX + 2) * lay.Tp]);
                right = mk(sh[(RT_SH_RX + 0) * lay.Tp], sh[(RT_SH_RX + 1) * lay.Tp], sh[(RT_SH_RX + 2) * lay.Tp]);
                fwd = mk(sh[(RT_SH_FX + 0) * lay.Tp], sh[(RT_SH_FX + 1) * lay.Tp], sh[(RT_SH_FX + 2) * lay.Tp]);
                diffuse = mk(sh[(RT_SH_DR + 0) * lay.Tp], sh[(RT_SH_DR + 1) * lay.Tp], sh[(RT_SH_DR + 2) * lay.Tp]);
            } else {
                uint32_t j = id - lay.T;
                const float *sh = sm + lay.sshade + j;
                if (sh[RT_SSH_FLAG * lay.Sp] != 0.0f) {
                    if (!MIS || b == 0) acc = mk(sh[(RT_SSH_ER + 0) * lay.Sp], sh[(RT_SSH_ER + 1) * lay.Sp], sh[(RT_SSH_ER + 2) * lay.Sp]);
                    break;
                }
                const float *S = sm + lay.sphere + RT_SPHERE_WORDS * j;
                f3 hp = ro + rd * dist;
                normal = normalize(hp - mk(S[0], S[1], S[2]));
                right = normalize(cross(normal, mk(0.0072f, 1.0f, 0.0034f)));
                fwd = cross(right, normal);
                diffuse = mk(sh[(RT_SSH_DR + 0) * lay.Sp], sh[(RT_SSH_DR + 1) * lay.Sp], sh[(RT_SSH_DR + 2) * lay.Sp]);
            }
            f3 P = (ro + rd * dist) + normal * 1e-3f;                                           /* :67 */
            if (COUNT) { cnt.shaded++; cnt.shadow += numLights; }

            float wx = halton_slot<0>(i, b), wy = halton_slot<1>(i, b);                         /* :72-73 */
            color = color * diffuse;                                                            /* :76 */
            for (uint32_t l = 0; l < numLights; l++) {
                const float *L = sm + lay.light + RT_LIGHT_WORDS * l;
                f3 ldir; float ldist;
                f3 lightColor = sample_area_light(L, wx, wy, P, ldir, ldist);
                lightColor = lightColor * saturate(dot(normal, ldir));                          /* :75 */
                if (MIS) lightColor = lightColor * misA_light_weight(L, P, normal, ldir);
                /* shadow ray: default-constructed => min_distance 0; max = dist - 1e-3   :79-85 */
                if (!any_hit<CONST_BANK, UNROLL>(sc, P, ldir, 0.0f, ldist - 1e-3f))
                    acc = acc + lightColor * color;                                             /* :88 */
            }
            if (!MIS && b + 1 >= nb) break;
            float ux = halton_slot<2>(i, b), uy = halton_slot<3>(i, b);                         /* :93-94 */
            f3 sdir = sample_cosine_hemisphere(ux, uy);
            rd = (right * sdir.x + normal * sdir.y) + fwd * sdir.z;                             /* sampling.metal:65 */
            ro = P;                                                                             /* :99-100 */
            if (COUNT) cnt.closest++;
            hit = closest_hit<CONST_BANK, false, UNROLL>(sc, ro, rd, 0.001f, 1000.0f, dist, id);        /* :49 */
            if (MIS) {      /* the cosine strategy's term of every light for this continuation ray (traced after the last bounce too) */
                for (uint32_t l = 0; l < numLights; l++)
                    acc = acc + misA_cosine_term(sm + lay.light + RT_LIGHT_WORDS * l, P, normal, rd, hit, dist, color);
                if (b + 1 >= nb) break;
            }
        }
        lum = lum + acc;                                                /* raytrace.metal:103 */
    }

    float4 o;
    if (p.accumulate) {
        o = make_float4(lum.x, lum.y, lum.z, (float)(p.firstSample + p.spp));
    } else {
        float samples = (float)p.spp;                                   /* raytrace.metal:106-109 */
        o = make_float4(lum.x / samples, lum.y / samples, lum.z / samples, 1.0f);
    }
    out[pix] = o;
    if (COUNT) {   /* exact work counters for the FLOP accounting; per-thread totals fit 32 bits */
        atomicAdd(&counters[0], (unsigned long long)p.spp);
        atomicAdd(&counters[1], (unsigned long long)cnt.closest);
        atomicAdd(&counters[2], (unsigned long long)cnt.shadow);
        atomicAdd(&counters[3], (unsigned long long)cnt.shaded);
    }
}

#define RT_PATHTRACE_ENTRY(name, CONST_BANK, COUNT, UNROLL, MINB)                                                \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters) {           \
        extern __shared__ __align__(16) float sm[];                                                              \
        pathtrace_body<CONST_BANK, COUNT, UNROLL>(scene, offsets, out, p, lay, sm, counters);                    \
    }
#ifndef RT_NOALLOC_BUILD   /* noalloc.cu includes this file for rt_prep_scene, the intersector and the wave body only: the default build is unchanged */
RT_PATHTRACE_ENTRY(pathTrace_smem, false, false, 4, 4)
RT_PATHTRACE_ENTRY(pathTrace_const, true, false, 4, 4)
RT_PATHTRACE_ENTRY(pathTrace_smem_count, false, true, 4, 4)
RT_PATHTRACE_ENTRY(pathTrace_const_count, true, true, 4, 4)
#endif
#ifndef RT_NOALLOC_BUILD
/* RT_VARIANT_MIS (mis.cuh): the per-pixel kernel with the two-strategy light term */
extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)
pathTrace_smem_mis(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,
                   const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters) {
    extern __shared__ __align__(16) float sm[];
    pathtrace_body<false, false, 4, true>(scene, offsets, out, p, lay, sm, counters);
}
#endif
#include "wave.cuh"
#define RT_WAVE_ENTRY(name, COUNT, UNROLL, R, MINB, SPATIAL)                                                            \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,             \
         const RTHaltonTables ht) {                                                                              \
        extern __shared__ __align__(16) float sm[];                                                              \
        pathtrace_wave_body<COUNT, UNROLL, R, false, SPATIAL>(scene, offsets, out, p, lay, sm, counters, RTTileSched{}, ht); \
    }
/* ticketed form: one block per (tile, sample chunk) unit, named by a ticket from an atomic counter (wave.cuh: RTTileSched) */
#define RT_WAVE_PERSIST_ENTRY(name, COUNT, UNROLL, R, MINB)                                                      \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,             \
         const RTHaltonTables ht, const RTTileSched ts) {                                                        \
        extern __shared__ __align__(16) float sm[];                                                              \
        pathtrace_wave_body<COUNT, UNROLL, R, true, false>(scene, offsets, out, p, lay, sm, counters, ts, ht);   \
    }
/* warp-level path compaction, R samples per pixel per round; rtcuda.cpp adds the queue memory.
 * R = 6 with four blocks per SM is the default (46 KB of queues per block): measured 5 615 Mpaths/s against
 * 5 392 for R = 8 with three blocks, 5 482 for R = 5, 4 889 for R = 12 with two (profiles/r02c_variants_R.txt).
 * Earlier in the kernel's life occupancy did not matter (r01c, r01h); it does since the primitive loops shrank
 * and the latency-bound code between them became a larger share. */
#ifndef RT_NOALLOC_BUILD   /* noalloc.cu includes this file for rt_prep_scene, the intersector and the wave body only: the default build is unchanged */
RT_WAVE_ENTRY(pathTrace_wave, false, 4, 6, 4, false)
RT_WAVE_ENTRY(pathTrace_wave_count, true, 4, 6, 4, false)
/* R = 4: two thirds of the queue memory; the library picks it when a large scene leaves room for fewer than
 * four R = 6 blocks per SM (256-sphere scene) */
/* fewer samples per launch than a round has slots: G = R / spp tiles side by side per warp (wave.cuh: SPATIAL) */
RT_WAVE_ENTRY(pathTrace_wide, false, 4, 6, 4, true)
RT_WAVE_ENTRY(pathTrace_wide4, false, 4, 4, 4, true)
RT_WAVE_PERSIST_ENTRY(pathTrace_tiles, false, 4, 6, 4)
RT_WAVE_PERSIST_ENTRY(pathTrace_tiles4, false, 4, 4, 4)
RT_WAVE_ENTRY(pathTrace_wave4, false, 4, 4, 4, false)
RT_WAVE_ENTRY(pathTrace_wave4_count, true, 4, 4, 4, false)
/* RT_VARIANT_MIS (mis.cuh): the compacting kernel with the two-strategy light term, R = 6 and R = 4 */
#define RT_WAVE_MIS_ENTRY(name, R)                                                                               \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)                                            \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,             \
         const RTHaltonTables ht) {                                                                              \
        extern __shared__ __align__(16) float sm[];                                                              \
        pathtrace_wave_body<false, 4, R, false, false, true>(scene, offsets, out, p, lay, sm, counters, RTTileSched{}, ht); \
    }
RT_WAVE_MIS_ENTRY(pathTrace_wave_mis, 6)
RT_WAVE_MIS_ENTRY(pathTrace_wave4_mis, 4)
#ifdef RT_TUNING_VARIANTS   /* queue depth x occupancy sweep (variants 5, 6, 10, 12 of tools/time_variants.py): tools/build_exp.sh */
RT_WAVE_ENTRY(pathTrace_exp10, false, 4, 2, 4, false)
RT_WAVE_ENTRY(pathTrace_exp8, false, 4, 8, 3, false)
/* (also measured: R = 4 with five blocks per SM at 102 registers, no spills: 5 168; the same with the loops unrolled
 * by 2 instead of 4: 4 620) */
RT_WAVE_ENTRY(pathTrace_exp5, false, 4, 5, 4, false)
RT_WAVE_ENTRY(pathTrace_exp12, false, 4, 12, 2, false)
#endif

#include "sorted.cuh"
/* rays sorted by the object groups they can reach (sorted.cuh); a block that finds the scene ungrouped
 * (rt_prep_scene: RT_GRP_OK clear) runs the compacting kernel with R = 4 in the same shared memory */
#define RT_SORTED_ENTRY(name, R, MINB)                                                                           \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,             \
         const RTHaltonTables ht) {                                                                              \
        extern __shared__ __align__(16) float sm[];                                                              \
        if (((const uint32_t *)scene)[lay.group + RT_GRP_OK])                                                    \
            pathtrace_sorted_body<R, false>(scene, offsets, out, p, lay, sm, RTTileSched{}, ht);                 \
        else                                                                                                     \
            pathtrace_wave_body<false, 4, 4, false, false>(scene, offsets, out, p, lay, sm, counters, RTTileSched{}, ht); \
    }
#define RT_SORTED_PERSIST_ENTRY(name, R, MINB)                                                                   \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,             \
         const RTHaltonTables ht, const RTTileSched ts) {                                                        \
        extern __shared__ __align__(16) float sm[];                                                              \
        if (((const uint32_t *)scene)[lay.group + RT_GRP_OK])                                                    \
            pathtrace_sorted_body<R, true>(scene, offsets, out, p, lay, sm, ts, ht);                             \
        else                                                                                                     \
            pathtrace_wave_body<false, 4, 4, true, false>(scene, offsets, out, p, lay, sm, counters, ts, ht);    \
    }
/* 15 words per slot: R = 6 fits four blocks per SM beside the Cornell scene (53 KB per block) */
RT_SORTED_ENTRY(pathTrace_sorted4, 4, 4)
RT_SORTED_ENTRY(pathTrace_sorted5, 5, 4)
RT_SORTED_ENTRY(pathTrace_sorted6, 6, 4)
RT_SORTED_ENTRY(pathTrace_sorted8, 8, 3)
/* ticketed (tile, sample chunk) units for launches with few blocks per SM (wave.cuh: RTTileSched) */
RT_SORTED_PERSIST_ENTRY(pathTrace_sorted_tiles6, 6, 4)

#include "sortedblk.cuh"
#ifndef RT_EXP_NO_FALLBACK
#define RT_EXP_NO_FALLBACK 0     /* A/B builds only: the block-sorted entry without the compacting kernel as its fallback */
#endif
/* the sort taken over the block's four warps (sortedblk.cuh) */
#define RT_SORTEDBLK_ENTRY(name, R, MINB)                                                                        \
    extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)                                         \
    name(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,        \
         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,             \
         const RTHaltonTables ht) {                                                                              \
        extern __shared__ __align__(16) float sm[];                                                              \
        if (RT_EXP_NO_FALLBACK || ((const uint32_t *)scene)[lay.group + RT_GRP_OK])                             \
            pathtrace_sortedblk_body<R, false>(scene, offsets, out, p, lay, sm, RTTileSched{}, ht);              \
        else if (!RT_EXP_NO_FALLBACK)                                                                            \
            pathtrace_wave_body<false, 4, 4, false, false>(scene, offsets, out, p, lay, sm, counters, RTTileSched{}, ht); \
    }
RT_SORTEDBLK_ENTRY(pathTrace_sortedblk6, 6, 4)
/* RT_VARIANT_MIS (mis.cuh) on the sorted path; an ungrouped scene falls to the compacting kernel's MIS form */
extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)
pathTrace_sortedblk6_mis(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,
                         const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,
                         const RTHaltonTables ht) {
    extern __shared__ __align__(16) float sm[];
    if (((const uint32_t *)scene)[lay.group + RT_GRP_OK])
        pathtrace_sortedblk_body<6, false, false, true>(scene, offsets, out, p, lay, sm, RTTileSched{}, ht);
    else
        pathtrace_wave_body<false, 4, 4, false, false, true>(scene, offsets, out, p, lay, sm, counters, RTTileSched{}, ht);
}
/* fewer samples per launch than a round has slots (progressive rendering): G = 6 / spp tiles side by side per warp */
extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)
pathTrace_sortedblk_wide6(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,
                          const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,
                          const RTHaltonTables ht) {
    extern __shared__ __align__(16) float sm[];
    if (((const uint32_t *)scene)[lay.group + RT_GRP_OK])
        pathtrace_sortedblk_body<6, false, true>(scene, offsets, out, p, lay, sm, RTTileSched{}, ht);
    else
        pathtrace_wave_body<false, 4, 6, false, true>(scene, offsets, out, p, lay, sm, counters, RTTileSched{}, ht);
}
#ifdef RT_TUNING_VARIANTS
RT_SORTEDBLK_ENTRY(pathTrace_exp14, 4, 5)      /* 512 slots, five blocks per SM */
RT_SORTEDBLK_ENTRY(pathTrace_exp15, 5, 4)
#endif
extern "C" __global__ void __launch_bounds__(RT_BLOCK_THREADS, 4)
pathTrace_sortedblk_tiles6(const float *__restrict__ scene, const uint32_t *__restrict__ offsets, float4 *__restrict__ out,
                           const RTLaunchParams p, const RTSceneLayout lay, unsigned long long *__restrict__ counters,
                           const RTHaltonTables ht, const RTTileSched ts) {
    extern __shared__ __align__(16) float sm[];
    if (((const uint32_t *)scene)[lay.group + RT_GRP_OK])
        pathtrace_sortedblk_body<6, true>(scene, offsets, out, p, lay, sm, ts, ht);
    else
        pathtrace_wave_body<false, 4, 4, true, false>(scene, offsets, out, p, lay, sm, counters, ts, ht);
}

#include "bvh.cuh"
#include "bvh_build.cuh"

#include "modeb.cuh"

/* --------------------------------------------------------- Halton tables ---- */
/* rt_device_math.cuh: RTHaltonTables.  bounce[b * n + i] = halton_slot<0..3>(i, b), camY[i] = halton<1>(i): the very
 * functions the kernels would otherwise call, so a looked-up value is the computed one bit for bit. */
extern "C" __global__ void rt_build_halton(float4 *__restrict__ bounce, float *__restrict__ camY, uint32_t n, uint32_t bounces) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    camY[i] = halton<1>(i);
    for (uint32_t b = 0; b < bounces; b++)
        bounce[(size_t)b * n + i] = make_float4(halton_slot<0>(i, b), halton_slot<1>(i, b), halton_slot<2>(i, b), halton_slot<3>(i, b));
}

/* ----------------------------------------------------------------- resolve ---- */
/* RTrace/image.swift:25-65: the [Float16] read-back of the render target, x2 exposure, Reinhard, gamma 1/2.2, clamp,
 * *255 -> u8.  powf here is CUDA's (<= 2 ulp); the result is compared with +-1 LSB. */
extern "C" __global__ void rt_resolve_rgba8(const float4 *__restrict__ in, uchar4 *__restrict__ out,
                                            unsigned long long n, int divideByAlpha) {
    unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = in[i];
    float c[3] = { v.x, v.y, v.z };
    unsigned char r[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float value = c[k];
        if (divideByAlpha) value = value / v.w;
        /* the reference tone-maps what it reads back from its rgba16Float target (image.swift:25-38): the value
         * rounded to half, widened again */
        value = __half2float(__float2half_rn(value));
        value *= 2.0f;
        value = value / (value + 1.0f);
        value = powf(value, 1.0f / 2.2f);
        value = fmaxf(0.0f, fminf(1.0f, value));
        r[k] = (unsigned char)(value * 255.0f);
    }
    out[i] = make_uchar4(r[0], r[1], r[2], 255);
}

/* The reference's render target is rgba16Float (RTrace/renderer.swift:74-82: the shader's write rounds fp32 to
 * fp16) and saveTextureToImage reads it back as [Float16] (image.swift:25-38).  This pass gives a host that keeps
 * that code the same bytes: round-to-nearest-even per channel, alpha as stored (1 after a mean; after
 * divideByAlpha the resolved progressive image, alpha 1). */
extern "C" __global__ void rt_resolve_rgba16f(const float4 *__restrict__ in, ushort4 *__restrict__ out,
                                              unsigned long long n, int divideByAlpha) {
    unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = in[i];
    if (divideByAlpha) { v.x = v.x / v.w; v.y = v.y / v.w; v.z = v.z / v.w; v.w = 1.0f; }
    out[i] = make_ushort4(__half_as_ushort(__float2half_rn(v.x)), __half_as_ushort(__float2half_rn(v.y)),
                          __half_as_ushort(__float2half_rn(v.z)), __half_as_ushort(__float2half_rn(v.w)));
}

/* -------------------------------------------------- FP32 peak microbenchmark ---- */
/* "an FFMA-saturating microbenchmark on the box": the measured FP32 peak bench.py reports beside the
 * nominal SMs x 128 x 2 x clock.  Eight independent chains per thread; mode 0 = scalar FFMA (fma.rn.f32), mode 1 = the
 * packed fma.rn.f32x2 of sm_100 (tools/fp32_peak.cu has the same loops plus the ALU mixes; 70.2 / 74.1 TFLOP/s in r01v). */
extern "C" __global__ void __launch_bounds__(256) rt_fp32_peak(const float *__restrict__ in, float *__restrict__ out, int iters, int mode) {
    float a[8];
    const float b = in[0], c = in[1];
    unsigned long long pk[8], pb, pc;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        a[k] = in[2 + k] + (float)threadIdx.x;
        asm("mov.b64 %0, {%1,%2};" : "=l"(pk[k]) : "f"(a[k]), "f"(a[k] + 1.0f));
    }
    asm("mov.b64 %0, {%1,%2};" : "=l"(pb) : "f"(b), "f"(b));
    asm("mov.b64 %0, {%1,%2};" : "=l"(pc) : "f"(c), "f"(c));
    if (mode == 0) {
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int k = 0; k < 8; k++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[k]) : "f"(b), "f"(c));
        }
    } else {
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int k = 0; k < 8; k++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pk[k]) : "l"(pb), "l"(pc));
        }
    }
    float sum = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        float lo, hi;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(pk[k]));
        sum += a[k] + lo + hi;
    }
    if (sum == 123.456f) out[0] = sum;
}

/* ------------------------------------------------------------------- debug ---- */
/* what: 0 halton(i,d) [in: u32 pairs]          -> float  (compile-time-base path)
 *       1 sincos(x)   [in: float]              -> float2
 *       2 hash(x)     [in: u32]                -> u32
 *       3 halton_rt(i,d)                       -> float  (generic loop)
 *       4 halton_slot<s>(i,b), in: (i, 4*b+s)  -> float  (the path tracer's table-driven form)
 *       5 one triangle test; 6 rt_prep_scene's object groups of a scene (see below) */
template <int D>
__device__ float halton_dispatch(uint32_t i, uint32_t d) {
    if (d == D) return halton<D>(i);
    if constexpr (D < 23) return halton_dispatch<D + 1>(i, d); else return 0.0f;
}

extern "C" __global__ void rt_debug_eval(int what, const void *__restrict__ in, void *__restrict__ out, uint32_t n) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (what == 6) {
        /* the object groups rt_prep_scene makes of a scene: in = T, L, paired, - (uint32), then 48*T bytes of
         * vertices, then 64*L bytes of lights; out = the RT_GROUP_WORDS words of the "group" section */
        __shared__ uint32_t s_adj[32];
        if (blockIdx.x) return;
        const uint32_t *h = (const uint32_t *)in;
        RTSceneLayout lay = rtSceneLayout(h[0], 0u, h[1]);
        lay.group = 0u;
        const unsigned char *v = (const unsigned char *)in + 16;
        if (k < 32) s_adj[k] = 0u;
        __syncthreads();
        if (h[2] != 0u && h[0] >= 2u && h[0] <= 64u) prep_adjacency(v, h[0] / 2u, k & 31u, k >> 5, blockDim.x >> 5, s_adj);
        __syncthreads();
        if (k < 32) prep_groups(v, v + 48 * (size_t)h[0], (float *)out, lay, h[2] != 0u, k, s_adj);
        return;
    }
    if (k >= n) return;
    if (what == 0) {
        const uint32_t *p = (const uint32_t *)in;
        ((float *)out)[k] = halton_dispatch<0>(p[2 * k], p[2 * k + 1]);
    } else if (what == 1) {
        float s, c;
        sincos_det(((const float *)in)[k], s, c);
        ((float *)out)[2 * k] = s; ((float *)out)[2 * k + 1] = c;
    } else if (what == 2) {
        ((uint32_t *)out)[k] = hash_u32(((const uint32_t *)in)[k]);
    } else if (what == 3) {
        const uint32_t *p = (const uint32_t *)in;
        ((float *)out)[k] = halton_rt(p[2 * k], p[2 * k + 1]);
    } else if (what == 4) {
        const uint32_t *p = (const uint32_t *)in;
        const uint32_t i = p[2 * k], b = p[2 * k + 1] >> 2, sl = p[2 * k + 1] & 3u;
        float r = sl == 0 ? halton_slot<0>(i, b) : sl == 1 ? halton_slot<1>(i, b) : sl == 2 ? halton_slot<2>(i, b) : halton_slot<3>(i, b);
        ((float *)out)[k] = r;
    } else if (what == 5) {
        /* one triangle test, the isect record built as rt_prep_scene builds it: in = v0 v1 v2 o d tmin tmax
         * (17 floats), out = hit, distance, bestD (3 words).  For the exact-zero cases of the sign rule. */
        const float *p = (const float *)in + 17 * k;
        const f3 v0 = mk(p[0], p[1], p[2]), v1 = mk(p[3], p[4], p[5]), v2 = mk(p[6], p[7], p[8]);
        const f3 o = mk(p[9], p[10], p[11]), d = mk(p[12], p[13], p[14]);
        const f3 e1 = v1 - v0, e2 = v2 - v0, ni = cross(e2, e1), e1n = v0 - v1;
        (void)e1;
        const float4 A = make_float4(v0.x, v0.y, v0.z, ni.x), B = make_float4(ni.y, ni.z, e2.x, e2.y), C = make_float4(e2.z, e1n.x, e1n.y, e1n.z);
        float N = 0.0f, a = 1.0f;
        const bool hit = tri_test(A, B, C, o, d, p[15], p[16], 1.0f, N, a);
        ((uint32_t *)out)[3 * k] = hit ? 1u : 0u;
        ((float *)out)[3 * k + 1] = hit ? N / a : 0.0f;
        ((float *)out)[3 * k + 2] = hit ? a : 0.0f;
    }
}
#endif /* RT_NOALLOC_BUILD */
