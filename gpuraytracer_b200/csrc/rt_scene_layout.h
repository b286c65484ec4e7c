/*
 * rt_scene_layout.h -- layout of the PREPARED scene: what the prep kernel writes
 * into HBM once per command and what every pathTrace block stages into shared
 * memory (and, for the constant-bank variant, what is copied into c_isect).
 * Shared by the kernels (pathtrace.cu) and the host shim (rtcuda.cpp).
 *
 * All sections are arrays of 32-bit words inside one flat buffer:
 *
 *   [0, RT_HDR_WORDS)               header: camera frame + counts   (floats / uints)
 *   --- intersection part (mirrored into the constant bank by the "const" variant) ---
 *   isect  : 16 words per triangle  v0.xyz n.x | n.yz e2.xy | e2.z e1n.xyz | w.xyz Nn
 *            e1n = v0 - v1, e2 = v2 - v0, n = cross(v1 - v0, e2)  (not normalised);
 *            w = v0 - camera.position, Nn = dot(w, n): the ray-independent part of
 *            the test for CAMERA rays, which read slots 1..3 (v0 is inside w);
 *            read with the SAME index by all lanes -> broadcast 128-bit loads
 *   sphere : 8 words per sphere     center.xyz radius | m.xyz C
 *            m = camera.position - center, C = dot(m,m) - r*r (camera rays)
 *   --- shading part ---
 *   shade  : RT_SHADE_ROWS rows of Tp words (structure of arrays, Tp = T rounded
 *            up to 4): looked up with the per-lane primitive id after a hit,
 *            so SoA keeps different ids on different banks
 *   sshade : RT_SSHADE_ROWS rows of Sp words
 *   light  : RT_LIGHT_WORDS per light (see the define below)
 *   group  : RT_GROUP_WORDS words: object groups for the sorted kernel (sorted.cuh), built by rt_prep_scene
 *            when every triangle 2k+1 shares v0 with 2k (pairs), T <= 64 and S == 0:
 *            [0] K = number of cullable groups (0..4)   [1] pair mask that is always tested
 *            [2] 1 when the sorted kernel applies       [3] -
 *            then K records of 8 words: box centre xyz, pair mask (uint) | inflated half extent xyz, -
 *            A group is a connected set of quads (connected = some vertex bit-equal), its box the bounds of
 *            its vertices grown by 1/128 of the scene's largest extent.
 *            [3] the SHELL: those always-tested quads that lie in a face plane of the always-tested quads' own bounds
 *            (all six vertices bit-equal to the bound on one axis: the walls of a room).  After the K records, 8 words:
 *            centre xyz of those bounds, the groups' margin (1/128 of the extent) | half extent xyz shrunk by 1/16384 of the
 *            scene's largest extent, that shrink.  Then the shell quads face by face (RT_GRP_FACES), for the EXIT RULE: a
 *            closest-hit ray that starts inside the shrunk box can be accepted by a shell quad only where it crosses that
 *            quad's face plane ahead of it, and only if the crossing point lies within the face grown by the margin; so
 *            the ray is tested against the shell quads of the faces it can leave the box through -- one, as a rule --
 *            instead of all of them (sortedblk.cuh: part of the sort key).
 *            RT_GRP_LIGHT_PLANE: with one light (or several whose centres share one y), normal (0,-1,0), the quads whose six vertices have the light centre's y,
 *            bit for bit -- the light's own geometry.  Every light sample point has that y too, so a shadow ray crosses
 *            the plane AT its sample point, at t = ldist up to 1e-6 x ldist, and the shadow test ends at ldist - 1e-3
 *            (raytrace.metal:81): such a quad cannot be accepted by any shadow ray of the scene and is left out of their
 *            lists (sortedblk.cuh, sorted.cuh).  Only scenes smaller than 64 units across get the mask.
 *            RT_GRP_LIGHTS_IN: every light's sample rectangle (taken with the larger of its two possible half extents,
 *            0.25 and half its width / depth; lights with a normal of their own: never) lies inside the shrunk shell box.
 *            Then the shell rule needs the shading point alone: the kernels for several lights (wave.cuh) skip the shell
 *            triangles for all shadow rays of a pass when every lane's point is inside.  The shell, its box and this
 *            word are built for scenes with spheres too (K, the boxes and [2] are not used there: [2] stays 0).
 *            A shadow ray whose two ends both lie inside the shrunk box cannot be accepted by a shell quad (the
 *            segment stays 3e-4 x extent clear of its plane; the test's own error is 1e-6 x extent), so those
 *            any-hit tests are skipped (sortedblk.cuh / sorted.cuh, pass B).
 */
#ifndef RT_SCENE_LAYOUT_H
#define RT_SCENE_LAYOUT_H

#include <stdint.h>

#define RT_HDR_WORDS     32
/* header word indices */
#define RT_HDR_CAM_O      0   /* 3: camera.position                                  */
#define RT_HDR_CAM_U      3   /* 3: u = normalize(cross(up, w))                      */
#define RT_HDR_CAM_V      6   /* 3: v = normalize(cross(w, u))                       */
#define RT_HDR_CAM_W      9   /* 3: w = -normalize(direction)                        */
#define RT_HDR_HALF_W    12   /* tan(fov/2)                                          */
#define RT_HDR_HALF_H    13   /* halfWidth / float(res.x / res.y)                    */
#define RT_HDR_RES_X     14   /* float(camera.resolution.x)                          */
#define RT_HDR_RES_Y     15   /* float(camera.resolution.y)                          */
#define RT_HDR_EXPOSURE  16   /* mode B: 1 / (1.2 * 2^ev100)                         */
#define RT_HDR_PAIRED    18   /* uint: every triangle 2k+1 shares v0 with triangle 2k (rt_prep_scene)   */
#define RT_HDR_TILE_COUNTER 17 /* uint, zeroed by rt_prep_scene: next 8x4 tile of pathTrace_tiles */

#define RT_ISECT_WORDS   16
#define RT_SPHERE_WORDS   8

/* shade rows (per triangle) */
#define RT_SH_NX          0   /* geometric normal, getTriangleNormal                 */
#define RT_SH_RX          3   /* right   = normalize(cross(n, (0.0072,1,0.0034)))    */
#define RT_SH_FX          6   /* forward = cross(right, n)                           */
#define RT_SH_DR          9   /* diffuse.rgb                                         */
#define RT_SH_ER         12   /* emissive.rgb                                        */
#define RT_SH_FLAG       15   /* 1.0f when length(emissive) > 0                      */
#define RT_SH_METAL      16   /* mode B */
#define RT_SH_ROUGH      17   /* mode B */
#define RT_SH_TX         18   /* mode B: buildOrthonormalBasis(normal).tangent       */
#define RT_SH_BX         21   /* mode B: buildOrthonormalBasis(normal).bitangent     */
#define RT_SHADE_ROWS    24

#define RT_SSH_DR         0
#define RT_SSH_ER         3
#define RT_SSH_FLAG       6
#define RT_SSHADE_ROWS    7

#define RT_LIGHT_WORDS   28   /* center.xyz hx | color.xyz hz | emittedRadiance.xyz general | width depth 0 0 | normal.xyz 0 |
                                 tangent.xyz 0 | bitangent.xyz 0.   (hx,hz) = half extent of the rectangle sampleAreaLight samples:
                                 0.25, 0.25 (sampling.metal:207-209) or width/2, depth/2 (RT_LAYOUT_LIGHT_EXTENT).
                                 normal / tangent / bitangent: (0,-1,0), (1,0,0), (0,0,1) -- the reference's literals -- and
                                 general = 0, unless a normal table is bound (RT_LAYOUT_LIGHT_NORMALS): then the table's
                                 normal, normalised, with buildOrthonormalBasis of it (sampling.metal:158-171), general = 1 */
#define RT_LIGHT_GENERAL 11
#define RT_LIGHT_NORMAL  16
#define RT_LIGHT_TANGENT 20
#define RT_LIGHT_BITANGENT 24
#define RT_LAYOUT_LIGHT_EXTENT 1u
#define RT_LAYOUT_LIGHT_NORMALS 2u   /* a float4-per-light normal table is bound at RT_INDEX_LIGHT_NORMALS */
#define RT_LAYOUT_NO_SGRID 4u        /* the prepared scene has no room for the sphere grid (sgrid.cuh): brute force (rtPathTraceNoAlloc) */
#define RT_MAX_GROUPS     4   /* cullable object groups (16 sort keys)                                 */
#define RT_GRP_K          0
#define RT_GRP_ALWAYS     1
#define RT_GRP_OK         2
#define RT_GRP_SHELL      3
#define RT_GRP_SHELL_BOX (4 + 8 * RT_MAX_GROUPS)
#define RT_GRP_FACES     (RT_GRP_SHELL_BOX + 8)     /* 6 pair masks: the shell quads in face 2 * axis + (1 = high side) */
#define RT_GRP_LIGHT_PLANE (RT_GRP_FACES + 6)      /* pair mask: quads in the plane of the light(s) */
#define RT_GRP_LIGHTS_IN   (RT_GRP_FACES + 7)      /* 1: the sample rectangle of EVERY light lies inside the shrunk shell box */
#define RT_GROUP_WORDS   (RT_GRP_FACES + 8)

/* constant-bank mirror of the intersection part: float4 slots (56 KB of the 64 KB bank) */
#define RT_CONST_SLOTS   3584
#define RT_MAX_TRIANGLES 1024            /* brute-force kernels (scene in shared memory) */
#define RT_MAX_TRIANGLES_BVH (1u << 22)  /* bvh.cuh (scene in global memory) */
#define RT_MAX_SPHERES   1024
#define RT_MAX_LIGHTS    16

typedef struct RTSceneLayout {
    uint32_t T, S, L;
    uint32_t Tp, Sp;          /* padded to a multiple of 4 (the unrolled loops' step) */
    uint32_t isect, sphere;                         /* intersection part, section starts in words */
    uint32_t shade, sshade, light;                  /* shading part                               */
    uint32_t words;           /* total size                                          */
    uint32_t group;           /* object groups (sorted.cuh), RT_GROUP_WORDS words    */
    uint32_t flags;           /* RT_LAYOUT_*: how rt_prep_scene fills the sections   */
    uint32_t sgrid;           /* sphere grid (scenes with spheres), RT_SGRID_WORDS words, in GLOBAL memory only: it starts at
                                 `words` and is not staged into shared memory                                          */
    uint32_t wordsGlobal;     /* size of the prepared scene in global memory: words (+ RT_SGRID_WORDS when S > 0)      */
} RTSceneLayout;

/* Sphere grid (sgrid.cuh): RT_SGRID_N^3 cells over the bounds of the spheres, each cell a list of up to RT_SGRID_CELL sphere
 * numbers (16 bits each, 0xffff = empty) -- every sphere whose bounding cube, grown by 1/64 of a cell, overlaps the cell.
 *   [0..2] lo.xyz of the grid   [3] ok (uint: 1 = usable; 0 = a cell overflowed or the bounds are degenerate: brute force)
 *   [4..6] cell size xyz        [7] -
 *   [8..10] 1 / cell size xyz   [11] -
 *   [16 ...] RT_SGRID_N^3 cells of RT_SGRID_CELL / 2 words */
#define RT_SGRID_N      8
#define RT_SGRID_CELL   8
#define RT_SGRID_HDR   16
#define RT_SGRID_WORDS (RT_SGRID_HDR + RT_SGRID_N * RT_SGRID_N * RT_SGRID_N * (RT_SGRID_CELL / 2))

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ inline
#else
#define RT_HD static inline
#endif

RT_HD RTSceneLayout rtSceneLayout(uint32_t T, uint32_t S, uint32_t L) {
    RTSceneLayout l;
    l.T = T; l.S = S; l.L = L;
    l.Tp = (T + 3u) & ~3u;
    l.Sp = (S + 3u) & ~3u;
    l.isect   = RT_HDR_WORDS;
    l.sphere  = l.isect + RT_ISECT_WORDS * l.Tp;
    l.shade   = l.sphere + RT_SPHERE_WORDS * l.Sp;
    l.sshade  = l.shade + RT_SHADE_ROWS * l.Tp;
    l.light   = l.sshade + RT_SSHADE_ROWS * l.Sp;
    l.group   = l.light + RT_LIGHT_WORDS * L;
    l.words  = l.group + RT_GROUP_WORDS;
    l.words  = (l.words + 3u) & ~3u;
    l.flags  = 0u;
    l.sgrid  = l.words;
    l.wordsGlobal = l.words + (S ? RT_SGRID_WORDS : 0u);
    return l;
}

#endif
