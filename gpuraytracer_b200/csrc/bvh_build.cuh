/*
 * bvh_build.cuh -- the acceleration structure of bvh.cuh built ON THE DEVICE (included by pathtrace.cu).
 *
 * Reference: setupAccelerationStructures (RTrace/computeShader.swift:45-97) hands the triangles to Metal's builder,
 * which builds on the GPU at init (encoder.build, :87-90).  Here, on the stream the caller names, with no host
 * round trip (round 1 copied the vertices back and built on one CPU thread inside rtCommit):
 *
 *   rt_bvh_bounds   the scene's bounds (ordered-integer atomicMin / atomicMax)
 *   rt_bvh_morton   key = 30-bit Morton code of the triangle's box centre << 32 | triangle id (unique keys), padded to a
 *                   power of two with all-ones
 *   rt_bvh_sort_*   bitonic sort of the 64-bit keys: steps with a partner less than 2048 apart run in shared memory
 *                   (2048 keys per block), the others one launch each -- 21 launches for 36 864 triangles
 *   rt_bvh_karras   the binary radix tree over the sorted keys (Karras 2012): node i's range and split from the length
 *                   of common key prefixes; internal nodes 0 .. T-2, primitive p is node T-1+p
 *   rt_bvh_fit      bottom-up boxes: every primitive walks to the root, the second visitor of a node unites its
 *                   children's boxes (and heights: the root's height bounds the traversal stack)
 *   rt_bvh_emit     the nodes bvh.cuh walks (host_bvh.h: RTBvhNode, both children's boxes grown by 1/1024 of the scene's
 *                   extent): a child whose range holds <= RT_BVH_LEAF triangles becomes a leaf reference (first slot,
 *                   count), so the radix tree below it is never visited; order[slot] = triangle id of leaf slot `slot`
 *
 * The image does not depend on the tree: any tree whose boxes contain their triangles gives the same hits
 * (tests/test_bvh_cpu.py states that property and checks it on the host builder; tests/test_gpu_bvh_build.py runs the
 * same checker over the trees this file builds).
 */
#pragma once

#define RT_BVH_SORT_TILE 2048u
#define RT_BVH_SORT_THREADS 1024u
#define RT_BVH_MAX_LEAF 12u                 /* = RT_BVH_LEAF of host_bvh.h */
#define RT_BVH_MAX_HEIGHT 46u               /* radix-tree height beyond which the traversal stack (RT_BVH_STACK 48) may overflow */
#define RT_STATUS_BVH_TOO_DEEP 2u           /* bit of the status word (wave.cuh: RT_STATUS_CHUNK_WAIT_TIMEOUT is 1) */

__device__ __forceinline__ uint32_t bvh_ord(float f) {          /* order-preserving map float -> uint */
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float bvh_unord(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

/* bounds[0..2] = min xyz, bounds[3..5] = max xyz, as ordered uints; the host sets them to 0xff.. / 0 first */
extern "C" __global__ void rt_bvh_bounds(const float4 *__restrict__ vertices, uint32_t T, uint32_t *__restrict__ bounds) {
    const float INF = __int_as_float(0x7f800000);
    float lo[3] = { INF, INF, INF }, hi[3] = { -INF, -INF, -INF };
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < 3u * T; i += gridDim.x * blockDim.x) {
        const float4 v = vertices[i];
        lo[0] = fminf(lo[0], v.x); lo[1] = fminf(lo[1], v.y); lo[2] = fminf(lo[2], v.z);
        hi[0] = fmaxf(hi[0], v.x); hi[1] = fmaxf(hi[1], v.y); hi[2] = fmaxf(hi[2], v.z);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int s = 16; s; s >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], s));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], s));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) { atomicMin(&bounds[k], bvh_ord(lo[k])); atomicMax(&bounds[3 + k], bvh_ord(hi[k])); }
    }
}

__device__ __forceinline__ uint32_t bvh_spread10(uint32_t v) {   /* 10 bits -> every third bit */
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

extern "C" __global__ void rt_bvh_morton(const float4 *__restrict__ vertices, uint32_t T, uint32_t n2,
                                         const uint32_t *__restrict__ bounds, unsigned long long *__restrict__ keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n2) return;
    if (i >= T) { keys[i] = ~0ull; return; }
    const float4 a = vertices[3 * (size_t)i], b = vertices[3 * (size_t)i + 1], c = vertices[3 * (size_t)i + 2];
    const float cx = 0.5f * fminf(fminf(a.x, b.x), c.x) + 0.5f * fmaxf(fmaxf(a.x, b.x), c.x);
    const float cy = 0.5f * fminf(fminf(a.y, b.y), c.y) + 0.5f * fmaxf(fmaxf(a.y, b.y), c.y);
    const float cz = 0.5f * fminf(fminf(a.z, b.z), c.z) + 0.5f * fmaxf(fmaxf(a.z, b.z), c.z);
    const float cen[3] = { cx, cy, cz };
    uint32_t q[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float lo = bvh_unord(bounds[k]), ext = bvh_unord(bounds[3 + k]) - lo;
        const float t = ext > 0.0f ? (cen[k] - lo) / ext : 0.0f;                  /* (a NaN vertex lands in cell 0) */
        q[k] = (uint32_t)fminf(fmaxf(t * 1024.0f, 0.0f), 1023.0f);
    }
    const uint32_t code = (bvh_spread10(q[0]) << 2) | (bvh_spread10(q[1]) << 1) | bvh_spread10(q[2]);
    keys[i] = ((unsigned long long)code << 32) | i;
}

/* one bitonic step (k, j) with j >= RT_BVH_SORT_TILE: partners live in different tiles */
extern "C" __global__ void rt_bvh_sort_global(unsigned long long *__restrict__ keys, uint32_t n2, uint32_t k, uint32_t j) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n2 / 2u) return;
    const uint32_t i = 2u * j * (t / j) + (t % j), p = i + j;
    const unsigned long long a = keys[i], b = keys[p];
    if ((a > b) == ((i & k) == 0u)) { keys[i] = b; keys[p] = a; }
}
/* all steps of the stages kFrom .. kTo (powers of two) whose partner distance is below the tile: j = min(k / 2, TILE / 2) .. 1.
 * kFrom = 2, kTo = TILE sorts every tile outright; for a later stage k the global steps run first, then this with kFrom = kTo = k. */
extern "C" __global__ void __launch_bounds__(RT_BVH_SORT_THREADS)
rt_bvh_sort_shared(unsigned long long *__restrict__ keys, uint32_t n2, uint32_t kFrom, uint32_t kTo) {
    __shared__ unsigned long long s[RT_BVH_SORT_TILE];
    const uint32_t base = blockIdx.x * RT_BVH_SORT_TILE, t = threadIdx.x;
    const uint32_t tile = min(RT_BVH_SORT_TILE, n2);                              /* n2 is a power of two: tile divides it */
    if (base >= n2) return;
    for (uint32_t e = t; e < tile; e += RT_BVH_SORT_THREADS) s[e] = keys[base + e];
    __syncthreads();
    for (uint32_t k = kFrom; k <= kTo; k <<= 1) {
        for (uint32_t j = min(k >> 1, tile >> 1); j >= 1u; j >>= 1) {
            if (t < tile / 2u) {
                const uint32_t i = 2u * j * (t / j) + (t % j), p = i + j;
                const unsigned long long a = s[i], b = s[p];
                if ((a > b) == (((base + i) & k) == 0u)) { s[i] = b; s[p] = a; }
            }
            __syncthreads();
        }
    }
    for (uint32_t e = t; e < tile; e += RT_BVH_SORT_THREADS) keys[base + e] = s[e];
}

/* the radix tree's arrays, T triangles: internal nodes 0 .. T-2, primitive p = node T-1+p */
struct RTBvhBuild {
    unsigned long long *keys;      /* n2 sorted keys (Morton << 32 | triangle id) */
    uint32_t *left, *right;        /* [T-1] children (node numbers) */
    uint32_t *parent;              /* [2T-1] */
    uint32_t *first, *last;        /* [T-1] range of sorted positions under the node */
    float4 *boxLo, *boxHi;         /* [2T-1]; boxHi.w = the subtree's height as float */
    uint32_t *visits;              /* [T-1], zeroed by the host */
    uint32_t *bounds;              /* 6 ordered uints */
};

__device__ __forceinline__ int bvh_delta(const unsigned long long *keys, int T, int i, int j) {
    if (j < 0 || j >= T) return -1;
    return __clzll((long long)(keys[i] ^ keys[j]));                               /* keys are unique: never 64 */
}

extern "C" __global__ void rt_bvh_karras(const RTBvhBuild B, uint32_t T) {
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x), n = (int)T;
    if (i >= n - 1) return;
    const unsigned long long *keys = B.keys;
    const int d = bvh_delta(keys, n, i, i + 1) - bvh_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = bvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (bvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1) if (bvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = bvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1; ; t = (t + 1) >> 1) {
        if (bvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t lc = lo == gamma ? (uint32_t)(n - 1 + gamma) : (uint32_t)gamma;
    const uint32_t rc = hi == gamma + 1 ? (uint32_t)(n - 1 + gamma + 1) : (uint32_t)(gamma + 1);
    B.left[i] = lc; B.right[i] = rc; B.first[i] = (uint32_t)lo; B.last[i] = (uint32_t)hi;
    B.parent[lc] = (uint32_t)i; B.parent[rc] = (uint32_t)i;
    if (i == 0) B.parent[0] = 0xffffffffu;
}

extern "C" __global__ void rt_bvh_fit(const RTBvhBuild B, const float4 *__restrict__ vertices, uint32_t T, unsigned int *status) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= T) return;
    const uint32_t tri = (uint32_t)B.keys[p];
    const float4 a = vertices[3 * (size_t)tri], b = vertices[3 * (size_t)tri + 1], c = vertices[3 * (size_t)tri + 2];
    float4 lo = make_float4(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z), 0.0f);
    float4 hi = make_float4(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z), 0.0f);
    uint32_t node = T - 1u + p;
    B.boxLo[node] = lo; B.boxHi[node] = hi;
    if (T < 2u) return;
    for (;;) {
        __threadfence();
        const uint32_t up = B.parent[node];
        if (up == 0xffffffffu) {                                                  /* `node` is the root: its height bounds the stack */
            if (hi.w > (float)RT_BVH_MAX_HEIGHT && status) atomicOr_system(status, RT_STATUS_BVH_TOO_DEEP);
            return;
        }
        if (atomicAdd(&B.visits[up], 1u) == 0u) return;                            /* the sibling's subtree is not done yet */
        __threadfence();
        const uint32_t l = B.left[up], r = B.right[up];
        const float4 l0 = __ldcg(&B.boxLo[l]), l1 = __ldcg(&B.boxHi[l]), r0 = __ldcg(&B.boxLo[r]), r1 = __ldcg(&B.boxHi[r]);
        lo = make_float4(fminf(l0.x, r0.x), fminf(l0.y, r0.y), fminf(l0.z, r0.z), 0.0f);
        hi = make_float4(fmaxf(l1.x, r1.x), fmaxf(l1.y, r1.y), fmaxf(l1.z, r1.z), fmaxf(l1.w, r1.w) + 1.0f);
        __stcg(&B.boxLo[up], lo); __stcg(&B.boxHi[up], hi);
        node = up;
    }
}

/* nodes: 4 float4 per node (host_bvh.h: RTBvhNode).  One thread per internal node; thread p < T also writes order[p]. */
extern "C" __global__ void rt_bvh_emit(const RTBvhBuild B, uint32_t T, float4 *__restrict__ nodes, uint32_t *__restrict__ order) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < T) order[i] = (uint32_t)B.keys[i];
    float ext = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; k++) ext = fmaxf(ext, bvh_unord(B.bounds[3 + k]) - bvh_unord(B.bounds[k]));
    const float margin = ext * (1.0f / 1024.0f);
    const float INF = __int_as_float(0x7f800000);
    if (T <= RT_BVH_MAX_LEAF) {
        /* one leaf: a root whose second child is an empty box (as host_bvh.cpp) */
        if (i == 0) {
            const float lo[3] = { bvh_unord(B.bounds[0]), bvh_unord(B.bounds[1]), bvh_unord(B.bounds[2]) };
            const float hi[3] = { bvh_unord(B.bounds[3]), bvh_unord(B.bounds[4]), bvh_unord(B.bounds[5]) };
            nodes[0] = make_float4(lo[0] - margin, lo[1] - margin, lo[2] - margin, __uint_as_float(RT_BVH_LEAF_FLAG | ((T - 1u) << 27)));
            nodes[1] = make_float4(hi[0] + margin, hi[1] + margin, hi[2] + margin, __uint_as_float(RT_BVH_LEAF_FLAG));
            nodes[2] = make_float4(INF, INF, INF, 0.0f);
            nodes[3] = make_float4(-INF, -INF, -INF, 0.0f);
        }
        return;
    }
    if (i >= T - 1u) return;
    uint32_t ref[2];
    float4 lo[2], hi[2];
    const uint32_t child[2] = { B.left[i], B.right[i] };
#pragma unroll
    for (int c = 0; c < 2; c++) {
        const uint32_t ch = child[c];
        lo[c] = B.boxLo[ch]; hi[c] = B.boxHi[ch];
        if (ch >= T - 1u) ref[c] = RT_BVH_LEAF_FLAG | (ch - (T - 1u));                                   /* one triangle */
        else {
            const uint32_t f = B.first[ch], n = B.last[ch] - f + 1u;
            ref[c] = n <= RT_BVH_MAX_LEAF ? (RT_BVH_LEAF_FLAG | ((n - 1u) << 27) | f) : ch;
        }
    }
    nodes[4 * (size_t)i]     = make_float4(lo[0].x - margin, lo[0].y - margin, lo[0].z - margin, __uint_as_float(ref[0]));
    nodes[4 * (size_t)i + 1] = make_float4(hi[0].x + margin, hi[0].y + margin, hi[0].z + margin, __uint_as_float(ref[1]));
    nodes[4 * (size_t)i + 2] = make_float4(lo[1].x - margin, lo[1].y - margin, lo[1].z - margin, 0.0f);
    nodes[4 * (size_t)i + 3] = make_float4(hi[1].x + margin, hi[1].y + margin, hi[1].z + margin, 0.0f);
}
