/*
 * rt_fixtures.c -- TEST INFRASTRUCTURE.  The reference's input data restated in C.
 *
 *   rt_fixture_cornell   RTrace/scene.swift:14-240 (initCornellBox,
 *                        createCornellBoxScene, createRotatedBoxVertices,
 *                        createBoxTriangles), converted exactly as
 *                        RTrace/computeShader.swift:13-41,242-253 and
 *                        RTrace/renderer.swift:49-57 do (materials[i] and
 *                        vertices[3i..3i+2] belong to triangle i).
 *   rt_fixture_offsets   seeded replacement for RTrace/renderer.swift:96-101.
 *   rt_fixture_spheres   config 4 of BASELINE.json (not in the reference).
 *
 * Swift evaluates these literals in Float; the same operation order is kept
 * here.  cos/sin of the box rotations come from rt_sincosf so that the bytes do
 * not depend on the host libm.
 */
#include "rt_oracle.h"
#include "rt_detmath.h"

#define RT_PI_F 3.14159265358979323846264338327950288f

static simd_float3 v3(float x, float y, float z) { simd_float3 r; r.x = x; r.y = y; r.z = z; r._pad = 0.0f; return r; }

static MaterialGPU material(float r, float g, float b, float a, float metallic, float roughness,
                            float ex, float ey, float ez) {
    MaterialGPU m;
    memset(&m, 0, sizeof m);
    m.diffuse.x = r; m.diffuse.y = g; m.diffuse.z = b; m.diffuse.w = a;
    m.metallic = metallic; m.roughness = roughness;
    m.emissive = v3(ex, ey, ez);
    return m;
}

typedef struct { simd_float3 *v; MaterialGPU *m; int n; } tri_sink;

static void push(tri_sink *s, simd_float3 a, simd_float3 b, simd_float3 c, MaterialGPU m) {
    s->v[s->n * 3 + 0] = a; s->v[s->n * 3 + 1] = b; s->v[s->n * 3 + 2] = c;
    s->m[s->n] = m;
    s->n++;
}

/* scene.swift:177-210 */
static void rotated_box(simd_float3 center, float width, float height, float depth, float rotationY,
                        simd_float3 out[8]) {
    float halfW = width / 2, halfH = height / 2, halfD = depth / 2;
    const float base[8][3] = {
        { -halfW, -halfH, -halfD }, {  halfW, -halfH, -halfD }, {  halfW,  halfH, -halfD }, { -halfW,  halfH, -halfD },
        { -halfW, -halfH,  halfD }, {  halfW, -halfH,  halfD }, {  halfW,  halfH,  halfD }, { -halfW,  halfH,  halfD },
    };
    float sinY, cosY;
    rt_sincosf(rotationY, &sinY, &cosY);
    /* simd_float4x4(c0,c1,c2,c3) takes COLUMNS (scene.swift:197-202):
     * c0 = (cos,0,sin,0), c2 = (-sin,0,cos,0)  =>  x' = cos*x - sin*z, z' = sin*x + cos*z */
    for (int i = 0; i < 8; i++) {
        float x = base[i][0], y = base[i][1], z = base[i][2];
        float rx = cosY * x - sinY * z;
        float rz = sinY * x + cosY * z;
        out[i] = v3(rx + center.x, y + center.y, rz + center.z);
    }
}

/* scene.swift:212-240 */
static void box_triangles(tri_sink *s, const simd_float3 v[8], MaterialGPU m) {
    push(s, v[0], v[2], v[1], m); push(s, v[0], v[3], v[2], m);   /* back   */
    push(s, v[4], v[5], v[6], m); push(s, v[4], v[6], v[7], m);   /* front  */
    push(s, v[0], v[4], v[7], m); push(s, v[0], v[7], v[3], m);   /* left   */
    push(s, v[1], v[6], v[5], m); push(s, v[1], v[2], v[6], m);   /* right  */
    push(s, v[0], v[5], v[4], m); push(s, v[0], v[1], v[5], m);   /* bottom */
    push(s, v[3], v[6], v[2], m); push(s, v[3], v[7], v[6], m);   /* top    */
}

/* scene.swift:80-138: back, left (red), right (green), floor, ceiling */
static void room_walls(tri_sink *s, float half) {
    MaterialGPU red   = material(0.9f, 0.0f, 0.0f, 1, 0.05f, 0.3f, 0, 0, 0);
    MaterialGPU green = material(0.0f, 0.7f, 0.0f, 1, 0.05f, 0.8f, 0, 0, 0);
    MaterialGPU white = material(0.9f, 0.9f, 0.9f, 1, 0.05f, 0.8f, 0, 0, 0);
    float h = half;
    push(s, v3(-h, -h, -h), v3( h,  h, -h), v3(-h,  h, -h), white);
    push(s, v3(-h, -h, -h), v3( h, -h, -h), v3( h,  h, -h), white);
    push(s, v3(-h, -h, -h), v3(-h,  h,  h), v3(-h, -h,  h), red);
    push(s, v3(-h, -h, -h), v3(-h,  h, -h), v3(-h,  h,  h), red);
    push(s, v3( h, -h, -h), v3( h,  h,  h), v3( h,  h, -h), green);
    push(s, v3( h, -h, -h), v3( h, -h,  h), v3( h,  h,  h), green);
    push(s, v3(-h, -h, -h), v3( h, -h,  h), v3( h, -h, -h), white);
    push(s, v3(-h, -h, -h), v3(-h, -h,  h), v3( h, -h,  h), white);
    push(s, v3(-h,  h, -h), v3( h,  h,  h), v3(-h,  h,  h), white);
    push(s, v3(-h,  h, -h), v3( h,  h, -h), v3( h,  h,  h), white);
}

static void fill_camera(CameraGPU *camera, int width, int height) {
    /* scene.swift:16-18, 290-297; direction = normalize((0,0,-2.5)-(0,0,9)) = (0,0,-1) exactly */
    memset(camera, 0, sizeof *camera);
    camera->position = v3(0, 0, 9);
    camera->direction = v3(0, 0, -1);
    camera->up = v3(0, 1, 0);
    camera->resolution.x = width;
    camera->resolution.y = height;
    camera->horizontalFov = RT_PI_F / 4.0f;
    camera->ev100 = 5.0f;
}

/* scene.swift:23-53, 257-270 + computeShader.swift:33-41 */
static void fill_light(SquareLightGPU *light, simd_float3 center, MaterialGPU lm, float w, float d) {
    memset(light, 0, sizeof *light);
    light->center = center;
    light->color = lm.diffuse;
    float luminousFlux = 100.0f * 12.0f;
    float area = w * d;
    float exitance = luminousFlux / area;
    float luminance = exitance / RT_PI_F;
    light->emittedRadiance = v3(lm.diffuse.x * luminance, lm.diffuse.y * luminance, lm.diffuse.z * luminance);
    light->width = w;
    light->depth = d;
}

static void light_quad(tri_sink *s, simd_float3 c, float w, float d, MaterialGPU lm) {
    float halfW = w / 2, halfD = d / 2;
    simd_float3 v0 = v3(c.x - halfW, c.y, c.z - halfD);
    simd_float3 v1 = v3(c.x + halfW, c.y, c.z - halfD);
    simd_float3 v2 = v3(c.x + halfW, c.y, c.z + halfD);
    simd_float3 v3_ = v3(c.x - halfW, c.y, c.z + halfD);
    push(s, v0, v1, v2, lm);          /* scene.swift:58 */
    push(s, v0, v2, v3_, lm);         /* scene.swift:59 */
}

void rt_fixture_cornell(int width, int height, float lightSize, CameraGPU *camera,
                        MaterialGPU *materials, SquareLightGPU *light, simd_float3 *vertices) {
    tri_sink s = { vertices, materials, 0 };
    const float roomSize = 5.0f;
    const float half = roomSize / 2.0f;

    fill_camera(camera, width, height);
    room_walls(&s, half);

    MaterialGPU diffuseBox = material(0.9f, 0.9f, 0.9f, 1, 0.05f, 0.3f, 0, 0, 0);
    simd_float3 box[8];
    /* tall box, scene.swift:141-155 */
    {
        float w = 1.2f, h = 2.8f, d = 1.2f;
        simd_float3 pos = v3(-1.0f, -half + h / 2 - 0.05f, -1.5f);
        rotated_box(pos, w, h, d, RT_PI_F / 2.4f, box);
        box_triangles(&s, box, diffuseBox);
    }
    /* short box, scene.swift:158-172 */
    {
        float size = 1.2f, h = 1.2f;
        simd_float3 pos = v3(0.7f, -half + h / 2 - 0.05f, 1.2f);
        rotated_box(pos, size, h, size, -RT_PI_F / 2.5f, box);
        box_triangles(&s, box, diffuseBox);
    }
    /* light, scene.swift:23-59: appended last => primitive ids 34, 35 */
    MaterialGPU lm = material(1.0f, 0.95f, 0.9f, 1.0f, 0.0f, 0.0f, 1.0f, 1.0f, 1.0f);
    float lightY = half - 0.01f;
    simd_float3 center = v3(0, lightY, 0);
    light_quad(&s, center, lightSize, lightSize, lm);
    fill_light(light, center, lm, lightSize, lightSize);
}

/* randomFloat, sampling.metal:77-79: hash(seed) / 2^32 */
static float random_float(uint32_t seed) {
    return (float)rt_oracle_hash(seed) / ((float)0xffffffffU + 1.0f);
}

void rt_fixture_spheres(int width, int height, uint32_t seed, uint32_t numSpheres,
                        CameraGPU *camera, MaterialGPU *materials, SquareLightGPU *lights,
                        simd_float3 *vertices, SphereGPU *spheres) {
    tri_sink s = { vertices, materials, 0 };
    const float half = 2.5f;
    fill_camera(camera, width, height);
    room_walls(&s, half);
    /* four 0.5 x 0.5 ceiling lights (the extent sampleAreaLight hard-codes) */
    MaterialGPU lm = material(1.0f, 0.95f, 0.9f, 1.0f, 0.0f, 0.0f, 1.0f, 1.0f, 1.0f);
    const float lx[4] = { -1.25f, 1.25f, -1.25f, 1.25f };
    const float lz[4] = { -1.25f, -1.25f, 1.25f, 1.25f };
    for (int l = 0; l < 4; l++) {
        simd_float3 c = v3(lx[l], half - 0.01f, lz[l]);
        light_quad(&s, c, 0.5f, 0.5f, lm);
        fill_light(&lights[l], c, lm, 0.5f, 0.5f);
    }
    /* jittered 8 x 8 x 4 lattice (x, z, y), cell 0.55, radius in [0.10, 0.20) */
    for (uint32_t k = 0; k < numSpheres; k++) {
        uint32_t cell = k % 256u;
        uint32_t gx = cell % 8u, gz = (cell / 8u) % 8u, gy = cell / 64u;
        uint32_t h = seed * 0x9E3779B9u + k * 16u;
        float jx = random_float(h + 0), jy = random_float(h + 1), jz = random_float(h + 2);
        float rr = random_float(h + 3);
        SphereGPU *sp = &spheres[k];
        memset(sp, 0, sizeof *sp);
        sp->radius = 0.10f + 0.10f * rr;
        sp->center = v3(-1.925f + 0.55f * (float)gx + 0.15f * (jx - 0.5f),
                        -2.1f + 0.9f * (float)gy + 0.3f * (jy - 0.5f),
                        -1.925f + 0.55f * (float)gz + 0.15f * (jz - 0.5f));
        sp->material = material(0.2f + 0.7f * random_float(h + 4), 0.2f + 0.7f * random_float(h + 5),
                                0.2f + 0.7f * random_float(h + 6), 1.0f, 0.05f, 0.5f, 0, 0, 0);
    }
}

void rt_fixture_offsets(uint32_t *offsets, int width, int row0, int rows, uint32_t seed) {
    for (int ly = 0; ly < rows; ly++)
        for (int x = 0; x < width; x++) {
            uint32_t p = (uint32_t)(row0 + ly) * (uint32_t)width + (uint32_t)x;
            offsets[(size_t)ly * width + x] = rt_oracle_hash(p + 0x9E3779B9u * seed) & 0xFFFFFu;
        }
}
