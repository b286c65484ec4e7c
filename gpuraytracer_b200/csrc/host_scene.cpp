/*
 * host_scene.cpp -- host-side mirror of the reference's scene construction and GPU
 * marshalling, above the C ABI (the reference host is Swift; no Swift toolchain
 * exists on the build or GPU machines, so the mirror is C++ with the same names):
 *
 *   initCornellBox / createCornellBoxScene / createRotatedBoxVertices /
 *   createBoxTriangles                    RTrace/scene.swift:14-240
 *   convertMaterial / convertSquareLight / convertCameras / convertSpheres
 *                                         RTrace/computeShader.swift:13-41,232-253
 *   SquareLight.emittedLuminance          RTrace/scene.swift:257-270
 *   the per-pixel offset texture          RTrace/renderer.swift:96-110 (seeded here)
 *
 * Swift evaluates the scene literals in Float, operation by operation; the same
 * order is kept (this file is built with -ffp-contract=off).  cos/sin of the box
 * rotations use the library's own polynomial so the bytes do not depend on libm.
 */
#include "host_scene.h"
#include <math.h>
#include <string.h>

namespace rthost {

static const float kPi = 3.14159265358979323846264338327950288f;

/* same arithmetic as the device sincos (rt_device_math.cuh) */
static void sincosDet(float x, float *s_out, float *c_out) {
    const float TWO_OVER_PI = 0.6366197466850281f, PIO2_HI = 1.5707963705062866f, PIO2_LO = -4.371138828673793e-08f;
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
    float ss = (k & 1) ? c : s, cc = (k & 1) ? s : c;
    *s_out = (k & 2) ? -ss : ss;
    *c_out = ((k + 1) & 2) ? -cc : cc;
}

static simd_float3 f3(float x, float y, float z) { simd_float3 v; v.x = x; v.y = y; v.z = z; v._pad = 0.0f; return v; }
static simd_float4 f4(float x, float y, float z, float w) { simd_float4 v; v.x = x; v.y = y; v.z = z; v.w = w; return v; }

/* scene.swift:177-210 */
static std::vector<simd_float3> createRotatedBoxVertices(simd_float3 center, float width, float height, float depth,
                                                         float rotationY) {
    float halfW = width / 2, halfH = height / 2, halfD = depth / 2;
    const simd_float3 base[8] = {
        f3(-halfW, -halfH, -halfD), f3(halfW, -halfH, -halfD), f3(halfW, halfH, -halfD), f3(-halfW, halfH, -halfD),
        f3(-halfW, -halfH, halfD),  f3(halfW, -halfH, halfD),  f3(halfW, halfH, halfD),  f3(-halfW, halfH, halfD) };
    float sinY, cosY;
    sincosDet(rotationY, &sinY, &cosY);
    std::vector<simd_float3> out;
    for (const simd_float3 &v : base) {
        /* simd_float4x4 takes columns: x' = cos*x - sin*z, z' = sin*x + cos*z (scene.swift:197-202) */
        float rx = cosY * v.x - sinY * v.z;
        float rz = sinY * v.x + cosY * v.z;
        out.push_back(f3(rx + center.x, v.y + center.y, rz + center.z));
    }
    return out;
}

/* scene.swift:212-240 */
static void createBoxTriangles(std::vector<Triangle> &t, const std::vector<simd_float3> &v, const Material &m) {
    const int idx[12][3] = { {0,2,1},{0,3,2}, {4,5,6},{4,6,7}, {0,4,7},{0,7,3}, {1,6,5},{1,2,6}, {0,5,4},{0,1,5}, {3,6,2},{3,7,6} };
    for (const int *i : idx) t.push_back(Triangle{ { v[i[0]], v[i[1]], v[i[2]] }, m });
}

static void roomWalls(std::vector<Triangle> &t, float h) {
    Material red{ f4(0.9f, 0.0f, 0.0f, 1), 0.05f, 0.3f, f3(0, 0, 0) };
    Material green{ f4(0.0f, 0.7f, 0.0f, 1), 0.05f, 0.8f, f3(0, 0, 0) };
    Material white{ f4(0.9f, 0.9f, 0.9f, 1), 0.05f, 0.8f, f3(0, 0, 0) };
    auto tri = [&](simd_float3 a, simd_float3 b, simd_float3 c, const Material &m) { t.push_back(Triangle{ { a, b, c }, m }); };
    tri(f3(-h,-h,-h), f3( h, h,-h), f3(-h, h,-h), white);  tri(f3(-h,-h,-h), f3( h,-h,-h), f3( h, h,-h), white);   /* back    */
    tri(f3(-h,-h,-h), f3(-h, h, h), f3(-h,-h, h), red);    tri(f3(-h,-h,-h), f3(-h, h,-h), f3(-h, h, h), red);     /* left    */
    tri(f3( h,-h,-h), f3( h, h, h), f3( h, h,-h), green);  tri(f3( h,-h,-h), f3( h,-h, h), f3( h, h, h), green);   /* right   */
    tri(f3(-h,-h,-h), f3( h,-h, h), f3( h,-h,-h), white);  tri(f3(-h,-h,-h), f3(-h,-h, h), f3( h,-h, h), white);   /* floor   */
    tri(f3(-h, h,-h), f3( h, h, h), f3(-h, h, h), white);  tri(f3(-h, h,-h), f3( h, h,-h), f3( h, h, h), white);   /* ceiling */
}

/* scene.swift:64-175 */
std::vector<Triangle> createCornellBoxScene() {
    std::vector<Triangle> t;
    const float roomSize = 5.0f, half = roomSize / 2.0f;
    roomWalls(t, half);
    Material diffuseBox{ f4(0.9f, 0.9f, 0.9f, 1), 0.05f, 0.3f, f3(0, 0, 0) };
    const float tallH = 2.8f;
    createBoxTriangles(t, createRotatedBoxVertices(f3(-1.0f, -half + tallH / 2 - 0.05f, -1.5f), 1.2f, tallH, 1.2f, kPi / 2.4f), diffuseBox);
    const float shortH = 1.2f;
    createBoxTriangles(t, createRotatedBoxVertices(f3(0.7f, -half + shortH / 2 - 0.05f, 1.2f), 1.2f, shortH, 1.2f, -kPi / 2.5f), diffuseBox);
    return t;
}

static SquareLight makeLight(simd_float3 center, float w, float d, std::vector<Triangle> *tris) {
    float halfW = w / 2, halfD = d / 2;
    simd_float3 v0 = f3(center.x - halfW, center.y, center.z - halfD), v1 = f3(center.x + halfW, center.y, center.z - halfD);
    simd_float3 v2 = f3(center.x + halfW, center.y, center.z + halfD), v3 = f3(center.x - halfW, center.y, center.z + halfD);
    Material lm{ f4(1.0f, 0.95f, 0.9f, 1.0f), 0.0f, 0.0f, f3(1.0f, 1.0f, 1.0f) };
    SquareLight l{ center, { v0, v1, v2, v3 }, lm, 100.0f, 12.0f, w, d };
    if (tris) { tris->push_back(Triangle{ { v0, v1, v2 }, lm }); tris->push_back(Triangle{ { v0, v2, v3 }, lm }); }
    return l;
}

static Camera defaultCamera(int width, int height) {
    Camera c;
    c.position = f3(0, 0, 9);
    c.direction = f3(0, 0, -1);        /* normalize((0,0,-2.5) - (0,0,9)) exactly */
    c.up = f3(0, 1, 0);
    c.resolution.x = width; c.resolution.y = height;
    c.horizontalFov = kPi / 4.0f;
    c.ev100 = 5.0f;
    return c;
}

/* scene.swift:14-62; lightSize 1.0 there, 1.5 in Sources/gpuRaytracer/main.swift:30-31 */
Scene initCornellBox(int width, int height, float lightSize) {
    Scene s;
    s.camera = defaultCamera(width, height);
    s.triangles = createCornellBoxScene();
    const float half = 5.0f / 2.0f;
    s.lights.push_back(makeLight(f3(0, half - 0.01f, 0), lightSize, lightSize, &s.triangles));
    return s;
}

/* sampling.metal:68-79 */
uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
static float randomFloat(uint32_t seed) { return (float)hash(seed) / ((float)0xffffffffU + 1.0f); }

/* BASELINE.json config 4: room + four 0.5x0.5 ceiling lights + spheres on a jittered lattice */
Scene initSphereScene(int width, int height, uint32_t seed, uint32_t numSpheres) {
    Scene s;
    s.camera = defaultCamera(width, height);
    const float half = 2.5f;
    roomWalls(s.triangles, half);
    const float lx[4] = { -1.25f, 1.25f, -1.25f, 1.25f }, lz[4] = { -1.25f, -1.25f, 1.25f, 1.25f };
    for (int l = 0; l < 4; l++) s.lights.push_back(makeLight(f3(lx[l], half - 0.01f, lz[l]), 0.5f, 0.5f, &s.triangles));
    for (uint32_t k = 0; k < numSpheres; k++) {
        uint32_t cell = k % 256u, gx = cell % 8u, gz = (cell / 8u) % 8u, gy = cell / 64u;
        uint32_t h = seed * 0x9E3779B9u + k * 16u;
        float jx = randomFloat(h + 0), jy = randomFloat(h + 1), jz = randomFloat(h + 2), rr = randomFloat(h + 3);
        Sphere sp;
        sp.radius = 0.10f + 0.10f * rr;
        sp.center = f3(-1.925f + 0.55f * (float)gx + 0.15f * (jx - 0.5f),
                       -2.1f + 0.9f * (float)gy + 0.3f * (jy - 0.5f),
                       -1.925f + 0.55f * (float)gz + 0.15f * (jz - 0.5f));
        sp.material = Material{ f4(0.2f + 0.7f * randomFloat(h + 4), 0.2f + 0.7f * randomFloat(h + 5),
                                   0.2f + 0.7f * randomFloat(h + 6), 1.0f), 0.05f, 0.5f, f3(0, 0, 0) };
        s.spheres.push_back(sp);
    }
    return s;
}

/* ---- converters, computeShader.swift:13-41,232-253 --------------------------- */
MaterialGPU convertMaterial(const Material &m) {
    MaterialGPU g; memset(&g, 0, sizeof g);
    g.diffuse = m.diffuse; g.metallic = m.metallic; g.roughness = m.roughness; g.emissive = m.emissive;
    return g;
}
simd_float3 SquareLight::emittedLuminance() const {       /* scene.swift:257-270 */
    float luminousFlux = luminousEfficacy * watts;
    float area = width * depth;
    float luminanceExitance = luminousFlux / area;
    float luminance = luminanceExitance / kPi;
    return f3(material.diffuse.x * luminance, material.diffuse.y * luminance, material.diffuse.z * luminance);
}
SquareLightGPU convertSquareLight(const SquareLight &l) {
    SquareLightGPU g; memset(&g, 0, sizeof g);
    g.center = l.center; g.color = l.material.diffuse; g.emittedRadiance = l.emittedLuminance();
    g.width = l.width; g.depth = l.depth;
    return g;
}
CameraGPU convertCamera(const Camera &c) {
    CameraGPU g; memset(&g, 0, sizeof g);
    g.position = c.position; g.direction = c.direction; g.up = c.up; g.resolution = c.resolution;
    g.horizontalFov = c.horizontalFov; g.ev100 = c.ev100;
    return g;
}
SphereGPU convertSphere(const Sphere &s) {
    SphereGPU g; memset(&g, 0, sizeof g);
    g.center = s.center; g.material = convertMaterial(s.material); g.radius = s.radius;
    return g;
}

void fillOffsets(uint32_t *offsets, int width, int row0, int rows, uint32_t seed) {
    for (int ly = 0; ly < rows; ly++)
        for (int x = 0; x < width; x++) {
            uint32_t p = (uint32_t)(row0 + ly) * (uint32_t)width + (uint32_t)x;
            offsets[(size_t)ly * width + x] = hash(p + 0x9E3779B9u * seed) & 0xFFFFFu;
        }
}

}  // namespace rthost

/* ---- flat C entry points (ctypes) --------------------------------------------- */
using namespace rthost;

static int flatten(const Scene &s, RTHostSceneBytes *out) {
    if (!out) return RT_ERR_INVALID_ARGUMENT;
    uint32_t T = (uint32_t)s.triangles.size(), S = (uint32_t)s.spheres.size(), L = (uint32_t)s.lights.size();
    if (out->capTriangles < T || out->capSpheres < S || out->capLights < L) return RT_ERR_OUT_OF_RANGE;
    *out->camera = convertCamera(s.camera);
    for (uint32_t i = 0; i < T; i++) {
        out->materials[i] = convertMaterial(s.triangles[i].material);          /* renderer.swift:51 */
        for (int k = 0; k < 3; k++) out->vertices[3 * i + k] = s.triangles[i].vertices[k];   /* :54-57 */
    }
    for (uint32_t l = 0; l < L; l++) out->lights[l] = convertSquareLight(s.lights[l]);
    for (uint32_t j = 0; j < S; j++) out->spheres[j] = convertSphere(s.spheres[j]);
    out->numTriangles = T; out->numSpheres = S; out->numLights = L;
    return RT_OK;
}

extern "C" int rtHostInitCornellBox(int width, int height, float lightSize, RTHostSceneBytes *out) {
    return flatten(initCornellBox(width, height, lightSize), out);
}
extern "C" int rtHostInitSphereScene(int width, int height, uint32_t seed, uint32_t numSpheres, RTHostSceneBytes *out) {
    return flatten(initSphereScene(width, height, seed, numSpheres), out);
}
extern "C" void rtHostFillOffsets(uint32_t *offsets, int width, int row0, int rows, uint32_t seed) {
    fillOffsets(offsets, width, row0, rows, seed);
}
