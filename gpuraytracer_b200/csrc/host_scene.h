/*
 * host_scene.h -- C++ mirror of the reference's host-side scene types
 * (RTrace/scene.swift:8-12,242-301) and the flat C entry points Python uses.
 */
#pragma once
#include <stdint.h>
#include <vector>
#include "../../include/rtcuda.h"
#include "../../include/shaderTypes.h"

namespace rthost {

struct Material {                    /* scene.swift:277-282 */
    simd_float4 diffuse;
    float metallic, roughness;
    simd_float3 emissive;
};
struct Triangle {                    /* scene.swift:242-245 */
    simd_float3 vertices[3];
    Material material;
};
struct SquareLight {                 /* scene.swift:248-271; LightType.bulb(luminousEfficacy:watts:) */
    simd_float3 center;
    simd_float3 vertices[4];
    Material material;
    float luminousEfficacy, watts;
    float width, depth;
    simd_float3 emittedLuminance() const;
};
struct Sphere {                      /* scene.swift:284-288 */
    simd_float3 center;
    Material material;
    float radius;
};
struct Camera {                      /* scene.swift:290-301 */
    simd_float3 position, direction, up;
    simd_int2 resolution;
    float horizontalFov, ev100;
    int pixelCount() const { return resolution.x * resolution.y; }
};
struct Scene {                       /* scene.swift:8-12 (+ spheres / several lights for config 4) */
    Camera camera;
    std::vector<SquareLight> lights;
    std::vector<Triangle> triangles;
    std::vector<Sphere> spheres;
};

Scene initCornellBox(int width = 800, int height = 600, float lightSize = 1.0f);
std::vector<Triangle> createCornellBoxScene();
Scene initSphereScene(int width, int height, uint32_t seed, uint32_t numSpheres);

MaterialGPU convertMaterial(const Material &m);
SquareLightGPU convertSquareLight(const SquareLight &l);
CameraGPU convertCamera(const Camera &c);
SphereGPU convertSphere(const Sphere &s);
uint32_t hash(uint32_t x);
void fillOffsets(uint32_t *offsets, int width, int row0, int rows, uint32_t seed);

}  // namespace rthost

extern "C" {
/* caller-provided arrays (capacities in elements); counts are written back */
typedef struct RTHostSceneBytes {
    CameraGPU *camera;
    MaterialGPU *materials;   uint32_t capTriangles;
    simd_float3 *vertices;    /* 3 * capTriangles */
    SquareLightGPU *lights;   uint32_t capLights;
    SphereGPU *spheres;       uint32_t capSpheres;
    uint32_t numTriangles, numSpheres, numLights;
} RTHostSceneBytes;

int  rtHostInitCornellBox(int width, int height, float lightSize, RTHostSceneBytes *out);
int  rtHostInitSphereScene(int width, int height, uint32_t seed, uint32_t numSpheres, RTHostSceneBytes *out);
void rtHostFillOffsets(uint32_t *offsets, int width, int row0, int rows, uint32_t seed);
}
