/*
 * host_renderer.h -- C++ mirror of the reference's host objects, written against the C ABI only
 * (include/rtcuda.h): what RTrace/renderer.swift and RTrace/image.swift would look like with every Metal
 * call replaced by its rt* entry.  The Swift toolchain is absent here, so this is the compiled host of the
 * build; INTEGRATION.md shows the same calls in Swift.
 *
 *   class Renderer             RTrace/renderer.swift:9-146      (same members, same order of calls)
 *   saveTextureToImage         RTrace/image.swift:15-100        (tone map on the device, PNG on the host)
 *   savePixelArrayToImage      RTrace/image.swift:102-157
 *
 * Errors: the Swift code force-unwraps / fatalError()s; here every failed rt* call throws std::runtime_error
 * carrying rtLastErrorString().
 */
#pragma once
#include <stdexcept>
#include <string>
#include <vector>
#include "host_scene.h"

namespace rthost {

/* what the reference hard-codes (raytrace.metal:24-25, scene.swift:16-18) plus the knobs BASELINE.json's configs need */
struct RenderOptions {
    int width = 800, height = 600;
    uint32_t spp = 400;              /* raytrace.metal:24 */
    int bounces = 3;                 /* raytrace.metal:25 */
    int ordinal = 0;                 /* CUDA device */
    uint32_t seed = 1;               /* per-pixel Halton offsets: hash(p + 0x9E3779B9*seed) & 0xFFFFF replaces
                                        arc4random() % 2^20 (renderer.swift:96-101) so that runs are repeatable */
    bool sphereScene = false;        /* BASELINE config 4 instead of initCornellBox() */
    uint32_t numSpheres = 256;
    bool lightExtent = false;        /* sample the light's own width x depth (RT_VARIANT_LIGHT_EXTENT) instead of the
                                        0.5 x 0.5 sampling.metal:207-209 hard-codes */
};

class Renderer {
public:
    /* renderer.swift:29-113: device, pipeline "pathTrace", queue, scene -> buffers, textures, accel */
    explicit Renderer(const RenderOptions &opt = RenderOptions());
    ~Renderer();
    Renderer(const Renderer &) = delete;
    Renderer &operator=(const Renderer &) = delete;

    /* renderer.swift:115-146: encode, dispatchThreads(width x height, 8 x 8), commit, wait, save the image */
    void draw(const std::string &outputFileName);
    /* the same without writing a file */
    void render();

    double gpuMilliseconds() const { return gpuMs_; }
    const Camera &cameraRef() const { return camera; }
    RTTexture *renderTextureHandle() const { return renderTexture; }
    RTQueue *queueHandle() const { return commandQueue; }
    RTDevice *deviceHandle() const { return device; }

private:
    /* the members of renderer.swift:10-27, in its order */
    RTDevice *device = nullptr;
    RTPipeline *computePipeline = nullptr;
    RTQueue *commandQueue = nullptr;
    RTBuffer *cameraBuffer = nullptr, *materialsBuffer = nullptr, *lightsBuffer = nullptr, *vertexBuffer = nullptr,
             *debugBuffer = nullptr, *spheresBuffer = nullptr;
    RTTexture *randomTexture = nullptr, *renderTexture = nullptr;
    RTAccel *accelerationStructure = nullptr;
    Camera camera;

    RenderOptions opt_;
    uint32_t numTriangles_ = 0, numSpheres_ = 0, numLights_ = 0;
    double gpuMs_ = 0.0;
    void release();
};

/* image.swift:15-100: x2 exposure, Reinhard, gamma 1/2.2, 8 bit (on the device: rtResolveToRGBA8), then a PNG */
void saveTextureToImage(RTDevice *device, RTQueue *queue, RTTexture *texture, int width, int height, const std::string &fileName);
/* image.swift:102-157: RGBA8 pixels -> PNG (uncompressed deflate blocks; no libpng / CoreGraphics needed) */
void savePixelArrayToImage(const std::vector<uint8_t> &pixels, int width, int height, const std::string &fileName);

}  // namespace rthost

extern "C" {
/* flat entry for the Python tests: 0 on success, -1 (and rtLastErrorString-free stderr text) on failure */
int rtHostSavePixelArrayToImage(const uint8_t *pixels, int width, int height, const char *fileName);
}
