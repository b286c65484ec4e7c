/*
 * host_renderer.cpp -- see host_renderer.h.  Every line that talks to the GPU is an rt* call of
 * include/rtcuda.h; the comments name the Metal call of RTrace/renderer.swift it stands for.
 */
#include "host_renderer.h"

#include <stdio.h>
#include <string.h>

namespace rthost {

static void check(int rc, const char *what) {
    if (rc != RT_OK) throw std::runtime_error(std::string(what) + ": " + rtLastErrorString());
}

Renderer::Renderer(const RenderOptions &opt) : opt_(opt) {
    try {
        check(rtDeviceCreate(opt.ordinal, &device), "rtDeviceCreate");                       /* :30 MTLCreateSystemDefaultDevice */
        check(rtPipelineCreate(device, "pathTrace", &computePipeline), "rtPipelineCreate");    /* :32-43 library, function, pipeline state */
        check(rtQueueCreate(device, &commandQueue), "rtQueueCreate");                         /* :44 makeCommandQueue */

        Scene scene = opt.sphereScene ? initSphereScene(opt.width, opt.height, 12345u, opt.numSpheres)
                                      : initCornellBox(opt.width, opt.height);                /* :46 */
        CameraGPU cameraGPU = convertCamera(scene.camera);                                    /* :49 convertCameras */
        std::vector<MaterialGPU> materialsGPU;                                                /* :51 */
        std::vector<simd_float3> allVertices;                                                 /* :54-57 */
        for (const Triangle &t : scene.triangles) {
            materialsGPU.push_back(convertMaterial(t.material));
            for (int k = 0; k < 3; k++) allVertices.push_back(t.vertices[k]);
        }
        std::vector<SquareLightGPU> squareLightsGPU;                                          /* :52 (one light in the reference) */
        for (const SquareLight &l : scene.lights) squareLightsGPU.push_back(convertSquareLight(l));
        std::vector<SphereGPU> spheresGPU;
        for (const Sphere &s : scene.spheres) spheresGPU.push_back(convertSphere(s));
        camera = scene.camera;                                                                /* :59 */
        numTriangles_ = (uint32_t)scene.triangles.size();
        numSpheres_ = (uint32_t)spheresGPU.size();
        numLights_ = (uint32_t)squareLightsGPU.size();

        /* :61-72 makeBuffer(bytes:length:options:.storageModeShared) */
        check(rtBufferCreate(device, &cameraGPU, sizeof cameraGPU, &cameraBuffer), "cameraBuffer");
        check(rtBufferCreate(device, materialsGPU.data(), materialsGPU.size() * sizeof(MaterialGPU), &materialsBuffer), "materialsBuffer");
        check(rtBufferCreate(device, squareLightsGPU.data(), squareLightsGPU.size() * sizeof(SquareLightGPU), &lightsBuffer), "lightsBuffer");
        check(rtBufferCreate(device, allVertices.data(), allVertices.size() * sizeof(simd_float3), &vertexBuffer), "vertexBuffer");
        check(rtBufferCreate(device, nullptr, (size_t)camera.pixelCount() * sizeof(simd_float3), &debugBuffer), "debugBuffer");
        if (numSpheres_)
            check(rtBufferCreate(device, spheresGPU.data(), spheresGPU.size() * sizeof(SphereGPU), &spheresBuffer), "spheresBuffer");

        /* :74-93 render target (rgba16Float there, float4 here) and the r32Uint random texture */
        const int width = camera.resolution.x, height = camera.resolution.y;
        check(rtTextureCreate(device, RT_FORMAT_RGBA32_FLOAT, width, height, &renderTexture), "renderTexture");
        check(rtTextureCreate(device, RT_FORMAT_R32_UINT, width, height, &randomTexture), "randomTexture");
        std::vector<uint32_t> randomData((size_t)width * height);                             /* :95-101 */
        fillOffsets(randomData.data(), width, 0, height, opt.seed);
        check(rtTextureReplaceRegion(randomTexture, 0, 0, width, height, randomData.data(),
                                     (size_t)width * sizeof(uint32_t)), "randomTexture.replace");   /* :103-110 */

        /* :112 setupAccelerationStructures(device:triangles:) */
        check(rtAccelerationStructureCreate(device, vertexBuffer, 0, sizeof(simd_float3), numTriangles_, &accelerationStructure),
              "setupAccelerationStructures");
    } catch (...) {
        release();
        throw;
    }
}

void Renderer::release() {
    rtAccelerationStructureRelease(accelerationStructure);
    rtTextureRelease(randomTexture); rtTextureRelease(renderTexture);
    rtBufferRelease(spheresBuffer); rtBufferRelease(debugBuffer); rtBufferRelease(vertexBuffer);
    rtBufferRelease(lightsBuffer); rtBufferRelease(materialsBuffer); rtBufferRelease(cameraBuffer);
    rtQueueRelease(commandQueue); rtPipelineRelease(computePipeline); rtDeviceRelease(device);
    accelerationStructure = nullptr; randomTexture = renderTexture = nullptr;
    spheresBuffer = debugBuffer = vertexBuffer = lightsBuffer = materialsBuffer = cameraBuffer = nullptr;
    commandQueue = nullptr; computePipeline = nullptr; device = nullptr;
}

Renderer::~Renderer() { release(); }

void Renderer::render() {
    RTCommand *commandBuffer = nullptr;
    check(rtCommandBegin(commandQueue, computePipeline, &commandBuffer), "makeCommandBuffer");   /* :118-121 */
    try {
        check(rtSetBuffer(commandBuffer, cameraBuffer, 0, 0), "setBuffer 0");                      /* :122 */
        check(rtSetBuffer(commandBuffer, materialsBuffer, 0, 1), "setBuffer 1");                   /* :123 */
        check(rtSetBuffer(commandBuffer, lightsBuffer, 0, 2), "setBuffer 2");                      /* :124 */
        check(rtSetBuffer(commandBuffer, vertexBuffer, 0, 3), "setBuffer 3");                      /* :125 */
        check(rtSetAccelerationStructure(commandBuffer, accelerationStructure, 5), "setAccelerationStructure");   /* :127 */
        check(rtSetBuffer(commandBuffer, debugBuffer, 0, 6), "setBuffer 6");                       /* :128 */
        if (spheresBuffer) check(rtSetBuffer(commandBuffer, spheresBuffer, 0, RT_INDEX_SPHERES), "setBuffer 7");
        check(rtSetTexture(commandBuffer, renderTexture, 0), "setTexture 0");                      /* :129 */
        check(rtSetTexture(commandBuffer, randomTexture, 1), "setTexture 1");                      /* :130 */

        const int width = camera.resolution.x, height = camera.resolution.y;
        RTLaunchParams p;
        check(rtDefaultParams(width, height, &p), "rtDefaultParams");     /* the shader's literals ... */
        p.spp = opt_.spp; p.bounces = opt_.bounces;                       /* ... as parameters */
        if (opt_.lightExtent) p.variant |= RT_VARIANT_LIGHT_EXTENT;       /* sampling.metal:207-209 with the light's own size */
        p.numTriangles = numTriangles_; p.numSpheres = numSpheres_; p.numLights = numLights_;
        check(rtSetParams(commandBuffer, &p), "rtSetParams");
        check(rtDispatchThreads(commandBuffer, width, height, 1, 8, 8, 1), "dispatchThreads");     /* :137-141 */
        check(rtCommit(commandBuffer), "commit");                                                  /* :142-143 */
        check(rtWaitUntilCompleted(commandBuffer), "waitUntilCompleted");                          /* :144 */
        check(rtCommandGetGpuTimeMs(commandBuffer, &gpuMs_), "gpu time");
    } catch (...) {
        rtCommandRelease(commandBuffer);
        throw;
    }
    rtCommandRelease(commandBuffer);
}

void Renderer::draw(const std::string &outputFileName) {
    render();
    saveTextureToImage(device, commandQueue, renderTexture, camera.resolution.x, camera.resolution.y, outputFileName);   /* :146 */
}

/* ------------------------------------------------------------------ image.swift ---- */
void saveTextureToImage(RTDevice *device, RTQueue *queue, RTTexture *texture, int width, int height, const std::string &fileName) {
    /* image.swift:40-65 runs on the device (rt_resolve_rgba8: x2, Reinhard, gamma 1/2.2, clamp, 8 bit, alpha 255) */
    RTBuffer *storage = rtTextureBuffer(texture);
    if (!storage) throw std::runtime_error("saveTextureToImage: texture has no storage");
    const uint64_t n = (uint64_t)width * height;
    RTBuffer *rgba = nullptr;
    check(rtBufferCreate(device, nullptr, n * 4, &rgba), "rgba buffer");
    std::vector<uint8_t> pixels(n * 4);
    int rc = rtResolveToRGBA8(queue, storage, 0, rgba, 0, n, 0);
    const void *host = rc == RT_OK ? rtBufferContents(rgba) : nullptr;      /* .contents(): waits for the queue */
    if (host) memcpy(pixels.data(), host, pixels.size());
    rtBufferRelease(rgba);
    check(rc, "rtResolveToRGBA8");
    if (!host) throw std::runtime_error(std::string("rtBufferContents: ") + rtLastErrorString());
    savePixelArrayToImage(pixels, width, height, fileName);
}

/* PNG: 8-bit RGBA, one IDAT holding a zlib stream of stored (uncompressed) deflate blocks */
static uint32_t crc32_update(uint32_t c, const uint8_t *p, size_t n) {
    static uint32_t table[256];
    if (!table[1]) for (uint32_t i = 0; i < 256; i++) { uint32_t v = i; for (int k = 0; k < 8; k++) v = (v >> 1) ^ (0xEDB88320u & (0u - (v & 1u))); table[i] = v; }
    for (size_t i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xff] ^ (c >> 8);
    return c;
}
static void put32(std::vector<uint8_t> &v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
static void chunk(std::vector<uint8_t> &png, const char type[4], const std::vector<uint8_t> &data) {
    put32(png, (uint32_t)data.size());
    const size_t start = png.size();
    png.insert(png.end(), type, type + 4);
    png.insert(png.end(), data.begin(), data.end());
    put32(png, crc32_update(0xffffffffu, png.data() + start, png.size() - start) ^ 0xffffffffu);
}

void savePixelArrayToImage(const std::vector<uint8_t> &pixels, int width, int height, const std::string &fileName) {
    if (width <= 0 || height <= 0 || pixels.size() < (size_t)width * height * 4)
        throw std::runtime_error("savePixelArrayToImage: pixel array smaller than width*height*4");   /* image.swift:103-106 */
    std::vector<uint8_t> raw;                                   /* filter byte 0 + the row */
    raw.reserve(((size_t)width * 4 + 1) * height);
    for (int y = 0; y < height; y++) {
        raw.push_back(0);
        raw.insert(raw.end(), pixels.begin() + (size_t)y * width * 4, pixels.begin() + (size_t)(y + 1) * width * 4);
    }
    std::vector<uint8_t> z = { 0x78, 0x01 };
    uint32_t a = 1, b = 0;                                      /* Adler-32 */
    for (size_t off = 0; off < raw.size(); off += 65535) {
        const size_t len = raw.size() - off < 65535 ? raw.size() - off : 65535;
        z.push_back(off + len == raw.size() ? 1 : 0);
        z.push_back(len & 0xff); z.push_back(len >> 8); z.push_back(~len & 0xff); z.push_back((~len >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + len);
        for (size_t i = 0; i < len; i++) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
    }
    put32(z, (b << 16) | a);
    std::vector<uint8_t> png = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
    std::vector<uint8_t> ihdr;
    put32(ihdr, (uint32_t)width); put32(ihdr, (uint32_t)height);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(png, "IHDR", ihdr);
    chunk(png, "IDAT", z);
    chunk(png, "IEND", {});
    FILE *f = fopen(fileName.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + fileName);
    const size_t w = fwrite(png.data(), 1, png.size(), f);
    fclose(f);
    if (w != png.size()) throw std::runtime_error("short write to " + fileName);
    printf("Image saved to %s\n", fileName.c_str());            /* image.swift:97 */
}

}  // namespace rthost

extern "C" int rtHostSavePixelArrayToImage(const uint8_t *pixels, int width, int height, const char *fileName) {
    try {
        if (!pixels || !fileName || width <= 0 || height <= 0) return -1;
        std::vector<uint8_t> v(pixels, pixels + (size_t)width * height * 4);
        rthost::savePixelArrayToImage(v, width, height, fileName);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "rtHostSavePixelArrayToImage: %s\n", e.what());
        return -1;
    }
}
