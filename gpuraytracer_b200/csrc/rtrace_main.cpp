/*
 * rtrace_main.cpp -- the `RTrace` executable of the reference (RTrace/main.swift:13-30) over librtcuda.so:
 *     outputFileName = argv[1] or "output.png";  let renderer = try! Renderer();  renderer.draw()
 * plus the knobs BASELINE.json's configs need.  Built as gpuraytracer_b200/rtrace by the Makefile.
 *
 *   rtrace [output.png] [--width W] [--height H] [--spp N] [--bounces B] [--device K] [--seed S] [--spheres]
 *          [--light-extent]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include "host_renderer.h"

int main(int argc, char **argv) {
    std::string outputFileName = "output.png";                 /* main.swift:13-26 */
    rthost::RenderOptions opt;
    bool named = false;
    for (int i = 1; i < argc; i++) {
        const char *a = argv[i];
        auto value = [&](const char *flag) -> long {
            if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", flag); exit(2); }
            return strtol(argv[++i], nullptr, 10);
        };
        if (!strcmp(a, "--width")) opt.width = (int)value(a);
        else if (!strcmp(a, "--height")) opt.height = (int)value(a);
        else if (!strcmp(a, "--spp")) opt.spp = (uint32_t)value(a);
        else if (!strcmp(a, "--bounces")) opt.bounces = (int)value(a);
        else if (!strcmp(a, "--device")) opt.ordinal = (int)value(a);
        else if (!strcmp(a, "--seed")) opt.seed = (uint32_t)value(a);
        else if (!strcmp(a, "--spheres")) opt.sphereScene = true;
        else if (!strcmp(a, "--light-extent")) opt.lightExtent = true;
        else if (a[0] == '-') { fprintf(stderr, "unknown option %s\n", a); return 2; }
        else if (!named) { outputFileName = a; named = true; }
    }
    try {
        const auto t0 = std::chrono::steady_clock::now();
        rthost::Renderer renderer(opt);                         /* main.swift:28 */
        renderer.draw(outputFileName);                          /* main.swift:30 */
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        const double paths = (double)opt.width * opt.height * opt.spp;
        printf("Render completed in %.2f seconds (device %.3f ms, %.1f Mpaths/s)\n", s, renderer.gpuMilliseconds(),
               paths / renderer.gpuMilliseconds() / 1e3);       /* Sources/gpuRaytracer/main.swift:86-91 */
    } catch (const std::exception &e) {
        fprintf(stderr, "rtrace: %s\n", e.what());              /* Swift: fatalError / try! */
        return 1;
    }
    return 0;
}
