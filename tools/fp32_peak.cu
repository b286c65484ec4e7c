// FP32 issue/pipe microbenchmark for sm_100a ("an FFMA-saturating microbenchmark on the box").
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/fp32_peak tools/fp32_peak.cu && gpurun_out/fp32_peak
// Measures, per SM and clock: scalar FFMA, packed FFMA2 (fma.rn.f32x2, new on sm_100), and each of them
// interleaved with ALU-pipe work (LOP3) -- the question the path tracer's triangle loop asks: the kernel is
// issue-slot bound (profiles/r01v), so does a packed FMA free an issue slot for the compare/select work?
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ unsigned lop(unsigned a, unsigned b, unsigned c) { unsigned r; asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ float fsel(float a, float b, float c) { float r; asm volatile("{.reg .pred p; setp.lt.f32 p, %1, %2; selp.f32 %0, %3, %1, p;}" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

// MODE 0: FFMA x8   1: FFMA2 x8   2: FFMA x8 + LOP3 x4   3: FFMA2 x4 + LOP3 x4   4: LOP3 x8
// MODE 5: FFMA2 x4 + LOP3 x8      6: FFMA x8 + LOP3 x8   7: FFMA2 x4 + (FSETP+FSEL) x4
template <int MODE>
__global__ void __launch_bounds__(256) kern(const float* __restrict__ in, float* out, int iters) {
    float a[8], b = in[0], c = in[1];
    u64 p[8], pb, pc;
    unsigned q[8], qb = __float_as_uint(in[2]);
    for (int k = 0; k < 8; k++) {
        a[k] = in[3 + k] + threadIdx.x;
        q[k] = __float_as_uint(a[k]);
        asm("mov.b64 %0, {%1,%2};" : "=l"(p[k]) : "f"(a[k]), "f"(a[k] + 1.0f));
    }
    asm("mov.b64 %0, {%1,%2};" : "=l"(pb) : "f"(b), "f"(b));
    asm("mov.b64 %0, {%1,%2};" : "=l"(pc) : "f"(c), "f"(c));
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            if (MODE == 0 || MODE == 2 || MODE == 6) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = fma1(a[k], b, c);
            }
            if (MODE == 1) {
#pragma unroll
                for (int k = 0; k < 8; k++) p[k] = fma2(p[k], pb, pc);
            }
            if (MODE == 3 || MODE == 5 || MODE == 7) {
#pragma unroll
                for (int k = 0; k < 4; k++) p[k] = fma2(p[k], pb, pc);
            }
            if (MODE == 2 || MODE == 3) {
#pragma unroll
                for (int k = 0; k < 4; k++) q[k] = lop(q[k], qb, q[(k + 1) & 3]);
            }
            if (MODE == 4 || MODE == 5 || MODE == 6) {
#pragma unroll
                for (int k = 0; k < 8; k++) q[k] = lop(q[k], qb, q[(k + 1) & 7]);
            }
            if (MODE == 7) {
#pragma unroll
                for (int k = 4; k < 8; k++) a[k] = fsel(a[k], b, c);
            }
        }
    }
    float s = 0;
    for (int k = 0; k < 8; k++) {
        float lo, hi;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[k]));
        s += a[k] + lo + hi + __uint_as_float(q[k]);
    }
    if (s == 123.456f) out[0] = s;
}

struct Mix { const char* name; int ffma, ffma2, alu; };
static const Mix MIX[8] = {
    {"FFMA x8", 8, 0, 0},           {"FFMA2 x8", 0, 8, 0},          {"FFMA x8 + LOP3 x4", 8, 0, 4},  {"FFMA2 x4 + LOP3 x4", 0, 4, 4},
    {"LOP3 x8", 0, 0, 8},           {"FFMA2 x4 + LOP3 x8", 0, 4, 8}, {"FFMA x8 + LOP3 x8", 8, 0, 8},  {"FFMA2 x4 + FSETP/FSEL x4", 0, 4, 8}};

template <int MODE>
static void run(const float* in, float* out, int sms, double ghz, int json) {
    const int iters = 4096, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<MODE><<<blocks, 256>>>(in, out, 64);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        kern<MODE><<<blocks, 256>>>(in, out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const Mix& m = MIX[MODE];
    double warps = (double)blocks * 8, per = (double)iters * 4;
    double inst = warps * per * (m.ffma + m.ffma2 + m.alu);
    double flop = (double)blocks * 256 * per * (2.0 * m.ffma + 4.0 * m.ffma2);
    double clk = best * 1e-3 * ghz * 1e9;
    printf(json ? "{\"mix\": \"%s\", \"ms\": %.4f, \"tflops\": %.3f, \"warp_inst_per_clk_per_sm\": %.3f, \"fma_lane_ops_per_clk_per_sm\": %.2f}\n"
                : "%-28s %8.3f ms  %7.2f TFLOP/s  %5.2f warp-inst/clk/SM  %6.1f FMA lanes/clk/SM\n",
           m.name, best, flop / best / 1e9, inst / clk / sms, flop / 2 / clk / sms);
}

int main(int argc, char** argv) {
    int json = argc > 1;
    cudaDeviceProp pr;
    if (cudaGetDeviceProperties(&pr, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double ghz = khz * 1e-6;
    if (!json) printf("%s: %d SMs, %.3f GHz nominal -> %.2f TFLOP/s nominal FP32\n", pr.name, pr.multiProcessorCount, ghz,
                      pr.multiProcessorCount * 128 * 2 * ghz * 1e-3);
    float h[16]; for (int i = 0; i < 16; i++) h[i] = 1.0f + i * 1e-3f;
    h[0] = 0.999f; h[1] = 1e-3f;
    float *in, *out;
    cudaMalloc(&in, sizeof h); cudaMalloc(&out, 64);
    cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
    run<0>(in, out, pr.multiProcessorCount, ghz, json); run<1>(in, out, pr.multiProcessorCount, ghz, json);
    run<2>(in, out, pr.multiProcessorCount, ghz, json); run<3>(in, out, pr.multiProcessorCount, ghz, json);
    run<4>(in, out, pr.multiProcessorCount, ghz, json); run<5>(in, out, pr.multiProcessorCount, ghz, json);
    run<6>(in, out, pr.multiProcessorCount, ghz, json); run<7>(in, out, pr.multiProcessorCount, ghz, json);
    return 0;
}
