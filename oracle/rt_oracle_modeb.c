/* placeholder until mode B lands (rt_oracle_drawtriangle) */
#include "rt_oracle.h"
int rt_oracle_drawtriangle(const CameraGPU *cameras, const MaterialGPU *materials,
                           const SquareLightGPU *lights, const simd_float3 *vertices,
                           uint8_t *pixels, float *text, const RTLaunchParams *params,
                           uint64_t counters[RT_CNT_COUNT], int nthreads) {
    (void)cameras; (void)materials; (void)lights; (void)vertices; (void)pixels; (void)text;
    (void)params; (void)counters; (void)nthreads;
    return RT_ERR_UNSUPPORTED;
}
