/*
 * rt_plan.h -- which kernel a command runs: a pure function of the launch parameters and of three device numbers
 * (SM count, shared memory per block and per SM), driven by the rule table in rt_plan.cpp.  No CUDA call in it,
 * so the choice is testable without a GPU (rtDebugChooseKernel; tests/test_plan_cpu.py pins it for every BASELINE
 * config at one GPU and at one rank's share of eight).
 *
 * The reference has no counterpart: its one kernel is named in makeFunction(name:) (RTrace/renderer.swift:39) and
 * dispatched over the image (renderer.swift:137-141); the forms here are schedules of that kernel, all giving the
 * same image bit for bit.
 */
#pragma once
#include <stdint.h>
#include <stddef.h>
#include "../../include/rtcuda.h"
#include "rt_scene_layout.h"

enum RTKernelForm {
    RT_FORM_SMEM = 0,          /* pathTrace_smem[_count]: one thread walks one pixel's samples                */
    RT_FORM_CONST,             /* pathTrace_const[_count]                                                     */
    RT_FORM_WAVE,              /* pathTrace_wave / _wave4 [_count]: warp-level path compaction, R = 6 / 4     */
    RT_FORM_TILES,             /* pathTrace_tiles / _tiles4: WAVE, one block per ticketed (tile, chunk) unit  */
    RT_FORM_WIDE,              /* pathTrace_wide / _wide4: WAVE for spp < R                                   */
    RT_FORM_SORTED,            /* pathTrace_sorted<R>: rays sorted by reachable object groups, per warp       */
    RT_FORM_SORTED_TILES,      /* pathTrace_sorted_tiles6                                                     */
    RT_FORM_SORTEDBLK,         /* pathTrace_sortedblk6: the sort taken over the block                         */
    RT_FORM_SORTEDBLK_TILES,   /* pathTrace_sortedblk_tiles6                                                  */
    RT_FORM_SORTEDBLK_WIDE,    /* pathTrace_sortedblk_wide6: spp < 6                                          */
    RT_FORM_BVH,               /* pathTrace_bvh: T > 1024                                                     */
    RT_FORM_MODEB,             /* drawTriangle_mis                                                            */
    RT_FORM_EXP,               /* pathTrace_exp<N> of a tuning build                                          */
    RT_FORM_COUNT
};

/* knobs for A/B measurements; read from the environment ONCE, at rtDeviceCreate (RT_SORTED_R, RT_NO_SORTED,
 * RT_NO_SORTED_BLOCK, RT_TILES_CHUNKS, RT_ALWAYS_PREP, RT_NO_HALTON_TABLE, RT_BVH_FORM), and settable per device with
 * rtDebugSetTuning (tests) */
struct RTTuning {
    int sortedR;               /* 4, 5, 6 or 8; default 6                                            */
    int noSorted, noSortedBlock, alwaysPrep, noHaltonTable;
    int tilesChunks;           /* > 0: force this many sample chunks in the ticketed forms           */
    char bvhForm[32];          /* tuning builds: pathTrace_bvh_<form>                                */
    uint32_t expMask;          /* bit v: the module has pathTrace_exp<v>                             */
};
void rtTuningFromEnvironment(RTTuning *t);

struct RTDeviceShape { int smCount; size_t maxSmemOptin, smemPerSM; };

struct RTKernelPlan {
    int form;                  /* RTKernelForm                                                       */
    int R;                     /* samples per pixel per round (queue depth); 0 where it has no meaning */
    int count;                 /* the counting twin of the form                                      */
    int expVariant;            /* RT_FORM_EXP: N                                                     */
    char kernel[48];           /* the __global__ function's name                                     */
    RTSceneLayout lay;
    size_t sceneBytes, extraBytes;   /* dynamic shared memory = sceneBytes + extraBytes (BVH / mode B: extraBytes only) */
    unsigned gridX, gridY;     /* blocks of 128 threads; the ticketed forms launch gridX * gridY * chunks */
    int ticketed;              /* RT_FORM_*TILES                                                     */
    int copyConst;             /* RT_FORM_CONST: the intersection part goes to the constant bank first */
    int mis;                   /* RT_VARIANT_MIS: the form's two-strategy twin (kernel name ends in _mis)   */
};

/* p: the command's parameters with the counts resolved (numTriangles, numLights != 0).  kind: 0 pathTrace, 1 drawTriangle.
 * Returns RT_OK or an RT_ERR_* with the text in err. */
/* layoutFlags: RT_LAYOUT_* the caller knows beyond p->variant (RT_LAYOUT_LIGHT_NORMALS when a normal table is bound). */
int rtPlanKernel(const RTLaunchParams *p, int kind, const RTTuning *tune, const RTDeviceShape *dev, size_t constBankBytes,
                 uint32_t layoutFlags, RTKernelPlan *out, char *err, size_t errCap);

/* how the ticketed forms cut the samples into chunks: blockTiles x chunks units for `resident` block slots */
void rtPlanChunks(const RTLaunchParams *p, const RTKernelPlan *plan, const RTTuning *tune, unsigned long long resident,
                  unsigned *chunks, unsigned *chunkSpp);
