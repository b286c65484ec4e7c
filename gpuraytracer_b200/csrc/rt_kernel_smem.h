/*
 * rt_kernel_smem.h -- how much dynamic shared memory each pathTrace kernel form takes AFTER the prepared scene
 * (rt_scene_layout.h), in 32-bit words.  ONE statement of it, shared by the kernels (wave.cuh, sorted.cuh,
 * sortedblk.cuh carve their arrays out of these sizes) and by the host's kernel choice (rt_plan.cpp sizes the launch
 * with them), so that the two cannot drift.
 */
#ifndef RT_KERNEL_SMEM_H
#define RT_KERNEL_SMEM_H

#include <stdint.h>
#include "rt_scene_layout.h"

#define RT_BLOCK_WARPS   4      /* 128 threads: a block is 16 x 8 pixels, a warp 8 x 4                                  */
#define RT_WAVE_FIELDS  12      /* wave.cuh queue entry: hit point, colour, radiance (3 each), id, Halton index, tag    */
#define RT_WAVE_RES      3      /* result slots per path                                                              */
#define RT_SORT_FIELDS  11      /* sorted.cuh slot: X.xyz (hit point, then P)  D.xyz  colour.xyz  light sample x, z   */
#define RT_SORT_KEYS    (1 << RT_MAX_GROUPS)
#define RT_SORTBLK_KEYS (8 << RT_MAX_GROUPS)   /* sortedblk.cuh: the groups' bits above three bits of exit face (rt_scene_layout.h) */
/* ... except in the wide form (a block lives for ONE round there and four blocks fit an SM only without the 1.3 KB the
 * longer tables take: profiles/r04z_c5_*.csv, three blocks: 0.338 ms a frame against 0.326), which keeps the groups' keys */
#if defined(__CUDACC__)
__host__ __device__
#endif
constexpr uint32_t rtSortedBlkKeys(bool wide) { return wide ? (uint32_t)RT_SORT_KEYS : (uint32_t)RT_SORTBLK_KEYS; }

#if defined(__CUDACC__)
#define RT_CX __host__ __device__ constexpr
#else
#define RT_CX constexpr
#endif

/* per warp one candidate bit per primitive (camera_candidates), then the block's unit word of the ticketed forms
 * (padded to four words) */
RT_CX uint32_t rtCullWords(uint32_t T, uint32_t S) { return RT_BLOCK_WARPS * ((T + 31u) / 32u + (S + 31u) / 32u) + 4u; }

/* wave.cuh: per warp a queue of 32 R entries + their result slots */
RT_CX uint32_t rtWaveWordsPerWarp(uint32_t R) { return (RT_WAVE_FIELDS + RT_WAVE_RES) * 32u * R; }
RT_CX uint32_t rtWaveWords(uint32_t R, uint32_t T, uint32_t S) { return RT_BLOCK_WARPS * rtWaveWordsPerWarp(R) + rtCullWords(T, S); }

/* sorted.cuh: per warp 32 R slots of fields + res.xyz + four byte lists, the lanes' offsets, the key counters;
 * after the warps: candidate bits, unit word, the key -> quads table */
RT_CX uint32_t rtSortedWordsPerWarp(uint32_t R) { return (RT_SORT_FIELDS + 3u + 1u) * 32u * R + 32u + RT_SORT_KEYS; }
RT_CX uint32_t rtSortedWords(uint32_t R, uint32_t T) { return RT_BLOCK_WARPS * rtSortedWordsPerWarp(R) + rtCullWords(T, 0u) + RT_SORT_KEYS; }

/* sortedblk.cuh: 128 R slots, two 16-bit lists per slot, the lanes' offsets (the wide form: every slot's Halton index), key
 * counters (count, running, start), four block counters; then as above.  A slot is
 *   RT_SORTBLK_F4 = 0 (default)  RT_SORT_FIELDS + 3 separate words and two byte arrays (id, key): 14.5 words
 *   RT_SORTBLK_F4 = 1            three float4 -- (X.xyz, light x) (D.xyz, light z) (colour.xyz, id | key << 8) -- + res.xyz: 15 words,
 *                                a third of the shared-memory instructions.  MEASURED SLOWER (profiles/r04o_ab_slot_layout.log:
 *                                16.24 against 15.94 ms at 64 spp; the wide form falls to three blocks per SM, config 5
 *                                6 477 -> 5 582 Mpaths/s): scattered 128-bit accesses cost more wavefronts than they save
 *                                instructions.  Kept as a switch so that it is not built a second time. */
#ifndef RT_SORTBLK_F4
#define RT_SORTBLK_F4 0
#endif
RT_CX uint32_t rtSortedBlkSlots(uint32_t R) { return RT_BLOCK_WARPS * 32u * R; }
RT_CX uint32_t rtSortedBlkCoreWords(uint32_t R, bool wide) {
    return (RT_SORTBLK_F4 ? 15u * rtSortedBlkSlots(R) + (4u * rtSortedBlkSlots(R)) / 4u
                          : (RT_SORT_FIELDS + 3u) * rtSortedBlkSlots(R) + (6u * rtSortedBlkSlots(R)) / 4u) +
           (wide ? rtSortedBlkSlots(R) : RT_BLOCK_WARPS * 32u) + 3u * rtSortedBlkKeys(wide) + 4u;
}
/* (after the core: candidate bits, unit word, the key -> quads tables of the continuation ray and of the shadow ray) */
RT_CX uint32_t rtSortedBlkWords(uint32_t R, bool wide, uint32_t T) { return rtSortedBlkCoreWords(R, wide) + rtCullWords(T, 0u) + rtSortedBlkKeys(wide) + RT_SORT_KEYS; }

#endif
