/*
 * rt_plan.cpp -- the kernel choice of rtCommit as a rule table (rt_plan.h).  Pure host logic: no CUDA call.
 *
 * Every pathTrace form renders the same image bit for bit (tests/test_gpu_parity.py runs them against each other and
 * the oracle); the choice is about speed only, and each row of the table cites the measurement it rests on
 * (profiles/).
 */
#include "rt_plan.h"
#include "rt_kernel_smem.h"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

void rtTuningFromEnvironment(RTTuning *t) {
    memset(t, 0, sizeof *t);
    t->sortedR = 6;
    if (const char *e = getenv("RT_SORTED_R")) { const int r = atoi(e); if (r == 4 || r == 5 || r == 6 || r == 8) t->sortedR = r; }
    auto flag = [](const char *name) { const char *e = getenv(name); return e && *e && *e != '0' ? 1 : 0; };
    t->noSorted = flag("RT_NO_SORTED");
    t->noSortedBlock = flag("RT_NO_SORTED_BLOCK");
    t->alwaysPrep = flag("RT_ALWAYS_PREP");
    t->noHaltonTable = flag("RT_NO_HALTON_TABLE");
    if (const char *e = getenv("RT_TILES_CHUNKS")) t->tilesChunks = atoi(e) > 0 ? atoi(e) : 0;
    if (const char *e = getenv("RT_BVH_FORM")) snprintf(t->bvhForm, sizeof t->bvhForm, "%s", e);
}

static int planFail(char *err, size_t cap, int code, const char *fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

/* ---- the facts a rule may ask about ------------------------------------------------------------------------------ */
enum : uint32_t {
    IS_PT         = 1u << 0,    /* the pipeline is pathTrace (else drawTriangle)                                        */
    IS_COUNT      = 1u << 1,    /* RT_VARIANT_COUNT_RAYS: the exact work counters are wanted                             */
    USE_BVH       = 1u << 2,    /* more triangles than shared memory takes, or the BVH asked for by name                 */
    WAVE_FITS     = 1u << 3,    /* scene + wave queues fit a block's shared memory                                       */
    CONST_FITS    = 1u << 4,    /* the intersection part fits the constant bank                                          */
    SPP_GE_R      = 1u << 5,    /* a round of the compacting kernel (R samples per pixel) can be filled                  */
    SPP_GE_4      = 1u << 6,
    SPP_GE_64     = 1u << 7,    /* enough samples to cut into chunks of >= 32                                            */
    FEW_BLOCKS    = 1u << 8,    /* the launch has fewer than 6 (sortable scene: 8) resident waves of blocks              */
    R_IS_6        = 1u << 9,    /* four R = 6 blocks fit an SM beside the scene (else R = 4)                             */
    GROUPABLE     = 1u << 10,   /* quads, T <= 64, no spheres, one light: rt_prep_scene may find object groups           */
    SORTED_OK     = 1u << 11,   /* GROUPABLE, spp >= the sorted kernel's R, its slots fit                                */
    SORTED6_FITS  = 1u << 12,
    BLK_FITS      = 1u << 13,   /* the block-wide sort's slots fit                                                       */
    SPP_GE_6      = 1u << 14,
    BLKWIDE_FITS  = 1u << 15,
    NO_BLK        = 1u << 16,   /* tuning: keep to the per-warp sort                                                     */
    EXP_EXISTS    = 1u << 17,   /* the module has pathTrace_exp<request>                                                 */
};

struct Rule {
    int request;        /* RT_VARIANT_* asked for (low byte of RTLaunchParams.variant), -1 = any */
    uint32_t need, forbid;
    int form;           /* RTKernelForm, or -1: "unknown kernel variant"                          */
    const char *why;
};

/* First matching row wins. */
static const Rule RULES[] = {
    { -1, USE_BVH, 0, RT_FORM_BVH, "T > 1024 (or asked for): the scene lives in global memory behind a BVH, bvh.cuh" },
    { -1, 0, IS_PT, RT_FORM_MODEB, "drawTriangle has one kernel" },

    /* ---- RT_VARIANT_DEFAULT ---- */
    { RT_VARIANT_DEFAULT, 0, WAVE_FITS, RT_FORM_SMEM, "no room for path queues beside the scene" },
    { RT_VARIANT_DEFAULT, IS_COUNT | SPP_GE_4, 0, RT_FORM_WAVE, "counting twin of the compacting kernel" },
    { RT_VARIANT_DEFAULT, IS_COUNT, 0, RT_FORM_SMEM, "counting twin of the per-pixel kernel" },
    { RT_VARIANT_DEFAULT, R_IS_6 | GROUPABLE | BLKWIDE_FITS, SPP_GE_R, RT_FORM_SORTEDBLK_WIDE,
      "fewer samples than a round has slots (progressive rendering): 6 / spp tiles per warp, sorted; config 5 0.469 -> 0.350 ms per frame (r03c)" },
    { RT_VARIANT_DEFAULT, 0, SPP_GE_R, RT_FORM_WIDE, "the same for scenes that cannot be grouped; 1 spp frame 0.625 -> 0.522 ms (r02a)" },
    { RT_VARIANT_DEFAULT, SPP_GE_64 | FEW_BLOCKS | SORTED_OK | R_IS_6 | SORTED6_FITS | BLK_FITS, NO_BLK, RT_FORM_SORTEDBLK_TILES,
      "one rank's share of a multi-GPU frame: ticketed (tile, sample chunk) units keep the last wave full; 1080p / 8: 52.1 -> 49.8 ms (r01x), sorted form ahead to 6.9 waves (r03a)" },
    { RT_VARIANT_DEFAULT, SPP_GE_64 | FEW_BLOCKS | SORTED_OK | R_IS_6 | SORTED6_FITS, 0, RT_FORM_SORTED_TILES, "the same with the per-warp sort" },
    { RT_VARIANT_DEFAULT, SPP_GE_64 | FEW_BLOCKS, 0, RT_FORM_TILES, "the same for scenes that cannot be grouped" },
    { RT_VARIANT_DEFAULT, SORTED_OK | BLK_FITS | SPP_GE_6, NO_BLK, RT_FORM_SORTEDBLK,
      "rays sorted by the object groups they can reach, block-wide: 23.64 (wave) -> 18.55 ms at 64 spp (r03b), 15.99 with the Halton tables and the shell rule (r04d)" },
    { RT_VARIANT_DEFAULT, SORTED_OK, 0, RT_FORM_SORTED, "per-warp sort: 18.93 ms (r03a)" },
    { RT_VARIANT_DEFAULT, 0, 0, RT_FORM_WAVE, "warp-level path compaction: 38.8 -> 35.2 ms over the per-pixel kernel (r01j)" },

    /* ---- forms asked for by name (profiling, tests) ---- */
    { RT_VARIANT_SMEM, 0, 0, RT_FORM_SMEM, "" },
    { RT_VARIANT_CONST, CONST_FITS, 0, RT_FORM_CONST, "" },
    { RT_VARIANT_CONST, 0, 0, RT_FORM_SMEM, "intersection part larger than the constant bank" },
    { RT_VARIANT_WAVE, 0, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_TILES, IS_COUNT, 0, RT_FORM_WAVE, "the ticketed forms have no counting twin" },
    { RT_VARIANT_TILES, 0, 0, RT_FORM_TILES, "" },
    { RT_VARIANT_WIDE, IS_COUNT, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_WIDE, SPP_GE_R, 0, RT_FORM_WAVE, "the wide form is for spp < R only" },
    { RT_VARIANT_WIDE, 0, 0, RT_FORM_WIDE, "" },
    { RT_VARIANT_SORTED, IS_COUNT, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_SORTED, SORTED_OK, 0, RT_FORM_SORTED, "" },
    { RT_VARIANT_SORTED, 0, 0, RT_FORM_WAVE, "scene cannot be grouped" },
    { RT_VARIANT_SORTED_BLOCK, IS_COUNT, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_SORTED_BLOCK, SORTED_OK | BLK_FITS | SPP_GE_6, 0, RT_FORM_SORTEDBLK, "" },
    { RT_VARIANT_SORTED_BLOCK, SORTED_OK, 0, RT_FORM_SORTED, "" },
    { RT_VARIANT_SORTED_BLOCK, 0, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_SORTED_TILES, IS_COUNT, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_SORTED_TILES, SORTED_OK | R_IS_6 | SORTED6_FITS, 0, RT_FORM_SORTED_TILES, "" },
    { RT_VARIANT_SORTED_TILES, 0, 0, RT_FORM_TILES, "" },
    { RT_VARIANT_SORTED_BLOCK_TILES, IS_COUNT, 0, RT_FORM_WAVE, "" },
    { RT_VARIANT_SORTED_BLOCK_TILES, SORTED_OK | R_IS_6 | SORTED6_FITS | BLK_FITS, 0, RT_FORM_SORTEDBLK_TILES, "" },
    { RT_VARIANT_SORTED_BLOCK_TILES, SORTED_OK | R_IS_6 | SORTED6_FITS, 0, RT_FORM_SORTED_TILES, "" },
    { RT_VARIANT_SORTED_BLOCK_TILES, 0, 0, RT_FORM_TILES, "" },
    { RT_VARIANT_BVH, 0, 0, -1, "the BVH kernel takes triangles only, no counting twin" },
};

int rtPlanKernel(const RTLaunchParams *pp, int kind, const RTTuning *tune, const RTDeviceShape *dev, size_t constBankBytes,
                 uint32_t layoutFlags, RTKernelPlan *out, char *err, size_t errCap) {
    const RTLaunchParams &p = *pp;
    memset(out, 0, sizeof *out);
    const uint32_t request = p.variant & 0xffu;
    const bool count = (p.variant & RT_VARIANT_COUNT_RAYS) != 0;
    const uint32_t T = p.numTriangles, S = p.numSpheres, L = p.numLights;

    const bool useBvh = kind == 0 && S == 0 && !count && T >= 1 && T <= RT_MAX_TRIANGLES_BVH &&
                        (T > RT_MAX_TRIANGLES || request == RT_VARIANT_BVH);
    if ((T > RT_MAX_TRIANGLES && !useBvh) || S > RT_MAX_SPHERES || L > RT_MAX_LIGHTS)
        return planFail(err, errCap, RT_ERR_UNSUPPORTED, "scene too large for the brute-force kernel (T<=%d, S<=%d, L<=%d)",
                        RT_MAX_TRIANGLES, RT_MAX_SPHERES, RT_MAX_LIGHTS);

    RTSceneLayout lay = rtSceneLayout(T, S, L);
    if (p.variant & RT_VARIANT_LIGHT_EXTENT) lay.flags |= RT_LAYOUT_LIGHT_EXTENT;
    lay.flags |= layoutFlags & RT_LAYOUT_LIGHT_NORMALS;
    out->lay = lay;
    const bool mis = kind == 0 && (p.variant & RT_VARIANT_MIS) != 0;
    if (mis && count) return planFail(err, errCap, RT_ERR_UNSUPPORTED, "RT_VARIANT_MIS has no counting twin (COUNT_RAYS)");
    const size_t sceneBytes = (size_t)lay.words * 4, room = dev->maxSmemOptin;
    out->sceneBytes = sceneBytes;
    if (!useBvh && sceneBytes + 1024 > room)
        return planFail(err, errCap, RT_ERR_UNSUPPORTED, "prepared scene (%zu bytes) does not fit shared memory (%zu)", sceneBytes, room);
    auto fits = [&](size_t extra) { return sceneBytes + extra + 1024 <= room; };

    /* queue depth of the compacting kernel: R = 6 when four such blocks fit an SM beside the scene (5 617 against 5 392
     * Mpaths/s for R = 8 at three, profiles/r02c_variants_R.txt), else R = 4 (the 256-sphere scene) */
    const size_t wave6 = (size_t)rtWaveWords(6, T, S) * 4, wave4 = (size_t)rtWaveWords(4, T, S) * 4;
    const int waveR = dev->smemPerSM / (sceneBytes + wave6 + 1024) >= 4 ? 6 : 4;
    const size_t waveBytes = waveR == 6 ? wave6 : wave4;
    const int sortedR = tune->sortedR;
    const size_t sortedBytes = (size_t)rtSortedWords((uint32_t)sortedR, T) * 4, sorted6 = (size_t)rtSortedWords(6, T) * 4;
    const size_t blkBytes = (size_t)rtSortedBlkWords(6, false, T) * 4, blkWideBytes = (size_t)rtSortedBlkWords(6, true, T) * 4;

    /* (a light with a normal of its own is not for the sorted kernels: they keep x and z of the light sample point only) */
    const bool groupable = kind == 0 && T >= 2 && T <= 64 && T % 2 == 0 && S == 0 && L == 1 && !tune->noSorted &&
                           !(lay.flags & RT_LAYOUT_LIGHT_NORMALS);
    const bool sortedOk = groupable && p.spp >= (uint32_t)sortedR && fits(sortedBytes);
    const unsigned gx = (unsigned)((p.width + 15) / 16), gy = (unsigned)((p.rows + 7) / 8);
    const unsigned long long blocks = (unsigned long long)gx * gy;
    const unsigned long long resident = (unsigned long long)(dev->smemPerSM / (sceneBytes + waveBytes + 1024)) * (unsigned)dev->smCount;

    uint32_t facts = 0;
    if (kind == 0) facts |= IS_PT;
    if (count) facts |= IS_COUNT;
    if (useBvh) facts |= USE_BVH;
    if (kind == 0 && fits(waveBytes)) facts |= WAVE_FITS;
    if ((size_t)(lay.shade - lay.isect) * 4 <= constBankBytes) facts |= CONST_FITS;
    if (p.spp >= (uint32_t)waveR) facts |= SPP_GE_R;
    if (p.spp >= 4) facts |= SPP_GE_4;
    if (p.spp >= 6) facts |= SPP_GE_6;
    if (p.spp >= 64) facts |= SPP_GE_64;
    if (blocks < (sortedOk ? 8ull : 6ull) * resident) facts |= FEW_BLOCKS;
    if (waveR == 6) facts |= R_IS_6;
    if (groupable) facts |= GROUPABLE;
    if (sortedOk) facts |= SORTED_OK;
    if (fits(sorted6)) facts |= SORTED6_FITS;
    if (fits(blkBytes)) facts |= BLK_FITS;
    if (fits(blkWideBytes)) facts |= BLKWIDE_FITS;
    if (tune->noSortedBlock) facts |= NO_BLK;
    const bool isExp = request >= 4 && request < 16;
    if (isExp && ((tune->expMask >> request) & 1u) && !count) facts |= EXP_EXISTS;

    int form = -1;
    if (isExp && !(facts & USE_BVH) && (facts & IS_PT)) {
        form = (facts & EXP_EXISTS) ? RT_FORM_EXP : -1;
    } else {
        for (const Rule &r : RULES) {
            if (r.request != -1 && (uint32_t)r.request != request) continue;
            if ((facts & r.need) != r.need || (facts & r.forbid)) continue;
            form = r.form;
            break;
        }
    }
    if (form < 0) return planFail(err, errCap, RT_ERR_INVALID_ARGUMENT, "unknown kernel variant %u", p.variant);
    if (mis) {
        /* RT_VARIANT_MIS (mis.cuh) exists as the per-pixel, the compacting, the block-sorted and the BVH kernel; every other
         * form falls to the nearest of those (the image is the same whichever runs) */
        switch (form) {
        case RT_FORM_SMEM: case RT_FORM_CONST: form = RT_FORM_SMEM; break;
        case RT_FORM_SORTEDBLK: case RT_FORM_SORTEDBLK_TILES: break;
        case RT_FORM_SORTEDBLK_WIDE: case RT_FORM_BVH: break;
        case RT_FORM_EXP: return planFail(err, errCap, RT_ERR_UNSUPPORTED, "tuning variants have no MIS twin");
        default: form = (facts & WAVE_FITS) ? RT_FORM_WAVE : RT_FORM_SMEM; break;
        }
        if (form == RT_FORM_SORTEDBLK_TILES) form = RT_FORM_SORTEDBLK;
        if (form == RT_FORM_SORTEDBLK_WIDE) form = RT_FORM_WAVE;            /* spp < 6: the compacting kernel takes any sample count */
        out->mis = 1;
    }

    out->form = form; out->count = count ? 1 : 0; out->gridX = gx; out->gridY = gy;
    const char *cnt = count ? "_count" : "";
    switch (form) {
    case RT_FORM_SMEM:  snprintf(out->kernel, sizeof out->kernel, "pathTrace_smem%s", cnt); break;
    case RT_FORM_CONST: snprintf(out->kernel, sizeof out->kernel, "pathTrace_const%s", cnt); out->copyConst = 1; break;
    case RT_FORM_WAVE:
        out->R = waveR; out->extraBytes = waveBytes;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_wave%s%s", waveR == 6 ? "" : "4", cnt);
        break;
    case RT_FORM_TILES:
        out->R = waveR; out->extraBytes = waveBytes; out->ticketed = 1;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_tiles%s", waveR == 6 ? "" : "4");
        break;
    case RT_FORM_WIDE:
        out->R = waveR; out->extraBytes = waveBytes;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_wide%s", waveR == 6 ? "" : "4");
        break;
    case RT_FORM_SORTED:
        out->R = sortedR; out->extraBytes = sortedBytes;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_sorted%d", sortedR);
        break;
    case RT_FORM_SORTED_TILES:
        out->R = 6; out->extraBytes = sorted6; out->ticketed = 1;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_sorted_tiles6");
        break;
    case RT_FORM_SORTEDBLK:
        out->R = 6; out->extraBytes = blkBytes;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_sortedblk6");
        break;
    case RT_FORM_SORTEDBLK_TILES:
        out->R = 6; out->extraBytes = blkBytes; out->ticketed = 1;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_sortedblk_tiles6");
        break;
    case RT_FORM_SORTEDBLK_WIDE:
        out->R = 6; out->extraBytes = blkWideBytes;
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_sortedblk_wide6");
        break;
    case RT_FORM_EXP:
        out->expVariant = (int)request;
        /* tools/build_exp.sh -DRT_TUNING_VARIANTS: exp14 / exp15 = sortedblk with R = 4 / 5; exp<R> = wave with queue depth R (exp10: 2) */
        if (request >= 14) { out->R = (int)request - 10; out->extraBytes = (size_t)rtSortedBlkWords((uint32_t)out->R, false, T) * 4; }
        else { out->R = request == 10 ? 2 : (int)request; out->extraBytes = (size_t)rtWaveWords((uint32_t)out->R, T, S) * 4; }
        snprintf(out->kernel, sizeof out->kernel, "pathTrace_exp%u", request);
        break;
    case RT_FORM_BVH:
        /* header + lights only; the geometry stays in global memory */
        out->sceneBytes = 0; out->extraBytes = (size_t)(RT_HDR_WORDS + RT_LIGHT_WORDS * L) * 4;
        if (tune->bvhForm[0]) snprintf(out->kernel, sizeof out->kernel, "pathTrace_bvh_%s", tune->bvhForm);
        else snprintf(out->kernel, sizeof out->kernel, "pathTrace_bvh");
        break;
    case RT_FORM_MODEB:
        if (p.cameraRays == 0 || p.misSamples < 3)
            return planFail(err, errCap, RT_ERR_INVALID_ARGUMENT, "drawTriangle needs cameraRays >= 1 and misSamples >= 3");
        if (S) return planFail(err, errCap, RT_ERR_UNSUPPORTED, "drawTriangle has no sphere primitives");
        out->extraBytes = (size_t)5 * (p.misSamples / 3u) * 8;      /* five Halton streams, float2 each */
        if (!fits(out->extraBytes))
            return planFail(err, errCap, RT_ERR_UNSUPPORTED, "misSamples = %u: Halton tables do not fit shared memory", p.misSamples);
        snprintf(out->kernel, sizeof out->kernel, "drawTriangle_mis");
        break;
    }
    if (out->mis) {
        const size_t n = strlen(out->kernel);
        snprintf(out->kernel + n, sizeof out->kernel - n, "_mis");
    }
    if (form == RT_FORM_WIDE || form == RT_FORM_SORTEDBLK_WIDE) {       /* a block covers 16 G x 8 pixels, G = R / spp (wave.cuh) */
        const unsigned G = (unsigned)out->R / (p.spp ? p.spp : 1u);
        if (G > 1) out->gridX = (unsigned)((p.width + 16 * (int)G - 1) / (16 * (int)G));
    }
    if (form != RT_FORM_BVH && !fits(out->extraBytes))
        return planFail(err, errCap, RT_ERR_UNSUPPORTED, "scene + path queues (%zu bytes) do not fit shared memory", sceneBytes + out->extraBytes);
    return RT_OK;
}

/* The samples are cut into chunks only as far as needed to give every resident block slot ~24 units, so that the last
 * wave of units is a small part of the frame, and a chunk keeps >= 32 samples (wave.cuh: RTTileSched). */
void rtPlanChunks(const RTLaunchParams *p, const RTKernelPlan *plan, const RTTuning *tune, unsigned long long resident,
                  unsigned *chunksOut, unsigned *chunkSppOut) {
    const unsigned long long blockTiles = (unsigned long long)plan->gridX * plan->gridY;
    const unsigned R = (unsigned)(plan->R > 0 ? plan->R : 1);
    unsigned chunks = 1;
    if (tune->tilesChunks > 0) chunks = (unsigned)tune->tilesChunks;
    else while (blockTiles * chunks < 24ull * resident && p->spp / (chunks * 2) >= 32) chunks *= 2;
    const unsigned chunkSpp = ((p->spp + chunks - 1) / chunks + R - 1) / R * R;     /* whole rounds */
    *chunkSppOut = chunkSpp;
    *chunksOut = (p->spp + chunkSpp - 1) / chunkSpp;
}
