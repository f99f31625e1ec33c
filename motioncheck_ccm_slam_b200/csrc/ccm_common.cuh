// Shared device/host declarations for the sm_100a ORB front-end kernels.
// Data layout in HBM: for every pyramid level l one array [batch][h_l][pitch_l] of u8
// (pitch_l = w_l rounded up to 32 B so every row starts on a sector boundary), twice: the level itself and its
// 7x7-blurred copy. Per (frame, level): a packed FAST-corner list (u32: x | y<<12 | score<<24), a u16 quadtree
// node id per corner, and a fixed-capacity slot list of selected keypoints. No 19-px apron is stored: no stage
// reads it (SURVEY §8a note 1); ccm_orb_get_pyramid_level() synthesises it on request.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define CCM_MAX_LEVELS 16
#define CCM_EDGE 19          // EDGE_THRESHOLD, cslam/src/ORBextractor.cpp:65
#define CCM_HALF_PATCH 15    // HALF_PATCH_SIZE, :64
#define CCM_MIN_BORDER 16    // EDGE_THRESHOLD-3, :941

struct LevelDesc {
    uint8_t* img;            // level image, frame 0
    uint8_t* blur;           // blurred level, frame 0
    size_t img_stride;       // bytes between consecutive frames of img
    size_t blur_stride;      // bytes between consecutive frames of blur
    int img_pitch, blur_pitch;
    int w, h;
    // cell grid of ComputeKeyPointsOctTree (:941-955)
    int nCols, nRows, wCell, hCell;
    int cell_base;           // first cell id of this level in the flattened (all levels) cell list
    int blur_tile_base;      // first blur tile (128 x 128 px) of this level in the flattened tile list
    int blur_tiles_x;
    // quadtree (:707-726)
    int quota;               // mnFeaturesPerLevel[l]
    int nIni;
    float hX;
    int node_cap;            // max list length: max(quota+3, 4*nIni)
    // per-frame scratch offsets
    int corner_off;          // into the per-frame corner array (u32 units)
    int corner_cap;
    int kp_off;              // into the per-frame level-keypoint slot list
    int desc_block_base;     // first CTA of this level in k_orient_describe's grid (8 slots per CTA, a CTA never spans two levels)
    float scale;             // mvScaleFactor[l]
    float kp_size;           // (float)(int)(31*mvScaleFactor[l]), :1006
    // resize tables for (l-1) -> l, device pointers (null for level 0)
    const int* xofs; const short* xcoef;   // w entries, 2 coefs each
    const int* yofs; const short* ycoef;   // h entries
    const uint4* xgrp;                     // per group of 4 outputs, 2 x uint4 (k_resize_level); null: use the generic kernel
    int resize_mlp;                        // 1: every 4-row strip taps rows s..s+5 with row j starting at s+j or s+j+1 (k_resize_level_mlp)
};

struct LevelTable {
    LevelDesc l[CCM_MAX_LEVELS];
    int nlevels;
    int total_cells;
    int total_blur_tiles;
    int total_desc_blocks;
    int corners_per_frame;   // u32 units
    int kps_per_frame;       // slot-list length per frame (sum of node_cap)
    int ini_th, min_th;
};

struct ccm_keypoint_dev { float x, y, size, angle, response; int octave; };   // == ccm_keypoint (include/ccm_orb.h)

__device__ __forceinline__ int ccm_reflect101(int p, int len) {
    // valid for -len < p < 2*len-1, which covers the 3-px blur halo and the 19-px apron on every supported size
    if (p < 0) p = -p;
    if (p >= len) p = 2 * (len - 1) - p;
    return p;
}

// Programmatic dependent launch (sm_90+): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// while its predecessor on the stream is still draining; pdl_wait() blocks until that predecessor has completed and its writes
// are visible, pdl_launch() lets the successor's CTAs take the SM slots this grid no longer needs. Both are no-ops for a
// normal launch. Everything a kernel reads before pdl_wait() must be data no earlier kernel of the chain writes (host tables).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Debug build only (-DCCM_DEBUG_TS): the first thread of a kernel's first CTA notes globaltimer when it is past its dependency
// wait and the last thread to leave... is not known, so only starts are kept; slot = kind * 16 + level.
#ifdef CCM_DEBUG_TS
__device__ unsigned long long g_dbg_ts[128];
__device__ __forceinline__ void dbg_ts(int slot) {
    unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_dbg_ts[slot] = t;
}
#define CCM_TS(cond, slot) do { if (cond) dbg_ts(slot); } while (0)
#else
#define CCM_TS(cond, slot) do { } while (0)
#endif
