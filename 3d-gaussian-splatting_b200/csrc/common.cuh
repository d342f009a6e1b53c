// common.cuh — shared definitions for the sm_100a splatting kernels (see include/b200splat.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/b200splat.h"

#define ALPHA_THRESHOLD (1.0f / 255.0f)
#define MAX_ALPHA 0.999f
#define T_STOP 1e-4f
#define TILE B200_TILE
#define SPLAT_F B200_SPLAT_FLOATS

// Camera constants, passed by value in kernel parameter space.
struct Cam {
    float R[9];      // world->camera rotation, row-major
    float t[3];
    float campos[3]; // -R^T t
    float fx, fy, cx, cy;
    float near_plane, far_plane, eps2d, radius_clip;
    int W, H, tile_w, tile_h;
};

void b200_set_error(const char* fmt, ...);
int b200_check_launch(const char* what);
Cam b200_make_cam(const float* cam_host, int W, int H);

// cudaFuncSetAttribute is per DEVICE: a process that renders on several GPUs must set it on each.
// Returns true exactly once per (call site's flag array, current device).
#define B200_MAX_DEVICES 64
static inline bool b200_first_use_on_device(bool* flags) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= B200_MAX_DEVICES) return true;  // unknown slot: set the attribute every time (harmless)
    if (flags[dev]) return false;
    flags[dev] = true;
    return true;
}

#define B200_REQUIRE(cond, msg)                         \
    do {                                                \
        if (!(cond)) {                                  \
            b200_set_error("%s: %s", __func__, msg);    \
            return B200_EARG;                           \
        }                                               \
    } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
