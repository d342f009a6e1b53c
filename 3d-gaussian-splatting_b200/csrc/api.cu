// api.cu — error reporting, camera block decoding, stage 5 (densification statistics) and version.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

static thread_local char g_err[512] = "";

void b200_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int b200_check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        b200_set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
        return B200_ECUDA;
    }
    return B200_OK;
}

Cam b200_make_cam(const float* c, int W, int H) {
    Cam cam;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) cam.R[i * 3 + j] = c[i * 4 + j];
        cam.t[i] = c[i * 4 + 3];
    }
    for (int j = 0; j < 3; ++j)
        cam.campos[j] = -(cam.R[0 * 3 + j] * cam.t[0] + cam.R[1 * 3 + j] * cam.t[1] + cam.R[2 * 3 + j] * cam.t[2]);
    cam.fx = c[16]; cam.fy = c[17]; cam.cx = c[18]; cam.cy = c[19];
    cam.near_plane = c[20]; cam.far_plane = c[21]; cam.eps2d = c[22]; cam.radius_clip = c[23];
    cam.W = W; cam.H = H;
    cam.tile_w = (W + TILE - 1) / TILE;
    cam.tile_h = (H + TILE - 1) / TILE;
    return cam;
}

extern "C" int b200_version(void) { return 200; }  // round 2: coarse binning, auxiliary projection outputs
extern "C" const char* b200_last_error(void) { return g_err; }

// accum[N,3] += ||g||_2 ; count[N] += 1   (reference src/training/trainer.py:786-791)
__global__ void __launch_bounds__(256)
densify_stats_kernel(const float* __restrict__ grad, int64_t N, float* __restrict__ accum, float* __restrict__ count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float gx = grad[3 * i], gy = grad[3 * i + 1], gz = grad[3 * i + 2];
    float n = sqrtf(gx * gx + gy * gy + gz * gz);
    accum[3 * i] += n; accum[3 * i + 1] += n; accum[3 * i + 2] += n;
    count[i] += 1.f;
}

extern "C" int b200_densify_stats(const float* xyz_grad, int64_t N, float* accum, float* count, void* stream) {
    B200_REQUIRE(N >= 0, "bad size");
    if (N == 0) return B200_OK;
    B200_REQUIRE(xyz_grad && accum && count, "null pointer");
    densify_stats_kernel<<<(unsigned)((N + 255) / 256), 256, 0, (cudaStream_t)stream>>>(xyz_grad, N, accum, count);
    return b200_check_launch("densify_stats");
}
