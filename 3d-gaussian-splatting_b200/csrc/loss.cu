// loss.cu — SURVEY.md §8 f3: the reference's per-view photometric loss, forward AND gradient in one kernel.
//
// Reference (src/training/trainer.py:485-492 with _stclamp :102-107 and _compute_ssim :354-371):
//     x    = clamp(render, 0, 1)            (straight-through: gradient of the clamp is the identity)
//     l1   = mean |x - gt|
//     ssim = mean over pixels and channels of the 3x3 box-window SSIM (avg_pool2d(k=3, s=1, padding=1),
//            zero padding counted in the divisor 9), C1 = 0.01^2, C2 = 0.03^2
//     loss = (1 - lambda) l1 + lambda (1 - ssim)
// which costs the reference five avg_pool2d passes, a dozen elementwise kernels and their autograd mirror per
// view.  Here one CTA owns a 32x16 pixel tile: it stages render/gt with a 2-pixel halo in shared memory,
// evaluates the window statistics for the tile plus a 1-pixel ring, and — because the derivative of a window's
// SSIM w.r.t. one of its pixels p is (alpha_w + beta_w gt_p + gamma_w x_p) / 9 — obtains the gradient of every
// tile pixel from 3x3 sums of the three per-window coefficients.  HBM traffic: read render + gt, write
// v_render (36 B/pixel), once.
#include "common.cuh"

#define LT_X 32
#define LT_Y 16
#define LOSS_THREADS 256
#define IN_W ((LT_X + 4) * 3)   // floats per staged row (interleaved RGB)
#define IN_H (LT_Y + 4)
#define CO_W ((LT_X + 2) * 3)
#define CO_H (LT_Y + 2)

namespace {

__global__ void __launch_bounds__(LOSS_THREADS)
l1_ssim_kernel(const float* __restrict__ render, const float* __restrict__ gt, int W, int H, float lambda,
               float grad_scale, float* __restrict__ v_render, float* __restrict__ sums) {
    __shared__ float sX[IN_H][IN_W], sY[IN_H][IN_W];
    __shared__ float sA[CO_H][CO_W], sB[CO_H][CO_W], sG[CO_H][CO_W];
    __shared__ float s_red[3][LOSS_THREADS / 32];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LT_X, y0 = blockIdx.y * LT_Y;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;

    // ---- stage clamp(render) and gt with a 2-pixel halo, zeros outside the image (= avg_pool2d's padding)
    for (int e = tid; e < IN_H * IN_W; e += LOSS_THREADS) {
        const int r = e / IN_W, c = e - r * IN_W;
        const int iy = y0 - 2 + r, ix = x0 - 2 + c / 3;
        float xv = 0.f, yv = 0.f;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) {
            const size_t o = ((size_t)iy * W + (x0 - 2)) * 3 + c;
            xv = fminf(fmaxf(__ldg(render + o), 0.f), 1.f);
            yv = __ldg(gt + o);
        }
        sX[r][c] = xv;
        sY[r][c] = yv;
    }
    __syncthreads();

    // ---- window statistics for the tile plus a 1-pixel ring (window centres)
    float acc_ssim = 0.f;
    for (int e = tid; e < CO_H * CO_W; e += LOSS_THREADS) {
        const int r = e / CO_W, c = e - r * CO_W;
        const int iy = y0 - 1 + r, ix = x0 - 1 + c / 3;
        float a = 0.f, b = 0.f, g = 0.f;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) {
            float sx = 0.f, sy = 0.f, sxx = 0.f, syy = 0.f, sxy = 0.f;
#pragma unroll
            for (int dr = 0; dr < 3; ++dr)
#pragma unroll
                for (int dc = 0; dc < 3; ++dc) {
                    const float xv = sX[r + dr][c + 3 * dc], yv = sY[r + dr][c + 3 * dc];
                    sx += xv; sy += yv;
                    sxx = fmaf(xv, xv, sxx); syy = fmaf(yv, yv, syy); sxy = fmaf(xv, yv, sxy);
                }
            const float k = 1.f / 9.f;
            const float mu1 = sx * k, mu2 = sy * k;
            const float s11 = sxx * k - mu1 * mu1, s22 = syy * k - mu2 * mu2, s12 = sxy * k - mu1 * mu2;
            const float A = 2.f * mu1 * mu2 + C1, Bq = 2.f * s12 + C2;
            const float Cq = mu1 * mu1 + mu2 * mu2 + C1, Dq = s11 + s22 + C2;
            const float rCD = 1.f / (Cq * Dq);
            const float ssim = A * Bq * rCD;
            // d ssim / d x_p = (alpha + beta * y_p + gamma * x_p) / 9 for every pixel p of the window
            b = 2.f * A * rCD;
            g = -2.f * ssim / Dq;
            a = 2.f * mu2 * (Bq - A) * rCD - 2.f * mu1 * ssim / Cq - mu1 * g;
            const bool interior = r >= 1 && r <= LT_Y && c >= 3 && c < 3 + LT_X * 3;
            if (interior) acc_ssim += ssim;  // every in-image window is interior to exactly one CTA
        }
        sA[r][c] = a; sB[r][c] = b; sG[r][c] = g;
    }
    __syncthreads();

    // ---- per-pixel loss terms and gradient for the tile
    float acc_l1 = 0.f, acc_sq = 0.f;
    const float n_inv = 1.f / (3.f * (float)W * (float)H);
    const float w_l1 = (1.f - lambda) * n_inv * grad_scale, w_ss = lambda * n_inv * grad_scale * (1.f / 9.f);
    for (int e = tid; e < LT_Y * LT_X * 3; e += LOSS_THREADS) {
        const int r = e / (LT_X * 3), c = e - r * (LT_X * 3);
        const int iy = y0 + r, ix = x0 + c / 3;
        if (iy < H && ix < W) {
            const float xv = sX[r + 2][c + 6], yv = sY[r + 2][c + 6];
            const float d = xv - yv;
            acc_l1 += fabsf(d);
            acc_sq = fmaf(d, d, acc_sq);
            if (v_render != nullptr) {
                float a = 0.f, b = 0.f, g = 0.f;
#pragma unroll
                for (int dr = 0; dr < 3; ++dr)
#pragma unroll
                    for (int dc = 0; dc < 3; ++dc) {
                        a += sA[r + dr][c + 3 * dc]; b += sB[r + dr][c + 3 * dc]; g += sG[r + dr][c + 3 * dc];
                    }
                const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
                v_render[((size_t)iy * W + x0) * 3 + c] = w_l1 * sgn - w_ss * (a + b * yv + g * xv);
            }
        }
    }
    // ---- block reduction -> three atomics
    float v[3] = {acc_l1, acc_ssim, acc_sq};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        v[i] = warp_sum(v[i]);
        if ((tid & 31) == 0) s_red[i][tid >> 5] = v[i];
    }
    __syncthreads();
    if (tid < 3) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < LOSS_THREADS / 32; ++w) t += s_red[tid][w];
        atomicAdd(sums + tid, t);
    }
    // ---- the last CTA to arrive turns the three sums into (loss, l1, ssim, mse): no scalar kernels afterwards
    __shared__ bool s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned arrived = atomicAdd(reinterpret_cast<unsigned*>(sums + 3), 1u);
        s_last = arrived == gridDim.x * gridDim.y - 1;
    }
    __syncthreads();
    if (s_last && tid == 0) {
        __threadfence();
        const float l1 = __ldcg(sums + 0) * n_inv, ssim = __ldcg(sums + 1) * n_inv, mse = __ldcg(sums + 2) * n_inv;
        sums[4] = (1.f - lambda) * l1 + lambda * (1.f - ssim);
        sums[5] = l1;
        sums[6] = ssim;
        sums[7] = mse;
    }
}

}  // namespace

// render[H,W,3], gt[H,W,3] -> sums[8]: [0..2] = (sum |x-gt|, sum of window SSIMs, sum (x-gt)^2) over n = 3*H*W
// elements, [3] = arrival counter (both cleared by this call), [4..7] = (loss, l1, ssim, mse) with
// loss = (1-lambda) * sums[0]/n + lambda * (1 - sums[1]/n), written by the last CTA; and, if v_render != NULL,
// v_render[H,W,3] = grad_scale * d loss / d render.
extern "C" int b200_l1_ssim_fwd_bwd(const float* render, const float* gt, int width, int height, float lambda_dssim,
                                    float grad_scale, float* v_render, float* sums, void* stream) {
    B200_REQUIRE(width > 0 && height > 0 && render && gt && sums, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    cudaMemsetAsync(sums, 0, 4 * sizeof(float), st);
    dim3 grid((width + LT_X - 1) / LT_X, (height + LT_Y - 1) / LT_Y);
    l1_ssim_kernel<<<grid, LOSS_THREADS, 0, st>>>(render, gt, width, height, lambda_dssim, grad_scale, v_render, sums);
    return b200_check_launch("l1_ssim_fwd_bwd");
}
