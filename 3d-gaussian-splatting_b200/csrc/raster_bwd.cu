// raster_bwd.cu — stage 4a: reverse replay of the blend and accumulation of the splat-record
// cotangents (gsplat rasterize_to_pixels backward, triggered by the reference at
// src/training/trainer.py:505-511).  One CTA per tile; batches are walked back to front from the
// tile's deepest composited index; each Gaussian's ten partial sums are reduced across the warp
// with shuffles and then added to the gradient record with one RED per warp and slot.
#include "common.cuh"

#define RB_THREADS (TILE * TILE)

namespace {

template <int CHSET> struct ChInfo;
template <> struct ChInfo<B200_CH_RGB>  { static constexpr int CH = 3; };
template <> struct ChInfo<B200_CH_D>    { static constexpr int CH = 1; };
template <> struct ChInfo<B200_CH_RGBD> { static constexpr int CH = 4; };

struct Bg { float v[4]; };

template <int CHSET>
__global__ void __launch_bounds__(RB_THREADS)
raster_bwd_kernel(const float4* __restrict__ splats, const int32_t* __restrict__ flatten_ids,
                  const int32_t* __restrict__ offsets, int M, int W, int H, int tile_w, int n_tiles, Bg bg,
                  const float* __restrict__ alpha_in, const int32_t* __restrict__ last_ids,
                  const float* __restrict__ v_render, const float* __restrict__ v_alpha,
                  float* __restrict__ v_splats) {
    constexpr int CH = ChInfo<CHSET>::CH;
    constexpr bool HAS_RGB = (CHSET != B200_CH_D);
    constexpr bool HAS_D = (CHSET != B200_CH_RGB);
    __shared__ float4 s_a[RB_THREADS];
    __shared__ float4 s_b[RB_THREADS];
    __shared__ float2 s_c[RB_THREADS];
    __shared__ int s_id[RB_THREADS];
    __shared__ int s_max[RB_THREADS / 32];

    const int tile = blockIdx.x;
    const int ty = tile / tile_w, tx = tile - ty * tile_w;
    const int lx = threadIdx.x & (TILE - 1), ly = threadIdx.x >> 4;
    const int ix = tx * TILE + lx, iy = ty * TILE + ly;
    const bool inside = ix < W && iy < H;
    const float px = (float)ix + 0.5f, py = (float)iy + 0.5f;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;

    const int start = offsets[tile];
    const size_t pix = (size_t)iy * W + ix;
    int bin_final = -1;
    float T_final = 1.f, va = 0.f;
    float vr[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) vr[c] = 0.f;
    if (inside) {
        bin_final = last_ids[pix];
        T_final = 1.f - alpha_in[pix];
#pragma unroll
        for (int c = 0; c < CH; ++c) vr[c] = v_render[pix * CH + c];
        if (v_alpha != nullptr) va = v_alpha[pix];
    }
    float bgdot = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) bgdot += bg.v[c] * vr[c];
    const float va_eff = T_final * (va - bgdot);  // multiplies ra below

    // tile-wide deepest composited index
    int mx = bin_final;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_max[wid] = mx;
    __syncthreads();
    int tile_last = s_max[0];
#pragma unroll
    for (int w = 1; w < RB_THREADS / 32; ++w) tile_last = max(tile_last, s_max[w]);
    if (tile_last < start) return;

    float T = T_final;
    float buf[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) buf[c] = 0.f;

    for (int hi = tile_last; hi >= start; hi -= RB_THREADS) {
        __syncthreads();
        const int idx_l = hi - (int)threadIdx.x;
        if (idx_l >= start) {
            const int g = __ldg(flatten_ids + idx_l);
            const float4* rec = splats + 3 * (size_t)g;
            float4 a = __ldg(rec), b = __ldg(rec + 1), c = __ldg(rec + 2);
            s_a[threadIdx.x] = a;
            s_b[threadIdx.x] = b;
            s_c[threadIdx.x] = make_float2(c.x, c.y);
            s_id[threadIdx.x] = g;
        }
        __syncthreads();
        const int nb = min(RB_THREADS, hi - start + 1);
        for (int j = 0; j < nb; ++j) {
            const int idx = hi - j;
            bool active = idx <= bin_final;  // bin_final = -1 for outside / empty pixels
            float4 a, b;
            float dx = 0.f, dy = 0.f, vis = 0.f, al = 0.f;
            if (active) {
                a = s_a[j];
                b = s_b[j];
                dx = a.x - px; dy = a.y - py;
                const float sigma = 0.5f * (a.z * dx * dx + b.x * dy * dy) + a.w * dx * dy;
                vis = __expf(-sigma);
                al = fminf(MAX_ALPHA, b.y * vis);
                if (sigma < 0.f || al < ALPHA_THRESHOLD) active = false;
            }
            if (!__any_sync(0xffffffffu, active)) continue;
            float v_col[CH];
#pragma unroll
            for (int c = 0; c < CH; ++c) v_col[c] = 0.f;
            float v_A = 0.f, v_B = 0.f, v_C = 0.f, v_mx = 0.f, v_my = 0.f, v_o = 0.f;
            if (active) {
                const float ra = 1.f / (1.f - al);
                T *= ra;
                const float fac = al * T;
                float col[CH];
                if (HAS_RGB) { col[0] = b.z; col[1] = b.w; col[2] = s_c[j].x; }
                if (HAS_D) col[CH - 1] = s_c[j].y;
                float v_al = 0.f;
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    v_col[c] = fac * vr[c];
                    v_al += (col[c] * T - buf[c] * ra) * vr[c];
                    buf[c] += col[c] * fac;
                }
                v_al += va_eff * ra;
                if (b.y * vis <= MAX_ALPHA) {
                    const float v_sigma = -b.y * vis * v_al;
                    v_A = 0.5f * v_sigma * dx * dx;
                    v_B = v_sigma * dx * dy;
                    v_C = 0.5f * v_sigma * dy * dy;
                    v_mx = v_sigma * (a.z * dx + a.w * dy);
                    v_my = v_sigma * (a.w * dx + b.x * dy);
                    v_o = vis * v_al;
                }
            }
#pragma unroll
            for (int c = 0; c < CH; ++c) v_col[c] = warp_sum(v_col[c]);
            v_A = warp_sum(v_A); v_B = warp_sum(v_B); v_C = warp_sum(v_C);
            v_mx = warp_sum(v_mx); v_my = warp_sum(v_my); v_o = warp_sum(v_o);
            if (lane == 0) {
                float* out = v_splats + (size_t)s_id[j] * SPLAT_F;
                atomicAdd(out + 0, v_mx); atomicAdd(out + 1, v_my);
                atomicAdd(out + 2, v_A); atomicAdd(out + 3, v_B); atomicAdd(out + 4, v_C);
                atomicAdd(out + 5, v_o);
                if (HAS_RGB) { atomicAdd(out + 6, v_col[0]); atomicAdd(out + 7, v_col[1]); atomicAdd(out + 8, v_col[2]); }
                if (HAS_D) atomicAdd(out + 9, v_col[CH - 1]);
            }
        }
    }
}

}  // namespace

int b200_channels(int chset);

extern "C" int b200_raster_bwd(const float* splats, const int32_t* flatten_ids, const int32_t* offsets, int64_t M,
                               int width, int height, int channels, const float* background_host,
                               const float* alpha, const int32_t* last_ids, const float* v_render,
                               const float* v_alpha, float* v_splats, void* stream) {
    B200_REQUIRE(width > 0 && height > 0 && M >= 0 && M < ((int64_t)1 << 31), "bad sizes");
    B200_REQUIRE(b200_channels(channels) > 0, "bad channel set");
    if (M == 0) return B200_OK;
    B200_REQUIRE(splats && flatten_ids && offsets && alpha && last_ids && v_render && v_splats, "null pointer");
    B200_REQUIRE(((uintptr_t)splats & 15) == 0, "misaligned splats");
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const int n_tiles = tile_w * tile_h;
    Bg bg;
    for (int c = 0; c < 4; ++c) bg.v[c] = (background_host && c < b200_channels(channels)) ? background_host[c] : 0.f;
    cudaStream_t st = (cudaStream_t)stream;
    const float4* sp = reinterpret_cast<const float4*>(splats);
#define LAUNCH(CS) raster_bwd_kernel<CS><<<n_tiles, RB_THREADS, 0, st>>>(sp, flatten_ids, offsets, (int)M, width, height, tile_w, n_tiles, bg, alpha, last_ids, v_render, v_alpha, v_splats)
    switch (channels) {
        case B200_CH_RGB: LAUNCH(B200_CH_RGB); break;
        case B200_CH_D: LAUNCH(B200_CH_D); break;
        default: LAUNCH(B200_CH_RGBD); break;
    }
#undef LAUNCH
    return b200_check_launch("raster_bwd");
}
