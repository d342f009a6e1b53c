// raster_fwd.cu — stage 3: per-tile front-to-back alpha blend (gsplat
// rasterize_to_pixels forward as reached from gsplat_backend.py:168-182).
// One CTA per 16x16 tile; each batch of 256 sorted Gaussians is gathered (three 16-B vectors per
// 48-B splat record, L2-resident: N*48 B < 126 MB up to 2.6 M Gaussians) into shared memory and
// broadcast to the 256 pixel threads.
#include "common.cuh"

#define RF_THREADS (TILE * TILE)

namespace {

template <int CHSET> struct ChInfo;
template <> struct ChInfo<B200_CH_RGB>  { static constexpr int CH = 3; };
template <> struct ChInfo<B200_CH_D>    { static constexpr int CH = 1; };
template <> struct ChInfo<B200_CH_RGBD> { static constexpr int CH = 4; };

struct Bg { float v[4]; };

template <int CHSET>
__global__ void __launch_bounds__(RF_THREADS)
raster_fwd_kernel(const float4* __restrict__ splats, const int32_t* __restrict__ flatten_ids,
                  const int32_t* __restrict__ offsets, int M, int W, int H, int tile_w, int n_tiles, Bg bg,
                  float* __restrict__ render, float* __restrict__ alpha_out, int32_t* __restrict__ last_ids) {
    constexpr int CH = ChInfo<CHSET>::CH;
    __shared__ float4 s_a[RF_THREADS];  // mx, my, A, B
    __shared__ float4 s_b[RF_THREADS];  // C, opacity, r, g
    __shared__ float2 s_c[RF_THREADS];  // b, depth

    const int tile = blockIdx.x;
    const int ty = tile / tile_w, tx = tile - ty * tile_w;
    const int lx = threadIdx.x & (TILE - 1), ly = threadIdx.x >> 4;
    const int ix = tx * TILE + lx, iy = ty * TILE + ly;
    const bool inside = ix < W && iy < H;
    const float px = (float)ix + 0.5f, py = (float)iy + 0.5f;

    const int start = offsets[tile];
    const int end = (tile + 1 < n_tiles) ? offsets[tile + 1] : M;

    float T = 1.f;
    float acc[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c] = 0.f;
    int last = -1;
    bool done = !inside;

    for (int b0 = start; b0 < end; b0 += RF_THREADS) {
        if (__syncthreads_count(done) == RF_THREADS) break;
        const int idx = b0 + threadIdx.x;
        if (idx < end) {
            const int g = __ldg(flatten_ids + idx);
            const float4* rec = splats + 3 * (size_t)g;
            float4 a = __ldg(rec), b = __ldg(rec + 1), c = __ldg(rec + 2);
            s_a[threadIdx.x] = a;
            s_b[threadIdx.x] = b;
            s_c[threadIdx.x] = make_float2(c.x, c.y);
        }
        __syncthreads();
        const int nb = min(RF_THREADS, end - b0);
        if (!done) {
            for (int j = 0; j < nb; ++j) {
                const float4 a = s_a[j];
                const float4 b = s_b[j];
                const float dx = a.x - px, dy = a.y - py;
                const float sigma = 0.5f * (a.z * dx * dx + b.x * dy * dy) + a.w * dx * dy;
                const float vis = __expf(-sigma);
                const float al = fminf(MAX_ALPHA, b.y * vis);
                if (sigma < 0.f || al < ALPHA_THRESHOLD) continue;
                const float Tn = T * (1.f - al);
                if (Tn <= T_STOP) { done = true; break; }
                const float w = al * T;
                if (CHSET == B200_CH_RGB || CHSET == B200_CH_RGBD) {
                    acc[0] += w * b.z; acc[1] += w * b.w; acc[2] += w * s_c[j].x;
                    if (CHSET == B200_CH_RGBD) acc[3] += w * s_c[j].y;
                } else {
                    acc[0] += w * s_c[j].y;
                }
                last = b0 + j;
                T = Tn;
            }
        }
    }
    if (inside) {
        const size_t pix = (size_t)iy * W + ix;
#pragma unroll
        for (int c = 0; c < CH; ++c) render[pix * CH + c] = acc[c] + T * bg.v[c];
        alpha_out[pix] = 1.f - T;
        last_ids[pix] = last;
    }
}

}  // namespace

int b200_channels(int chset) {
    switch (chset) {
        case B200_CH_RGB: return 3;
        case B200_CH_D: return 1;
        case B200_CH_RGBD: return 4;
        default: return -1;
    }
}

extern "C" int b200_raster_fwd(const float* splats, const int32_t* flatten_ids, const int32_t* offsets, int64_t M,
                               int width, int height, int channels, const float* background_host, float* render,
                               float* alpha, int32_t* last_ids, void* stream) {
    B200_REQUIRE(width > 0 && height > 0 && M >= 0 && M < ((int64_t)1 << 31), "bad sizes");
    B200_REQUIRE(b200_channels(channels) > 0, "bad channel set");
    B200_REQUIRE(offsets && render && alpha && last_ids, "null pointer");
    B200_REQUIRE(M == 0 || (splats && flatten_ids), "null pointer");
    B200_REQUIRE(((uintptr_t)splats & 15) == 0, "misaligned splats");
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const int n_tiles = tile_w * tile_h;
    Bg bg;
    for (int c = 0; c < 4; ++c) bg.v[c] = (background_host && c < b200_channels(channels)) ? background_host[c] : 0.f;
    cudaStream_t st = (cudaStream_t)stream;
    const float4* sp = reinterpret_cast<const float4*>(splats);
#define LAUNCH(CS) raster_fwd_kernel<CS><<<n_tiles, RF_THREADS, 0, st>>>(sp, flatten_ids, offsets, (int)M, width, height, tile_w, n_tiles, bg, render, alpha, last_ids)
    switch (channels) {
        case B200_CH_RGB: LAUNCH(B200_CH_RGB); break;
        case B200_CH_D: LAUNCH(B200_CH_D); break;
        default: LAUNCH(B200_CH_RGBD); break;
    }
#undef LAUNCH
    return b200_check_launch("raster_fwd");
}
