// raster_bwd.cu — stage 4a: reverse replay of the blend and accumulation of the splat-record
// cotangents (gsplat rasterize_to_pixels backward, triggered by the reference at
// src/training/trainer.py:505-511).  Work decomposition: raster_common.cuh.  Each lane sums a
// Gaussian's ten partial sums over its four pixels in registers and parks them in shared memory; every
// RB_GROUP Gaussians the warp reduces the parked block with a transposed read (flush_pending) and issues
// three 16-byte vector REDs per Gaussian into the 48-B gradient record — 3 L2 atomic operations per
// (Gaussian, half tile) instead of 10 per warp of 32 px, and no shuffle trees.
#include <stdlib.h>
#include "raster_common.cuh"

namespace {

#ifndef RB_MIN_WARPS
#define RB_MIN_WARPS 24          // resident warps per SM the register allocation must allow (<= 85 registers)
#endif
#define RB_GROUP 4            // Gaussians whose per-lane partial sums wait in shared memory for one reduction
#define RB_ROW (32 * 3)       // float4 per waiting Gaussian: 32 lanes x {q0, q1, q2}

struct __align__(16) Pending { int id; float A, B, C; };

#ifdef B200_TRACE_UNITS  // tuning aid (profiles/unit_trace.py): start / end time and SM of every work unit
__device__ unsigned long long g_unit_trace[3 * 65536];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned smid() { unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }
#define TRACE_BEGIN() do { if (lane == 0 && unit < 65536) { g_unit_trace[3 * unit] = gtime(); g_unit_trace[3 * unit + 2] = smid(); } } while (0)
#define TRACE_END() do { if (lane == 0 && unit < 65536) g_unit_trace[3 * unit + 1] = gtime(); } while (0)
#else
#define TRACE_BEGIN()
#define TRACE_END()
#endif

// Reduction of RB_GROUP x 32 lanes x 12 partial sums through shared memory.  Lane (g, q, h) = (lane>>3,
// (lane>>1)&3, lane&1) adds the float4 group q of Gaussian g over the 16 source lanes 2i+h — 16 LDS.128 +
// 64 FADD for four Gaussians, one shuffle level to join the halves — and the 12 lanes with h == 0, q < 3 each
// issue ONE 16-byte vector RED.  ~30 warp instructions per Gaussian instead of ~105 for ten shuffle trees.
// Bank check (float4 units, 8 per wavefront): quarter-warp = one g, index (2i+h)*3 + q = 6i + 3h + q ->
// distinct for the six active (q, h) pairs.
__device__ __forceinline__ void flush_pending(const float4* __restrict__ part, const Pending* __restrict__ pend, int np,
                                              int lane, float* __restrict__ v_splats) {
    __syncwarp();
    const int g = lane >> 3, q = (lane >> 1) & 3, h = lane & 1;
    const bool live = q < 3 && g < np;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
        const float4* row = part + g * RB_ROW + 3 * h + q;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float4 v = row[6 * i];
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    }
    acc.x += __shfl_xor_sync(0xffffffffu, acc.x, 1);
    acc.y += __shfl_xor_sync(0xffffffffu, acc.y, 1);
    acc.z += __shfl_xor_sync(0xffffffffu, acc.z, 1);
    acc.w += __shfl_xor_sync(0xffffffffu, acc.w, 1);
    if (live && h == 0) {
        const Pending p = pend[g];
        float4 o = acc;
        if (q == 0) {  // (Sx, Sy, Sxx, Sxy) -> (v_mx, v_my, v_A, v_B)
            o.x = p.A * acc.x + p.B * acc.y;
            o.y = p.B * acc.x + p.C * acc.y;
            o.z = 0.5f * acc.z;
        } else if (q == 1) {  // (Syy, v_opacity, v_r, v_g)
            o.x = 0.5f * acc.x;
        }
        red_add_v4(v_splats + (size_t)p.id * SPLAT_F + 4 * q, o.x, o.y, o.z, o.w);
    }
    __syncwarp();
}

template <int CHSET, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, RB_MIN_WARPS / WARPS)
raster_bwd_kernel(const float4* __restrict__ splats, const int32_t* __restrict__ flatten_ids,
                  const int32_t* __restrict__ offsets, int M, int W, int H, int tile_w, int n_tiles,
                  const float* __restrict__ background, int n_bg, const float* __restrict__ alpha_in,
                  const int32_t* __restrict__ last_ids, const float* __restrict__ v_render,
                  const float* __restrict__ v_alpha, float* __restrict__ v_splats) {
    constexpr int CH = ChInfo<CHSET>::CH;
    constexpr bool HAS_RGB = (CHSET != B200_CH_D);
    constexpr bool HAS_D = (CHSET != B200_CH_RGB);
    __shared__ StagedSplat s_splat[WARPS][32];
    __shared__ float4 s_part[WARPS][RB_GROUP * RB_ROW];
    __shared__ Pending s_pend[WARPS][RB_GROUP];

    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int unit = blockIdx.x * WARPS + wid;
    if (unit >= 2 * n_tiles) return;
    TRACE_BEGIN();
    const int tile = unit >> 1, half = unit & 1;
    const int ty = tile / tile_w, tx = tile - ty * tile_w;
    const int lx = lane & 7, ly = lane >> 3;
    const int x_base = tx * TILE, y_base = ty * TILE + half * 8;
    StagedSplat* stage = s_splat[wid];
    float4* part = s_part[wid];
    Pending* pend = s_pend[wid];
    const int start = offsets[tile];
    Bg bg;
#pragma unroll
    for (int c = 0; c < CH; ++c) bg.v[c] = (c < n_bg) ? __ldg(background + c) : 0.f;

    // per-lane state of four pixels: transmittance, cotangents, and D = sum_c buffer_c * v_render_c (the only
    // form in which the accumulated colour behind the current Gaussian is needed)
    float T[RU_SLOTS], va_eff[RU_SLOTS], vr[RU_SLOTS][CH], D[RU_SLOTS];
    int bin_final[RU_SLOTS];
    int warp_last = -1;
#pragma unroll
    for (int s = 0; s < RU_SLOTS; ++s) {
        const int ix = x_base + (s & 1) * 8 + lx, iy = y_base + (s >> 1) * 4 + ly;
        bin_final[s] = -1;
        T[s] = 1.f;
        va_eff[s] = 0.f;
        D[s] = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) vr[s][c] = 0.f;
        if (ix < W && iy < H) {
            const size_t pix = (size_t)iy * W + ix;
            bin_final[s] = last_ids[pix];
            const float T_final = 1.f - alpha_in[pix];
            T[s] = T_final;
            float bgdot = 0.f;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                vr[s][c] = v_render[pix * CH + c];
                bgdot += bg.v[c] * vr[s][c];
            }
            const float va = (v_alpha != nullptr) ? v_alpha[pix] : 0.f;
            va_eff[s] = T_final * (va - bgdot);
        }
        warp_last = max(warp_last, bin_final[s]);
    }
    warp_last = __reduce_max_sync(0xffffffffu, warp_last);
    if (warp_last < start) { TRACE_END(); return; }
    const float pxc = (float)(x_base + lx) + 4.5f, pyc = (float)(y_base + ly) + 2.5f;  // centroid of the 4 pixels
    const float bx0 = (float)x_base + 0.5f, bx1 = (float)min(x_base + 15, W - 1) + 0.5f;
    const float by0 = (float)y_base + 0.5f, by1 = (float)min(y_base + 7, H - 1) + 0.5f;
    int np = 0;  // Gaussians waiting in `part` (warp-uniform)

    for (int hi = warp_last; hi >= start; hi -= 32) {
        const int idx_l = hi - lane;
        bool keep = false;
        if (idx_l >= start) {
            const int g = __ldg(flatten_ids + idx_l);
            const float4* rec = splats + 3 * (size_t)g;
            const float4 a = __ldg(rec), b = __ldg(rec + 1), c = __ldg(rec + 2);
            keep = splat_reaches_box(a, b, bx0, bx1, by0, by1);
            if (keep) stage_splat(stage[lane], a, b, c, g, idx_l);
        }
        unsigned mask = __ballot_sync(0xffffffffu, keep);
        __syncwarp();
        while (mask) {
            const int j = __ffs(mask) - 1;  // ascending lane == descending sorted index
            mask &= mask - 1;
            const float4 a = stage[j].a;
            const float4 b = stage[j].b;
            const int idx = stage[j].idx;
            const float u = a.x - pxc, v = a.y - pyc;
            float sg[RU_SLOTS], vis[RU_SLOTS], al[RU_SLOTS];
            eval_sigma4(a, b, u, v, sg);
            bool ok[RU_SLOTS];
#pragma unroll
            for (int s = 0; s < RU_SLOTS; ++s) {
                vis[s] = fast_ex2(-sg[s]);
                al[s] = fminf(MAX_ALPHA, b.y * vis[s]);
                ok[s] = al[s] >= ALPHA_THRESHOLD && sg[s] >= 0.f && idx <= bin_final[s];
            }
            if (!__any_sync(0xffffffffu, ok[0] || ok[1] || ok[2] || ok[3])) continue;
            const float4 c4 = stage[j].c;
            float col[CH];
            if (HAS_RGB) { col[0] = c4.x; col[1] = c4.y; col[2] = c4.z; }
            if (HAS_D) col[CH - 1] = c4.w;
            // per-lane partial sums over this lane's pixels: colour cotangents, opacity, and the moments
            // Sx = sum v_sigma dx, Sxx = sum v_sigma dx^2, ... from which conic / mean cotangents follow
            float g_col[CH], g_o = 0.f, Sx = 0.f, Sy = 0.f, Sxx = 0.f, Sxy = 0.f, Syy = 0.f;
#pragma unroll
            for (int c = 0; c < CH; ++c) g_col[c] = 0.f;
#pragma unroll
            for (int s = 0; s < RU_SLOTS; ++s) {
                if (ok[s]) {  // a slot no lane reaches is one untaken branch
                    const float ra = fast_rcp(1.f - al[s]);
                    T[s] *= ra;
                    const float fac = al[s] * T[s];
                    float cv = 0.f;  // colour . v_render
#pragma unroll
                    for (int c = 0; c < CH; ++c) {
                        g_col[c] = fmaf(fac, vr[s][c], g_col[c]);
                        cv = fmaf(col[c], vr[s][c], cv);
                    }
                    // v_alpha = sum_c (col_c T - buffer_c / (1 - alpha)) v_render_c + T_final (v_a - bg.v_render) / (1 - alpha)
                    const float v_al = fmaf(T[s], cv, ra * (va_eff[s] - D[s]));
                    D[s] = fmaf(fac, cv, D[s]);
                    const float ov = b.y * vis[s];
                    if (ov <= MAX_ALPHA) {
                        const float dx = (s & 1) ? u - 4.f : u + 4.f, dy = (s >> 1) ? v - 2.f : v + 2.f;
                        const float v_sigma = -ov * v_al;
                        const float vx = v_sigma * dx, vy = v_sigma * dy;
                        Sx += vx; Sy += vy;
                        Sxx = fmaf(vx, dx, Sxx);
                        Sxy = fmaf(vx, dy, Sxy);
                        Syy = fmaf(vy, dy, Syy);
                        g_o = fmaf(vis[s], v_al, g_o);
                    }
                }
            }
            // park the lane's partials; the warp reduces RB_GROUP Gaussians at a time (flush_pending)
            float4* row = part + np * RB_ROW + 3 * lane;
            row[0] = make_float4(Sx, Sy, Sxx, Sxy);
            row[1] = make_float4(Syy, g_o, HAS_RGB ? g_col[0] : 0.f, HAS_RGB ? g_col[1] : 0.f);
            row[2] = make_float4(HAS_RGB ? g_col[2] : 0.f, HAS_D ? g_col[CH - 1] : 0.f, 0.f, 0.f);
            if (lane == 0) {
                Pending p;
                p.id = stage[j].id;
                p.A = a.z * (2.f * LN2_F); p.B = a.w * LN2_F; p.C = b.x * (2.f * LN2_F);  // undo the staging scale
                pend[np] = p;
            }
            if (++np == RB_GROUP) {
                flush_pending(part, pend, np, lane, v_splats);
                np = 0;
            }
        }
        __syncwarp();
    }
    if (np) flush_pending(part, pend, np, lane, v_splats);
    TRACE_END();
}

}  // namespace

#ifdef B200_TRACE_UNITS
extern "C" int b200_debug_unit_trace(unsigned long long* host_out, int n_units) {
    cudaDeviceSynchronize();
    return cudaMemcpyFromSymbol(host_out, g_unit_trace, sizeof(unsigned long long) * 3 * (size_t)n_units) == cudaSuccess ? 0 : 1;
}
#endif

extern "C" int b200_raster_bwd(const float* splats, const int32_t* flatten_ids, const int32_t* offsets, int64_t M,
                               int width, int height, int channels, const float* background, int n_background,
                               const float* alpha, const int32_t* last_ids, const float* v_render,
                               const float* v_alpha, float* v_splats, void* stream) {
    B200_REQUIRE(width > 0 && height > 0 && M >= 0 && M < ((int64_t)1 << 31), "bad sizes");
    B200_REQUIRE(b200_channels(channels) > 0, "bad channel set");
    if (M == 0) return B200_OK;
    B200_REQUIRE(splats && flatten_ids && offsets && alpha && last_ids && v_render && v_splats, "null pointer");
    B200_REQUIRE(((uintptr_t)splats & 15) == 0 && ((uintptr_t)v_splats & 15) == 0, "misaligned splats");
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const int n_tiles = tile_w * tile_h;
    B200_REQUIRE(n_background >= 0 && n_background <= b200_channels(channels) && (n_background == 0 || background), "bad background");
    cudaStream_t st = (cudaStream_t)stream;
    const float4* sp = reinterpret_cast<const float4*>(splats);
    const int warps = b200_ru_warps("B200_BWD_WARPS", RU_WARPS);
    const int grid = (2 * n_tiles + warps - 1) / warps;
#define LAUNCH1(CS, WP) raster_bwd_kernel<CS, WP><<<grid, WP * 32, 0, st>>>(sp, flatten_ids, offsets, (int)M, width, height, tile_w, n_tiles, background, n_background, alpha, last_ids, v_render, v_alpha, v_splats)
#define LAUNCH(CS) do { if (warps == 1) LAUNCH1(CS, 1); else if (warps == 2) LAUNCH1(CS, 2); else LAUNCH1(CS, 4); } while (0)
    switch (channels) {
        case B200_CH_RGB: LAUNCH(B200_CH_RGB); break;
        case B200_CH_D: LAUNCH(B200_CH_D); break;
        default: LAUNCH(B200_CH_RGBD); break;
    }
#undef LAUNCH
#undef LAUNCH1
    return b200_check_launch("raster_bwd");
}
