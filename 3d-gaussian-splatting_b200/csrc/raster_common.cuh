// raster_common.cuh — shared pieces of the forward/backward blend kernels.
//
// Work decomposition (both directions): one WARP owns one half tile (16x8 px) of a 16x16 tile and
// walks that tile's sorted list on its own — no block barriers.  The half tile is four 8x4 "slots";
// lane l owns pixel (l & 7, l >> 3) of every slot, so a lane carries four pixels' state in registers
// and a slot whose 32 pixels a Gaussian does not reach costs one untaken branch.  Batches of 32 Gaussians are
// gathered by the 32 lanes (three 16-B vectors per 48-B splat record), screened against the half
// tile with an exact convex-quadratic bound, and only survivors are broadcast through shared memory.
#pragma once
#include "common.cuh"

#define RU_WARPS 4                   // warps (= half-tile work units) per CTA
#define RU_THREADS (RU_WARPS * 32)
#define RU_SLOTS 4                   // 8x4 sub-tiles per half tile

template <int CHSET> struct ChInfo;
template <> struct ChInfo<B200_CH_RGB>  { static constexpr int CH = 3; };
template <> struct ChInfo<B200_CH_D>    { static constexpr int CH = 1; };
template <> struct ChInfo<B200_CH_RGBD> { static constexpr int CH = 4; };

struct Bg { float v[4]; };

#define LOG2E_F 1.4426950408889634f
#define LN2_F 0.6931471805599453f

// Staged form of a splat record (64 B, four 16-B vectors).  The staging lane pre-multiplies the conic by
// log2(e) (and halves A, C) so the 32 consumers evaluate sigma in base-2 units and feed it straight to
// MUFU.EX2, and pre-computes the two constants of the four-slot expansion (eval_sigma4).
struct __align__(16) StagedSplat {
    float4 a;  // mx, my, hA = 0.5*A*log2e, B' = B*log2e
    float4 b;  // hC = 0.5*C*log2e, opacity, c0 = 16 hA + 4 hC, B8 = 8 B'
    float4 c;  // r, g, b, depth
    int id;    // Gaussian index
    int idx;   // position in the sorted list
    int pad0, pad1;
};

// in: a = (mx, my, A, B), b = (C, opacity, r, g), c = (blue, depth, ., .) as stored in the splat record
__device__ __forceinline__ void stage_splat(StagedSplat& st, float4 a, float4 b, const float4 c, int id, int idx) {
    const float hA = a.z * (0.5f * LOG2E_F), Bp = a.w * LOG2E_F, hC = b.x * (0.5f * LOG2E_F);
    st.a = make_float4(a.x, a.y, hA, Bp);
    st.b = make_float4(hC, b.y, __fmaf_rn(16.f, hA, 4.f * hC), 8.f * Bp);
    st.c = make_float4(b.z, b.w, c.x, c.y);
    *reinterpret_cast<int2*>(&st.id) = make_int2(id, idx);
}

// 2^x, flush-to-zero: one MUFU.EX2 without the denormal range fix-up (results below 2^-126 are far under
// the 1/255 alpha threshold anyway).  Forward and backward share it, so they agree bit for bit.
__device__ __forceinline__ float fast_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// sigma (base-2 units) at a lane's four pixels.  The pixels sit at (+-4, +-2) around their centroid, so
// with (u, v) = mean - centroid:
//   sigma(u + 4 sx, v + 2 sy) = [hA u^2 + hC v^2 + B u v] + c0 + sx 4(2 hA u + B v) + sy 2(2 hC v + B u) + sx sy 8B
// 21 operations for four values instead of 4 x 9.  Slot s has sx = (s & 1) ? -1 : +1, sy = (s >> 1) ? -1 : +1
// (d = mean - pixel, and the pixel of slot s is 8 (s & 1) to the right / 4 (s >> 1) below slot 0's).
// Explicit round-to-nearest intrinsics pin ONE operation order (no contraction differences between the
// forward and the backward kernel), so both make the same skip decision for a (pixel, Gaussian) pair.
__device__ __forceinline__ void eval_sigma4(const float4 a, const float4 b, float u, float v, float (&s)[4]) {
    const float hA = a.z, B = a.w, hC = b.x, c0 = b.z, B8 = b.w;
    const float p = __fmul_rn(hA, u), q = __fmul_rn(hC, v), w = __fmul_rn(B, u);
    const float sc = __fmaf_rn(p, u, __fmaf_rn(q, v, __fmul_rn(w, v)));
    const float a4 = __fmaf_rn(2.f, p, __fmul_rn(B, v));  // (2 hA u + B v)
    const float b2 = __fmaf_rn(2.f, q, w);                // (2 hC v + B u)
    const float base = __fadd_rn(sc, c0);
    const float P = __fadd_rn(base, B8), M = __fsub_rn(base, B8);
    const float bb = __fmul_rn(2.f, b2);
    const float t1 = __fmaf_rn(4.f, a4, bb), t2 = __fmaf_rn(4.f, a4, -bb);
    s[0] = __fadd_rn(P, t1);
    s[1] = __fsub_rn(M, t2);
    s[2] = __fadd_rn(M, t2);
    s[3] = __fsub_rn(P, t1);
}

// sigma = hA dx^2 + hC dy^2 + B dx dy at one point (used by the half-tile screen below).
__device__ __forceinline__ float eval_sigma(float hA, float B, float hC, float dx, float dy) {
    float s = (hA * dx) * dx;
    s = fmaf(hC * dy, dy, s);
    s = fmaf(B * dx, dy, s);
    return s;
}

// Conservative screen: can alpha = opacity * exp(-sigma) reach 1/255 anywhere on the pixel-centre
// rectangle [x0,x1] x [y0,y1]?  The conic is positive definite, so sigma's minimum over the box is 0 if
// the mean is inside, else it lies on one of the four edges where it is a clamped 1-D quadratic minimum.
// A 1e-4 relative + absolute slack keeps every pair the exact per-pixel test could accept.
__device__ __forceinline__ bool splat_reaches_box(const float4 a, const float4 b, float x0, float x1, float y0, float y1) {
    const float mx = a.x, my = a.y, A = a.z, B = a.w, C = b.x, o = b.y;
    const float dxl = mx - x1, dxh = mx - x0, dyl = my - y1, dyh = my - y0;  // d = mean - pixel
    if (dxl <= 0.f && dxh >= 0.f && dyl <= 0.f && dyh >= 0.f) return true;
    const float hA = 0.5f * A, hC = 0.5f * C;
    const float rA = __fdividef(1.f, A), rC = __fdividef(1.f, C);
    float smin;
    {
        float dy = fminf(fmaxf(-B * dxl * rC, dyl), dyh);
        smin = eval_sigma(hA, B, hC, dxl, dy);
        dy = fminf(fmaxf(-B * dxh * rC, dyl), dyh);
        smin = fminf(smin, eval_sigma(hA, B, hC, dxh, dy));
        float dx = fminf(fmaxf(-B * dyl * rA, dxl), dxh);
        smin = fminf(smin, eval_sigma(hA, B, hC, dx, dyl));
        dx = fminf(fmaxf(-B * dyh * rA, dxl), dxh);
        smin = fminf(smin, eval_sigma(hA, B, hC, dx, dyh));
    }
    const float tau = __logf(255.f * o);  // alpha >= 1/255  <=>  sigma <= ln(255 o)
    return !(smin > tau * 1.0001f + 1e-4f);
}

__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// 16-byte vector reduction into global memory (sm_90+): one RED for four adjacent floats.
__device__ __forceinline__ void red_add_v4(float* addr, float x, float y, float z, float w) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ void red_add_v2(float* addr, float x, float y) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(x), "f"(y) : "memory");
}

// Reduce eight per-lane values across the warp with a transposing butterfly: 7 shuffles for the first
// three levels (instead of 8 x 3), then two plain levels.  On return lane l holds the warp total of
// v[(l >> 2) & 7].  32 instructions instead of 80 for eight plain shuffle trees.
__device__ __forceinline__ float warp_reduce8_transposed(const float (&v)[8], int lane) {
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    float w[4], u[2], r;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float send = b4 ? v[i] : v[i + 4];
        const float keep = b4 ? v[i + 4] : v[i];
        w[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float send = b3 ? w[i] : w[i + 2];
        const float keep = b3 ? w[i + 2] : w[i];
        u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    {
        const float send = b2 ? u[0] : u[1];
        const float keep = b2 ? u[1] : u[0];
        r = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    r += __shfl_xor_sync(0xffffffffu, r, 2);
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    return r;
}

int b200_channels(int chset);
int b200_ru_warps(const char* env, int dflt);
