// project_sh.cu — stage 1 (EWA projection + SH colour + tile count, fused) and stage 4b (its VJP).
//
// Semantics follow gsplat 1.5.x as the reference calls it (src/backends/gsplat_backend.py:168-182;
// SURVEY.md Appendix A.1, A.2, A.6).  One thread per Gaussian; the 48-B splat record is the only
// per-Gaussian float output so the blend stages gather three 16-B vectors per Gaussian.
#include "common.cuh"

#define PROJ_THREADS 256

// The camera block is read from DEVICE memory (the caller's packed block, b200_cam_pack: the pose never visits the
// host) and staged once per CTA in shared memory.  No global slot is shared between launches, so projection kernels
// of different views may be in flight on different streams at the same time (streams.ViewStreams).  A HOST camera
// block is first copied into c_cam_staging, which is one slot: such launches on a different stream wait for the
// last kernel that read it (load_cam / cam_used below).
__device__ Cam c_cam_staging;

#define STAGE_CAM(cam_g)                                                                                           \
    __shared__ Cam s_cam;                                                                                          \
    if (threadIdx.x < sizeof(Cam) / 4)                                                                             \
        reinterpret_cast<uint32_t*>(&s_cam)[threadIdx.x] = __ldg(reinterpret_cast<const uint32_t*>(cam_g) + threadIdx.x); \
    __syncthreads();                                                                                               \
    const Cam& cam = s_cam

namespace {

constexpr float SH_C0 = 0.2820947917738781f;
constexpr float SH_C1 = 0.48860251190292f;
constexpr float SH_C2A = 0.5462742152960395f;
constexpr float SH_C2B = 1.092548430592079f;
constexpr float SH_C2C = 0.9461746957575601f;
constexpr float SH_C2D = 0.3153915652525201f;
constexpr float SH_C3A = 0.5900435899266435f;
constexpr float SH_C3B = 1.445305721320277f;
constexpr float SH_C3C = 2.285228997322329f;
constexpr float SH_C3D = 0.4570457994644658f;
constexpr float SH_C3E = 1.865881662950577f;
constexpr float SH_C3F = 1.119528997770346f;

struct Proj {
    float pcx, pcy, pcz;   // camera-space mean
    float Rq[9];           // rotation from the normalised quaternion
    float qn[4];           // normalised quaternion
    float inv_norm;        // 1/|q|
    float Wm[9];           // R * Rq * diag(s)
    float J00, J02, J11, J12;
    float T[6];            // J * Wm (2x3)
    float tx, ty, rz;
    bool clamp_x, clamp_y;
    float a, b, c, det;    // blurred 2-D covariance and determinant
    float mx, my;
};

// Forward projection up to the conic; returns false if culled by depth or determinant.
__device__ __forceinline__ bool project_core(const Cam& cam, float px, float py, float pz, float4 q,
                                             float sx, float sy, float sz, Proj& P) {
    const float* R = cam.R;
    P.pcx = R[0] * px + R[1] * py + R[2] * pz + cam.t[0];
    P.pcy = R[3] * px + R[4] * py + R[5] * pz + cam.t[1];
    P.pcz = R[6] * px + R[7] * py + R[8] * pz + cam.t[2];
    if (!(P.pcz >= cam.near_plane) || !(P.pcz <= cam.far_plane)) return false;

    float n2 = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
    P.inv_norm = rsqrtf(n2);
    // quaternion stored (w,x,y,z) == (q.x,q.y,q.z,q.w)
    float w = q.x * P.inv_norm, x = q.y * P.inv_norm, y = q.z * P.inv_norm, z = q.w * P.inv_norm;
    P.qn[0] = w; P.qn[1] = x; P.qn[2] = y; P.qn[3] = z;
    float* Rq = P.Rq;
    Rq[0] = 1.f - 2.f * (y * y + z * z); Rq[1] = 2.f * (x * y - w * z); Rq[2] = 2.f * (x * z + w * y);
    Rq[3] = 2.f * (x * y + w * z); Rq[4] = 1.f - 2.f * (x * x + z * z); Rq[5] = 2.f * (y * z - w * x);
    Rq[6] = 2.f * (x * z - w * y); Rq[7] = 2.f * (y * z + w * x); Rq[8] = 1.f - 2.f * (x * x + y * y);
    const float s[3] = {sx, sy, sz};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            P.Wm[i * 3 + j] = (R[i * 3 + 0] * Rq[0 * 3 + j] + R[i * 3 + 1] * Rq[1 * 3 + j] +
                               R[i * 3 + 2] * Rq[2 * 3 + j]) * s[j];

    float tanx = 0.5f * cam.W / cam.fx, tany = 0.5f * cam.H / cam.fy;
    float lim_xp = (cam.W - cam.cx) / cam.fx + 0.3f * tanx;
    float lim_xn = cam.cx / cam.fx + 0.3f * tanx;
    float lim_yp = (cam.H - cam.cy) / cam.fy + 0.3f * tany;
    float lim_yn = cam.cy / cam.fy + 0.3f * tany;
    float rz = 1.f / P.pcz;
    float rz2 = rz * rz;
    float xr = P.pcx * rz, yr = P.pcy * rz;
    P.clamp_x = !(xr <= lim_xp && xr >= -lim_xn);
    P.clamp_y = !(yr <= lim_yp && yr >= -lim_yn);
    P.tx = P.pcz * fminf(lim_xp, fmaxf(-lim_xn, xr));
    P.ty = P.pcz * fminf(lim_yp, fmaxf(-lim_yn, yr));
    P.rz = rz;
    P.J00 = cam.fx * rz;
    P.J02 = -cam.fx * P.tx * rz2;
    P.J11 = cam.fy * rz;
    P.J12 = -cam.fy * P.ty * rz2;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        P.T[j] = P.J00 * P.Wm[j] + P.J02 * P.Wm[6 + j];
        P.T[3 + j] = P.J11 * P.Wm[3 + j] + P.J12 * P.Wm[6 + j];
    }
    float s00 = P.T[0] * P.T[0] + P.T[1] * P.T[1] + P.T[2] * P.T[2];
    float s01 = P.T[0] * P.T[3] + P.T[1] * P.T[4] + P.T[2] * P.T[5];
    float s11 = P.T[3] * P.T[3] + P.T[4] * P.T[4] + P.T[5] * P.T[5];
    P.a = s00 + cam.eps2d;
    P.b = s01;
    P.c = s11 + cam.eps2d;
    P.det = P.a * P.c - P.b * P.b;
    P.mx = cam.fx * P.pcx * rz + cam.cx;
    P.my = cam.fy * P.pcy * rz + cam.cy;
    return P.det > 0.f;
}

// Opacity-aware rectangular extent and screen cull.
__device__ __forceinline__ bool radius_and_cull(const Cam& cam, const Proj& P, float opac, int& rx, int& ry) {
    if (!(opac >= ALPHA_THRESHOLD)) return false;
    float extend = fminf(3.33f, sqrtf(2.0f * logf(opac / ALPHA_THRESHOLD)));
    float frx = ceilf(extend * sqrtf(P.a));
    float fry = ceilf(extend * sqrtf(P.c));
    if (frx <= cam.radius_clip && fry <= cam.radius_clip) return false;
    if (P.mx + frx <= 0.f || P.mx - frx >= (float)cam.W || P.my + fry <= 0.f || P.my - fry >= (float)cam.H)
        return false;
    rx = (int)frx;
    ry = (int)fry;
    return rx > 0 && ry > 0;
}

// Real SH basis up to `DEG` for unit direction (x,y,z): B[(DEG+1)^2].
template <int DEG>
__device__ __forceinline__ void sh_basis(float x, float y, float z, float* B) {
    B[0] = SH_C0;
    if (DEG >= 1) {
        B[1] = -SH_C1 * y; B[2] = SH_C1 * z; B[3] = -SH_C1 * x;
    }
    if (DEG >= 2) {
        float z2 = z * z, fT0B = -SH_C2B * z, fC1 = x * x - y * y, fS1 = 2.f * x * y;
        B[4] = SH_C2A * fS1; B[5] = fT0B * y; B[6] = SH_C2C * z2 - SH_C2D; B[7] = fT0B * x; B[8] = SH_C2A * fC1;
        if (DEG >= 3) {
            float fT0C = -SH_C3C * z2 + SH_C3D, fT1B = SH_C3B * z;
            float fC2 = x * fC1 - y * fS1, fS2 = x * fS1 + y * fC1;
            B[9] = -SH_C3A * fS2; B[10] = fT1B * fS1; B[11] = fT0C * y;
            B[12] = z * (SH_C3E * z2 - SH_C3F); B[13] = fT0C * x; B[14] = fT1B * fC1; B[15] = -SH_C3A * fC2;
        }
    }
}

// Coefficient fetch: NV = number of 16-B vectors holding the first 3*(DEG+1)^2 floats.
template <int DEG>
struct ShCount {
    static constexpr int NB = (DEG + 1) * (DEG + 1);
    static constexpr int NF = 3 * NB;
    static constexpr int NV = (NF + 3) / 4;
};

#ifndef SH_LDG256
#define SH_LDG256 1  // 32-byte loads of the SH row where its start is 32-B aligned (sm_100: LDG.E.256)
#endif

// One 32-byte read-only load (sm_100a).  A lane that walks its own 192-B SH row with 16-byte loads touches every
// 32-B sector twice and relies on L1 to keep it between the two; with 32-byte loads each sector is one request.
struct F8 { float v[8]; };
__device__ __forceinline__ F8 ldg256(const float* p) {
    F8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
    return r;
}

// vec_ok: 0 = scalar loads, 1 = rows start 16-B aligned, 2 = rows start 32-B aligned
template <int DEG>
__device__ __forceinline__ void load_sh(const float* __restrict__ shs, int64_t g, int sh_coeffs, int vec_ok,
                                        float* c /* NV*4 */) {
    constexpr int NV = ShCount<DEG>::NV;
    constexpr int NF = ShCount<DEG>::NF;
    const float* base = shs + g * (int64_t)sh_coeffs * 3;
#if SH_LDG256
    if (vec_ok == 2) {
#pragma unroll
        for (int i = 0; i < NV / 2; ++i) {
            const F8 v = ldg256(base + 8 * i);
#pragma unroll
            for (int k = 0; k < 8; ++k) c[8 * i + k] = v.v[k];
        }
        if (NV & 1) {
            float4 v = __ldg(reinterpret_cast<const float4*>(base) + (NV - 1));
            c[4 * NV - 4] = v.x; c[4 * NV - 3] = v.y; c[4 * NV - 2] = v.z; c[4 * NV - 1] = v.w;
        }
    } else
#endif
    if (vec_ok) {
        const float4* b4 = reinterpret_cast<const float4*>(base);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float4 v = __ldg(b4 + i);
            c[4 * i + 0] = v.x; c[4 * i + 1] = v.y; c[4 * i + 2] = v.z; c[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < NF; ++i) c[i] = __ldg(base + i);
    }
}

// ---- bulk-async (TMA engine) staging of SH rows -------------------------------------------------------------------
// A lane that fetches its own 192-B SH row with vector loads sends six to twelve single-sector requests down the
// LSU -> L1 -> L2 path, next to the kernel's other scattered traffic (record stores, the four tile-grid REDs).  In BULK
// mode every visible lane instead issues ONE cp.async.bulk (SASS UBLKCP) of its row into a padded shared-memory row:
// the copy engine moves whole rows, completion is signalled on one mbarrier per warp (expect_tx = rows x bytes), and
// the lane then reads its row with conflict-free LDS.128 (row stride 13 x 16 B: 13 l mod 8 is a permutation of the
// quarter warp).  Culled Gaussians never touch their SH row, as before.
#ifndef SH_BULK
#define SH_BULK 1
#endif
#ifndef SH_BULK_EARLY
#define SH_BULK_EARLY 0
#endif
#define SH_ROW_F 52  // padded shared-memory row stride in floats (48 used at degree 3)
#define SH_BULK_SMEM (PROJ_THREADS * SH_ROW_F * 4 + (PROJ_THREADS / 32) * 8)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// RAW: the inputs are the reference's raw parameters (gaussian_model.py:65-76) and the getters' activations
// (gaussian_model.py:90-108: exp / sigmoid; F.normalize is the quaternion normalisation done here anyway) are
// applied in registers — no activated copies of the parameter store are ever materialised.
// BULK: SH rows staged by cp.async.bulk (above); needs 16-B aligned rows and SH_BULK_SMEM bytes of dynamic shared memory.
template <int DEG, bool RAW, bool BULK>
__global__ void __launch_bounds__(PROJ_THREADS)
project_sh_fwd_kernel(const float* __restrict__ means, const float* __restrict__ quats,
                      const float* __restrict__ scales, const float* __restrict__ opacities,
                      const float* __restrict__ shs, int sh_coeffs, int vec_ok, int64_t N,
                      int32_t* __restrict__ radii, float* __restrict__ splats,
                      int32_t* __restrict__ tiles_per_gauss, int64_t* __restrict__ block_sums,
                      uint32_t* __restrict__ depth_keys, int32_t* __restrict__ gauss_ids,
                      int32_t* __restrict__ tile_grid, const Cam* __restrict__ cam_g,
                      uchar2* __restrict__ visibility, float4* __restrict__ grad_records, uint2* __restrict__ tile_rects) {
    extern __shared__ __align__(128) unsigned char s_dyn[];
    float* s_sh = reinterpret_cast<float*>(s_dyn);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_dyn + PROJ_THREADS * SH_ROW_F * 4) + (threadIdx.x >> 5);
    if (BULK) {
        if ((threadIdx.x & 31) == 0) mbar_init(s_bar, 1);
    }
    STAGE_CAM(cam_g);  // (block barrier: orders the mbarrier initialisation as well)
    int64_t g = (int64_t)blockIdx.x * PROJ_THREADS + threadIdx.x;
    if (BULK && SH_BULK_EARLY) {  // probe: fetch every row before the cull decision
        constexpr int D = DEG < 0 ? 0 : DEG;
        constexpr uint32_t ROW_BYTES = ShCount<D>::NV * 16;
        const unsigned vmask = __ballot_sync(0xffffffffu, g < N);
        if ((threadIdx.x & 31) == 0) mbar_arrive_expect_tx(s_bar, ROW_BYTES * (uint32_t)__popc(vmask));
        if (g < N) bulk_g2s(s_sh + threadIdx.x * SH_ROW_F, shs + g * (int64_t)sh_coeffs * 3, ROW_BYTES, s_bar);
    }
    int ntiles = 0;
    float px = 0.f, py = 0.f, pz = 0.f, opac = 0.f;
    Proj P;
    int rx = 0, ry = 0;
    bool ok = false;
    if (g < N) {
        px = __ldg(means + 3 * g); py = __ldg(means + 3 * g + 1); pz = __ldg(means + 3 * g + 2);
        float4 q = __ldg(reinterpret_cast<const float4*>(quats) + g);
        float sx = __ldg(scales + 3 * g), sy = __ldg(scales + 3 * g + 1), sz = __ldg(scales + 3 * g + 2);
        opac = __ldg(opacities + g);
        if (RAW) {
            sx = expf(sx); sy = expf(sy); sz = expf(sz);
            opac = 1.f / (1.f + expf(-opac));
        }
        ok = project_core(cam, px, py, pz, q, sx, sy, sz, P) && radius_and_cull(cam, P, opac, rx, ry);
    }
    if (BULK && !SH_BULK_EARLY) {  // every lane of the warp gets here
        constexpr int D = DEG < 0 ? 0 : DEG;
        constexpr uint32_t ROW_BYTES = ShCount<D>::NV * 16;
        const unsigned vmask = __ballot_sync(0xffffffffu, ok);
        if (vmask != 0u) {
            if ((threadIdx.x & 31) == 0) mbar_arrive_expect_tx(s_bar, ROW_BYTES * (uint32_t)__popc(vmask));
            if (ok) bulk_g2s(s_sh + threadIdx.x * SH_ROW_F, shs + g * (int64_t)sh_coeffs * 3, ROW_BYTES, s_bar);
        }
    }
    if (g < N) {
        // Everything that depends only on the projection goes out BEFORE the wait on the SH row — the tile-grid REDs, the
        // small per-Gaussian outputs and the first record vector travel while the copy engine fetches the row instead
        // of queueing behind it.
        float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0;
        float inv_det = 0.f;
        uint2 rect = make_uint2(0u, 0u);
        if (ok) {
            // tile rectangle — same fp32 operations as the oracle's tile_bounds
            float tile_x = P.mx / (float)TILE, tile_y = P.my / (float)TILE;
            float trx = (float)rx / (float)TILE, try_ = (float)ry / (float)TILE;
            float fw = (float)cam.tile_w, fh = (float)cam.tile_h;
            int minx = (int)fminf(fmaxf(floorf(tile_x - trx), 0.f), fw);
            int maxx = (int)fminf(fmaxf(ceilf(tile_x + trx), 0.f), fw);
            int miny = (int)fminf(fmaxf(floorf(tile_y - try_), 0.f), fh);
            int maxy = (int)fminf(fmaxf(ceilf(tile_y + try_), 0.f), fh);
            ntiles = (maxy - miny) * (maxx - minx);
            // packed tile rectangle: the binning kernels gather ONE 8-byte word per Gaussian behind the depth permutation
            // instead of radii + means2d (two 32-B sectors)
            if (ntiles > 0) rect = make_uint2((uint32_t)minx | ((uint32_t)maxx << 16), (uint32_t)miny | ((uint32_t)maxy << 16));
            if (tile_grid != nullptr && ntiles > 0) {
                // corner differences of the tile rectangle: a 2-D prefix sum (b200_tile_prep) turns the grid
                // into per-tile intersection counts, i.e. the tile ranges and the radix digit histograms,
                // without another pass over the M intersections
                const int gs = cam.tile_w + 1;
                atomicAdd(tile_grid + miny * gs + minx, 1);
                atomicAdd(tile_grid + miny * gs + maxx, -1);
                atomicAdd(tile_grid + maxy * gs + minx, -1);
                atomicAdd(tile_grid + maxy * gs + maxx, 1);
            }
            inv_det = 1.f / P.det;
            r0 = make_float4(P.mx, P.my, P.c * inv_det, -P.b * inv_det);
        } else {
            rx = ry = 0;
        }
        reinterpret_cast<int2*>(radii)[g] = make_int2(rx, ry);
        // what the caller would otherwise produce with two more launches per view: `visibility_filter = radii > 0`
        // (gsplat_backend.py:206-208) and the zeroed 48-B cotangent record the backward blend accumulates into
        if (visibility != nullptr) visibility[g] = make_uchar2(rx > 0, ry > 0);
        if (tile_rects != nullptr) tile_rects[g] = rect;
        if (grad_records != nullptr) {
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            grad_records[3 * g] = z; grad_records[3 * g + 1] = z; grad_records[3 * g + 2] = z;
        }
        float4* rec = reinterpret_cast<float4*>(splats) + 3 * g;
        rec[0] = r0;
        tiles_per_gauss[g] = ntiles;
        if (depth_keys != nullptr) {  // split-key sort: depth digits are sorted over Gaussians, before duplication
            depth_keys[g] = ntiles > 0 ? __float_as_uint(P.pcz) : 0xFFFFFFFFu;
            gauss_ids[g] = (int32_t)g;
        }
        if (ok) {
            float col[3];
            if (DEG >= 0) {
                constexpr int D = DEG < 0 ? 0 : DEG;
                constexpr int NB = ShCount<D>::NB;
                float c[ShCount<D>::NV * 4];
                float ux = px - cam.campos[0], uy = py - cam.campos[1], uz = pz - cam.campos[2];
                float inv = rsqrtf(ux * ux + uy * uy + uz * uz);
                float B[NB];
                sh_basis<D>(ux * inv, uy * inv, uz * inv, B);
                if (BULK) {
                    mbar_wait(s_bar, 0);
                    const float4* row = reinterpret_cast<const float4*>(s_sh + threadIdx.x * SH_ROW_F);
#pragma unroll
                    for (int i = 0; i < ShCount<D>::NV; ++i) {
                        const float4 v = row[i];
                        c[4 * i + 0] = v.x; c[4 * i + 1] = v.y; c[4 * i + 2] = v.z; c[4 * i + 3] = v.w;
                    }
                } else {
                    load_sh<D>(shs, g, sh_coeffs, vec_ok, c);
                }
                float r = 0.f, gg = 0.f, b = 0.f;
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    r += B[k] * c[3 * k];
                    gg += B[k] * c[3 * k + 1];
                    b += B[k] * c[3 * k + 2];
                }
                col[0] = fmaxf(r + 0.5f, 0.f); col[1] = fmaxf(gg + 0.5f, 0.f); col[2] = fmaxf(b + 0.5f, 0.f);
            } else {  // plain colours [N,3]
                col[0] = __ldg(shs + 3 * g); col[1] = __ldg(shs + 3 * g + 1); col[2] = __ldg(shs + 3 * g + 2);
            }
            r1 = make_float4(P.a * inv_det, opac, col[0], col[1]);
            r2 = make_float4(col[2], P.pcz, 0.f, 0.f);
        }
        rec[1] = r1; rec[2] = r2;
    }
    // block sum of tile counts (feeds b200_bin_scan on the 64-bit-key path; the split-key path recounts in depth order
    // and passes no buffer, which spares the CTA its closing barrier)
    if (block_sums == nullptr) return;
    __shared__ int warp_tot[PROJ_THREADS / 32];
    int v = ntiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t tot = 0;
#pragma unroll
        for (int i = 0; i < PROJ_THREADS / 32; ++i) tot += warp_tot[i];
        block_sums[blockIdx.x] = tot;
    }
}

// MODE 0: activated inputs, gradients written.  MODE 1: raw parameters (see the forward kernel), gradients
// w.r.t. the RAW parameters written.  MODE 2: raw parameters, gradients ACCUMULATED into the caller's buffers
// (the flat gradient store): views after the first add in place, culled Gaussians and SH coefficients above
// the active degree are not touched at all.
//
// The 192-B SH gradient row of a Gaussian is produced by one thread but stored by the WARP: rows are staged in
// shared memory (two halves of six 16-B vectors, padded to 7 so that staging is conflict-free) and written —
// or, in MODE 2, read-modified-written with all loads of a half issued back to back — as runs of 96 contiguous
// bytes by consecutive lanes.  A per-thread "+=" over its own row would serialise twelve dependent round trips.
template <int NB, bool ACC>
__device__ __forceinline__ void store_sh_rows(float4* __restrict__ s_buf, const float* Bk, const float* vr, unsigned vis_mask,
                                              int lane, int64_t g0, int64_t N, float* __restrict__ v_shs) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            float o4[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int e = 4 * (6 * h + i) + j, k = e / 3, ch = e - 3 * k;
                o4[j] = (k < NB) ? Bk[k < NB ? k : 0] * vr[ch] : 0.f;
            }
            s_buf[lane * 7 + i] = make_float4(o4[0], o4[1], o4[2], o4[3]);
        }
        __syncwarp();
        float4* out[6];
        float4 old[6];
        bool on[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const int f = j * 32 + lane, r = f / 6, c = f - 6 * r;
            const int64_t gg = g0 + r;
            on[j] = gg < N && (!ACC || (((vis_mask >> r) & 1u) && 4 * (6 * h + c) < 3 * NB));
            out[j] = reinterpret_cast<float4*>(v_shs) + gg * 12 + 6 * h + c;
            if (ACC && on[j]) old[j] = *out[j];
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const int f = j * 32 + lane, r = f / 6, c = f - 6 * r;
            if (on[j]) {
                float4 v = s_buf[r * 7 + c];
                if (ACC) { v.x += old[j].x; v.y += old[j].y; v.z += old[j].z; v.w += old[j].w; }
                *out[j] = v;
            }
        }
        __syncwarp();
    }
}

template <int DEG, int MODE>
__global__ void __launch_bounds__(PROJ_THREADS, 2)
project_sh_bwd_kernel(const float* __restrict__ means, const float* __restrict__ quats,
                      const float* __restrict__ scales, const float* __restrict__ opacities,
                      const float* __restrict__ shs, int sh_coeffs, int vec_ok, int64_t N,
                      const int32_t* __restrict__ radii, const float* __restrict__ v_splats,
                      float* __restrict__ v_means, float* __restrict__ v_quats, float* __restrict__ v_scales,
                      float* __restrict__ v_opacities, float* __restrict__ v_shs, const Cam* __restrict__ cam_g) {
    STAGE_CAM(cam_g);
    constexpr bool RAW = MODE != 0, ACC = MODE == 2;
    constexpr int D = DEG < 0 ? 0 : DEG;
    constexpr int NB = ShCount<D>::NB;
    __shared__ float4 s_sh[PROJ_THREADS / 32][32 * 7];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t g = (int64_t)blockIdx.x * PROJ_THREADS + threadIdx.x;
    const bool in = g < N;
    bool vis = false;
    if (in) {
        const int2 rad = __ldg(reinterpret_cast<const int2*>(radii) + g);
        vis = rad.x > 0 && rad.y > 0;
    }
    const bool coop = DEG >= 0 && vec_ok && sh_coeffs == 16;  // block-uniform
    float vm[3] = {0.f, 0.f, 0.f}, vq[4] = {0.f, 0.f, 0.f, 0.f}, vs[3] = {0.f, 0.f, 0.f}, vo = 0.f;
    const int ncoef_f = sh_coeffs * 3;
    float* vsh = (DEG >= 0) ? v_shs + g * (int64_t)ncoef_f : v_shs + 3 * g;
    float px = 0.f, py = 0.f, pz = 0.f;
    float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0, g2 = g0;
    float B[NB], vr[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < NB; ++k) B[k] = 0.f;

    // ---- phase 1: SH colour VJP (A.2) ----
    if (vis) {
        px = __ldg(means + 3 * g); py = __ldg(means + 3 * g + 1); pz = __ldg(means + 3 * g + 2);
        const float4* vrec = reinterpret_cast<const float4*>(v_splats) + 3 * g;
        g0 = __ldg(vrec); g1 = __ldg(vrec + 1); g2 = __ldg(vrec + 2);
        const float v_col[3] = {g1.z, g1.w, g2.x};
        if (DEG >= 0) {
            float c[ShCount<D>::NV * 4];
            load_sh<D>(shs, g, sh_coeffs, vec_ok, c);
            float ux = px - cam.campos[0], uy = py - cam.campos[1], uz = pz - cam.campos[2];
            float inv = rsqrtf(ux * ux + uy * uy + uz * uz);
            float x = ux * inv, y = uy * inv, z = uz * inv;
            sh_basis<D>(x, y, z, B);
            float rgb[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < NB; ++k) {
                rgb[0] += B[k] * c[3 * k]; rgb[1] += B[k] * c[3 * k + 1]; rgb[2] += B[k] * c[3 * k + 2];
            }
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) vr[ch] = (rgb[ch] + 0.5f >= 0.f) ? v_col[ch] : 0.f;
            // coefficient gradients: v_c[k][ch] = B_k * v_rgb[ch], zero above the active degree
            if (!coop) {
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    if (ACC) {
                        vsh[3 * k] += B[k] * vr[0]; vsh[3 * k + 1] += B[k] * vr[1]; vsh[3 * k + 2] += B[k] * vr[2];
                    } else {
                        vsh[3 * k] = B[k] * vr[0]; vsh[3 * k + 1] = B[k] * vr[1]; vsh[3 * k + 2] = B[k] * vr[2];
                    }
                }
                if (!ACC) for (int i = 3 * NB; i < ncoef_f; ++i) vsh[i] = 0.f;
            }
            // direction gradient
            float gk[NB];
#pragma unroll
            for (int k = 0; k < NB; ++k) gk[k] = c[3 * k] * vr[0] + c[3 * k + 1] * vr[1] + c[3 * k + 2] * vr[2];
            float vdx = 0.f, vdy = 0.f, vdz = 0.f;
            if (D >= 1) {
                vdy += -SH_C1 * gk[1]; vdz += SH_C1 * gk[2]; vdx += -SH_C1 * gk[3];
            }
            if (D >= 2) {
                vdx += SH_C2A * 2.f * y * gk[4]; vdy += SH_C2A * 2.f * x * gk[4];
                vdy += -SH_C2B * z * gk[5]; vdz += -SH_C2B * y * gk[5];
                vdz += 2.f * SH_C2C * z * gk[6];
                vdx += -SH_C2B * z * gk[7]; vdz += -SH_C2B * x * gk[7];
                vdx += SH_C2A * 2.f * x * gk[8]; vdy += -SH_C2A * 2.f * y * gk[8];
            }
            if (D >= 3) {
                float z2 = z * z, fT0C = -SH_C3C * z2 + SH_C3D, x2my2 = x * x - y * y;
                vdx += -SH_C3A * 6.f * x * y * gk[9]; vdy += -SH_C3A * 3.f * x2my2 * gk[9];
                vdx += 2.f * SH_C3B * y * z * gk[10]; vdy += 2.f * SH_C3B * x * z * gk[10]; vdz += 2.f * SH_C3B * x * y * gk[10];
                vdy += fT0C * gk[11]; vdz += -2.f * SH_C3C * z * y * gk[11];
                vdz += (3.f * SH_C3E * z2 - SH_C3F) * gk[12];
                vdx += fT0C * gk[13]; vdz += -2.f * SH_C3C * z * x * gk[13];
                vdx += SH_C3B * 2.f * x * z * gk[14]; vdy += -SH_C3B * 2.f * y * z * gk[14]; vdz += SH_C3B * x2my2 * gk[14];
                vdx += -SH_C3A * 3.f * x2my2 * gk[15]; vdy += SH_C3A * 6.f * x * y * gk[15];
            }
            float dot = x * vdx + y * vdy + z * vdz;
            vm[0] += (vdx - x * dot) * inv; vm[1] += (vdy - y * dot) * inv; vm[2] += (vdz - z * dot) * inv;
        } else if (ACC) {
            vsh[0] += v_col[0]; vsh[1] += v_col[1]; vsh[2] += v_col[2];
        } else {
            vsh[0] = v_col[0]; vsh[1] = v_col[1]; vsh[2] = v_col[2];
        }
    } else if (in && !ACC && !(DEG >= 0 && coop)) {
        const int nz = (DEG >= 0) ? ncoef_f : 3;
        for (int i = 0; i < nz; ++i) vsh[i] = 0.f;
    }
    if (DEG >= 0 && coop) {
        const unsigned vis_mask = __ballot_sync(0xffffffffu, vis);
        if (!ACC || vis_mask)
            store_sh_rows<NB, ACC>(s_sh[wid], B, vr, vis_mask, lane, g - lane, N, v_shs);
    }
    if (!in || (ACC && !vis)) return;

    // ---- phase 2: projection VJP ----
    if (vis) {
        float4 q = __ldg(reinterpret_cast<const float4*>(quats) + g);
        float sx = __ldg(scales + 3 * g), sy = __ldg(scales + 3 * g + 1), sz = __ldg(scales + 3 * g + 2);
        if (RAW) { sx = expf(sx); sy = expf(sy); sz = expf(sz); }
        float v_mx = g0.x, v_my = g0.y, v_A = g0.z, v_B = g0.w, v_C = g1.x;
        vo = g1.y;
        float v_depth = g2.y;
        Proj P;
        project_core(cam, px, py, pz, q, sx, sy, sz, P);
        // ---- conic -> 2-D covariance (A.6) ----
        float inv_det = 1.f / P.det;
        float A = P.c * inv_det, Bc = -P.b * inv_det, C = P.a * inv_det;
        float hB = 0.5f * v_B;
        float xv00 = A * v_A + Bc * hB, xv01 = A * hB + Bc * v_C;
        float xv10 = Bc * v_A + C * hB, xv11 = Bc * hB + C * v_C;
        float G00 = -(xv00 * A + xv01 * Bc), G01 = -(xv00 * Bc + xv01 * C);
        float G10 = -(xv10 * A + xv11 * Bc), G11 = -(xv10 * Bc + xv11 * C);
        float Gs = 0.5f * (G01 + G10);
        // S2 = T T^T  ->  v_T = 2 G T
        float vT[6];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            vT[j] = 2.f * (G00 * P.T[j] + Gs * P.T[3 + j]);
            vT[3 + j] = 2.f * (Gs * P.T[j] + G11 * P.T[3 + j]);
        }
        // T = J Wm: v_J = v_T Wm^T (only 4 live entries), v_Wm = J^T v_T
        float vJ00 = 0.f, vJ02 = 0.f, vJ11 = 0.f, vJ12 = 0.f;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            vJ00 += vT[j] * P.Wm[j];
            vJ02 += vT[j] * P.Wm[6 + j];
            vJ11 += vT[3 + j] * P.Wm[3 + j];
            vJ12 += vT[3 + j] * P.Wm[6 + j];
        }
        float vW[9];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            vW[j] = P.J00 * vT[j];
            vW[3 + j] = P.J11 * vT[3 + j];
            vW[6 + j] = P.J02 * vT[j] + P.J12 * vT[3 + j];
        }
        // Wm = R Rq diag(s): v_(Rq diag s) = R^T v_Wm
        const float* R = cam.R;
        const float s[3] = {sx, sy, sz};
        float vRq[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                float vMij = R[0 * 3 + i] * vW[0 * 3 + j] + R[1 * 3 + i] * vW[1 * 3 + j] + R[2 * 3 + i] * vW[2 * 3 + j];
                vs[j] += P.Rq[i * 3 + j] * vMij;
                vRq[i * 3 + j] = vMij * s[j];
            }
        {
            float w = P.qn[0], x = P.qn[1], y = P.qn[2], z = P.qn[3];
            float vw = 2.f * (-z * vRq[1] + y * vRq[2] + z * vRq[3] - x * vRq[5] - y * vRq[6] + x * vRq[7]);
            float vx = 2.f * (y * vRq[1] + z * vRq[2] + y * vRq[3] - 2.f * x * vRq[4] - w * vRq[5] + z * vRq[6] + w * vRq[7] - 2.f * x * vRq[8]);
            float vy = 2.f * (-2.f * y * vRq[0] + x * vRq[1] + w * vRq[2] + x * vRq[3] + z * vRq[5] - w * vRq[6] + z * vRq[7] - 2.f * y * vRq[8]);
            float vz = 2.f * (-2.f * z * vRq[0] - w * vRq[1] + x * vRq[2] + w * vRq[3] - 2.f * z * vRq[4] + y * vRq[5] + x * vRq[6] + y * vRq[7]);
            float dot = w * vw + x * vx + y * vy + z * vz;
            vq[0] = (vw - w * dot) * P.inv_norm; vq[1] = (vx - x * dot) * P.inv_norm;
            vq[2] = (vy - y * dot) * P.inv_norm; vq[3] = (vz - z * dot) * P.inv_norm;
        }
        // ---- perspective projection VJP: J, mean2d, depth -> camera-space mean ----
        float rz = P.rz, rz2 = rz * rz, rz3 = rz2 * rz;
        float vpx = cam.fx * rz * v_mx;
        float vpy = cam.fy * rz * v_my;
        float vpz = -(cam.fx * P.pcx * v_mx + cam.fy * P.pcy * v_my) * rz2 + v_depth;
        vpz += -cam.fx * rz2 * vJ00 - cam.fy * rz2 * vJ11;
        if (!P.clamp_x) { vpx += -cam.fx * rz2 * vJ02; vpz += 2.f * cam.fx * P.tx * rz3 * vJ02; }
        else            { vpz += cam.fx * P.tx * rz3 * vJ02; }
        if (!P.clamp_y) { vpy += -cam.fy * rz2 * vJ12; vpz += 2.f * cam.fy * P.ty * rz3 * vJ12; }
        else            { vpz += cam.fy * P.ty * rz3 * vJ12; }
        vm[0] += R[0] * vpx + R[3] * vpy + R[6] * vpz;
        vm[1] += R[1] * vpx + R[4] * vpy + R[7] * vpz;
        vm[2] += R[2] * vpx + R[5] * vpy + R[8] * vpz;
        if (RAW) {  // chain through exp (scales) and sigmoid (opacity); vs[] was w.r.t. the activated scale
            vs[0] *= sx; vs[1] *= sy; vs[2] *= sz;
            const float o = 1.f / (1.f + expf(-__ldg(opacities + g)));
            vo *= o * (1.f - o);
        }
    }
    if (ACC) {  // all loads first: one round trip instead of eleven dependent ones
        const float m0 = v_means[3 * g], m1 = v_means[3 * g + 1], m2 = v_means[3 * g + 2];
        const float4 oq = reinterpret_cast<float4*>(v_quats)[g];
        const float s0 = v_scales[3 * g], s1 = v_scales[3 * g + 1], s2 = v_scales[3 * g + 2];
        const float o0 = v_opacities[g];
        v_means[3 * g] = m0 + vm[0]; v_means[3 * g + 1] = m1 + vm[1]; v_means[3 * g + 2] = m2 + vm[2];
        reinterpret_cast<float4*>(v_quats)[g] = make_float4(oq.x + vq[0], oq.y + vq[1], oq.z + vq[2], oq.w + vq[3]);
        v_scales[3 * g] = s0 + vs[0]; v_scales[3 * g + 1] = s1 + vs[1]; v_scales[3 * g + 2] = s2 + vs[2];
        v_opacities[g] = o0 + vo;
    } else {
        v_means[3 * g] = vm[0]; v_means[3 * g + 1] = vm[1]; v_means[3 * g + 2] = vm[2];
        reinterpret_cast<float4*>(v_quats)[g] = make_float4(vq[0], vq[1], vq[2], vq[3]);
        v_scales[3 * g] = vs[0]; v_scales[3 * g + 1] = vs[1]; v_scales[3 * g + 2] = vs[2];
        v_opacities[g] = vo;
    }
}

// Builds the packed camera block on the device from a device-resident pose and intrinsics.
__global__ void cam_pack_kernel(const float* __restrict__ viewmat, const float* __restrict__ K, float near_plane,
                                float far_plane, float eps2d, float radius_clip, int W, int H, Cam* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    Cam cam;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) cam.R[i * 3 + j] = viewmat[i * 4 + j];
        cam.t[i] = viewmat[i * 4 + 3];
    }
    for (int j = 0; j < 3; ++j)
        cam.campos[j] = -(cam.R[0 * 3 + j] * cam.t[0] + cam.R[1 * 3 + j] * cam.t[1] + cam.R[2 * 3 + j] * cam.t[2]);
    cam.fx = K[0]; cam.fy = K[4]; cam.cx = K[2]; cam.cy = K[5];
    cam.near_plane = near_plane; cam.far_plane = far_plane; cam.eps2d = eps2d; cam.radius_clip = radius_clip;
    cam.W = W; cam.H = H;
    cam.tile_w = (W + TILE - 1) / TILE;
    cam.tile_h = (H + TILE - 1) / TILE;
    *out = cam;
}

#define CAM_MAX_DEV 16
cudaEvent_t g_cam_event[CAM_MAX_DEV];
cudaStream_t g_cam_stream[CAM_MAX_DEV];
bool g_cam_event_ok[CAM_MAX_DEV];

// Device address of the camera block the next projection launch on `st` reads: the caller's device block as is,
// or the staging slot after a stream-ordered upload of a host block.  `*slot_dev` receives the device ordinal when
// the staging slot is in use (for cam_used), else -1.  Returns nullptr after recording an error.
const Cam* load_cam(const float* cam_host, const float* cam_dev, int W, int H, cudaStream_t st, int* slot_dev) {
    *slot_dev = -1;
    if (cam_dev) return reinterpret_cast<const Cam*>(cam_dev);
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= CAM_MAX_DEV) { b200_set_error("device ordinal %d out of range", dev); return nullptr; }
    if (!g_cam_event_ok[dev]) {
        if (cudaEventCreateWithFlags(&g_cam_event[dev], cudaEventDisableTiming) != cudaSuccess) { b200_check_launch("cam event"); return nullptr; }
        g_cam_event_ok[dev] = true;
        g_cam_stream[dev] = st;
    } else if (g_cam_stream[dev] != st) {
        cudaStreamWaitEvent(st, g_cam_event[dev], 0);  // the previous reader of the staging slot ran on another stream
    }
    Cam cam = b200_make_cam(cam_host, W, H);
    cudaMemcpyToSymbolAsync(c_cam_staging, &cam, sizeof(Cam), 0, cudaMemcpyHostToDevice, st);
    void* addr = nullptr;
    if (cudaGetSymbolAddress(&addr, c_cam_staging) != cudaSuccess) { b200_check_launch("cam staging address"); return nullptr; }
    *slot_dev = dev;
    return reinterpret_cast<const Cam*>(addr);
}

void cam_used(int dev, cudaStream_t st) {
    if (dev < 0) return;
    cudaEventRecord(g_cam_event[dev], st);
    g_cam_stream[dev] = st;
}

}  // namespace

static_assert(sizeof(Cam) <= B200_CAM_DEV_FLOATS * sizeof(float), "packed camera block too small");

extern "C" int b200_cam_pack(const float* viewmat, const float* K, float near_plane, float far_plane, float eps2d,
                             float radius_clip, int width, int height, float* cam_dev, void* stream) {
    B200_REQUIRE(viewmat && K && cam_dev && width > 0 && height > 0, "bad argument");
    B200_REQUIRE(((uintptr_t)cam_dev & 15) == 0, "misaligned cam_dev");
    cam_pack_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(viewmat, K, near_plane, far_plane, eps2d, radius_clip, width, height,
                                                        reinterpret_cast<Cam*>(cam_dev));
    return b200_check_launch("cam_pack");
}

extern "C" int64_t b200_project_blocks(int64_t N) { return (N + PROJ_THREADS - 1) / PROJ_THREADS; }

static bool sh_args_ok(int sh_coeffs, int sh_degree) {
    if (sh_degree < 0) return true;
    return sh_degree <= 3 && (sh_degree + 1) * (sh_degree + 1) <= sh_coeffs;
}

static int project_fwd_impl(bool raw, const float* means, const float* quats, const float* scales,
                                   const float* opacities, const float* shs, int sh_coeffs, int sh_degree,
                                   int64_t N, const float* cam_host, const float* cam_dev, int width, int height,
                                   int32_t* radii, float* splats, int32_t* tiles_per_gauss, int64_t* block_sums,
                                   uint32_t* depth_keys, int32_t* gauss_ids, int32_t* tile_grid, uint8_t* visibility,
                                   float* grad_records, void* tile_rects, void* stream) {
    B200_REQUIRE(N >= 0 && width > 0 && height > 0, "bad sizes");
    B200_REQUIRE(((uintptr_t)tile_rects & 7) == 0, "misaligned tile_rects");
    B200_REQUIRE(sh_args_ok(sh_coeffs, sh_degree), "sh_degree/sh_coeffs mismatch");
    B200_REQUIRE(((uintptr_t)grad_records & 15) == 0 && ((uintptr_t)visibility & 1) == 0, "misaligned auxiliary output");
    if (tile_grid != nullptr) {
        const size_t cells = (size_t)((width + TILE - 1) / TILE + 1) * (size_t)((height + TILE - 1) / TILE + 1);
        cudaMemsetAsync(tile_grid, 0, cells * sizeof(int32_t), (cudaStream_t)stream);
    }
    if (N == 0) return B200_OK;
    B200_REQUIRE(means && quats && scales && opacities && shs && radii && splats && tiles_per_gauss, "null pointer");
    B200_REQUIRE(block_sums != nullptr || depth_keys != nullptr, "block_sums may be omitted only on the split-key path (depth_keys given)");
    B200_REQUIRE((cam_host != nullptr) != (cam_dev != nullptr), "exactly one of cam_host / cam_dev");
    B200_REQUIRE(((uintptr_t)quats & 15) == 0 && ((uintptr_t)splats & 15) == 0 && ((uintptr_t)radii & 7) == 0, "misaligned pointer");
    B200_REQUIRE((depth_keys == nullptr) == (gauss_ids == nullptr), "depth_keys and gauss_ids go together");
    int vec_ok = sh_degree >= 0 && ((sh_coeffs * 12) % 16 == 0) && (((uintptr_t)shs & 15) == 0);
    if (vec_ok && ((sh_coeffs * 12) % 32 == 0) && (((uintptr_t)shs & 31) == 0)) vec_ok = 2;
    dim3 grid((unsigned)b200_project_blocks(N)), block(PROJ_THREADS);
    cudaStream_t st = (cudaStream_t)stream;
    int cdev = -1;
    const Cam* cam_g = load_cam(cam_host, cam_dev, width, height, st, &cdev);
    if (cam_g == nullptr) return B200_ECUDA;
    // BULK staging of the SH rows needs 16-B aligned rows; the dynamic shared-memory limit is raised once per instantiation
    const bool bulk = SH_BULK && sh_degree >= 0 && vec_ok != 0;
#define FWD_ARGS means, quats, scales, opacities, shs, sh_coeffs, vec_ok, N, radii, splats, tiles_per_gauss, block_sums, depth_keys, gauss_ids, tile_grid, cam_g, \
                 reinterpret_cast<uchar2*>(visibility), reinterpret_cast<float4*>(grad_records), reinterpret_cast<uint2*>(tile_rects)
#define LAUNCH_B(D, R) do { \
        static bool attr_set[B200_MAX_DEVICES] = {}; \
        if (b200_first_use_on_device(attr_set)) cudaFuncSetAttribute(project_sh_fwd_kernel<D, R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SH_BULK_SMEM); \
        project_sh_fwd_kernel<D, R, true><<<grid, block, SH_BULK_SMEM, st>>>(FWD_ARGS); } while (0)
#define LAUNCH(D) do { if (bulk && D >= 0) { if (raw) LAUNCH_B((D < 0 ? 0 : D), true); else LAUNCH_B((D < 0 ? 0 : D), false); } \
                       else if (raw) project_sh_fwd_kernel<D, true, false><<<grid, block, 0, st>>>(FWD_ARGS); \
                       else project_sh_fwd_kernel<D, false, false><<<grid, block, 0, st>>>(FWD_ARGS); } while (0)
    switch (sh_degree) {
        case 0: LAUNCH(0); break;
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 3: LAUNCH(3); break;
        default: LAUNCH(-1); break;
    }
#undef LAUNCH
#undef LAUNCH_B
#undef FWD_ARGS
    cam_used(cdev, st);
    return b200_check_launch("project_sh_fwd");
}

extern "C" int b200_project_sh_fwd(const float* means, const float* quats, const float* scales, const float* opacities,
                                   const float* shs, int sh_coeffs, int sh_degree, int64_t N, const float* cam_host,
                                   const float* cam_dev, int width, int height, int32_t* radii, float* splats,
                                   int32_t* tiles_per_gauss, int64_t* block_sums, uint32_t* depth_keys, int32_t* gauss_ids,
                                   int32_t* tile_grid, uint8_t* visibility, float* grad_records, void* tile_rects,
                                   void* stream) {
    return project_fwd_impl(false, means, quats, scales, opacities, shs, sh_coeffs, sh_degree, N, cam_host, cam_dev, width,
                            height, radii, splats, tiles_per_gauss, block_sums, depth_keys, gauss_ids, tile_grid, visibility,
                            grad_records, tile_rects, stream);
}

extern "C" int b200_project_sh_fwd_raw(const float* xyz, const float* rotation, const float* scaling, const float* opacity,
                                       const float* features, int sh_coeffs, int sh_degree, int64_t N, const float* cam_host,
                                       const float* cam_dev, int width, int height, int32_t* radii, float* splats,
                                       int32_t* tiles_per_gauss, int64_t* block_sums, uint32_t* depth_keys,
                                       int32_t* gauss_ids, int32_t* tile_grid, uint8_t* visibility, float* grad_records,
                                       void* tile_rects, void* stream) {
    return project_fwd_impl(true, xyz, rotation, scaling, opacity, features, sh_coeffs, sh_degree, N, cam_host, cam_dev, width,
                            height, radii, splats, tiles_per_gauss, block_sums, depth_keys, gauss_ids, tile_grid, visibility,
                            grad_records, tile_rects, stream);
}

static int project_bwd_impl(int mode, const float* means, const float* quats, const float* scales,
                                   const float* opacities, const float* shs, int sh_coeffs, int sh_degree,
                                   int64_t N, const float* cam_host, const float* cam_dev, int width, int height,
                                   const int32_t* radii, const float* v_splats, float* v_means, float* v_quats, float* v_scales,
                                   float* v_opacities, float* v_shs, void* stream) {
    B200_REQUIRE(N >= 0 && width > 0 && height > 0, "bad sizes");
    B200_REQUIRE(sh_args_ok(sh_coeffs, sh_degree), "sh_degree/sh_coeffs mismatch");
    if (N == 0) return B200_OK;
    B200_REQUIRE((cam_host != nullptr) != (cam_dev != nullptr), "exactly one of cam_host / cam_dev");
    B200_REQUIRE(means && quats && scales && opacities && shs && radii && v_splats && v_means && v_quats && v_scales && v_opacities && v_shs, "null pointer");
    B200_REQUIRE(((uintptr_t)quats & 15) == 0 && ((uintptr_t)v_splats & 15) == 0 && ((uintptr_t)v_quats & 15) == 0 && ((uintptr_t)radii & 7) == 0, "misaligned pointer");
    int vec_ok = sh_degree >= 0 && ((sh_coeffs * 12) % 16 == 0) && (((uintptr_t)shs & 15) == 0) && (((uintptr_t)v_shs & 15) == 0);
    if (vec_ok && ((sh_coeffs * 12) % 32 == 0) && (((uintptr_t)shs & 31) == 0)) vec_ok = 2;
    dim3 grid((unsigned)b200_project_blocks(N)), block(PROJ_THREADS);
    cudaStream_t st = (cudaStream_t)stream;
    int cdev = -1;
    const Cam* cam_g = load_cam(cam_host, cam_dev, width, height, st, &cdev);
    if (cam_g == nullptr) return B200_ECUDA;
#define ARGS means, quats, scales, opacities, shs, sh_coeffs, vec_ok, N, radii, v_splats, v_means, v_quats, v_scales, v_opacities, v_shs, cam_g
#define LAUNCH(D) do { if (mode == 0) project_sh_bwd_kernel<D, 0><<<grid, block, 0, st>>>(ARGS); \
                       else if (mode == 1) project_sh_bwd_kernel<D, 1><<<grid, block, 0, st>>>(ARGS); \
                       else project_sh_bwd_kernel<D, 2><<<grid, block, 0, st>>>(ARGS); } while (0)
    switch (sh_degree) {
        case 0: LAUNCH(0); break;
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 3: LAUNCH(3); break;
        default: LAUNCH(-1); break;
    }
#undef LAUNCH
#undef ARGS
    cam_used(cdev, st);
    return b200_check_launch("project_sh_bwd");
}

extern "C" int b200_project_sh_bwd(const float* means, const float* quats, const float* scales, const float* opacities,
                                   const float* shs, int sh_coeffs, int sh_degree, int64_t N, const float* cam_host,
                                   const float* cam_dev, int width, int height, const int32_t* radii, const float* v_splats,
                                   float* v_means, float* v_quats, float* v_scales, float* v_opacities, float* v_shs,
                                   void* stream) {
    return project_bwd_impl(0, means, quats, scales, opacities, shs, sh_coeffs, sh_degree, N, cam_host, cam_dev, width, height,
                            radii, v_splats, v_means, v_quats, v_scales, v_opacities, v_shs, stream);
}

extern "C" int b200_project_sh_bwd_raw(const float* xyz, const float* rotation, const float* scaling, const float* opacity,
                                       const float* features, int sh_coeffs, int sh_degree, int64_t N, const float* cam_host,
                                       const float* cam_dev, int width, int height, const int32_t* radii,
                                       const float* v_splats, float* v_xyz, float* v_rotation, float* v_scaling,
                                       float* v_opacity, float* v_features, int accumulate, void* stream) {
    return project_bwd_impl(accumulate ? 2 : 1, xyz, rotation, scaling, opacity, features, sh_coeffs, sh_degree, N, cam_host,
                            cam_dev, width, height, radii, v_splats, v_xyz, v_rotation, v_scaling, v_opacity, v_features, stream);
}
