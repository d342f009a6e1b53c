// densify.cu — SURVEY.md §8 f2: the data movement of the reference's densify_and_prune
// (src/gaussians/gaussian_model.py:110-366) as ONE gather over the flat store.
//
// The reference rebuilds every parameter tensor and both Adam moments with torch.cat / boolean masks once per
// clone, split-removal, split-append and prune (6 attributes x 3 buffers x 4 operations of full-size copies).
// Here the host-side logic (store.py) only computes, for every row of the NEW population, which old row it
// comes from and whether it keeps its optimiser state; this kernel then writes the new parameter buffer and
// both new moment buffers in one pass: 16 threads per destination row (12 x 16 B of SH coefficients, xyz,
// scaling, rotation, opacity).  Rows that are new (clones, split children) get zero moments, as the reference's
// _replace_param_in_optimizer does for appended rows.
#include "common.cuh"

struct StoreLayout { long long xyz, scaling, rotation, opacity, features; };  // block offsets in floats

namespace {

__device__ __forceinline__ void copy_row_part(const float* __restrict__ src, float* __restrict__ dst, const StoreLayout& ls,
                                              const StoreLayout& ld, long long s, long long j, int part, bool zero) {
    if (part < 12) {
        const float4 v = zero ? make_float4(0.f, 0.f, 0.f, 0.f)
                              : __ldg(reinterpret_cast<const float4*>(src + ls.features) + s * 12 + part);
        reinterpret_cast<float4*>(dst + ld.features)[j * 12 + part] = v;
    } else if (part == 12) {
#pragma unroll
        for (int k = 0; k < 3; ++k) dst[ld.xyz + 3 * j + k] = zero ? 0.f : __ldg(src + ls.xyz + 3 * s + k);
    } else if (part == 13) {
#pragma unroll
        for (int k = 0; k < 3; ++k) dst[ld.scaling + 3 * j + k] = zero ? 0.f : __ldg(src + ls.scaling + 3 * s + k);
    } else if (part == 14) {
        const float4 v = zero ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(reinterpret_cast<const float4*>(src + ls.rotation) + s);
        reinterpret_cast<float4*>(dst + ld.rotation)[j] = v;
    } else {
        dst[ld.opacity + j] = zero ? 0.f : __ldg(src + ls.opacity + s);
    }
}

__global__ void __launch_bounds__(256)
store_gather_kernel(const float* __restrict__ p_src, const float* __restrict__ m_src, const float* __restrict__ v_src,
                    float* __restrict__ p_dst, float* __restrict__ m_dst, float* __restrict__ v_dst, StoreLayout ls,
                    StoreLayout ld, const long long* __restrict__ src_row, const unsigned char* __restrict__ keep_state,
                    long long n_dst) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long j = t >> 4;
    const int part = (int)(t & 15);
    if (j >= n_dst) return;
    const long long s = __ldg(src_row + j);
    copy_row_part(p_src, p_dst, ls, ld, s, j, part, false);
    if (m_src != nullptr) {
        const bool zero = keep_state != nullptr && keep_state[j] == 0;
        copy_row_part(m_src, m_dst, ls, ld, s, j, part, zero);
        copy_row_part(v_src, v_dst, ls, ld, s, j, part, zero);
    }
}

}  // namespace

// layout_*_host: 5 int64 block offsets {xyz, scaling, rotation, opacity, features} (floats, multiples of 4).
extern "C" int b200_store_gather(const float* params_src, const float* exp_avg_src, const float* exp_avg_sq_src,
                                 const int64_t* layout_src_host, float* params_dst, float* exp_avg_dst,
                                 float* exp_avg_sq_dst, const int64_t* layout_dst_host, const int64_t* src_row,
                                 const uint8_t* keep_state, int64_t n_dst, void* stream) {
    B200_REQUIRE(n_dst >= 0 && layout_src_host && layout_dst_host, "bad argument");
    if (n_dst == 0) return B200_OK;
    B200_REQUIRE(params_src && params_dst && src_row, "null pointer");
    B200_REQUIRE((exp_avg_src == nullptr) == (exp_avg_sq_src == nullptr) && (exp_avg_src == nullptr) == (exp_avg_dst == nullptr) &&
                     (exp_avg_src == nullptr) == (exp_avg_sq_dst == nullptr), "moment buffers go together");
    StoreLayout ls{layout_src_host[0], layout_src_host[1], layout_src_host[2], layout_src_host[3], layout_src_host[4]};
    StoreLayout ld{layout_dst_host[0], layout_dst_host[1], layout_dst_host[2], layout_dst_host[3], layout_dst_host[4]};
    B200_REQUIRE(((ls.features | ls.rotation | ld.features | ld.rotation) & 3) == 0, "blocks must be 16-byte aligned");
    const long long threads = (long long)n_dst * 16;
    store_gather_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        params_src, exp_avg_src, exp_avg_sq_src, params_dst, exp_avg_dst, exp_avg_sq_dst, ls, ld,
        reinterpret_cast<const long long*>(src_row), keep_state, (long long)n_dst);
    return b200_check_launch("store_gather");
}
