// collective.cu — the gradient all-reduce of the view-sharded training step as THIS library's kernel over NVSwitch
// multicast memory (NVLink SHARP), instead of an NCCL call (BASELINE.json north_star: "parameter gradients ... reduced
// ... over NVLink").
//
// Every rank holds its replica of the flat gradient buffer in SYMMETRIC memory (same size on every rank, mapped into one
// multicast address range; the allocation and the rendezvous are host plumbing: torch.distributed._symmetric_memory).
// After a cross-rank barrier (all local backward passes have finished writing), rank r owns the slice
// [r * chunk, (r + 1) * chunk): it reads each 16-byte vector of the slice ONCE through the multicast address with
//   multimem.ld_reduce.add.v4.f32   — the switch fetches the vector from all N replicas and returns their sum,
// scales it (1 / global views: the MEAN loss of src/training/trainer.py:499-500) and writes it back with
//   multimem.st.v4.f32              — the switch stores it into all N replicas.
// Per GPU and step that is B / N bytes pulled and B / N bytes pushed per peer link-direction, against NCCL ring's
// 2 (N-1)/N B; the reduction is done once per element, by one rank, so the replicas are bit-identical by construction.
// A second barrier (host side) orders the stores before any rank reads its replica.
//
// b200_allreduce_adam_multimem fuses the optimiser into the same pass (reduce-scatter -> Adam on the rank's slice ->
// all-gather of the PARAMETERS): the reduced gradient never returns to memory, the 28 B/parameter optimiser pass is
// divided by N, and the moments are sharded (each rank keeps only its slice).
#include "common.cuh"

namespace {

// The ordering between the ranks' producers and these accesses comes from the barriers around the kernel (system-scope
// release / acquire on the signal pads), so the accesses themselves are weak.
__device__ __forceinline__ float4 mm_ld_reduce(const float* mc_addr) {
    float4 v;
    asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc_addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(float* mc_addr, float4 v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

#define MM_THREADS 512

template <int MM_UNROLL>
__global__ void __launch_bounds__(MM_THREADS)
multimem_allreduce_kernel(float* __restrict__ mc, long long v4_begin, long long v4_end, float scale) {
    const long long stride = (long long)gridDim.x * MM_THREADS;
    long long i = v4_begin + (long long)blockIdx.x * MM_THREADS + threadIdx.x;
    // MM_UNROLL independent multicast loads in flight per thread before the first store
    for (; i + (MM_UNROLL - 1) * stride < v4_end; i += MM_UNROLL * stride) {
        float4 v[MM_UNROLL];
#pragma unroll
        for (int k = 0; k < MM_UNROLL; ++k) v[k] = mm_ld_reduce(mc + 4 * (i + k * stride));
#pragma unroll
        for (int k = 0; k < MM_UNROLL; ++k) {
            v[k].x *= scale; v[k].y *= scale; v[k].z *= scale; v[k].w *= scale;
            mm_st(mc + 4 * (i + k * stride), v[k]);
        }
    }
    for (; i < v4_end; i += stride) {
        float4 v = mm_ld_reduce(mc + 4 * i);
        v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
        mm_st(mc + 4 * i, v);
    }
}

// ---- reduce-scatter -> Adam on the rank's slice -> all-gather of the parameters, one kernel ----------------------------
// Same segment table as adam.cu (per-segment learning rates, the dc / rest alternation inside the SH block, the cap on
// the log-scales).  For element i of this rank's slice:
//   g  = scale * multimem.ld_reduce(grads[i])          in-switch sum over the replicas: the reduce-scatter
//   m, v, p: this rank's LOCAL moments and parameter (the moments of other slices are never touched: sharded)
//   multimem.st(params[i]) <- updated p                 every replica receives the new parameter: the all-gather
// and, inside [keep_begin, keep_end) (the position block), multimem.st(grads[i]) <- g so that every rank can take the
// densification statistic (||reduced xyz.grad||, trainer.py:786-791) from its own replica afterwards.
struct MmAdamSeg {
    long long begin, end;
    float lr, lr_alt;
    int period, split;
    float clamp_max;
};
struct MmAdamSegs { MmAdamSeg s[8]; int n; };

template <int UNROLL>
__global__ void __launch_bounds__(MM_THREADS)
multimem_adam_kernel(float* __restrict__ mc_grads, float* __restrict__ mc_params, const float* __restrict__ params,
                     float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq, long long v4_begin, long long v4_end,
                     float scale, MmAdamSegs segs, float b1, float b2, float eps, float inv_bc1, float inv_sqrt_bc2,
                     long long keep_begin, long long keep_end) {
    const long long stride = (long long)gridDim.x * MM_THREADS;
    for (long long i0 = v4_begin + (long long)blockIdx.x * MM_THREADS + threadIdx.x; i0 < v4_end; i0 += UNROLL * stride) {
        float4 g[UNROLL];
#pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            const long long i4 = i0 + k * stride;
            if (i4 < v4_end) g[k] = mm_ld_reduce(mc_grads + 4 * i4);
        }
#pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            const long long i4 = i0 + k * stride;
            if (i4 >= v4_end) continue;
            const long long i = 4 * i4;
            float4 gg = g[k];
            gg.x *= scale; gg.y *= scale; gg.z *= scale; gg.w *= scale;
            if (i >= keep_begin && i < keep_end) mm_st(mc_grads + i, gg);
            int si = -1;
#pragma unroll
            for (int q = 0; q < 8; ++q)
                if (q < segs.n && i >= segs.s[q].begin && i < segs.s[q].end) si = q;
            if (si < 0) continue;  // padding between the blocks
            const MmAdamSeg sg = segs.s[si];
            float4 pp = reinterpret_cast<const float4*>(params)[i4];
            float4 mm = reinterpret_cast<float4*>(exp_avg)[i4], vv = reinterpret_cast<float4*>(exp_avg_sq)[i4];
            float* pa = reinterpret_cast<float*>(&pp);
            const float* ga = reinterpret_cast<const float*>(&gg);
            float* ma = reinterpret_cast<float*>(&mm);
            float* va = reinterpret_cast<float*>(&vv);
            const int phase = sg.period ? (int)((i - sg.begin) % sg.period) : 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float lr = sg.lr;
                if (sg.period) {
                    int ph = phase + j;
                    if (ph >= sg.period) ph -= sg.period;
                    if (ph >= sg.split) lr = sg.lr_alt;
                }
                const float gj = ga[j];
                ma[j] = b1 * ma[j] + (1.f - b1) * gj;       // identical arithmetic to adam_kernel (adam.cu)
                va[j] = b2 * va[j] + (1.f - b2) * gj * gj;
                const float denom = sqrtf(va[j]) * inv_sqrt_bc2 + eps;
                pa[j] = fminf(pa[j] - (lr * inv_bc1) * (ma[j] / denom), sg.clamp_max);
            }
            reinterpret_cast<float4*>(exp_avg)[i4] = mm;
            reinterpret_cast<float4*>(exp_avg_sq)[i4] = vv;
            mm_st(mc_params + i, pp);
        }
    }
}

// every replica of [v4_begin, v4_end) <- this rank's local copy (hands a rank's slice of a sharded buffer to all ranks)
__global__ void __launch_bounds__(MM_THREADS)
multimem_share_slice_kernel(float* __restrict__ mc, const float* __restrict__ local, long long v4_begin, long long v4_end) {
    const long long stride = (long long)gridDim.x * MM_THREADS;
    for (long long i4 = v4_begin + (long long)blockIdx.x * MM_THREADS + threadIdx.x; i4 < v4_end; i4 += stride)
        mm_st(mc + 4 * i4, reinterpret_cast<const float4*>(local)[i4]);
}

static inline void mm_slice(int64_t n_floats, int rank, int world, long long* b, long long* e) {
    const long long n4 = n_floats / 4;
    const long long chunk = (n4 + world - 1) / world;
    *b = (long long)rank * chunk;
    *e = (*b + chunk < n4) ? *b + chunk : n4;
    if (*b > n4) *b = n4;
}

}  // namespace

extern "C" void b200_multimem_slice(int64_t n_floats, int rank, int world, int64_t* begin_float, int64_t* end_float) {
    long long b, e;
    mm_slice(n_floats, rank, world, &b, &e);
    if (begin_float) *begin_float = 4 * b;
    if (end_float) *end_float = 4 * e;
}

extern "C" int b200_allreduce_adam_multimem(float* mc_grads, float* mc_params, const float* params, float* exp_avg,
                                            float* exp_avg_sq, int64_t n_floats, int rank, int world, float grad_scale,
                                            const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                                            int64_t keep_grad_begin, int64_t keep_grad_end, int max_ctas, void* stream) {
    B200_REQUIRE(n_floats >= 0 && n_floats % 4 == 0 && world >= 1 && rank >= 0 && rank < world && n_seg >= 0 && n_seg <= 8 && step >= 1,
                 "bad argument");
    if (n_floats == 0) return B200_OK;
    B200_REQUIRE(mc_grads && mc_params && params && exp_avg && exp_avg_sq && seg_host, "null pointer");
    B200_REQUIRE((((uintptr_t)mc_grads | (uintptr_t)mc_params | (uintptr_t)params | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
                 "misaligned buffer");
    B200_REQUIRE(keep_grad_begin % 4 == 0 && keep_grad_end % 4 == 0 && keep_grad_begin <= keep_grad_end, "bad keep range");
    MmAdamSegs segs;
    segs.n = n_seg;
    for (int k = 0; k < n_seg; ++k) {
        const double* r = seg_host + 7 * k;
        segs.s[k].begin = (long long)r[0]; segs.s[k].end = (long long)r[1];
        segs.s[k].lr = (float)r[2]; segs.s[k].lr_alt = (float)r[3];
        segs.s[k].period = (int)r[4]; segs.s[k].split = (int)r[5];
        segs.s[k].clamp_max = (float)r[6];
        B200_REQUIRE(segs.s[k].begin % 4 == 0 && segs.s[k].end <= n_floats, "bad segment");
    }
    long long b, e;
    mm_slice(n_floats, rank, world, &b, &e);
    if (b >= e) return B200_OK;
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    const long long cap = (max_ctas & 0xFFFF) > 0 ? (max_ctas & 0xFFFF) : 64;
    long long ctas = (e - b + MM_THREADS * 4 - 1) / (MM_THREADS * 4);
    if (ctas > cap) ctas = cap;
    multimem_adam_kernel<4><<<(unsigned)ctas, MM_THREADS, 0, (cudaStream_t)stream>>>(
        mc_grads, mc_params, params, exp_avg, exp_avg_sq, b, e, grad_scale, segs, beta1, beta2, eps, (float)(1.0 / bc1),
        (float)(1.0 / sqrt(bc2)), keep_grad_begin, keep_grad_end);
    return b200_check_launch("allreduce_adam_multimem");
}

extern "C" int b200_share_slice_multimem(float* mc_ptr, const float* local, int64_t n_floats, int rank, int world, void* stream) {
    B200_REQUIRE(n_floats >= 0 && n_floats % 4 == 0 && world >= 1 && rank >= 0 && rank < world, "bad argument");
    if (n_floats == 0) return B200_OK;
    B200_REQUIRE(mc_ptr && local && (((uintptr_t)mc_ptr | (uintptr_t)local) & 15) == 0, "null or misaligned pointer");
    long long b, e;
    mm_slice(n_floats, rank, world, &b, &e);
    if (b >= e) return B200_OK;
    long long ctas = (e - b + MM_THREADS - 1) / MM_THREADS;
    if (ctas > 64) ctas = 64;
    multimem_share_slice_kernel<<<(unsigned)ctas, MM_THREADS, 0, (cudaStream_t)stream>>>(mc_ptr, local, b, e);
    return b200_check_launch("share_slice_multimem");
}

// In place: every replica of the symmetric buffer <- scale * sum over ranks.  mc_ptr = multicast address of the buffer
// (n_floats floats, a multiple of 4, 16-byte aligned); this rank reduces and broadcasts its 1/world slice.  The caller
// brackets the call with cross-rank barriers on the same stream (before: all replicas written; after: all slices stored).
extern "C" int b200_allreduce_multimem(float* mc_ptr, int64_t n_floats, int rank, int world, float scale, int max_ctas,
                                       void* stream) {
    B200_REQUIRE(n_floats >= 0 && n_floats % 4 == 0 && world >= 1 && rank >= 0 && rank < world, "bad argument");
    if (n_floats == 0) return B200_OK;
    B200_REQUIRE(mc_ptr != nullptr && ((uintptr_t)mc_ptr & 15) == 0, "null or misaligned multicast pointer");
    const long long n4 = n_floats / 4;
    const long long chunk = (n4 + world - 1) / world;
    const long long b = (long long)rank * chunk, e = (b + chunk < n4) ? b + chunk : n4;
    if (b >= e) return B200_OK;
    // max_ctas encodes the tuning point: low 16 bits = CTAs (0: 148), bits 16.. = loads in flight per thread (0: 8)
    // (measured at 8 ranks, 236 MB: 32 CTAs x 4 loads in flight 0.548 ms ... 296 CTAs x 16: 0.647 ms — the switch, not the SMs, is
    // the bound; few CTAs also leave the SMs to whatever else is in flight)
    const int unroll = (max_ctas >> 16) > 0 ? (max_ctas >> 16) : 4;
    const long long cap = (max_ctas & 0xFFFF) > 0 ? (max_ctas & 0xFFFF) : 32;
    long long ctas = (e - b + MM_THREADS * unroll - 1) / (MM_THREADS * unroll);
    if (ctas > cap) ctas = cap;
    cudaStream_t st = (cudaStream_t)stream;
    if (unroll >= 16) multimem_allreduce_kernel<16><<<(unsigned)ctas, MM_THREADS, 0, st>>>(mc_ptr, b, e, scale);
    else if (unroll >= 8) multimem_allreduce_kernel<8><<<(unsigned)ctas, MM_THREADS, 0, st>>>(mc_ptr, b, e, scale);
    else multimem_allreduce_kernel<4><<<(unsigned)ctas, MM_THREADS, 0, st>>>(mc_ptr, b, e, scale);
    return b200_check_launch("allreduce_multimem");
}
