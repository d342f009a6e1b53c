// adam.cu — SURVEY.md §8 f4: the reference's optimiser step (torch.optim.Adam with six parameter groups,
// eps = 1e-15, src/training/trainer.py:331-340) as ONE pass over the flat parameter store.
//
// The store (3d-gaussian-splatting_b200/store.py) keeps parameters, gradients and both moments in flat fp32
// buffers of identical layout, so the whole step is p/g/m/v streamed once: 16 B read + 12 B written per
// element, no per-group launches, no foreach temporaries.  Learning rates are per SEGMENT of the flat buffer;
// inside the SH segment the rate alternates between the dc coefficient (first 3 of every 48 floats) and the
// rest, which is how the reference's separate features_dc / features_rest groups map onto the [N,16,3] layout.
// Arithmetic follows torch.optim.Adam (no amsgrad, no weight decay, not maximize):
//   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  p -= (lr / (1-b1^t)) * m / (sqrt(v) / sqrt(1-b2^t) + eps)
// A segment may carry an upper bound applied to the updated parameter in the same pass — the reference's
// per-iteration hard cap on the log-scales, `gaussians.scaling.data.clamp_(max=scale_ceiling)`
// (trainer.py:854-859), which otherwise is one more read+write of the scaling block per step.
//
// Stage 5 (K10) rides along when asked: the densification statistic `accum[N,3] += ||xyz.grad||_2; count[N] += 1`
// (trainer.py:786-791) reads the SAME position gradient the optimiser has just streamed (the reference takes it after
// optimizer.step(), which leaves .grad untouched), so the first N threads of the pass also update the two statistic
// buffers — 28 B/Gaussian of read-modify-write, the 12 B gradient read being an L2 hit — instead of a second launch.
#include "common.cuh"

#define ADAM_MAX_SEG 8

struct AdamSeg {
    long long begin, end;   // element range in the flat buffer (multiples of 4)
    float lr, lr_alt;       // lr_alt applies where (i - begin) % period >= split
    int period, split;      // period == 0: one rate for the whole segment
    float clamp_max;        // updated parameters are min(p, clamp_max); +inf = no bound
};
struct AdamSegs { AdamSeg s[ADAM_MAX_SEG]; int n; };

namespace {

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            long long n4, AdamSegs segs, float b1, float b2, float eps, float inv_bc1, float inv_sqrt_bc2,
            const float* __restrict__ xyz_grad, long long n_gauss, float* __restrict__ accum, float* __restrict__ count) {
    const long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (xyz_grad != nullptr && i4 < n_gauss) {  // stage 5 for Gaussian i4 (same arithmetic as densify_stats_kernel)
        const float gx = __ldg(xyz_grad + 3 * i4), gy = __ldg(xyz_grad + 3 * i4 + 1), gz = __ldg(xyz_grad + 3 * i4 + 2);
        const float nrm = sqrtf(gx * gx + gy * gy + gz * gz);
        accum[3 * i4] += nrm; accum[3 * i4 + 1] += nrm; accum[3 * i4 + 2] += nrm;
        count[i4] += 1.f;
    }
    if (i4 >= n4) return;
    const long long i = 4 * i4;
    int si = -1;
#pragma unroll
    for (int k = 0; k < ADAM_MAX_SEG; ++k)
        if (k < segs.n && i >= segs.s[k].begin && i < segs.s[k].end) si = k;
    if (si < 0) return;  // padding between segments
    const AdamSeg sg = segs.s[si];
    float4 pp = reinterpret_cast<float4*>(p)[i4];
    const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + i4);
    float4 mm = reinterpret_cast<float4*>(m)[i4], vv = reinterpret_cast<float4*>(v)[i4];
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
        ma[j] = b1 * ma[j] + (1.f - b1) * gj;           // torch: exp_avg.lerp_(grad, 1 - beta1)
        va[j] = b2 * va[j] + (1.f - b2) * gj * gj;      // exp_avg_sq.mul_(b2).addcmul_(grad, grad, 1 - b2)
        const float denom = sqrtf(va[j]) * inv_sqrt_bc2 + eps;
        pa[j] = fminf(pa[j] - (lr * inv_bc1) * (ma[j] / denom), sg.clamp_max);
    }
    reinterpret_cast<float4*>(p)[i4] = pp;
    reinterpret_cast<float4*>(m)[i4] = mm;
    reinterpret_cast<float4*>(v)[i4] = vv;
}

}  // namespace

// seg_host: n_seg rows of B200_ADAM_SEG_DOUBLES (7) doubles {begin, end, lr, lr_alt, period, split, clamp_max};
// clamp_max = +inf for "no bound"; step >= 1 is Adam's t.
static int adam_launch(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_floats,
                       const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                       const float* xyz_grad, int64_t n_gauss, float* accum, float* count, void* stream) {
    B200_REQUIRE(n_floats >= 0 && n_floats % 4 == 0 && n_seg >= 0 && n_seg <= ADAM_MAX_SEG && step >= 1, "bad argument");
    if (n_floats == 0 || n_seg == 0) return B200_OK;
    B200_REQUIRE(params && grads && exp_avg && exp_avg_sq && seg_host, "null pointer");
    B200_REQUIRE((((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0, "misaligned buffer");
    AdamSegs segs;
    segs.n = n_seg;
    for (int k = 0; k < n_seg; ++k) {
        const double* r = seg_host + 7 * k;
        segs.s[k].begin = (long long)r[0]; segs.s[k].end = (long long)r[1];
        segs.s[k].lr = (float)r[2]; segs.s[k].lr_alt = (float)r[3];
        segs.s[k].period = (int)r[4]; segs.s[k].split = (int)r[5];
        segs.s[k].clamp_max = (float)r[6];
        // begin may be negative: a rank stepping only the slice [lo, hi) of the flat buffer passes pointers offset by lo
        // and segments shifted by -lo, so that (i - begin) % period keeps the phase of the full buffer
        B200_REQUIRE(segs.s[k].begin % 4 == 0 && segs.s[k].end <= n_floats, "bad segment");
    }
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    const long long n4 = n_floats / 4;
    B200_REQUIRE(xyz_grad == nullptr || (n_gauss >= 0 && n_gauss <= n4 && accum && count), "bad statistic buffers");
    adam_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        params, grads, exp_avg, exp_avg_sq, n4, segs, beta1, beta2, eps, (float)(1.0 / bc1), (float)(1.0 / sqrt(bc2)),
        xyz_grad, n_gauss, accum, count);
    return b200_check_launch("adam_step");
}

extern "C" int b200_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_floats,
                              const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                              void* stream) {
    return adam_launch(params, grads, exp_avg, exp_avg_sq, n_floats, seg_host, n_seg, step, beta1, beta2, eps,
                       nullptr, 0, nullptr, nullptr, stream);
}

extern "C" int b200_adam_step_stats(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_floats,
                                    const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                                    const float* xyz_grad, int64_t n_gauss, float* accum, float* count, void* stream) {
    B200_REQUIRE(xyz_grad && accum && count && n_gauss >= 0, "null statistic buffers");
    return adam_launch(params, grads, exp_avg, exp_avg_sq, n_floats, seg_host, n_seg, step, beta1, beta2, eps,
                       xyz_grad, n_gauss, accum, count, stream);
}
