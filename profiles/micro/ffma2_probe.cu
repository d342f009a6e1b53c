// ffma2_probe.cu — issue-rate probe for the packed FP32 instructions of sm_100a (FFMA2 / FMUL2 / FADD2) against
// their scalar forms, alone and mixed with MUFU.EX2, to decide whether packing the blend kernels' per-slot
// arithmetic can relieve their binding resource (instruction issue).  Build + run: profiles/micro/run_ffma2.sh
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096

template <int MODE>
__global__ void __launch_bounds__(256) probe(float* out, float seed) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(seed + i + threadIdx.x, seed - i);
    const float2 m = make_float2(1.0000001f, 0.9999999f), c = make_float2(1e-7f, -1e-7f);
    float e0 = seed, e1 = seed * 0.5f;
    __shared__ float sm[64];
    if (threadIdx.x < 64) sm[threadIdx.x] = 0.f;
    __syncthreads();
    const unsigned saddr = (unsigned)__cvta_generic_to_shared(sm) + (threadIdx.x & 7) * 4;
    int q[4] = {(int)threadIdx.x, 1, 2, 3};
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if ((MODE & 1) == 0) {  // scalar: 16 FFMA per unrolled body
                a[i].x = __fmaf_rn(a[i].x, m.x, c.x);
                a[i].y = __fmaf_rn(a[i].y, m.y, c.y);
            } else {                        // packed: 8 FFMA2 (= 16 FMAs)
                a[i] = __ffma2_rn(a[i], m, c);
            }
        }
        if (MODE == 4 || MODE == 5) {  // + 16 integer LOP3 per body (dependent chains on four registers)
#pragma unroll
            for (int i = 0; i < 16; ++i) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(q[i & 3]) : "r"(it), "r"(i));
        }
        if (MODE == 6 || MODE == 7) {  // + 8 shared-memory loads per body
#pragma unroll
            for (int i = 0; i < 8; ++i) { float t; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(t) : "r"(saddr + 4 * i)); e0 += t; }
        }
        if (MODE == 2 || MODE == 3) {  // + 4 MUFU.EX2 per body (the blend kernels' ratio is about 1 MUFU : 10 FP32)
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e0));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e1));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e0));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e1));
        }
    }
    float s = e0 + e1 + (float)(q[0] ^ q[1] ^ q[2] ^ q[3]);
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static void run(const char* name, float* out, int sms) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * 8;
    probe<MODE><<<grid, 256>>>(out, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    probe<MODE><<<grid, 256>>>(out, 1.0f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fmas = (double)grid * 256 * ITERS * 16;
    printf("%-28s %8.3f ms  %7.2f TFMA/s  (%.1f FMA/clk/SM at 1.9 GHz)\n", name, ms, fmas / ms * 1e-9,
           fmas / (ms * 1e-3) / sms / 1.9e9);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    float* out;
    cudaMalloc(&out, (size_t)p.multiProcessorCount * 8 * 256 * sizeof(float));
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<0>("FFMA  (scalar)", out, p.multiProcessorCount);
    run<1>("FFMA2 (packed)", out, p.multiProcessorCount);
    run<2>("FFMA  + MUFU.EX2 (4:1)", out, p.multiProcessorCount);
    run<3>("FFMA2 + MUFU.EX2 (2:1 issue)", out, p.multiProcessorCount);
    run<4>("16 FFMA  + 16 LOP3", out, p.multiProcessorCount);
    run<5>(" 8 FFMA2 + 16 LOP3", out, p.multiProcessorCount);
    run<6>("16 FFMA  + 8 LDS + 8 FADD", out, p.multiProcessorCount);
    run<7>(" 8 FFMA2 + 8 LDS + 8 FADD", out, p.multiProcessorCount);
    return cudaGetLastError() != cudaSuccess;
}
