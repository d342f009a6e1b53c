// radix_sort.cu — stage 2c: stable LSD radix sort of (uint64 key, uint32 value) pairs, 8-bit digits,
// single-read "onesweep" passes with decoupled look-back between tiles.  Replaces the
// cub::DeviceRadixSort::SortPairs call inside gsplat isect_tiles(sort=True) that the reference reaches
// through gsplat_backend.py:168-182.  Stability (ties keep emission order) comes from
// in-order ranking inside a warp, warp-major order inside a tile and ticket order between tiles.
#include "common.cuh"

#define RS_BITS 8
#define RS_RADIX 256
#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_ITEMS 16
#define RS_TILE (RS_THREADS * RS_ITEMS)
#define RS_MAX_PASSES 8

#define FLAG_AGG 0x40000000u
#define FLAG_PREFIX 0x80000000u
#define FLAG_MASK 0xC0000000u
#define VAL_MASK 0x3FFFFFFFu

namespace {

// All per-pass digit histograms in one read of the keys.
__global__ void __launch_bounds__(256)
rs_hist_kernel(const uint64_t* __restrict__ keys, int64_t n, int npasses, int end_bit, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[RS_MAX_PASSES * RS_RADIX];
    for (int i = threadIdx.x; i < npasses * RS_RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n_round = (n + 31) / 32 * 32;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool valid = i < n;
        const uint64_t k = valid ? keys[i] : 0;
        const unsigned act = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            for (int p = 0; p < npasses; ++p) {
                int shift = p * RS_BITS;
                int bits = min(RS_BITS, end_bit - shift);
                uint32_t d = (uint32_t)(k >> shift) & ((1u << bits) - 1u);
                unsigned peers = __match_any_sync(act, d);
                if ((__ffs(peers) - 1) == lane) atomicAdd(&s_hist[p * RS_RADIX + d], __popc(peers));
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npasses * RS_RADIX; i += blockDim.x)
        if (s_hist[i]) atomicAdd(&hist[i], s_hist[i]);
}

// Exclusive scan of each pass's 256 bins, in place.
__global__ void __launch_bounds__(RS_RADIX)
rs_scan_kernel(uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_w[RS_RADIX / 32];
    uint32_t* h = hist + blockIdx.x * RS_RADIX;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t v = h[threadIdx.x], inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[wid] = inc;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < RS_RADIX / 32; ++w) off += (w < wid) ? s_w[w] : 0;
    h[threadIdx.x] = off + inc - v;
}

struct __align__(16) RsSmem {
    uint64_t keys[RS_TILE];
    uint32_t vals[RS_TILE];
    uint32_t warp_hist[RS_WARPS][RS_RADIX];  // per-warp digit counts -> exclusive offsets across warps
    uint32_t bin_excl[RS_RADIX];             // tile-local exclusive start of each digit
    int64_t global_off[RS_RADIX];            // global index of smem slot p with digit d = global_off[d] + p
    uint32_t warp_scan[RS_WARPS];
    uint32_t tile;
};

__global__ void __launch_bounds__(RS_THREADS)
rs_onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                   uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n, int shift, int bits,
                   const uint32_t* __restrict__ digit_base, volatile uint32_t* __restrict__ status,
                   uint32_t* __restrict__ ticket) {
    extern __shared__ __align__(16) unsigned char rs_smem_raw[];
    RsSmem& S = *reinterpret_cast<RsSmem*>(rs_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t dmask = (1u << bits) - 1u;

    if (tid == 0) S.tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&S.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = S.tile;
    const int64_t tile_base = (int64_t)tile * RS_TILE;
    const int n_valid = (int)min((int64_t)RS_TILE, n - tile_base);

    // ---- load (warp-striped so that order inside a warp is item-major, lane-minor) ----
    uint64_t key[RS_ITEMS];
    uint32_t val[RS_ITEMS];
    uint32_t pos[RS_ITEMS];
    const int warp_base = wid * (32 * RS_ITEMS);
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        int local = warp_base + i * 32 + lane;
        bool ok = local < n_valid;
        key[i] = ok ? keys_in[tile_base + local] : ~0ull;
        val[i] = ok ? vals_in[tile_base + local] : 0u;
    }
    // ---- rank inside the warp by digit, in order ----
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (lane == leader) {
            base = S.warp_hist[wid][d];
            S.warp_hist[wid][d] = base + __popc(peers);
        }
        base = __shfl_sync(0xffffffffu, base, leader);
        pos[i] = base + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    // ---- per digit (thread d): exclusive prefix over warps, tile count ----
    uint32_t count;
    {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            uint32_t c = S.warp_hist[w][tid];
            S.warp_hist[w][tid] = run;
            run += c;
        }
        count = run;
    }
    // exclusive scan of counts over digits
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) S.warp_scan[wid] = inc;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) woff += (w < wid) ? S.warp_scan[w] : 0;
    const uint32_t excl_local = woff + inc - count;
    S.bin_excl[tid] = excl_local;

    // ---- decoupled look-back over previous tiles for this digit ----
    {
        volatile uint32_t* my = status + (size_t)tile * RS_RADIX + tid;
        uint32_t excl = 0;
        if (tile == 0) {
            *my = count | FLAG_PREFIX;
        } else {
            *my = count | FLAG_AGG;
            int64_t t = (int64_t)tile - 1;
            while (true) {
                uint32_t s = status[(size_t)t * RS_RADIX + tid];
                if ((s & FLAG_MASK) == 0) continue;  // not published yet
                excl += s & VAL_MASK;
                if (s & FLAG_PREFIX) break;
                --t;
            }
            *my = (excl + count) | FLAG_PREFIX;
        }
        S.global_off[tid] = (int64_t)digit_base[tid] + (int64_t)excl - (int64_t)excl_local;
    }
    __syncthreads();
    // ---- scatter into tile-local sorted order ----
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        uint32_t p = S.bin_excl[d] + S.warp_hist[wid][d] + pos[i];
        S.keys[p] = key[i];
        S.vals[p] = val[i];
    }
    __syncthreads();
    // ---- coalesced write-out ----
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        int p = i * RS_THREADS + tid;
        if (p < n_valid) {
            uint64_t k = S.keys[p];
            uint32_t d = (uint32_t)(k >> shift) & dmask;
            int64_t o = S.global_off[d] + p;
            keys_out[o] = k;
            vals_out[o] = S.vals[p];
        }
    }
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" size_t b200_sort_scratch_bytes(int64_t M) {
    if (M < 0) M = 0;
    size_t ntiles = (size_t)((M + RS_TILE - 1) / RS_TILE);
    size_t hist = RS_MAX_PASSES * RS_RADIX * sizeof(uint32_t);
    size_t tickets = 256;
    size_t status = (size_t)RS_MAX_PASSES * ntiles * RS_RADIX * sizeof(uint32_t);
    return align_up(hist + tickets + status, 256);
}

extern "C" int b200_sort_pairs(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, int64_t M,
                               int end_bit, void* scratch, size_t scratch_bytes, int* result_in_b_host, void* stream) {
    B200_REQUIRE(M >= 0 && end_bit > 0 && end_bit <= 64, "bad sizes");
    B200_REQUIRE(M < ((int64_t)1 << 30), "M >= 2^30 not supported");
    const int npasses = (end_bit + RS_BITS - 1) / RS_BITS;
    if (result_in_b_host) *result_in_b_host = (M > 0) ? (npasses & 1) : 0;
    if (M == 0) return B200_OK;
    B200_REQUIRE(keys_a && vals_a && keys_b && vals_b && scratch, "null pointer");
    if (scratch_bytes < b200_sort_scratch_bytes(M)) {
        b200_set_error("b200_sort_pairs: scratch too small (%zu < %zu)", scratch_bytes, b200_sort_scratch_bytes(M));
        return B200_ESCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(rs_onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem));
        attr_set = true;
    }
    const size_t ntiles = (size_t)((M + RS_TILE - 1) / RS_TILE);
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* tickets = hist + RS_MAX_PASSES * RS_RADIX;
    uint32_t* status = tickets + 64;
    const size_t used = (RS_MAX_PASSES * RS_RADIX + 64) * sizeof(uint32_t) + (size_t)npasses * ntiles * RS_RADIX * sizeof(uint32_t);
    cudaMemsetAsync(scratch, 0, used, st);
    int hist_blocks = (int)min((int64_t)148 * 8, (M + 255) / 256);
    rs_hist_kernel<<<hist_blocks, 256, 0, st>>>(keys_a, M, npasses, end_bit, hist);
    rs_scan_kernel<<<npasses, RS_RADIX, 0, st>>>(hist);
    uint64_t* kin = keys_a; uint32_t* vin = vals_a; uint64_t* kout = keys_b; uint32_t* vout = vals_b;
    for (int p = 0; p < npasses; ++p) {
        int shift = p * RS_BITS;
        int bits = (end_bit - shift) < RS_BITS ? (end_bit - shift) : RS_BITS;
        rs_onesweep_kernel<<<(unsigned)ntiles, RS_THREADS, sizeof(RsSmem), st>>>(
            kin, vin, kout, vout, M, shift, bits, hist + p * RS_RADIX, status + (size_t)p * ntiles * RS_RADIX, tickets + p);
        uint64_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    return b200_check_launch("sort_pairs");
}
