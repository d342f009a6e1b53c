// radix_sort.cu — stage 2c: stable LSD radix sort of (key, uint32 value) pairs, 8-bit digits,
// single-read "onesweep" passes with decoupled look-back between tiles.  Replaces the
// cub::DeviceRadixSort::SortPairs call inside gsplat isect_tiles(sort=True) that the reference reaches
// through gsplat_backend.py:168-182.  Stability (ties keep emission order) comes from
// in-order ranking inside a warp, warp-major order inside a tile and ticket order between tiles.
//
// Three key widths share one kernel template: 64-bit (tile<<32 | depth) keys of the literal gsplat
// formulation, and the split form the fast path uses — 32-bit depth keys over Gaussians BEFORE tile
// duplication followed by 16-bit tile keys over intersections (the same LSD digit order, with the
// duplication moved between the depth digits and the tile digits; see DESIGN.md "Sort").
#include <stdlib.h>
#include "common.cuh"

#define RS_BITS 8
#define RS_RADIX 256
#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#ifndef RS_ITEMS
#define RS_ITEMS 16
#endif
#define RS_TILE (RS_THREADS * RS_ITEMS)
#define RS_MAX_PASSES 8
#define RS_LOOKBACK 8  // predecessor status words fetched per look-back step (memory-level parallelism)

#define FLAG_AGG 0x40000000u
#define FLAG_PREFIX 0x80000000u
#define FLAG_MASK 0xC0000000u
#define VAL_MASK 0x3FFFFFFFu

namespace {

// Lanes holding the same BITS-bit digit, three instructions per bit after the predicates (ptxas packs those into
// R2P): ballot, the lane's own bit widened to a mask (SELP), and ONE LOP3 folding "lanes that disagree with me in
// this bit" into x (x |= vote ^ mask).  The earlier form (two predicated LOP3) compiled to LOP3+SEL pairs, five
// ALU-pipe instructions per bit, and the ALU pipe (half rate) was what throttled the pass (math_pipe_throttle).
template <int BITS>
__device__ __forceinline__ unsigned digit_peers_n(uint32_t d) {
    unsigned x = 0;
#pragma unroll
    for (int b = 0; b < BITS; ++b) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .b32 v, m, t;\n\t"
            "and.b32 t, %1, %2;\n\t"
            "setp.ne.u32 p, t, 0;\n\t"
            "vote.sync.ballot.b32 v, p, 0xffffffff;\n\t"
            "selp.b32 m, 0xffffffff, 0, p;\n\t"
            "lop3.b32 %0, %0, v, m, 0xF6;\n\t}"
            : "+r"(x)
            : "r"(d), "r"(1u << b));
    }
    return ~x;
}

// All per-pass digit histograms in one read of the keys.  Each warp owns private histogram rows in
// shared memory and updates them with plain LDS/STS from the lowest lane of every digit group: shared
// ATOMICS cost ~2 cycles per distinct address (B300_MICROARCH.md) and with high-entropy digits that is
// 60 cycles per warp instruction — it, not HBM, bounded the first two versions (60 us for a 13 MB read).
// Block size is chosen by the launcher so that warps * npasses <= 32 rows (32 KB).
#define RS_HIST_ROWS 32
template <typename KeyT>
__global__ void __launch_bounds__(256)
rs_hist_kernel(const KeyT* __restrict__ keys, int64_t n, int npasses, int end_bit, uint32_t* __restrict__ hist) {
    constexpr int VEC = 16 / sizeof(KeyT);
    __shared__ uint32_t s_hist[RS_HIST_ROWS * RS_RADIX];
    const int nwarps = blockDim.x >> 5;
    for (int i = threadIdx.x; i < nwarps * npasses * RS_RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t* my_rows = s_hist + wid * npasses * RS_RADIX;
    const int64_t nchunks = (n + VEC - 1) / VEC;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int iters = (int)((nchunks + stride - 1) / stride);  // block-uniform trip count: ballots stay plain VOTEs
    const bool aligned = ((uintptr_t)keys & 15) == 0;
    for (int it = 0; it < iters; ++it) {
        const int64_t c = ((int64_t)it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        KeyT k[VEC];
        const int64_t base = c * VEC;
        if (aligned && base + VEC <= n) {
            const uint4 raw = __ldg(reinterpret_cast<const uint4*>(keys) + c);
            memcpy(k, &raw, 16);
        } else {
#pragma unroll
            for (int v = 0; v < VEC; ++v) k[v] = (base + v < n) ? keys[base + v] : (KeyT)0;
        }
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            const bool valid = base + v < n;
            const unsigned vmask = __ballot_sync(0xffffffffu, valid);
            for (int p = 0; p < npasses; ++p) {
                const int shift = p * RS_BITS;
                const int bits = min(RS_BITS, end_bit - shift);
                const uint32_t d = (uint32_t)((uint64_t)k[v] >> shift) & ((1u << bits) - 1u);
                const unsigned peers = digit_peers_n<RS_BITS>(d) & vmask;
                if (valid && (31 - __clz(peers)) == lane) my_rows[p * RS_RADIX + d] += __popc(peers);
                __syncwarp();
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npasses * RS_RADIX; i += blockDim.x) {
        uint32_t tot = 0;
        for (int w = 0; w < nwarps; ++w) tot += s_hist[w * npasses * RS_RADIX + i];
        if (tot) atomicAdd(&hist[i], tot);
    }
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

template <typename KeyT>
struct __align__(16) RsSmem {
    KeyT keys[RS_TILE];
    uint32_t vals[RS_TILE];
    uint32_t warp_hist[RS_WARPS][RS_RADIX];  // per-warp digit counts -> exclusive offsets across warps
    uint32_t bin_excl[RS_RADIX];             // tile-local exclusive start of each digit
    uint32_t global_off[RS_RADIX];           // global index of smem slot p with digit d = global_off[d] + p (mod 2^32)
    uint32_t warp_scan[RS_WARPS];
    uint32_t tile;
};

// One LSD pass.  dmask selects the digit bits of this pass, BITS >= popcount(dmask) is the number of ballots spent.  All indices are 32-bit
// (n < 2^30); full tiles take an unpredicated path.
template <typename KeyT, int BITS>
__global__ void __launch_bounds__(RS_THREADS)
rs_onesweep_kernel(const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                   KeyT* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift,
                   uint32_t dmask, const uint32_t* __restrict__ digit_base, volatile uint32_t* __restrict__ status,
                   uint32_t* __restrict__ ticket, const uint32_t* __restrict__ n_dev) {
    extern __shared__ __align__(16) unsigned char rs_smem_raw[];
    RsSmem<KeyT>& S = *reinterpret_cast<RsSmem<KeyT>*>(rs_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    // n_dev: the element count lives on the device (the coarse binning path sizes its launch by an upper bound known to
    // the host and lets the surplus CTAs leave here); tickets are handed out in order, so the leavers are the last ones
    if (n_dev != nullptr) n = __ldg(n_dev);
    if (tid == 0) S.tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&S.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = S.tile;
    if ((uint64_t)tile * RS_TILE >= n) return;
    const uint32_t tile_base = tile * RS_TILE;
    const uint32_t n_valid = min((uint32_t)RS_TILE, n - tile_base);
    const bool full = n_valid == RS_TILE;

    // ---- load (warp-striped so that order inside a warp is item-major, lane-minor) ----
    KeyT key[RS_ITEMS];
    uint32_t val[RS_ITEMS];
    uint32_t pos[RS_ITEMS];
    const uint32_t warp_base = wid * (32 * RS_ITEMS) + lane;
    const KeyT* kin = keys_in + tile_base + warp_base;
    const uint32_t* vin = vals_in + tile_base + warp_base;
    if (full) {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i) {
            key[i] = kin[i * 32];
            val[i] = vin[i * 32];
        }
    } else {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i) {
            const bool ok = warp_base + i * 32 < n_valid;
            key[i] = ok ? kin[i * 32] : (KeyT)~(KeyT)0;
            val[i] = ok ? vin[i * 32] : 0u;
        }
    }
    // ---- rank inside the warp by digit, in order ----
    // The warp's histogram row is private to it, so the lowest lane of each digit group updates it
    // with a plain load + store (no shared atomics: ~2 cycles per distinct address, i.e. up to 60 cycles
    // per item for high-entropy digits); __syncwarp orders the row between successive items.
    const unsigned lt_mask = (1u << lane) - 1u;
    uint32_t* my_row = S.warp_hist[wid];
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        const unsigned peers = digit_peers_n<BITS>(d);
        const int leader = 31 - __clz(peers);  // any one lane of the group may update the row; the highest costs one FLO
        uint32_t base = 0;
        if (lane == leader) {
            base = my_row[d];
            my_row[d] = base + __popc(peers);
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
    // publish this tile's aggregate as early as possible: successors can already add it
    volatile uint32_t* my = status + (size_t)tile * RS_RADIX + tid;
    *my = count | (tile == 0 ? FLAG_PREFIX : FLAG_AGG);

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

    // ---- decoupled look-back: RS_LOOKBACK predecessors per step, consumed nearest-first ----
    {
        uint32_t excl = 0;
        if (tile != 0) {
            int t = (int)tile - 1;
            bool found = false;
            while (!found) {
                uint32_t s[RS_LOOKBACK];
#pragma unroll
                for (int k = 0; k < RS_LOOKBACK; ++k) s[k] = (t - k >= 0) ? status[(size_t)(t - k) * RS_RADIX + tid] : FLAG_PREFIX;
#pragma unroll
                for (int k = 0; k < RS_LOOKBACK; ++k) {
                    if (found) break;
                    if ((s[k] & FLAG_MASK) == 0) break;  // not published yet: re-read from here
                    excl += s[k] & VAL_MASK;
                    --t;
                    if (s[k] & FLAG_PREFIX) found = true;
                }
            }
            *my = (excl + count) | FLAG_PREFIX;
        }
        S.global_off[tid] = digit_base[tid] + excl - excl_local;  // wraps mod 2^32; global_off[d] + p is exact
    }
    __syncthreads();
    // ---- scatter into tile-local sorted order ----
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        const uint32_t p = S.bin_excl[d] + my_row[d] + pos[i];
        S.keys[p] = key[i];
        S.vals[p] = val[i];
    }
    __syncthreads();
    // ---- coalesced write-out ----
    if (full) {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i) {
            const uint32_t p = i * RS_THREADS + tid;
            const KeyT k = S.keys[p];
            const uint32_t o = S.global_off[(uint32_t)(k >> shift) & dmask] + p;
            keys_out[o] = k;
            vals_out[o] = S.vals[p];
        }
    } else {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i) {
            const uint32_t p = i * RS_THREADS + tid;
            if (p < n_valid) {
                const KeyT k = S.keys[p];
                const uint32_t o = S.global_off[(uint32_t)(k >> shift) & dmask] + p;
                keys_out[o] = k;
                vals_out[o] = S.vals[p];
            }
        }
    }
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// begin_bit: lowest key bit that takes part (bits below ride along as payload).  n_dev: device-side element count
// <= M (M then only sizes the launch and the scratch).  prepared_base: exclusive digit offsets of every pass supplied
// by the caller (no histogram pass).
template <typename KeyT>
int sort_pairs_impl(KeyT* keys_a, uint32_t* vals_a, KeyT* keys_b, uint32_t* vals_b, int64_t M, int end_bit,
                    void* scratch, size_t scratch_bytes, int* result_in_b_host, void* stream, const char* what,
                    const uint32_t* prepared_base = nullptr, int begin_bit = 0, const uint32_t* n_dev = nullptr) {
    if (!(M >= 0 && begin_bit >= 0 && end_bit > begin_bit && end_bit <= (int)(8 * sizeof(KeyT)))) {
        b200_set_error("%s: bad sizes", what);
        return B200_EARG;
    }
    if (n_dev != nullptr && prepared_base == nullptr) {
        b200_set_error("%s: a device-side count needs prepared digit bases", what);
        return B200_EARG;
    }
    if (M >= ((int64_t)1 << 30)) {
        b200_set_error("%s: M >= 2^30 not supported", what);
        return B200_EARG;
    }
    const int npasses = (end_bit - begin_bit + RS_BITS - 1) / RS_BITS;
    if (result_in_b_host) *result_in_b_host = (M > 0) ? (npasses & 1) : 0;
    if (M == 0) return B200_OK;
    if (!(keys_a && vals_a && keys_b && vals_b && scratch)) {
        b200_set_error("%s: null pointer", what);
        return B200_EARG;
    }
    if (scratch_bytes < b200_sort_scratch_bytes(M)) {
        b200_set_error("%s: scratch too small (%zu < %zu)", what, scratch_bytes, b200_sort_scratch_bytes(M));
        return B200_ESCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    static bool attr_set[B200_MAX_DEVICES] = {};  // one flag array per template instantiation, one flag per device
    if (b200_first_use_on_device(attr_set)) {
        const int smem_max = (int)sizeof(RsSmem<KeyT>) + 160 * 1024;
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
    }
    const size_t ntiles = (size_t)((M + RS_TILE - 1) / RS_TILE);
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* tickets = hist + RS_MAX_PASSES * RS_RADIX;
    uint32_t* status = tickets + 64;
    const size_t used = (RS_MAX_PASSES * RS_RADIX + 64) * sizeof(uint32_t) + (size_t)npasses * ntiles * RS_RADIX * sizeof(uint32_t);
    if (prepared_base == nullptr) {
        cudaMemsetAsync(scratch, 0, used, st);
    } else {  // digit bases supplied by the caller (b200_tile_prep): only tickets + status need clearing
        cudaMemsetAsync(tickets, 0, used - RS_MAX_PASSES * RS_RADIX * sizeof(uint32_t), st);
    }
    static int pad = -1;
    if (pad < 0) { const char* e = getenv("B200_SORT_SMEM_PAD"); pad = e ? atoi(e) : 0; }
    static int hb = -1;
    if (hb < 0) { const char* e = getenv("B200_HIST_BLOCKS"); hb = e ? atoi(e) : 148 * 2; }
    int hist_blocks = (int)min((int64_t)hb, (M + 255) / 256);
    const int hist_threads = (npasses <= 4) ? 256 : 128;  // warps * npasses <= RS_HIST_ROWS
    if (prepared_base == nullptr) {
        rs_hist_kernel<KeyT><<<hist_blocks, hist_threads, 0, st>>>(keys_a, M, npasses, end_bit, hist);  // (begin_bit == 0 here)
        rs_scan_kernel<<<npasses, RS_RADIX, 0, st>>>(hist);
    }
    const uint32_t* bases = prepared_base ? prepared_base : hist;
    KeyT* kin = keys_a; uint32_t* vin = vals_a; KeyT* kout = keys_b; uint32_t* vout = vals_b;
    const size_t smem = sizeof(RsSmem<KeyT>) + pad;
    if (prepared_base == nullptr && begin_bit != 0) {
        b200_set_error("%s: begin_bit needs prepared digit bases", what);
        return B200_EARG;
    }
    for (int p = 0; p < npasses; ++p) {
        int shift = begin_bit + p * RS_BITS;
        int bits = (end_bit - shift) < RS_BITS ? (end_bit - shift) : RS_BITS;
#define RS_LAUNCH(B) rs_onesweep_kernel<KeyT, B><<<(unsigned)ntiles, RS_THREADS, smem, st>>>( \
        kin, vin, kout, vout, (uint32_t)M, shift, (1u << bits) - 1u, bases + p * RS_RADIX, status + (size_t)p * ntiles * RS_RADIX, tickets + p, n_dev)
        // digits narrower than 5 bits run the 5-bit kernel (the upper ballots see zeros): same result
        if (bits >= 8) RS_LAUNCH(8); else if (bits == 7) RS_LAUNCH(7); else if (bits == 6) RS_LAUNCH(6); else RS_LAUNCH(5);
#undef RS_LAUNCH
        KeyT* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    return b200_check_launch(what);
}

}  // namespace

// Used by the coarse binning path (binning.cu): (supertile << 16 | coverage mask) keys sorted on bits [16, end_bit) with
// caller-supplied digit bases and a device-side element count.
int b200_sort_pairs_coarse(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, int64_t cap, int end_bit,
                           const uint32_t* digit_base, const uint32_t* n_dev, void* scratch, size_t scratch_bytes,
                           int* result_in_b_host, void* stream) {
    return sort_pairs_impl<uint32_t>(keys_a, vals_a, keys_b, vals_b, cap, end_bit, scratch, scratch_bytes, result_in_b_host,
                                     stream, "b200_bin_coarse(sort)", digit_base, 16, n_dev);
}

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
    return sort_pairs_impl<uint64_t>(keys_a, vals_a, keys_b, vals_b, M, end_bit, scratch, scratch_bytes, result_in_b_host, stream, "b200_sort_pairs");
}

extern "C" int b200_sort_pairs_k32(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, int64_t M,
                                   int end_bit, void* scratch, size_t scratch_bytes, int* result_in_b_host, void* stream) {
    return sort_pairs_impl<uint32_t>(keys_a, vals_a, keys_b, vals_b, M, end_bit, scratch, scratch_bytes, result_in_b_host, stream, "b200_sort_pairs_k32");
}

extern "C" int b200_sort_pairs_k16(uint16_t* keys_a, uint32_t* vals_a, uint16_t* keys_b, uint32_t* vals_b, int64_t M,
                                   int end_bit, void* scratch, size_t scratch_bytes, int* result_in_b_host, void* stream) {
    return sort_pairs_impl<uint16_t>(keys_a, vals_a, keys_b, vals_b, M, end_bit, scratch, scratch_bytes, result_in_b_host, stream, "b200_sort_pairs_k16");
}

extern "C" int b200_sort_pairs_k16_prepared(uint16_t* keys_a, uint32_t* vals_a, uint16_t* keys_b, uint32_t* vals_b, int64_t M,
                                            int end_bit, const uint32_t* digit_base, void* scratch, size_t scratch_bytes,
                                            int* result_in_b_host, void* stream) {
    if (M > 0 && digit_base == nullptr) {
        b200_set_error("b200_sort_pairs_k16_prepared: null digit_base");
        return B200_EARG;
    }
    return sort_pairs_impl<uint16_t>(keys_a, vals_a, keys_b, vals_b, M, end_bit, scratch, scratch_bytes, result_in_b_host, stream,
                                     "b200_sort_pairs_k16_prepared", digit_base);
}

// ---- allocation-free variant (no counterpart in the reference) ---------
// b200_sort_pairs_k32 with its scratch (digit histograms, tile tickets, look-back status words) in a file-scope device
// array instead of a caller buffer, for callers that cannot allocate (CUDA graph capture, serving loops).  Every GPU that
// loads the library holds the array (2 105 600 B) until exit.  All calls share it: sort_pairs_impl clears the part a call
// uses on `stream` before the first kernel, so stale contents of an earlier call never matter, but two calls that overlap
// on the GPU (different streams, no ordering between them) race.
#define RS_STATIC_TILES ((B200_SORT_STATIC_MAX_KEYS + RS_TILE - 1) / RS_TILE)
#define RS_STATIC_BYTES ((((size_t)RS_MAX_PASSES * RS_RADIX * 4 + 256 + (size_t)RS_MAX_PASSES * RS_STATIC_TILES * RS_RADIX * 4) + 255) / 256 * 256)
static __device__ __align__(256) unsigned char g_sort_static_scratch[RS_STATIC_BYTES];

extern "C" int b200_sort_pairs_k32_static(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, int64_t M,
                                          int end_bit, int* result_in_b_host, void* stream) {
    if (M > (int64_t)B200_SORT_STATIC_MAX_KEYS) {
        b200_set_error("b200_sort_pairs_k32_static: M = %lld exceeds the static scratch (B200_SORT_STATIC_MAX_KEYS = %d keys)",
                       (long long)M, (int)B200_SORT_STATIC_MAX_KEYS);
        return B200_EARG;
    }
    // In host code the array's name is only a handle: the device address is asked for once per device, with that
    // device current, and cached.
    static void* dev_ptr[B200_MAX_DEVICES] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= B200_MAX_DEVICES) {
        b200_set_error("b200_sort_pairs_k32_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return B200_ECUDA;
    }
    if (dev_ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_sort_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            b200_set_error("b200_sort_pairs_k32_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return B200_ECUDA;
        }
        dev_ptr[dev] = p;
    }
    if (b200_sort_scratch_bytes(B200_SORT_STATIC_MAX_KEYS) > RS_STATIC_BYTES) {  // layout drifted from the array's size
        b200_set_error("b200_sort_pairs_k32_static: static scratch smaller than b200_sort_scratch_bytes(max)");
        return B200_ESCRATCH;
    }
    return sort_pairs_impl<uint32_t>(keys_a, vals_a, keys_b, vals_b, M, end_bit, dev_ptr[dev], RS_STATIC_BYTES,
                                     result_in_b_host, stream, "b200_sort_pairs_k32_static");
}
