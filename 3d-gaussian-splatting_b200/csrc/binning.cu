// binning.cu — stage 2a/2b/2d: scan of per-block tile counts, load-balanced key emission, per-tile
// ranges.  Integer work, bit-exact against oracle.bin_tiles / sort_and_ranges, which
// restate gsplat isect_tiles / isect_offset_encode as called through gsplat_backend.py:168-182.
#include "common.cuh"

#define EMIT_THREADS 256  // must equal PROJ_THREADS: block b owns Gaussians [256 b, 256 b + 256)

namespace {

__global__ void __launch_bounds__(1024)
bin_scan_kernel(const int64_t* __restrict__ block_sums, int64_t nblocks, int64_t* __restrict__ block_offsets,
                int64_t* __restrict__ total) {
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t carry_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int64_t base = 0; base < nblocks; base += 1024) {
        int64_t i = base + threadIdx.x;
        int64_t v = i < nblocks ? block_sums[i] : 0;
        int64_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) warp_tot[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int64_t w = warp_tot[lane];
            int64_t winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int64_t n = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += n;
            }
            warp_tot[lane] = winc - w;  // exclusive
        }
        __syncthreads();
        int64_t carry = carry_s;
        int64_t excl = carry + warp_tot[wid] + inc - v;
        if (i < nblocks) block_offsets[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) total[0] = carry_s;
}

__global__ void __launch_bounds__(EMIT_THREADS)
bin_emit_kernel(const float* __restrict__ splats, const int32_t* __restrict__ radii,
                const int32_t* __restrict__ tiles_per_gauss, const int64_t* __restrict__ block_offsets,
                int64_t N, int tile_w, int tile_h, int64_t M, int64_t* __restrict__ isect_ids,
                int32_t* __restrict__ flatten_ids, int64_t* __restrict__ cum_tiles) {
    __shared__ int s_excl[EMIT_THREADS + 1];
    __shared__ int s_minx[EMIT_THREADS], s_miny[EMIT_THREADS], s_bw[EMIT_THREADS];
    __shared__ uint32_t s_depth[EMIT_THREADS];
    __shared__ int s_warp[EMIT_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t g = (int64_t)blockIdx.x * EMIT_THREADS + tid;
    int cnt = 0;
    if (g < N) {
        cnt = tiles_per_gauss[g];
        if (cnt > 0) {
            const float4 r0 = __ldg(reinterpret_cast<const float4*>(splats) + 3 * g);
            const float depth = __ldg(splats + SPLAT_F * g + 9);
            const int2 rad = __ldg(reinterpret_cast<const int2*>(radii) + g);
            float tile_x = r0.x / (float)TILE, tile_y = r0.y / (float)TILE;
            float trx = (float)rad.x / (float)TILE, try_ = (float)rad.y / (float)TILE;
            float fw = (float)tile_w, fh = (float)tile_h;
            int minx = (int)fminf(fmaxf(floorf(tile_x - trx), 0.f), fw);
            int maxx = (int)fminf(fmaxf(ceilf(tile_x + trx), 0.f), fw);
            int miny = (int)fminf(fmaxf(floorf(tile_y - try_), 0.f), fh);
            s_minx[tid] = minx; s_miny[tid] = miny; s_bw[tid] = maxx - minx;
            s_depth[tid] = __float_as_uint(depth);
        }
    }
    // block exclusive scan of cnt
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    int woff = 0;
#pragma unroll
    for (int w = 0; w < EMIT_THREADS / 32; ++w) woff += (w < wid) ? s_warp[w] : 0;
    const int excl = woff + inc - cnt;
    s_excl[tid] = excl;
    if (tid == EMIT_THREADS - 1) s_excl[EMIT_THREADS] = excl + cnt;
    const int64_t boff = block_offsets[blockIdx.x];
    if (cum_tiles != nullptr && g < N) cum_tiles[g] = boff + excl + cnt;
    __syncthreads();
    const int total = s_excl[EMIT_THREADS];
    for (int k = tid; k < total; k += EMIT_THREADS) {
        // largest i with s_excl[i] <= k  (upper_bound - 1)
        int lo = 0, hi = EMIT_THREADS;  // invariant: s_excl[lo] <= k < s_excl[hi]
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (s_excl[mid] <= k) lo = mid; else hi = mid;
        }
        const int kk = k - s_excl[lo];
        const int bw = s_bw[lo];
        const int ty = s_miny[lo] + kk / bw, tx = s_minx[lo] + kk % bw;
        const int64_t tile_id = (int64_t)ty * tile_w + tx;
        const int64_t out = boff + k;
        if (out < M) {
            isect_ids[out] = (tile_id << 32) | (int64_t)s_depth[lo];
            flatten_ids[out] = (int32_t)((int64_t)blockIdx.x * EMIT_THREADS + lo);
        }
    }
}

__global__ void tile_ranges_kernel(const int64_t* __restrict__ ids, int64_t M, int n_tiles, int32_t* __restrict__ offsets) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M) return;
    int cur = (int)(ids[idx] >> 32);
    if (idx == 0) {
        for (int t = 0; t <= cur; ++t) offsets[t] = 0;
    } else {
        int prev = (int)(ids[idx - 1] >> 32);
        for (int t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)idx;
    }
    if (idx == M - 1)
        for (int t = cur + 1; t < n_tiles; ++t) offsets[t] = (int32_t)M;
}

// ---- split-key path: emission in DEPTH order (perm = Gaussian ids sorted by depth bits) -------------
__global__ void __launch_bounds__(EMIT_THREADS)
bin_count_sorted_kernel(const int32_t* __restrict__ tiles_per_gauss, const int32_t* __restrict__ perm, int64_t N,
                        int64_t* __restrict__ block_sums) {
    __shared__ int warp_tot[EMIT_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * EMIT_THREADS + threadIdx.x;
    int v = (i < N) ? __ldg(tiles_per_gauss + __ldg(perm + i)) : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t tot = 0;
#pragma unroll
        for (int w = 0; w < EMIT_THREADS / 32; ++w) tot += warp_tot[w];
        block_sums[blockIdx.x] = tot;
    }
}

__global__ void __launch_bounds__(EMIT_THREADS)
bin_emit_sorted_kernel(const float* __restrict__ splats, const int32_t* __restrict__ radii,
                       const int32_t* __restrict__ tiles_per_gauss, const int32_t* __restrict__ perm,
                       const int64_t* __restrict__ block_offsets, int64_t N, int tile_w, int tile_h, int64_t M,
                       uint16_t* __restrict__ tile_keys, int32_t* __restrict__ flatten_ids) {
    __shared__ int s_excl[EMIT_THREADS + 1];
    __shared__ int s_minx[EMIT_THREADS], s_miny[EMIT_THREADS], s_bw[EMIT_THREADS], s_g[EMIT_THREADS];
    __shared__ float s_rbw[EMIT_THREADS];
    __shared__ int s_warp[EMIT_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t i = (int64_t)blockIdx.x * EMIT_THREADS + tid;
    int cnt = 0;
    if (i < N) {
        const int g = __ldg(perm + i);
        cnt = __ldg(tiles_per_gauss + g);
        if (cnt > 0) {
            const float4 r0 = __ldg(reinterpret_cast<const float4*>(splats) + 3 * (size_t)g);
            const int2 rad = __ldg(reinterpret_cast<const int2*>(radii) + g);
            float tile_x = r0.x / (float)TILE, tile_y = r0.y / (float)TILE;
            float trx = (float)rad.x / (float)TILE, try_ = (float)rad.y / (float)TILE;
            float fw = (float)tile_w, fh = (float)tile_h;
            int minx = (int)fminf(fmaxf(floorf(tile_x - trx), 0.f), fw);
            int maxx = (int)fminf(fmaxf(ceilf(tile_x + trx), 0.f), fw);
            int miny = (int)fminf(fmaxf(floorf(tile_y - try_), 0.f), fh);
            s_minx[tid] = minx; s_miny[tid] = miny; s_bw[tid] = maxx - minx; s_g[tid] = g;
            s_rbw[tid] = 1.f / (float)(maxx - minx);
        }
    }
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    int woff = 0;
#pragma unroll
    for (int w = 0; w < EMIT_THREADS / 32; ++w) woff += (w < wid) ? s_warp[w] : 0;
    const int excl = woff + inc - cnt;
    s_excl[tid] = excl;
    if (tid == EMIT_THREADS - 1) s_excl[EMIT_THREADS] = excl + cnt;
    const int64_t boff = block_offsets[blockIdx.x];
    __syncthreads();
    const int total = s_excl[EMIT_THREADS];
    for (int k = tid; k < total; k += EMIT_THREADS) {
        int lo = 0, hi = EMIT_THREADS;
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (s_excl[mid] <= k) lo = mid; else hi = mid;
        }
        const int kk = k - s_excl[lo];
        const int bw = s_bw[lo];
        // row = kk / bw without an integer division: kk < 65536 tiles, so (kk + 0.5) / bw in fp32 floors exactly
        const int row = (int)(((float)kk + 0.5f) * s_rbw[lo]);
        const int ty = s_miny[lo] + row, tx = s_minx[lo] + (kk - row * bw);
        const int64_t out = boff + k;
        if (out < M) {
            tile_keys[out] = (uint16_t)(ty * tile_w + tx);
            flatten_ids[out] = s_g[lo];
        }
    }
}

__global__ void tile_ranges_k16_kernel(const uint16_t* __restrict__ keys, int64_t M, int n_tiles, int32_t* __restrict__ offsets) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M) return;
    int cur = (int)keys[idx];
    if (idx == 0) {
        for (int t = 0; t <= cur; ++t) offsets[t] = 0;
    } else {
        int prev = (int)keys[idx - 1];
        for (int t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)idx;
    }
    if (idx == M - 1)
        for (int t = cur + 1; t < n_tiles; ++t) offsets[t] = (int32_t)M;
}


// ---- tile_prep: per-tile counts from the corner-difference grid -> tile ranges + radix digit bases ----
// One CTA.  grid_g[(tile_h+1) x (tile_w+1)] holds +1/-1 at the corners of every visible Gaussian's tile
// rectangle (project_sh_fwd).  Its 2-D inclusive prefix sum is the number of intersections per tile; the
// row-major exclusive scan of that is isect_offsets (what tile_ranges derives from the SORTED keys), and the
// two 8-bit digit histograms of the tile-key sort follow from the same counts — so neither a histogram
// pass nor a ranges pass over the M intersections is needed.
#define PREP_THREADS 1024
__global__ void __launch_bounds__(PREP_THREADS)
tile_prep_kernel(const int32_t* __restrict__ grid_g, int tile_w, int tile_h, int32_t* __restrict__ offsets,
                 int64_t* __restrict__ total, uint32_t* __restrict__ digit_base) {
    extern __shared__ int s_grid[];
    __shared__ int s_wtot[PREP_THREADS / 32];
    __shared__ int s_h0[256];
    __shared__ int s_hw[8];
    __shared__ int s_total;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int stride = tile_w + 1, rows = tile_h + 1, cells = stride * rows, n_tiles = tile_w * tile_h;
    for (int i = tid; i < cells; i += PREP_THREADS) s_grid[i] = grid_g[i];
    __syncthreads();
    // inclusive scan along x, one warp per row
    for (int r = wid; r < rows; r += PREP_THREADS / 32) {
        int carry = 0;
        for (int x0 = 0; x0 < stride; x0 += 32) {
            const int x = x0 + lane;
            int v = (x < stride) ? s_grid[r * stride + x] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int n = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += n;
            }
            v += carry;
            if (x < stride) s_grid[r * stride + x] = v;
            carry = __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
    // inclusive scan along y, one thread per column
    for (int x = tid; x < stride; x += PREP_THREADS) {
        int acc = 0;
        for (int r = 0; r < rows; ++r) {
            acc += s_grid[r * stride + x];
            s_grid[r * stride + x] = acc;
        }
    }
    __syncthreads();
    // row-major exclusive scan of the per-tile counts: thread t owns `per` consecutive tiles
    const int per = (n_tiles + PREP_THREADS - 1) / PREP_THREADS;
    const int t0 = tid * per, t1 = min(t0 + per, n_tiles);
    int sum = 0;
    for (int t = t0; t < t1; ++t) {
        const int ty = t / tile_w, tx = t - ty * tile_w;
        sum += s_grid[ty * stride + tx];
    }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_wtot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int w = s_wtot[lane];
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += n;
        }
        s_wtot[lane] = winc - w;
        if (lane == 31) { total[0] = (int64_t)winc; s_total = winc; }
    }
    __syncthreads();
    int run = s_wtot[wid] + inc - sum;
    for (int t = t0; t < t1; ++t) {
        const int ty = t / tile_w, tx = t - ty * tile_w;
        offsets[t] = run;
        run += s_grid[ty * stride + tx];
    }
    // digit histogram of pass 0 (tile id & 255) and its exclusive scan
    if (tid < 256) {
        int h = 0;
        for (int t = tid; t < n_tiles; t += 256) {
            const int ty = t / tile_w, tx = t - ty * tile_w;
            h += s_grid[ty * stride + tx];
        }
        s_h0[tid] = h;
    }
    __syncthreads();  // also orders the global `offsets` writes above for the reads below (same CTA)
    if (tid < 256) {
        int h = s_h0[tid], hinc = h;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, hinc, o);
            if (lane >= o) hinc += n;
        }
        if (lane == 31) s_hw[wid] = hinc;
        asm volatile("bar.sync 1, 256;" ::: "memory");  // named barrier over the 8 warps taking part
        int woff = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) woff += (w < wid) ? s_hw[w] : 0;
        digit_base[tid] = (uint32_t)(woff + hinc - h);
        // pass 1 digit = tile id >> 8: tiles are already in digit order, so the base is the tile offset
        const int t = tid << 8;
        digit_base[256 + tid] = (uint32_t)((t < n_tiles) ? offsets[t] : s_total);
    }
}

// sorted 64-bit keys of the literal formulation, rebuilt on request (parity tests / gsplat meta)
__global__ void isect_ids_kernel(const uint16_t* __restrict__ tile_keys, const int32_t* __restrict__ flatten_ids,
                                 const float* __restrict__ splats, int64_t M, int64_t* __restrict__ isect_ids) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M) return;
    const uint32_t db = __float_as_uint(__ldg(splats + (size_t)flatten_ids[idx] * SPLAT_F + 9));
    isect_ids[idx] = ((int64_t)tile_keys[idx] << 32) | (int64_t)db;
}


// ---- two-level ("coarse-first") binning --------------------------------------------------------------------------
// The split-key path above still sorts the M (Gaussian, tile) intersections twice (two 8-bit passes of the tile key):
// 2 x 12 B x M of traffic and, at 6 M Gaussians / 4K, 1.2 ms of a 5.4 ms view.  Here the sorted unit is the
// (Gaussian, SUPERTILE) pair — a supertile is 4x4 tiles, 64x64 px — of which there are M' ~ M/3 .. M/5:
//   emit_coarse   Gaussians in depth order -> (supertile << 16 | 16-bit coverage mask, Gaussian id), M' pairs;
//                 output offsets by a decoupled look-back scan over 256-Gaussian chunks (no count / scan kernels),
//                 per-supertile counts in a CTA-private shared-memory histogram flushed once per CTA
//   coarse_prep   counts -> supertile ranges + the two digit bases of the sort (one CTA)
//   onesweep x2   stable sort of the M' pairs on the supertile bits (radix_sort.cu, device-side count)
//   expand        one CTA per supertile, one warp per tile: streams the supertile's depth-ordered list once and
//                 compacts, for its tile, the ids whose mask covers it straight into flatten_ids at
//                 isect_offsets[tile] (+ running count) — order inside a tile = order of the list = (depth, id)
// The result (flatten_ids, isect_offsets) is bit-identical to the stable sort of the 64-bit (tile, depth) keys: the
// coarse sort is stable, emission is in depth order with ties in ascending Gaussian index, and the compaction keeps
// list order.  Only ids are written per intersection (4 B x M); the tile of a position follows from isect_offsets.
#define ST_SHIFT 2
#define ST_SIDE (1 << ST_SHIFT)
#define ST_MAX_BINS 4096
#define EC_THREADS 1024
#define EC_CTAS_PER_SM 2

// Pass A: Gaussians in depth order -> packed tile rectangle (4 x u16, zero for culled) in depth order + the number of
// (Gaussian, supertile) pairs of every 1024-Gaussian chunk.  Embarrassingly parallel: the two dependent gathers
// (perm -> radii / means2d) of one CTA overlap with those of all the others; nothing waits on a neighbour.
__global__ void __launch_bounds__(EC_THREADS)
coarse_count_kernel(const float* __restrict__ splats, const int32_t* __restrict__ radii, const uint2* __restrict__ rects,
                    const int32_t* __restrict__ perm, int64_t N, int tile_w, int tile_h, uint2* __restrict__ rect_sorted,
                    uint32_t* __restrict__ chunk_sums) {
    __shared__ int s_warp[EC_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t i = (int64_t)blockIdx.x * EC_THREADS + tid;
    int cnt = 0;
    if (i < N) {
        const int g = __ldg(perm + i);
        uint2 packed = make_uint2(0u, 0u);
        if (rects != nullptr) {  // packed by the projection kernel: one 8-byte gather
            packed = __ldg(rects + g);
            if (packed.x != 0u || packed.y != 0u) {
                const int minx = (int)(packed.x & 0xFFFFu), maxx = (int)(packed.x >> 16);
                const int miny = (int)(packed.y & 0xFFFFu), maxy = (int)(packed.y >> 16);
                cnt = (((maxx - 1) >> ST_SHIFT) - (minx >> ST_SHIFT) + 1) * (((maxy - 1) >> ST_SHIFT) - (miny >> ST_SHIFT) + 1);
            }
            rect_sorted[i] = packed;
        }
        const int2 rad = (rects == nullptr) ? __ldg(reinterpret_cast<const int2*>(radii) + g) : make_int2(0, 0);
        if (rad.x > 0) {  // culled Gaussians carry radii (0, 0) (project_sh_fwd) and sort to the end of perm
            const float2 m2 = __ldg(reinterpret_cast<const float2*>(splats + SPLAT_F * (size_t)g));
            // the tile rectangle, with the same fp32 operations as the projection kernel / the oracle's tile_bounds
            float tile_x = m2.x / (float)TILE, tile_y = m2.y / (float)TILE;
            float trx = (float)rad.x / (float)TILE, try_ = (float)rad.y / (float)TILE;
            float fw = (float)tile_w, fh = (float)tile_h;
            const int minx = (int)fminf(fmaxf(floorf(tile_x - trx), 0.f), fw);
            const int maxx = (int)fminf(fmaxf(ceilf(tile_x + trx), 0.f), fw);
            const int miny = (int)fminf(fmaxf(floorf(tile_y - try_), 0.f), fh);
            const int maxy = (int)fminf(fmaxf(ceilf(tile_y + try_), 0.f), fh);
            if (maxx > minx && maxy > miny) {
                cnt = (((maxx - 1) >> ST_SHIFT) - (minx >> ST_SHIFT) + 1) * (((maxy - 1) >> ST_SHIFT) - (miny >> ST_SHIFT) + 1);
                packed = make_uint2((uint32_t)minx | ((uint32_t)maxx << 16), (uint32_t)miny | ((uint32_t)maxy << 16));
            }
        }
        if (rects == nullptr) rect_sorted[i] = packed;
    }
    int v = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_warp[wid] = v;
    __syncthreads();
    if (wid == 0) {
        int t = s_warp[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) chunk_sums[blockIdx.x] = (uint32_t)t;
    }
}

// Pass B: persistent, co-resident CTAs; CTA b takes chunks b, b + G, ... in increasing order and carries its output
// offset along (offset(c) = offset(c - G) + the chunk sums of [c - G, c): one block reduction, no waiting — every
// chunk sum was written by pass A).  Streams the packed rectangles (no gathers), emits the pairs
// load-balanced, counts them per supertile in shared memory, flushes the histogram once.
__global__ void __launch_bounds__(EC_THREADS, EC_CTAS_PER_SM)
emit_coarse_kernel(const uint2* __restrict__ rect_sorted, const int32_t* __restrict__ perm, int64_t N,
                   const uint32_t* __restrict__ chunk_sums, int st_w, int n_st, uint32_t cap,
                   uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ st_counts,
                   uint32_t* __restrict__ total_out) {
    extern __shared__ uint32_t s_hist[];  // n_st bins
    __shared__ int s_excl[EC_THREADS + 1];
    __shared__ uint2 s_rect[EC_THREADS];
    __shared__ short s_sx0[EC_THREADS], s_sy0[EC_THREADS], s_sw[EC_THREADS];
    __shared__ int s_g[EC_THREADS];
    __shared__ float s_rsw[EC_THREADS];
    __shared__ int s_warp[EC_THREADS / 32];
    __shared__ uint32_t s_red[EC_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int b = tid; b < n_st; b += EC_THREADS) s_hist[b] = 0;
    const uint32_t nchunks = (uint32_t)((N + EC_THREADS - 1) / EC_THREADS);
    const uint32_t G = gridDim.x;  // <= EC_THREADS (checked by the launcher)
    uint32_t boff = 0;
    for (uint32_t c = blockIdx.x; c < nchunks; c += G) {
        // output offset of chunk c: add the chunk sums of [c - G, c) (first trip: of [0, c))
        {
            const uint32_t first = (c >= G && c != blockIdx.x) ? c - G : 0u;
            const uint32_t j = first + tid;
            uint32_t v = (j < c) ? __ldg(chunk_sums + j) : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            __syncthreads();  // previous chunk's shared arrays (and s_red) are free; the histogram is zeroed on the first trip
            if (lane == 0) s_red[wid] = v;
        }
        const int64_t i = (int64_t)c * EC_THREADS + tid;
        int cnt = 0;
        if (i < N) {
            const uint2 r = __ldg(rect_sorted + i);
            if (r.x != 0u || r.y != 0u) {  // (maxx, maxy >= 1 for a visible Gaussian)
                const int minx = (int)(r.x & 0xFFFFu), maxx = (int)(r.x >> 16), miny = (int)(r.y & 0xFFFFu), maxy = (int)(r.y >> 16);
                const int sx0 = minx >> ST_SHIFT, sx1 = (maxx - 1) >> ST_SHIFT;
                const int sy0 = miny >> ST_SHIFT, sy1 = (maxy - 1) >> ST_SHIFT;
                const int sw = sx1 - sx0 + 1;
                cnt = sw * (sy1 - sy0 + 1);
                s_rect[tid] = r;
                s_sx0[tid] = (short)sx0; s_sy0[tid] = (short)sy0; s_sw[tid] = (short)sw;
                s_rsw[tid] = 1.f / (float)sw;
                s_g[tid] = __ldg(perm + i);
            }
        }
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        if (wid == 0) {  // scan of the 32 warp totals
            const int w = s_warp[lane];
            int winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int n = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += n;
            }
            s_warp[lane] = winc - w;
            if (lane == 31) s_excl[EC_THREADS] = winc;
        }
        {
            uint32_t t = s_red[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            boff += t;
        }
        __syncthreads();
        const int excl = s_warp[wid] + inc - cnt;
        s_excl[tid] = excl;
        __syncthreads();
        const uint32_t total = (uint32_t)s_excl[EC_THREADS];
        if (tid == 0 && c == nchunks - 1) total_out[0] = boff + total;
        for (uint32_t k = tid; k < total; k += EC_THREADS) {
            int lo = 0, hi = EC_THREADS;  // largest lo with s_excl[lo] <= k
#pragma unroll
            for (int step = 0; step < 10; ++step) {  // log2(EC_THREADS)
                const int mid = (lo + hi) >> 1;
                if (s_excl[mid] <= (int)k) lo = mid; else hi = mid;
            }
            const int kk = (int)k - s_excl[lo];
            const int sw = s_sw[lo];
            const int row = (int)(((float)kk + 0.5f) * s_rsw[lo]);  // kk / sw (kk < 4096: exact in fp32)
            const int sx = s_sx0[lo] + (kk - row * sw), sy = s_sy0[lo] + row;
            const uint2 r = s_rect[lo];
            const int minx = (int)(r.x & 0xFFFFu), maxx = (int)(r.x >> 16), miny = (int)(r.y & 0xFFFFu), maxy = (int)(r.y >> 16);
            // coverage of the supertile's 4x4 tiles by the Gaussian's tile rectangle: bit (4*ly + lx)
            const int x0 = max(minx - (sx << ST_SHIFT), 0), x1 = min(maxx - (sx << ST_SHIFT), ST_SIDE);
            const int y0 = max(miny - (sy << ST_SHIFT), 0), y1 = min(maxy - (sy << ST_SHIFT), ST_SIDE);
            const uint32_t rowmask = ((1u << x1) - 1u) & ~((1u << x0) - 1u);
            const uint32_t rows = ((1u << (4 * y1)) - 1u) & ~((1u << (4 * y0)) - 1u);  // y1 <= 4: 1u << 16 is fine
            const uint32_t mask = (rowmask * 0x1111u) & rows;
            const uint32_t st = (uint32_t)(sy * st_w + sx);
            const uint32_t out = boff + k;
            if (out < cap) {
                keys[out] = (st << 16) | mask;
                vals[out] = (uint32_t)s_g[lo];
            }
            atomicAdd(&s_hist[st], 1u);
        }
    }
    __syncthreads();
    for (int b = tid; b < n_st; b += EC_THREADS) {
        const uint32_t h = s_hist[b];
        if (h) atomicAdd(st_counts + b, h);
    }
}

// counts[n_st] -> st_offsets[n_st + 1] (exclusive scan; last = M') and the digit bases of the two sort passes.
__global__ void __launch_bounds__(1024)
coarse_prep_kernel(const uint32_t* __restrict__ st_counts, int n_st, uint32_t* __restrict__ st_offsets,
                   uint32_t* __restrict__ digit_base) {
    __shared__ uint32_t s_cnt[ST_MAX_BINS];
    __shared__ uint32_t s_off[ST_MAX_BINS + 1];
    __shared__ uint32_t s_wtot[32];
    __shared__ uint32_t s_h0[256];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i < ST_MAX_BINS; i += 1024) s_cnt[i] = (i < n_st) ? st_counts[i] : 0u;
    __syncthreads();
    // thread t owns 4 consecutive bins
    uint32_t v[4], sum = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) { v[j] = s_cnt[4 * tid + j]; sum += v[j]; }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_wtot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const uint32_t w = s_wtot[lane];
        uint32_t winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += n;
        }
        s_wtot[lane] = winc - w;
        if (lane == 31) s_off[ST_MAX_BINS] = winc;
    }
    __syncthreads();
    uint32_t run = s_wtot[wid] + inc - sum;
#pragma unroll
    for (int j = 0; j < 4; ++j) { s_off[4 * tid + j] = run; run += v[j]; }
    __syncthreads();
    const uint32_t total = s_off[ST_MAX_BINS];
    for (int i = tid; i <= n_st; i += 1024) st_offsets[i] = (i < n_st) ? s_off[i] : total;
    // pass 0 digit = supertile id & 255: histogram over residue classes, then exclusive scan
    if (tid < 256) {
        uint32_t h = 0;
        for (int sidx = tid; sidx < n_st; sidx += 256) h += s_cnt[sidx];
        s_h0[tid] = h;
    }
    __syncthreads();
    if (tid < 256) {
        uint32_t h = s_h0[tid], hinc = h;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, hinc, o);
            if (lane >= o) hinc += n;
        }
        if (lane == 31) s_wtot[wid] = hinc;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        uint32_t woff = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) woff += (w < wid) ? s_wtot[w] : 0u;
        digit_base[tid] = woff + hinc - h;
        // pass 1 digit = supertile id >> 8: supertiles are already in digit order, the base is the range start
        const int sidx = tid << 8;
        digit_base[256 + tid] = (sidx < n_st) ? s_off[sidx] : total;
    }
}

#define EX_THREADS 512
#define EX_CHUNK 2048
__global__ void __launch_bounds__(EX_THREADS)
expand_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals, const uint32_t* __restrict__ st_offsets,
              const int32_t* __restrict__ isect_offsets, int tile_w, int tile_h, int st_w, int32_t* __restrict__ flatten_ids) {
    __shared__ __align__(8) uint16_t s_mask[EX_CHUNK];
    __shared__ __align__(8) uint32_t s_val[EX_CHUNK];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;  // warp w <-> tile (w & 3, w >> 2) of the supertile
    const int st = blockIdx.x, sy = st / st_w, sx = st - sy * st_w;
    const int tx = (sx << ST_SHIFT) + (w & 3), ty = (sy << ST_SHIFT) + (w >> 2);
    const bool active = tx < tile_w && ty < tile_h;
    uint32_t out = active ? (uint32_t)__ldg(isect_offsets + ty * tile_w + tx) : 0u;
    const uint32_t b = __ldg(st_offsets + st), e = __ldg(st_offsets + st + 1);
    const unsigned lt_mask = (1u << lane) - 1u;
    const uint32_t bit = 1u << w, bit_hi = bit << 16;
    for (uint32_t base = b; base < e; base += EX_CHUNK) {
        const int n = (int)min((uint32_t)EX_CHUNK, e - base);
        __syncthreads();
        for (int j = tid; j < n; j += EX_THREADS) {
            s_mask[j] = (uint16_t)(__ldg(keys + base + j) & 0xFFFFu);
            s_val[j] = __ldg(vals + base + j);
        }
        if (tid < 64 && n + tid < EX_CHUNK) s_mask[n + tid] = 0;  // the pair loop below may read one pair past n
        __syncthreads();
        if (active) {
            // two list entries per lane and step (entries 2*lane, 2*lane + 1 of a run of 64): one LDS.32 of masks, one
            // LDS.64 of ids, two ballots; list order is kept (rank = hits in lower lanes, then the lane's even entry)
            for (int j = 0; j < n; j += 64) {
                const int idx = j + 2 * lane;
                const uint32_t m2 = *reinterpret_cast<const uint32_t*>(s_mask + idx);
                const bool h0 = (m2 & bit) != 0u, h1 = (m2 & bit_hi) != 0u;
                const unsigned b0 = __ballot_sync(0xffffffffu, h0), b1 = __ballot_sync(0xffffffffu, h1);
                if ((b0 | b1) == 0u) continue;
                const uint2 v2 = *reinterpret_cast<const uint2*>(s_val + idx);
                const uint32_t r = out + __popc(b0 & lt_mask) + __popc(b1 & lt_mask);
                if (h0) flatten_ids[r] = (int32_t)v2.x;
                if (h1) flatten_ids[r + (h0 ? 1u : 0u)] = (int32_t)v2.y;
                out += __popc(b0) + __popc(b1);
            }
        }
    }
}


// ---- tile_count: the corner-difference grid, off the projection kernel's critical path ----------------------------------
// Four scattered REDs per visible Gaussian inside the projection kernel cost it ~18 us per million Gaussians (one L2
// request per lane, adding to — not overlapping with — its other scattered traffic).  The same +1/-1 corner updates are
// made here instead, from the stored means2d / radii (identical fp32 operations, hence identical rectangles), by a few
// persistent CTAs that accumulate into a SHARED-MEMORY copy of the grid and flush it once: n_CTAs x cells REDs reach L2
// instead of 4 x N_visible.  The launch goes on the side stream with tile_prep and runs under the depth sort.
#define TC_THREADS 512
__global__ void __launch_bounds__(TC_THREADS)
tile_count_kernel(const float* __restrict__ splats, const int32_t* __restrict__ radii, int64_t N, int tile_w, int tile_h,
                  int32_t* __restrict__ grid_g, int use_smem) {
    extern __shared__ int s_cgrid[];
    const int stride = tile_w + 1, cells = stride * (tile_h + 1);
    if (use_smem) {
        for (int i = threadIdx.x; i < cells; i += TC_THREADS) s_cgrid[i] = 0;
        __syncthreads();
    }
    int* G = use_smem ? s_cgrid : grid_g;
    const float fw = (float)tile_w, fh = (float)tile_h;
    for (int64_t g = (int64_t)blockIdx.x * TC_THREADS + threadIdx.x; g < N; g += (int64_t)gridDim.x * TC_THREADS) {
        const int2 rad = __ldg(reinterpret_cast<const int2*>(radii) + g);
        if (rad.x <= 0) continue;
        const float2 m2 = __ldg(reinterpret_cast<const float2*>(splats + SPLAT_F * (size_t)g));
        const float tile_x = m2.x / (float)TILE, tile_y = m2.y / (float)TILE;
        const float trx = (float)rad.x / (float)TILE, try_ = (float)rad.y / (float)TILE;
        const int minx = (int)fminf(fmaxf(floorf(tile_x - trx), 0.f), fw);
        const int maxx = (int)fminf(fmaxf(ceilf(tile_x + trx), 0.f), fw);
        const int miny = (int)fminf(fmaxf(floorf(tile_y - try_), 0.f), fh);
        const int maxy = (int)fminf(fmaxf(ceilf(tile_y + try_), 0.f), fh);
        if (!(maxx > minx && maxy > miny)) continue;
        atomicAdd(G + miny * stride + minx, 1);
        atomicAdd(G + miny * stride + maxx, -1);
        atomicAdd(G + maxy * stride + minx, -1);
        atomicAdd(G + maxy * stride + maxx, 1);
    }
    if (use_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < cells; i += TC_THREADS) {
            const int v = s_cgrid[i];
            if (v != 0) atomicAdd(grid_g + i, v);
        }
    }
}

// sorted 64-bit keys rebuilt from (flatten_ids, isect_offsets) on request: the tile of position idx is the last tile
// whose range starts at or before idx
__global__ void isect_ids_from_offsets_kernel(const int32_t* __restrict__ flatten_ids, const int32_t* __restrict__ isect_offsets,
                                              int n_tiles, const float* __restrict__ splats, int64_t M,
                                              int64_t* __restrict__ isect_ids) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M) return;
    int lo = 0, hi = n_tiles;  // largest lo with isect_offsets[lo] <= idx
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if ((int64_t)__ldg(isect_offsets + mid) <= idx) lo = mid; else hi = mid;
    }
    const uint32_t db = __float_as_uint(__ldg(splats + (size_t)flatten_ids[idx] * SPLAT_F + 9));
    isect_ids[idx] = ((int64_t)lo << 32) | (int64_t)db;
}

}  // namespace

extern "C" int b200_bin_scan(const int64_t* block_sums, int64_t nblocks, int64_t* block_offsets, int64_t* total, void* stream) {
    B200_REQUIRE(nblocks >= 0 && total, "bad args");
    B200_REQUIRE(nblocks == 0 || (block_sums && block_offsets), "null pointer");
    bin_scan_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(block_sums, nblocks, block_offsets, total);
    return b200_check_launch("bin_scan");
}

extern "C" int b200_bin_emit(const float* splats, const int32_t* radii, const int32_t* tiles_per_gauss,
                             const int64_t* block_offsets, int64_t N, int width, int height, int64_t M,
                             int64_t* isect_ids, int32_t* flatten_ids, int64_t* cum_tiles, void* stream) {
    B200_REQUIRE(N >= 0 && M >= 0 && width > 0 && height > 0, "bad sizes");
    B200_REQUIRE(M < ((int64_t)1 << 30), "too many intersections (M >= 2^30)");
    if (N == 0) return B200_OK;
    B200_REQUIRE(splats && radii && tiles_per_gauss && block_offsets, "null pointer");
    B200_REQUIRE(M == 0 || (isect_ids && flatten_ids), "null output");
    int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    bin_emit_kernel<<<(unsigned)b200_project_blocks(N), EMIT_THREADS, 0, (cudaStream_t)stream>>>(
        splats, radii, tiles_per_gauss, block_offsets, N, tile_w, tile_h, M, isect_ids, flatten_ids, cum_tiles);
    return b200_check_launch("bin_emit");
}

extern "C" int b200_tile_ranges(const int64_t* sorted_isect_ids, int64_t M, int n_tiles, int32_t* offsets, void* stream) {
    B200_REQUIRE(M >= 0 && n_tiles > 0 && offsets, "bad args");
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 0) {
        cudaMemsetAsync(offsets, 0, sizeof(int32_t) * (size_t)n_tiles, st);
        return b200_check_launch("tile_ranges(memset)");
    }
    B200_REQUIRE(sorted_isect_ids, "null pointer");
    tile_ranges_kernel<<<(unsigned)((M + 255) / 256), 256, 0, st>>>(sorted_isect_ids, M, n_tiles, offsets);
    return b200_check_launch("tile_ranges");
}

extern "C" int b200_bin_count_sorted(const int32_t* tiles_per_gauss, const int32_t* perm, int64_t N, int64_t* block_sums,
                                     void* stream) {
    B200_REQUIRE(N >= 0, "bad size");
    if (N == 0) return B200_OK;
    B200_REQUIRE(tiles_per_gauss && perm && block_sums, "null pointer");
    bin_count_sorted_kernel<<<(unsigned)b200_project_blocks(N), EMIT_THREADS, 0, (cudaStream_t)stream>>>(tiles_per_gauss, perm, N, block_sums);
    return b200_check_launch("bin_count_sorted");
}

extern "C" int b200_bin_emit_sorted(const float* splats, const int32_t* radii, const int32_t* tiles_per_gauss,
                                    const int32_t* perm, const int64_t* block_offsets, int64_t N, int width, int height,
                                    int64_t M, uint16_t* tile_keys, int32_t* flatten_ids, void* stream) {
    B200_REQUIRE(N >= 0 && M >= 0 && width > 0 && height > 0, "bad sizes");
    B200_REQUIRE(M < ((int64_t)1 << 30), "too many intersections (M >= 2^30)");
    int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    B200_REQUIRE((int64_t)tile_w * tile_h <= 65536, "more than 65536 tiles: use the 64-bit key path");
    if (N == 0) return B200_OK;
    B200_REQUIRE(splats && radii && tiles_per_gauss && perm && block_offsets, "null pointer");
    B200_REQUIRE(M == 0 || (tile_keys && flatten_ids), "null output");
    bin_emit_sorted_kernel<<<(unsigned)b200_project_blocks(N), EMIT_THREADS, 0, (cudaStream_t)stream>>>(
        splats, radii, tiles_per_gauss, perm, block_offsets, N, tile_w, tile_h, M, tile_keys, flatten_ids);
    return b200_check_launch("bin_emit_sorted");
}

extern "C" int b200_tile_ranges_k16(const uint16_t* sorted_tile_keys, int64_t M, int n_tiles, int32_t* offsets, void* stream) {
    B200_REQUIRE(M >= 0 && n_tiles > 0 && n_tiles <= 65536 && offsets, "bad args");
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 0) {
        cudaMemsetAsync(offsets, 0, sizeof(int32_t) * (size_t)n_tiles, st);
        return b200_check_launch("tile_ranges_k16(memset)");
    }
    B200_REQUIRE(sorted_tile_keys, "null pointer");
    tile_ranges_k16_kernel<<<(unsigned)((M + 255) / 256), 256, 0, st>>>(sorted_tile_keys, M, n_tiles, offsets);
    return b200_check_launch("tile_ranges_k16");
}

extern "C" int b200_isect_ids(const uint16_t* sorted_tile_keys, const int32_t* flatten_ids, const float* splats, int64_t M,
                              int64_t* isect_ids, void* stream) {
    B200_REQUIRE(M >= 0, "bad size");
    if (M == 0) return B200_OK;
    B200_REQUIRE(sorted_tile_keys && flatten_ids && splats && isect_ids, "null pointer");
    isect_ids_kernel<<<(unsigned)((M + 255) / 256), 256, 0, (cudaStream_t)stream>>>(sorted_tile_keys, flatten_ids, splats, M, isect_ids);
    return b200_check_launch("isect_ids");
}

extern "C" size_t b200_tile_prep_max_cells(void) { return (size_t)(200 * 1024) / sizeof(int32_t); }

extern "C" int b200_tile_prep(const int32_t* tile_grid, int width, int height, int32_t* offsets, int64_t* total,
                              uint32_t* digit_base, void* stream) {
    B200_REQUIRE(width > 0 && height > 0 && tile_grid && offsets && total && digit_base, "bad args");
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const size_t cells = (size_t)(tile_w + 1) * (size_t)(tile_h + 1);
    B200_REQUIRE((int64_t)tile_w * tile_h <= 65536 && cells <= b200_tile_prep_max_cells(), "tile grid too large for tile_prep");
    static bool attr_set[B200_MAX_DEVICES] = {};
    if (b200_first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(tile_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    tile_prep_kernel<<<1, PREP_THREADS, cells * sizeof(int32_t), (cudaStream_t)stream>>>(tile_grid, tile_w, tile_h, offsets, total, digit_base);
    return b200_check_launch("tile_prep");
}

// ---- two-level binning: C ABI ----
int b200_sort_pairs_coarse(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, int64_t cap, int end_bit,
                           const uint32_t* digit_base, const uint32_t* n_dev, void* scratch, size_t scratch_bytes,
                           int* result_in_b_host, void* stream);  // radix_sort.cu

static inline size_t coarse_align(size_t x) { return (x + 255) / 256 * 256; }

extern "C" int b200_bin_coarse_supported(int width, int height) {
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const int st_w = (tile_w + ST_SIDE - 1) >> ST_SHIFT, st_h = (tile_h + ST_SIDE - 1) >> ST_SHIFT;
    return (int64_t)tile_w * tile_h <= 65536 && tile_w < 32768 && tile_h < 32768 && (int64_t)st_w * st_h <= ST_MAX_BINS;
}

extern "C" size_t b200_bin_coarse_scratch_bytes(int64_t N, int64_t M_cap) {
    if (N < 0) N = 0;
    if (M_cap < 0) M_cap = 0;
    const size_t head = 256                                   // ticket, M'
                        + coarse_align(ST_MAX_BINS * 4)       // st_counts
                        + coarse_align((ST_MAX_BINS + 1) * 4) // st_offsets
                        + coarse_align(512 * 4)               // digit bases
                        + coarse_align((size_t)((N + EC_THREADS - 1) / EC_THREADS) * 4)   // chunk sums
                        + coarse_align((size_t)N * 8);                                   // packed rectangles, depth order
    return head + b200_sort_scratch_bytes(M_cap);
}

extern "C" int b200_bin_coarse(const float* splats, const int32_t* radii, const void* tile_rects, const int32_t* perm,
                               int64_t N, int width, int height, const int32_t* isect_offsets, int64_t M,
                               uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b,
                               int32_t* flatten_ids, void* scratch, size_t scratch_bytes, void* stream) {
    B200_REQUIRE(N >= 0 && M >= 0 && width > 0 && height > 0, "bad sizes");
    B200_REQUIRE(M < ((int64_t)1 << 30), "too many intersections (M >= 2^30)");
    B200_REQUIRE(b200_bin_coarse_supported(width, height), "tile grid outside the coarse path's limits");
    if (N == 0 || M == 0) return B200_OK;
    B200_REQUIRE(splats && radii && perm && isect_offsets, "null pointer");
    B200_REQUIRE(keys_a && vals_a && keys_b && vals_b && flatten_ids && scratch, "null buffer");
    if (scratch_bytes < b200_bin_coarse_scratch_bytes(N, M)) {
        b200_set_error("b200_bin_coarse: scratch too small (%zu < %zu)", scratch_bytes, b200_bin_coarse_scratch_bytes(N, M));
        return B200_ESCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const int st_w = (tile_w + ST_SIDE - 1) >> ST_SHIFT, st_h = (tile_h + ST_SIDE - 1) >> ST_SHIFT;
    const int n_st = st_w * st_h;
    unsigned char* p = reinterpret_cast<unsigned char*>(scratch);
    uint32_t* ticket = reinterpret_cast<uint32_t*>(p);
    uint32_t* total = ticket + 16;
    p += 256;
    uint32_t* st_counts = reinterpret_cast<uint32_t*>(p); p += coarse_align(ST_MAX_BINS * 4);
    uint32_t* st_offsets = reinterpret_cast<uint32_t*>(p); p += coarse_align((ST_MAX_BINS + 1) * 4);
    uint32_t* digit_base = reinterpret_cast<uint32_t*>(p); p += coarse_align(512 * 4);
    const size_t nchunks = (size_t)((N + EC_THREADS - 1) / EC_THREADS);
    uint32_t* chunk_sums = reinterpret_cast<uint32_t*>(p); p += coarse_align(nchunks * 4);
    const size_t zeroed = (size_t)(p - reinterpret_cast<unsigned char*>(scratch));  // (the rectangles are fully written)
    uint2* rect_sorted = reinterpret_cast<uint2*>(p); p += coarse_align((size_t)N * 8);
    void* sort_scratch = p;
    const size_t head = (size_t)(p - reinterpret_cast<unsigned char*>(scratch));
    cudaMemsetAsync(scratch, 0, zeroed, st);
    static int sms[B200_MAX_DEVICES] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < B200_MAX_DEVICES && sms[dev] == 0) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    const int n_sm = (dev >= 0 && dev < B200_MAX_DEVICES && sms[dev] > 0) ? sms[dev] : 148;
    const unsigned grid = (unsigned)min((size_t)min(n_sm * EC_CTAS_PER_SM, EC_THREADS), nchunks);  // persistent CTAs
    (void)ticket;
    coarse_count_kernel<<<(unsigned)nchunks, EC_THREADS, 0, st>>>(splats, radii, reinterpret_cast<const uint2*>(tile_rects), perm, N,
                                                                  tile_w, tile_h, rect_sorted, chunk_sums);
    emit_coarse_kernel<<<grid, EC_THREADS, (size_t)n_st * 4, st>>>(rect_sorted, perm, N, chunk_sums, st_w, n_st, (uint32_t)M,
                                                                   keys_a, vals_a, st_counts, total);
    coarse_prep_kernel<<<1, 1024, 0, st>>>(st_counts, n_st, st_offsets, digit_base);
    int bits = 1;
    while ((1 << bits) < n_st) ++bits;  // bits carrying supertile ids 0 .. n_st-1
    int in_b = 0;
    const int rc = b200_sort_pairs_coarse(keys_a, vals_a, keys_b, vals_b, M, 16 + bits, digit_base, total, sort_scratch,
                                          scratch_bytes - head, &in_b, stream);
    if (rc != B200_OK) return rc;
    const uint32_t* ks = in_b ? keys_b : keys_a;
    const uint32_t* vs = in_b ? vals_b : vals_a;
    expand_kernel<<<(unsigned)n_st, EX_THREADS, 0, st>>>(ks, vs, st_offsets, isect_offsets, tile_w, tile_h, st_w, flatten_ids);
    return b200_check_launch("bin_coarse");
}

extern "C" int b200_isect_ids_from_offsets(const int32_t* flatten_ids, const int32_t* isect_offsets, int n_tiles,
                                           const float* splats, int64_t M, int64_t* isect_ids, void* stream) {
    B200_REQUIRE(M >= 0 && n_tiles > 0, "bad size");
    if (M == 0) return B200_OK;
    B200_REQUIRE(flatten_ids && isect_offsets && splats && isect_ids, "null pointer");
    isect_ids_from_offsets_kernel<<<(unsigned)((M + 255) / 256), 256, 0, (cudaStream_t)stream>>>(flatten_ids, isect_offsets, n_tiles,
                                                                                                  splats, M, isect_ids);
    return b200_check_launch("isect_ids_from_offsets");
}

// Corner-difference grid of the visible Gaussians' tile rectangles (zeroed here first): the input of b200_tile_prep when
// b200_project_sh_fwd was called with tile_grid == NULL.
extern "C" int b200_tile_count(const float* splats, const int32_t* radii, int64_t N, int width, int height, int32_t* tile_grid,
                               void* stream) {
    B200_REQUIRE(N >= 0 && width > 0 && height > 0 && tile_grid, "bad args");
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const size_t cells = (size_t)(tile_w + 1) * (size_t)(tile_h + 1);
    cudaStream_t st = (cudaStream_t)stream;
    cudaMemsetAsync(tile_grid, 0, cells * sizeof(int32_t), st);
    if (N == 0) return b200_check_launch("tile_count(memset)");
    B200_REQUIRE(splats && radii, "null pointer");
    static bool attr_set[B200_MAX_DEVICES] = {};
    if (b200_first_use_on_device(attr_set)) cudaFuncSetAttribute(tile_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const size_t smem = cells * sizeof(int32_t);
    const int use_smem = smem <= 200 * 1024;
    // one CTA per SM: every CTA flushes up to `cells` REDs, so fewer CTAs = fewer REDs reaching L2 (148 x 8281 = 1.2 M
    // against 4 x 857 k = 3.4 M at 1 M Gaussians / 1080p; 4.8 M against 20.6 M at 6 M / 4K); without the shared copy
    // (grids beyond 200 KB) the REDs go straight to L2 from more CTAs
    int64_t grid = (N + TC_THREADS - 1) / TC_THREADS;
    const int64_t cap = use_smem ? 148 : 148 * 4;
    if (grid > cap) grid = cap;
    tile_count_kernel<<<(unsigned)grid, TC_THREADS, use_smem ? smem : 0, st>>>(splats, radii, N, tile_w, tile_h, tile_grid, use_smem);
    return b200_check_launch("tile_count");
}
