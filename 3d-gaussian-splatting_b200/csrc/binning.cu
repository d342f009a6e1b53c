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
