/*
 * cpu_splat.c — TEST INFRASTRUCTURE ONLY: plain-C / OpenMP restatement of the rasterizer the reference
 * calls, used (a) as the checker at sizes the PyTorch oracle cannot reach and (b) as the timed CPU
 * baseline of bench.py (`cpu_baseline.kind = "port"`, and the `--impl reference` arm).
 *
 * Status: PARITY UNPINNED.  The arithmetic on this path lives in the un-vendored third-party wheel
 * gsplat==1.5.3 (reference requirements.txt:9; call sites src/backends/gsplat_backend.py:83-97 and
 * :168-182); the reference has no CPU implementation (src/backends/pytorch_backend.py:1-5 raises) and
 * no test of this path.  This file restates gsplat 1.5.x's published algorithm exactly as
 * oracle/torch_oracle.py does (SURVEY.md Appendix A is the spec) and is checked against that oracle's
 * autograd in tests/test_oracle_cpu.py.  The product path never links or loads this file.
 *
 * Each function names the spec section and the reference file:line whose behaviour it follows.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16
#define ALPHA_THRESHOLD (1.0f / 255.0f)
#define MAX_ALPHA 0.999f
#define T_STOP 1e-4f

typedef struct {
    float R[9], t[3], campos[3];
    float fx, fy, cx, cy, near_plane, far_plane, eps2d, radius_clip;
    int W, H, tile_w, tile_h;
} Cam;

/* cam24: 16 viewmat floats (row-major world->camera) + fx fy cx cy + near far eps2d radius_clip —
 * the arguments gsplat_backend.py:168-182 passes (viewmats, Ks) plus gsplat's defaults. */
static Cam make_cam(const float* c, int W, int H) {
    Cam cam;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) cam.R[i * 3 + j] = c[i * 4 + j];
        cam.t[i] = c[i * 4 + 3];
    }
    for (int j = 0; j < 3; ++j)
        cam.campos[j] = -(cam.R[j] * cam.t[0] + cam.R[3 + j] * cam.t[1] + cam.R[6 + j] * cam.t[2]);
    cam.fx = c[16]; cam.fy = c[17]; cam.cx = c[18]; cam.cy = c[19];
    cam.near_plane = c[20]; cam.far_plane = c[21]; cam.eps2d = c[22]; cam.radius_clip = c[23];
    cam.W = W; cam.H = H;
    cam.tile_w = (W + TILE - 1) / TILE;
    cam.tile_h = (H + TILE - 1) / TILE;
    return cam;
}

/* launchers such as torchrun export OMP_NUM_THREADS=1; the timed baseline asks for the host cores explicitly */
void cpu_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int cpu_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---------------------------------------------------------------------------------------------
 * A.1 projection (gsplat fully_fused_projection as called from gsplat_backend.py:168-182)
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    float pc[3], Rq[9], qn[4], inv_norm, Wm[9], J00, J02, J11, J12, T[6], tx, ty, rz;
    int clamp_x, clamp_y;
    float a, b, c, det, mx, my;
} Proj;

static int project_core(const Cam* cam, const float* p, const float* q, const float* s, Proj* P) {
    const float* R = cam->R;
    for (int i = 0; i < 3; ++i) P->pc[i] = R[i * 3] * p[0] + R[i * 3 + 1] * p[1] + R[i * 3 + 2] * p[2] + cam->t[i];
    float z = P->pc[2];
    if (!(z >= cam->near_plane) || !(z <= cam->far_plane)) return 0;
    float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    P->inv_norm = 1.0f / n;
    float w = q[0] / n, x = q[1] / n, y = q[2] / n, zz = q[3] / n;
    P->qn[0] = w; P->qn[1] = x; P->qn[2] = y; P->qn[3] = zz;
    float* Rq = P->Rq;
    Rq[0] = 1.f - 2.f * (y * y + zz * zz); Rq[1] = 2.f * (x * y - w * zz); Rq[2] = 2.f * (x * zz + w * y);
    Rq[3] = 2.f * (x * y + w * zz); Rq[4] = 1.f - 2.f * (x * x + zz * zz); Rq[5] = 2.f * (y * zz - w * x);
    Rq[6] = 2.f * (x * zz - w * y); Rq[7] = 2.f * (y * zz + w * x); Rq[8] = 1.f - 2.f * (x * x + y * y);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            P->Wm[i * 3 + j] = (R[i * 3] * Rq[j] + R[i * 3 + 1] * Rq[3 + j] + R[i * 3 + 2] * Rq[6 + j]) * s[j];
    float tanx = 0.5f * cam->W / cam->fx, tany = 0.5f * cam->H / cam->fy;
    float lim_xp = (cam->W - cam->cx) / cam->fx + 0.3f * tanx, lim_xn = cam->cx / cam->fx + 0.3f * tanx;
    float lim_yp = (cam->H - cam->cy) / cam->fy + 0.3f * tany, lim_yn = cam->cy / cam->fy + 0.3f * tany;
    float rz = 1.f / z, rz2 = rz * rz;
    float xr = P->pc[0] * rz, yr = P->pc[1] * rz;
    P->clamp_x = !(xr <= lim_xp && xr >= -lim_xn);
    P->clamp_y = !(yr <= lim_yp && yr >= -lim_yn);
    P->tx = z * fminf(lim_xp, fmaxf(-lim_xn, xr));
    P->ty = z * fminf(lim_yp, fmaxf(-lim_yn, yr));
    P->rz = rz;
    P->J00 = cam->fx * rz; P->J02 = -cam->fx * P->tx * rz2;
    P->J11 = cam->fy * rz; P->J12 = -cam->fy * P->ty * rz2;
    for (int j = 0; j < 3; ++j) {
        P->T[j] = P->J00 * P->Wm[j] + P->J02 * P->Wm[6 + j];
        P->T[3 + j] = P->J11 * P->Wm[3 + j] + P->J12 * P->Wm[6 + j];
    }
    float s00 = P->T[0] * P->T[0] + P->T[1] * P->T[1] + P->T[2] * P->T[2];
    float s01 = P->T[0] * P->T[3] + P->T[1] * P->T[4] + P->T[2] * P->T[5];
    float s11 = P->T[3] * P->T[3] + P->T[4] * P->T[4] + P->T[5] * P->T[5];
    P->a = s00 + cam->eps2d; P->b = s01; P->c = s11 + cam->eps2d;
    P->det = P->a * P->c - P->b * P->b;
    P->mx = cam->fx * P->pc[0] * rz + cam->cx;
    P->my = cam->fy * P->pc[1] * rz + cam->cy;
    return P->det > 0.f;
}

static int radius_and_cull(const Cam* cam, const Proj* P, float opac, int* rx, int* ry) {
    if (!(opac >= ALPHA_THRESHOLD)) return 0;
    float extend = fminf(3.33f, sqrtf(2.0f * logf(opac / ALPHA_THRESHOLD)));
    float frx = ceilf(extend * sqrtf(P->a)), fry = ceilf(extend * sqrtf(P->c));
    if (frx <= cam->radius_clip && fry <= cam->radius_clip) return 0;
    if (P->mx + frx <= 0.f || P->mx - frx >= (float)cam->W || P->my + fry <= 0.f || P->my - fry >= (float)cam->H) return 0;
    *rx = (int)frx; *ry = (int)fry;
    return *rx > 0 && *ry > 0;
}

/* A.2 real SH basis, degree <= 3 */
static void sh_basis(int deg, float x, float y, float z, float* B) {
    B[0] = 0.2820947917738781f;
    if (deg >= 1) { B[1] = -0.48860251190292f * y; B[2] = 0.48860251190292f * z; B[3] = -0.48860251190292f * x; }
    if (deg >= 2) {
        float z2 = z * z, fT0B = -1.092548430592079f * z, fC1 = x * x - y * y, fS1 = 2.f * x * y;
        B[4] = 0.5462742152960395f * fS1; B[5] = fT0B * y; B[6] = 0.9461746957575601f * z2 - 0.3153915652525201f;
        B[7] = fT0B * x; B[8] = 0.5462742152960395f * fC1;
        if (deg >= 3) {
            float fT0C = -2.285228997322329f * z2 + 0.4570457994644658f, fT1B = 1.445305721320277f * z;
            float fC2 = x * fC1 - y * fS1, fS2 = x * fS1 + y * fC1;
            B[9] = -0.5900435899266435f * fS2; B[10] = fT1B * fS1; B[11] = fT0C * y;
            B[12] = z * (1.865881662950577f * z2 - 1.119528997770346f); B[13] = fT0C * x; B[14] = fT1B * fC1;
            B[15] = -0.5900435899266435f * fC2;
        }
    }
}

static void tile_rect(const Cam* cam, float mx, float my, int rx, int ry, int* minx, int* miny, int* maxx, int* maxy) {
    float tile_x = mx / (float)TILE, tile_y = my / (float)TILE;
    float trx = (float)rx / (float)TILE, try_ = (float)ry / (float)TILE;
    float fw = (float)cam->tile_w, fh = (float)cam->tile_h;
    *minx = (int)fminf(fmaxf(floorf(tile_x - trx), 0.f), fw);
    *maxx = (int)fminf(fmaxf(ceilf(tile_x + trx), 0.f), fw);
    *miny = (int)fminf(fmaxf(floorf(tile_y - try_), 0.f), fh);
    *maxy = (int)fminf(fmaxf(ceilf(tile_y + try_), 0.f), fh);
}

/* Stage 1: radii[N,2], means2d[N,2], depths[N], conics[N,3], colors[N,3], tiles_per_gauss[N]. */
void cpu_project_sh_fwd(const float* means, const float* quats, const float* scales, const float* opac,
                        const float* shs, int sh_coeffs, int deg, int64_t N, const float* cam24, int W, int H,
                        int32_t* radii, float* means2d, float* depths, float* conics, float* colors,
                        int32_t* tiles_per_gauss) {
    Cam cam = make_cam(cam24, W, H);
    const int nb = (deg + 1) * (deg + 1);
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < N; ++g) {
        Proj P;
        int rx = 0, ry = 0;
        int ok = project_core(&cam, means + 3 * g, quats + 4 * g, scales + 3 * g, &P) && radius_and_cull(&cam, &P, opac[g], &rx, &ry);
        if (!ok) {
            radii[2 * g] = radii[2 * g + 1] = 0;
            means2d[2 * g] = means2d[2 * g + 1] = 0.f; depths[g] = 0.f;
            conics[3 * g] = conics[3 * g + 1] = conics[3 * g + 2] = 0.f;
            colors[3 * g] = colors[3 * g + 1] = colors[3 * g + 2] = 0.f;
            tiles_per_gauss[g] = 0;
            continue;
        }
        radii[2 * g] = rx; radii[2 * g + 1] = ry;
        means2d[2 * g] = P.mx; means2d[2 * g + 1] = P.my; depths[g] = P.pc[2];
        float inv_det = 1.f / P.det;
        conics[3 * g] = P.c * inv_det; conics[3 * g + 1] = -P.b * inv_det; conics[3 * g + 2] = P.a * inv_det;
        float u[3] = {means[3 * g] - cam.campos[0], means[3 * g + 1] - cam.campos[1], means[3 * g + 2] - cam.campos[2]};
        float inv = 1.f / sqrtf(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        float B[16];
        sh_basis(deg, u[0] * inv, u[1] * inv, u[2] * inv, B);
        const float* c = shs + g * (int64_t)sh_coeffs * 3;
        for (int ch = 0; ch < 3; ++ch) {
            float acc = 0.f;
            for (int k = 0; k < nb; ++k) acc += B[k] * c[3 * k + ch];
            colors[3 * g + ch] = fmaxf(acc + 0.5f, 0.f);
        }
        int minx, miny, maxx, maxy;
        tile_rect(&cam, P.mx, P.my, rx, ry, &minx, &miny, &maxx, &maxy);
        tiles_per_gauss[g] = (maxy - miny) * (maxx - minx);
    }
}

/* ---------------------------------------------------------------------------------------------
 * A.3 tile binning: emission in gsplat isect_tiles order, stable LSD radix sort (what
 * cub::DeviceRadixSort gives gsplat), isect_offset_encode.
 * ------------------------------------------------------------------------------------------- */
/* cum[N] = inclusive cumsum of tiles_per_gauss; returns M. */
int64_t cpu_cumsum(const int32_t* tiles_per_gauss, int64_t N, int64_t* cum) {
    int64_t run = 0;
    for (int64_t g = 0; g < N; ++g) { run += tiles_per_gauss[g]; cum[g] = run; }
    return run;
}

void cpu_bin_emit(const float* means2d, const int32_t* radii, const float* depths, const int64_t* cum, int64_t N,
                  int W, int H, int64_t* isect_ids, int32_t* flatten_ids) {
    Cam cam; memset(&cam, 0, sizeof cam);
    cam.tile_w = (W + TILE - 1) / TILE; cam.tile_h = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < N; ++g) {
        if (radii[2 * g] <= 0 || radii[2 * g + 1] <= 0) continue;
        int minx, miny, maxx, maxy;
        tile_rect(&cam, means2d[2 * g], means2d[2 * g + 1], radii[2 * g], radii[2 * g + 1], &minx, &miny, &maxx, &maxy);
        int64_t o = g ? cum[g - 1] : 0;
        uint32_t db; memcpy(&db, depths + g, 4);
        for (int i = miny; i < maxy; ++i)
            for (int j = minx; j < maxx; ++j) {
                isect_ids[o] = ((int64_t)(i * cam.tile_w + j) << 32) | (int64_t)db;
                flatten_ids[o] = (int32_t)g;
                ++o;
            }
    }
}

/* Stable LSD radix sort of (key,val) over bits [0,end_bit), 8 bits per pass, thread-chunked
 * histograms.  Result always ends in (keys, vals); (tk, tv) are caller-provided temporaries. */
void cpu_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t* tk, uint32_t* tv, int64_t M, int end_bit) {
    int npass = (end_bit + 7) / 8;
    int nt = cpu_num_threads();
    int64_t* hist = (int64_t*)malloc(sizeof(int64_t) * 256 * (size_t)nt);
    uint64_t *ki = keys, *ko = tk; uint32_t *vi = vals, *vo = tv;
    for (int p = 0; p < npass; ++p) {
        int shift = p * 8;
        memset(hist, 0, sizeof(int64_t) * 256 * (size_t)nt);
#pragma omp parallel num_threads(nt)
        {
#ifdef _OPENMP
            int t = omp_get_thread_num();
#else
            int t = 0;
#endif
            int64_t lo = M * t / nt, hi = M * (t + 1) / nt;
            int64_t* h = hist + 256 * t;
            for (int64_t i = lo; i < hi; ++i) h[(ki[i] >> shift) & 255]++;
#pragma omp barrier
#pragma omp single
            {
                int64_t run = 0;
                for (int d = 0; d < 256; ++d)
                    for (int tt = 0; tt < nt; ++tt) { int64_t c = hist[256 * tt + d]; hist[256 * tt + d] = run; run += c; }
            }
            for (int64_t i = lo; i < hi; ++i) {
                int64_t o = h[(ki[i] >> shift) & 255]++;
                ko[o] = ki[i]; vo[o] = vi[i];
            }
        }
        uint64_t* a = ki; ki = ko; ko = a;
        uint32_t* b = vi; vi = vo; vo = b;
    }
    if (ki != keys) { memcpy(keys, ki, sizeof(uint64_t) * (size_t)M); memcpy(vals, vi, sizeof(uint32_t) * (size_t)M); }
    free(hist);
}

void cpu_tile_ranges(const int64_t* ids, int64_t M, int n_tiles, int32_t* offsets) {
    if (M == 0) { memset(offsets, 0, sizeof(int32_t) * (size_t)n_tiles); return; }
    int64_t i = 0;
    for (int t = 0; t < n_tiles; ++t) {
        while (i < M && (ids[i] >> 32) < t) ++i;
        offsets[t] = (int32_t)i;
    }
}

/* ---------------------------------------------------------------------------------------------
 * A.4 forward blend (gsplat rasterize_to_pixels fwd); CH = 3 (rgb), 1 (depth) or 4 (rgb+depth):
 * feat[N,CH] is the per-Gaussian feature vector composited.
 * ------------------------------------------------------------------------------------------- */
/* P of SURVEY.md §8(d): (pixel, Gaussian) pairs evaluated by the last cpu_raster_fwd call (pairs visited before a
 * pixel terminates), and how many of them were composited. */
static long long g_pairs_evaluated = 0, g_pairs_blended = 0;
long long cpu_pairs_evaluated(void) { return g_pairs_evaluated; }
long long cpu_pairs_blended(void) { return g_pairs_blended; }

void cpu_raster_fwd(const float* means2d, const float* conics, const float* feat, int CH, const float* opac,
                    const int32_t* flatten_ids, const int32_t* offsets, int64_t M, int W, int H, const float* bg,
                    float* render, float* alpha, int32_t* last_ids) {
    int tile_w = (W + TILE - 1) / TILE, tile_h = (H + TILE - 1) / TILE, n_tiles = tile_w * tile_h;
    long long pairs = 0, blended = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : pairs, blended)
    for (int t = 0; t < n_tiles; ++t) {
        int start = offsets[t], end = (t + 1 < n_tiles) ? offsets[t + 1] : (int)M;
        int ty = t / tile_w, tx = t % tile_w;
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int ix = tx * TILE + lx, iy = ty * TILE + ly;
                if (ix >= W || iy >= H) continue;
                float px = ix + 0.5f, py = iy + 0.5f, T = 1.f, acc[4] = {0, 0, 0, 0};
                int last = -1;
                for (int idx = start; idx < end; ++idx) {
                    ++pairs;
                    int g = flatten_ids[idx];
                    float dx = means2d[2 * g] - px, dy = means2d[2 * g + 1] - py;
                    float sigma = 0.5f * (conics[3 * g] * dx * dx + conics[3 * g + 2] * dy * dy) + conics[3 * g + 1] * dx * dy;
                    float al = fminf(MAX_ALPHA, opac[g] * expf(-sigma));
                    if (sigma < 0.f || al < ALPHA_THRESHOLD) continue;
                    float Tn = T * (1.f - al);
                    if (Tn <= T_STOP) break;
                    float w = al * T;
                    for (int c = 0; c < CH; ++c) acc[c] += w * feat[(size_t)g * CH + c];
                    last = idx; T = Tn;
                    ++blended;
                }
                size_t pix = (size_t)iy * W + ix;
                for (int c = 0; c < CH; ++c) render[pix * CH + c] = acc[c] + T * (bg ? bg[c] : 0.f);
                alpha[pix] = 1.f - T;
                last_ids[pix] = last;
            }
    }
    g_pairs_evaluated = pairs;
    g_pairs_blended = blended;
}

/* ---------------------------------------------------------------------------------------------
 * A.5 backward blend.  Accumulates v_means2d[N,2], v_conics[N,3], v_feat[N,CH], v_opac[N]
 * (caller zeroes).  Per tile a private accumulator per listed Gaussian, flushed with atomics.
 * ------------------------------------------------------------------------------------------- */
void cpu_raster_bwd(const float* means2d, const float* conics, const float* feat, int CH, const float* opac,
                    const int32_t* flatten_ids, const int32_t* offsets, int64_t M, int W, int H, const float* bg,
                    const float* alpha, const int32_t* last_ids, const float* v_render, const float* v_alpha,
                    float* v_means2d, float* v_conics, float* v_feat, float* v_opac) {
    int tile_w = (W + TILE - 1) / TILE, tile_h = (H + TILE - 1) / TILE, n_tiles = tile_w * tile_h;
#pragma omp parallel
    {
        float* loc = NULL; size_t cap = 0;
#pragma omp for schedule(dynamic, 4)
        for (int t = 0; t < n_tiles; ++t) {
            int start = offsets[t];
            int ty = t / tile_w, tx = t % tile_w;
            int tile_last = -1;
            for (int ly = 0; ly < TILE; ++ly)
                for (int lx = 0; lx < TILE; ++lx) {
                    int ix = tx * TILE + lx, iy = ty * TILE + ly;
                    if (ix < W && iy < H && last_ids[(size_t)iy * W + ix] > tile_last) tile_last = last_ids[(size_t)iy * W + ix];
                }
            if (tile_last < start) continue;
            size_t cnt = (size_t)(tile_last - start + 1), stride = (size_t)(6 + CH);
            if (cnt * stride > cap) { free(loc); cap = cnt * stride * 2; loc = (float*)malloc(sizeof(float) * cap); }
            memset(loc, 0, sizeof(float) * cnt * stride);
            for (int ly = 0; ly < TILE; ++ly)
                for (int lx = 0; lx < TILE; ++lx) {
                    int ix = tx * TILE + lx, iy = ty * TILE + ly;
                    if (ix >= W || iy >= H) continue;
                    size_t pix = (size_t)iy * W + ix;
                    int bin_final = last_ids[pix];
                    if (bin_final < start) continue;
                    float px = ix + 0.5f, py = iy + 0.5f;
                    float T_final = 1.f - alpha[pix], T = T_final, buf[4] = {0, 0, 0, 0}, vr[4] = {0, 0, 0, 0};
                    float va = v_alpha ? v_alpha[pix] : 0.f, bgdot = 0.f;
                    for (int c = 0; c < CH; ++c) { vr[c] = v_render[pix * CH + c]; bgdot += (bg ? bg[c] : 0.f) * vr[c]; }
                    for (int idx = bin_final; idx >= start; --idx) {
                        int g = flatten_ids[idx];
                        float A = conics[3 * g], B = conics[3 * g + 1], Cc = conics[3 * g + 2], o = opac[g];
                        float dx = means2d[2 * g] - px, dy = means2d[2 * g + 1] - py;
                        float sigma = 0.5f * (A * dx * dx + Cc * dy * dy) + B * dx * dy;
                        float vis = expf(-sigma), al = fminf(MAX_ALPHA, o * vis);
                        if (sigma < 0.f || al < ALPHA_THRESHOLD) continue;
                        float ra = 1.f / (1.f - al);
                        T *= ra;
                        float fac = al * T, v_al = 0.f;
                        float* L = loc + (size_t)(idx - start) * stride;
                        for (int c = 0; c < CH; ++c) {
                            float col = feat[(size_t)g * CH + c];
                            L[6 + c] += fac * vr[c];
                            v_al += (col * T - buf[c] * ra) * vr[c];
                            buf[c] += col * fac;
                        }
                        v_al += T_final * ra * (va - bgdot);
                        if (o * vis <= MAX_ALPHA) {
                            float v_sigma = -o * vis * v_al;
                            L[0] += v_sigma * (A * dx + B * dy);
                            L[1] += v_sigma * (B * dx + Cc * dy);
                            L[2] += 0.5f * v_sigma * dx * dx;
                            L[3] += v_sigma * dx * dy;
                            L[4] += 0.5f * v_sigma * dy * dy;
                            L[5] += vis * v_al;
                        }
                    }
                }
            for (size_t k = 0; k < cnt; ++k) {
                int g = flatten_ids[start + (int)k];
                const float* L = loc + k * stride;
#pragma omp atomic
                v_means2d[2 * g] += L[0];
#pragma omp atomic
                v_means2d[2 * g + 1] += L[1];
#pragma omp atomic
                v_conics[3 * g] += L[2];
#pragma omp atomic
                v_conics[3 * g + 1] += L[3];
#pragma omp atomic
                v_conics[3 * g + 2] += L[4];
#pragma omp atomic
                v_opac[g] += L[5];
                for (int c = 0; c < CH; ++c) {
#pragma omp atomic
                    v_feat[(size_t)g * CH + c] += L[6 + c];
                }
            }
        }
        free(loc);
    }
}

/* ---------------------------------------------------------------------------------------------
 * A.6 projection + SH backward: cotangents of (means2d, conics, colors, depths) -> the five inputs.
 * v_opac_in is the opacity cotangent from the blend (passed through).
 * ------------------------------------------------------------------------------------------- */
void cpu_project_sh_bwd(const float* means, const float* quats, const float* scales, const float* shs, int sh_coeffs,
                        int deg, int64_t N, const float* cam24, int W, int H, const int32_t* radii,
                        const float* v_means2d, const float* v_conics, const float* v_colors, const float* v_depths,
                        float* v_means, float* v_quats, float* v_scales, float* v_shs) {
    Cam cam = make_cam(cam24, W, H);
    const int nb = (deg + 1) * (deg + 1);
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < N; ++g) {
        float* vm = v_means + 3 * g; float* vq = v_quats + 4 * g; float* vs = v_scales + 3 * g;
        float* vsh = v_shs + g * (int64_t)sh_coeffs * 3;
        vm[0] = vm[1] = vm[2] = 0.f; vq[0] = vq[1] = vq[2] = vq[3] = 0.f; vs[0] = vs[1] = vs[2] = 0.f;
        for (int i = 0; i < sh_coeffs * 3; ++i) vsh[i] = 0.f;
        if (radii[2 * g] <= 0 || radii[2 * g + 1] <= 0) continue;
        Proj P;
        const float* s = scales + 3 * g;
        project_core(&cam, means + 3 * g, quats + 4 * g, s, &P);
        /* SH */
        float u[3] = {means[3 * g] - cam.campos[0], means[3 * g + 1] - cam.campos[1], means[3 * g + 2] - cam.campos[2]};
        float inv = 1.f / sqrtf(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        float x = u[0] * inv, y = u[1] * inv, z = u[2] * inv, B[16], gk[16];
        sh_basis(deg, x, y, z, B);
        const float* c = shs + g * (int64_t)sh_coeffs * 3;
        float vr[3];
        for (int ch = 0; ch < 3; ++ch) {
            float acc = 0.f;
            for (int k = 0; k < nb; ++k) acc += B[k] * c[3 * k + ch];
            vr[ch] = (acc + 0.5f >= 0.f) ? v_colors[3 * g + ch] : 0.f;
        }
        for (int k = 0; k < nb; ++k) {
            for (int ch = 0; ch < 3; ++ch) vsh[3 * k + ch] = B[k] * vr[ch];
            gk[k] = c[3 * k] * vr[0] + c[3 * k + 1] * vr[1] + c[3 * k + 2] * vr[2];
        }
        float vd[3] = {0, 0, 0};
        const float C1 = 0.48860251190292f, C2A = 0.5462742152960395f, C2B = 1.092548430592079f, C2C = 0.9461746957575601f;
        const float C3A = 0.5900435899266435f, C3B = 1.445305721320277f, C3C = 2.285228997322329f, C3D = 0.4570457994644658f;
        const float C3E = 1.865881662950577f, C3F = 1.119528997770346f;
        if (deg >= 1) { vd[1] += -C1 * gk[1]; vd[2] += C1 * gk[2]; vd[0] += -C1 * gk[3]; }
        if (deg >= 2) {
            vd[0] += C2A * 2.f * y * gk[4]; vd[1] += C2A * 2.f * x * gk[4];
            vd[1] += -C2B * z * gk[5]; vd[2] += -C2B * y * gk[5];
            vd[2] += 2.f * C2C * z * gk[6];
            vd[0] += -C2B * z * gk[7]; vd[2] += -C2B * x * gk[7];
            vd[0] += C2A * 2.f * x * gk[8]; vd[1] += -C2A * 2.f * y * gk[8];
        }
        if (deg >= 3) {
            float z2 = z * z, fT0C = -C3C * z2 + C3D, d2 = x * x - y * y;
            vd[0] += -C3A * 6.f * x * y * gk[9]; vd[1] += -C3A * 3.f * d2 * gk[9];
            vd[0] += 2.f * C3B * y * z * gk[10]; vd[1] += 2.f * C3B * x * z * gk[10]; vd[2] += 2.f * C3B * x * y * gk[10];
            vd[1] += fT0C * gk[11]; vd[2] += -2.f * C3C * z * y * gk[11];
            vd[2] += (3.f * C3E * z2 - C3F) * gk[12];
            vd[0] += fT0C * gk[13]; vd[2] += -2.f * C3C * z * x * gk[13];
            vd[0] += C3B * 2.f * x * z * gk[14]; vd[1] += -C3B * 2.f * y * z * gk[14]; vd[2] += C3B * d2 * gk[14];
            vd[0] += -C3A * 3.f * d2 * gk[15]; vd[1] += C3A * 6.f * x * y * gk[15];
        }
        float dot = x * vd[0] + y * vd[1] + z * vd[2];
        vm[0] += (vd[0] - x * dot) * inv; vm[1] += (vd[1] - y * dot) * inv; vm[2] += (vd[2] - z * dot) * inv;
        /* conic -> covariance */
        float inv_det = 1.f / P.det, A = P.c * inv_det, Bc = -P.b * inv_det, C = P.a * inv_det;
        float v_A = v_conics[3 * g], hB = 0.5f * v_conics[3 * g + 1], v_C = v_conics[3 * g + 2];
        float xv00 = A * v_A + Bc * hB, xv01 = A * hB + Bc * v_C, xv10 = Bc * v_A + C * hB, xv11 = Bc * hB + C * v_C;
        float G00 = -(xv00 * A + xv01 * Bc), G01 = -(xv00 * Bc + xv01 * C), G10 = -(xv10 * A + xv11 * Bc), G11 = -(xv10 * Bc + xv11 * C);
        float Gs = 0.5f * (G01 + G10), vT[6], vW[9], vRq[9];
        for (int j = 0; j < 3; ++j) {
            vT[j] = 2.f * (G00 * P.T[j] + Gs * P.T[3 + j]);
            vT[3 + j] = 2.f * (Gs * P.T[j] + G11 * P.T[3 + j]);
        }
        float vJ00 = 0, vJ02 = 0, vJ11 = 0, vJ12 = 0;
        for (int j = 0; j < 3; ++j) {
            vJ00 += vT[j] * P.Wm[j]; vJ02 += vT[j] * P.Wm[6 + j];
            vJ11 += vT[3 + j] * P.Wm[3 + j]; vJ12 += vT[3 + j] * P.Wm[6 + j];
            vW[j] = P.J00 * vT[j]; vW[3 + j] = P.J11 * vT[3 + j]; vW[6 + j] = P.J02 * vT[j] + P.J12 * vT[3 + j];
        }
        const float* R = cam.R;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) {
                float vMij = R[i] * vW[j] + R[3 + i] * vW[3 + j] + R[6 + i] * vW[6 + j];
                vs[j] += P.Rq[i * 3 + j] * vMij;
                vRq[i * 3 + j] = vMij * s[j];
            }
        {
            float w = P.qn[0], qx = P.qn[1], qy = P.qn[2], qz = P.qn[3];
            float vw = 2.f * (-qz * vRq[1] + qy * vRq[2] + qz * vRq[3] - qx * vRq[5] - qy * vRq[6] + qx * vRq[7]);
            float vx = 2.f * (qy * vRq[1] + qz * vRq[2] + qy * vRq[3] - 2.f * qx * vRq[4] - w * vRq[5] + qz * vRq[6] + w * vRq[7] - 2.f * qx * vRq[8]);
            float vy = 2.f * (-2.f * qy * vRq[0] + qx * vRq[1] + w * vRq[2] + qx * vRq[3] + qz * vRq[5] - w * vRq[6] + qz * vRq[7] - 2.f * qy * vRq[8]);
            float vz = 2.f * (-2.f * qz * vRq[0] - w * vRq[1] + qx * vRq[2] + w * vRq[3] - 2.f * qz * vRq[4] + qy * vRq[5] + qx * vRq[6] + qy * vRq[7]);
            float d = w * vw + qx * vx + qy * vy + qz * vz;
            vq[0] = (vw - w * d) * P.inv_norm; vq[1] = (vx - qx * d) * P.inv_norm;
            vq[2] = (vy - qy * d) * P.inv_norm; vq[3] = (vz - qz * d) * P.inv_norm;
        }
        float rz = P.rz, rz2 = rz * rz, rz3 = rz2 * rz, v_mx = v_means2d[2 * g], v_my = v_means2d[2 * g + 1];
        float vpx = cam.fx * rz * v_mx, vpy = cam.fy * rz * v_my;
        float vpz = -(cam.fx * P.pc[0] * v_mx + cam.fy * P.pc[1] * v_my) * rz2 + (v_depths ? v_depths[g] : 0.f);
        vpz += -cam.fx * rz2 * vJ00 - cam.fy * rz2 * vJ11;
        if (!P.clamp_x) { vpx += -cam.fx * rz2 * vJ02; vpz += 2.f * cam.fx * P.tx * rz3 * vJ02; } else vpz += cam.fx * P.tx * rz3 * vJ02;
        if (!P.clamp_y) { vpy += -cam.fy * rz2 * vJ12; vpz += 2.f * cam.fy * P.ty * rz3 * vJ12; } else vpz += cam.fy * P.ty * rz3 * vJ12;
        vm[0] += R[0] * vpx + R[3] * vpy + R[6] * vpz;
        vm[1] += R[1] * vpx + R[4] * vpy + R[7] * vpz;
        vm[2] += R[2] * vpx + R[5] * vpy + R[8] * vpz;
    }
}

/* A.7 stage 5 (reference src/training/trainer.py:786-791) */
void cpu_densify_stats(const float* grad, int64_t N, float* accum, float* count) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        float n = sqrtf(grad[3 * i] * grad[3 * i] + grad[3 * i + 1] * grad[3 * i + 1] + grad[3 * i + 2] * grad[3 * i + 2]);
        accum[3 * i] += n; accum[3 * i + 1] += n; accum[3 * i + 2] += n;
        count[i] += 1.f;
    }
}
