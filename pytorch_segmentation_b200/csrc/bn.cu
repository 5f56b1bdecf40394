// Training-mode BatchNorm around the tcgen05 convs: per-channel statistics, scale/shift finalisation with the
// running-stat update, the normalise(+ReLU) apply pass that also places the result in a concat slice, and the
// two-pass backward.  All HBM-bound: 128-bit loads of 8 bf16 channels per thread, consecutive threads on
// consecutive channel groups of one pixel (coalesced over the NHWC row), fp32 math, one atomic per channel per
// block for the reductions.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

#include <stdlib.h>

namespace b200seg {
namespace {

constexpr int kRedThreads = 256;

// pixels per thread per trip of the streaming BatchNorm passes (1, 2 or 4): B200SEG_BN_K sets all three kernels,
// B200SEG_BN_K_APPLY / _BWD / _RED one of them (A/B runs: scripts/bench_bandwidth.py)
int bn_env_k(const char* name, int dflt) {
    const char* e = getenv(name);
    if (!e) e = getenv("B200SEG_BN_K");
    const int v = e ? atoi(e) : dflt;
    return v == 1 || v == 2 || v == 4 ? v : dflt;
}

int bn_env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    const int v = e ? atoi(e) : dflt;
    return v > 0 ? v : dflt;
}

// Geometry of a column reduction over a [pixels][c] slice: G = c/8 channel groups; a block covers
// `gpb` groups (<= 256) and 256/gpb pixel rows per iteration.
struct ColGeom {
    int groups, gpb, rows, gy;
    int64_t pixels_per_block;
    int gx;
};

// max_gpb: widest channel span of one block.  A REDUCTION ends with one atomic per channel per block, so total atomics =
// blocks x channels-per-block: spanning 32 groups (a warp reads 512 contiguous bytes of a pixel) instead of the whole
// row cuts them 4-8x on the 1024 / 2048-channel tensors and lets the grid's second dimension cover the rest.
// launch geometry of the two elementwise passes (bn_apply, bn_bwd_apply): blocks per SM the grid is sized for and the
// widest channel span of a block (A/B: B200SEG_BNAPPLY_CTAS / B200SEG_BNAPPLY_GPB)
int apply_ctas() {
    static const int v = bn_env_int("B200SEG_BNAPPLY_CTAS", 4);
    return v;
}
int apply_gpb() {
    static const int v = bn_env_int("B200SEG_BNAPPLY_GPB", 32);
    return v;
}

ColGeom col_geom(int c, int64_t pixels, int ctas_per_sm = 8, int max_gpb = kRedThreads) {
    ColGeom g;
    g.groups = c / 8;
    g.gpb = g.groups < max_gpb ? g.groups : max_gpb;
    g.rows = kRedThreads / g.gpb;
    g.gy = ceil_div(g.groups, g.gpb);
    int sms = sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = static_cast<int64_t>(sms) * ctas_per_sm / g.gy;
    if (want < 1) want = 1;
    int64_t ppb = ceil_div64(pixels, want);
    const int64_t min_ppb = static_cast<int64_t>(g.rows) * 4;
    if (ppb < min_ppb) ppb = min_ppb;
    ppb = ceil_div64(ppb, g.rows) * g.rows;
    g.pixels_per_block = ppb;
    g.gx = static_cast<int>(ceil_div64(pixels, ppb));
    return g;
}

// Shared tail of the column reductions: NV values per thread, reduce over the `rows` pixel lanes of the block
// and add to out[v * c + channel].
template <int NV>
__device__ __forceinline__ void block_col_reduce(float (&acc)[NV][8], int g_local, int row, int gpb, int rows, int group,
                                                 int groups, int c, float* out) {
    __shared__ float red[kRedThreads * 8];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        __syncthreads();
        if (row < rows) {
#pragma unroll
            for (int j = 0; j < 8; ++j) red[(row * gpb + g_local) * 8 + j] = acc[v][j];
        }
        __syncthreads();
        // thread t < gpb*8 sums channel (t) over rows
        for (int t = threadIdx.x; t < gpb * 8; t += blockDim.x) {
            const int gl = t >> 3, j = t & 7;
            const int gg = blockIdx.y * gpb + gl;
            if (gg < groups) {
                float s = 0.f;
                for (int r = 0; r < rows; ++r) s += red[(r * gpb + gl) * 8 + j];
                atomicAdd(out + v * c + gg * 8 + j, s);
            }
        }
    }
    (void)group;
}

__global__ void __launch_bounds__(kRedThreads) bn_stats_kernel(const __nv_bfloat16* __restrict__ x, int c_total, int c_off,
                                                                int c, int64_t pixels, int64_t ppb, int gpb, int rows,
                                                                float* __restrict__ stats) {
    pdl_prologue();
    const int groups = c >> 3;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    float acc[2][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
    if (row < rows && group < groups) {
        // strided sweep: all CTAs move through the tensor together (measured faster than one contiguous chunk per CTA)
        const int64_t p0 = static_cast<int64_t>(blockIdx.x) * rows;
        const int64_t p1 = pixels;
        const int64_t pstep = static_cast<int64_t>(gridDim.x) * rows;
        (void)ppb;
        const __nv_bfloat16* base = x + c_off + group * 8;
#pragma unroll 4
        for (int64_t p = p0 + row; p < p1; p += pstep) {
            float f[8];
            unpack8(ldg128(base + p * c_total), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                acc[0][j] += f[j];
                acc[1][j] = fmaf(f[j], f[j], acc[1][j]);
            }
        }
    }
    block_col_reduce<2>(acc, g_local, row, gpb, rows, group, groups, c, stats);
}

// optional second output: plain per-channel sum only
__global__ void __launch_bounds__(kRedThreads) channel_sum_kernel(const __nv_bfloat16* __restrict__ x, int c_total,
                                                                   int c_off, int c, int64_t pixels, int64_t ppb, int gpb,
                                                                   int rows, float* __restrict__ out) {
    pdl_prologue();
    const int groups = c >> 3;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    float acc[1][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[0][j] = 0.f;
    if (row < rows && group < groups) {
        // strided sweep: all CTAs move through the tensor together (measured faster than one contiguous chunk per CTA)
        const int64_t p0 = static_cast<int64_t>(blockIdx.x) * rows;
        const int64_t p1 = pixels;
        const int64_t pstep = static_cast<int64_t>(gridDim.x) * rows;
        (void)ppb;
        const __nv_bfloat16* base = x + c_off + group * 8;
        for (int64_t p = p0 + row; p < p1; p += pstep) {
            float f[8];
            unpack8(ldg128(base + p * c_total), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[0][j] += f[j];
        }
    }
    block_col_reduce<1>(acc, g_local, row, gpb, rows, group, groups, c, out);
}

__global__ void bn_finalize_kernel(const float* __restrict__ stats, int c, float inv_count, float unbias,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                   float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                                   float* save_mean, float* save_invstd) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    const float mean = stats[i] * inv_count;
    float var = fmaf(-mean, mean, stats[c + i] * inv_count);
    var = fmaxf(var, 0.f);
    const float invstd = rsqrtf(var + eps);
    const float g = gamma ? gamma[i] : 1.f, b = beta ? beta[i] : 0.f;
    const float sc = g * invstd;
    scale[i] = sc;
    shift[i] = fmaf(-mean, sc, b);
    if (save_mean) save_mean[i] = mean;
    if (save_invstd) save_invstd[i] = invstd;
    if (running_mean) running_mean[i] = fmaf(momentum, mean - running_mean[i], running_mean[i]);
    if (running_var) running_var[i] = fmaf(momentum, var * unbias - running_var[i], running_var[i]);
}

// eval mode: scale/shift from the running statistics
__global__ void bn_eval_affine_kernel(int c, const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ running_mean, const float* __restrict__ running_var,
                                      float eps, float* scale, float* shift, float* save_mean, float* save_invstd) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    const float invstd = rsqrtf(running_var[i] + eps);
    const float sc = (gamma ? gamma[i] : 1.f) * invstd;
    scale[i] = sc;
    shift[i] = fmaf(-running_mean[i], sc, beta ? beta[i] : 0.f);
    if (save_mean) save_mean[i] = running_mean[i];
    if (save_invstd) save_invstd[i] = invstd;
}

// Training-mode inputs of bn_apply_kernel when it also does bn_finalize's job (stats == nullptr: plain apply)
struct BnTrain {
    const float* stats;      // [2c] per-channel sum and sum of squares from the conv epilogue
    const float* gamma;
    const float* beta;
    float inv_count, unbias, eps, momentum;
    float* running_mean;
    float* running_var;
    float* vec;              // [4c] out: scale, shift, mean, invstd
    unsigned char* bits;     // optional [pixels][c/8]: bit j of a byte = (y[8g + j] > 0), the ReLU mask of the output in
                             // 1/16 of its bytes — what the residual gradient join reads instead of the output itself
};

// Elementwise kernels over a [pixels][c] slice use the column geometry too: a thread owns one 8-channel group for
// its whole pixel range, so the per-channel parameters live in registers (reloading them per element made these
// kernels L1-bound at ~65 % of HBM bandwidth).
// kResAffine: the residual carries its own BN affine (downsample branch); kBits: also write the sign mask.  Separate
// instantiations keep the common ones at 64 registers or fewer (4 resident CTAs per SM): at 73 the 134 MB passes ran
// 113 us instead of 85.
// K: pixels per thread per trip, their loads all issued before the first is consumed.  A streaming pass needs
// bandwidth x latency bytes in flight (6.5 TB/s x ~1 us = 44 KB per SM); one 16-byte load per thread at 1024 resident
// threads is 16 KB, which is what held these passes at ~3 TB/s.  The per-channel parameters live in SHARED memory
// (one LDS.128 pair per vector per trip, far below the LDS rate): in registers they left no room for K loads at four
// resident CTAs per SM.
// (volatile: the compiler would otherwise hoist the rows back into registers for the whole sweep and spill the loads)
__device__ __forceinline__ void lds8(const float* p, float (&f)[8]) {
    const unsigned a = static_cast<unsigned>(__cvta_generic_to_shared(p));
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[0]), "=f"(f[1]), "=f"(f[2]), "=f"(f[3]) : "r"(a));
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[4]), "=f"(f[5]), "=f"(f[6]), "=f"(f[7]) : "r"(a + 16));
}

// kRes: 0 = no residual, 1 = identity residual, 2 = residual with its own BN affine (downsample branch); compile-time so
// the K staged residual loads cost registers only where a residual exists
template <int kRes, bool kBits, int K>
__global__ void __launch_bounds__(kRedThreads, 4) bn_apply_kernel(
    const __nv_bfloat16* __restrict__ x, int x_ct, int x_off, __nv_bfloat16* __restrict__ y, int y_ct, int y_off, int groups,
    int64_t pixels, int64_t ppb, int gpb, int rows, const float* __restrict__ scale, const float* __restrict__ shift,
    int relu, const __nv_bfloat16* __restrict__ res, int r_ct, int r_off, const float* __restrict__ res_scale,
    const float* __restrict__ res_shift, const BnTrain tr) {
    pdl_prologue();
    extern __shared__ __align__(16) float bn_sp[];      // [2 or 4][gpb * 8]: scale, shift (, residual scale, shift)
    const int cw = gpb * 8;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    {
        // the block's parameter rows, one CHANNEL per thread (coalesced loads, every warp takes part: with one thread
        // per 8-channel group the 32-128 threads that did this ran ~50 dependent strided loads each while the rest of
        // the block waited — on the 1024 / 2048-channel tensors that prologue cost more than the 4-pixel sweep behind it)
        const int c = groups * 8;
        const int ch0 = blockIdx.y * cw;
        const bool publish = blockIdx.x == 0;
        for (int lc = threadIdx.x; lc < cw; lc += blockDim.x) {
            const int ch = ch0 + lc;
            if (ch >= c) break;
            if (tr.stats) {
                // training mode without a separate finalize launch: scale / shift from the conv epilogue's sums
                // (bn_finalize_kernel's arithmetic); block row 0 also publishes them for the backward and moves the
                // running statistics
                const float mean = __ldg(tr.stats + ch) * tr.inv_count;
                const float var = fmaxf(fmaf(-mean, mean, __ldg(tr.stats + c + ch) * tr.inv_count), 0.f);
                const float invstd = rsqrtf(var + tr.eps);
                const float scj = __ldg(tr.gamma + ch) * invstd, shj = fmaf(-mean, scj, __ldg(tr.beta + ch));
                bn_sp[lc] = scj;
                bn_sp[cw + lc] = shj;
                if (publish) {
                    tr.vec[ch] = scj;
                    tr.vec[c + ch] = shj;
                    tr.vec[2 * c + ch] = mean;
                    tr.vec[3 * c + ch] = invstd;
                    if (tr.running_mean) tr.running_mean[ch] = fmaf(tr.momentum, mean - tr.running_mean[ch], tr.running_mean[ch]);
                    if (tr.running_var)
                        tr.running_var[ch] = fmaf(tr.momentum, var * tr.unbias - tr.running_var[ch], tr.running_var[ch]);
                }
            } else {
                bn_sp[lc] = __ldg(scale + ch);
                bn_sp[cw + lc] = __ldg(shift + ch);
            }
            if (kRes == 2) {
                bn_sp[2 * cw + lc] = __ldg(res_scale + ch);
                bn_sp[3 * cw + lc] = __ldg(res_shift + ch);
            }
        }
    }
    __syncthreads();
    if (row >= rows || group >= groups) return;
    const float* psc = bn_sp + g_local * 8;
    const float* psh = psc + cw;
    const int64_t p0 = static_cast<int64_t>(blockIdx.x) * rows;
    const int64_t p1 = pixels;
    const int64_t pstep = static_cast<int64_t>(gridDim.x) * rows;
    (void)ppb;
    const __nv_bfloat16* xb = x + x_off + group * 8;
    const __nv_bfloat16* rb = kRes ? res + r_off + group * 8 : nullptr;
    __nv_bfloat16* yb = y + y_off + group * 8;
    // a char store may alias anything: without the restrict copy the next iteration's loads could not move above it
    unsigned char* __restrict__ bits = kBits ? tr.bits : nullptr;
    // running pointers and a 32-bit trip count instead of 64-bit index products per access (registers)
    const int64_t first = p0 + row;
    const int cnt = first < p1 ? static_cast<int>((p1 - first + pstep - 1) / pstep) : 0;
    const int xs = static_cast<int>(pstep * x_ct), rs = static_cast<int>(pstep * r_ct), ys = static_cast<int>(pstep * y_ct),
              bs = static_cast<int>(pstep * groups);        // launcher-checked to fit 32 bits
    const __nv_bfloat16* xp = xb + first * x_ct;
    const __nv_bfloat16* rp = kRes ? rb + first * r_ct : nullptr;
    __nv_bfloat16* yp = yb + first * y_ct;
    unsigned char* bp = kBits ? bits + first * groups + group : nullptr;
#pragma unroll 1
    for (int i = 0; i < cnt; i += K, xp += K * xs, yp += K * ys) {
        if (kRes) rp += (i ? K * rs : 0);
        if (kBits) bp += (i ? K * bs : 0);
        uint4 xv[K], rv[kRes ? K : 1];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (i + k < cnt) {
                xv[k] = ldcs128(xp + k * xs);
                if (kRes) rv[k] = ldcs128(rp + k * rs);
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (i + k >= cnt) break;
            float f[8], sc[8], sh[8];
            lds8(psc, sc);
            lds8(psh, sh);
            unpack8(xv[k], f);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaf(f[j], sc[j], sh[j]);
            if (kRes) {
                // residual branch: the identity, or the downsample conv's raw output with its own BN affine
                float r[8];
                unpack8(rv[k], r);
                if (kRes == 2) {
                    float rsc[8], rsh[8];
                    lds8(psc + 2 * cw, rsc);
                    lds8(psc + 3 * cw, rsh);
#pragma unroll
                    for (int j = 0; j < 8; ++j) r[j] = fmaf(r[j], rsc[j], rsh[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] += r[j];
            }
            if (relu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
            }
            st128(yp + k * ys, pack8(f));
            if (kBits) {
                unsigned int m = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) m |= (f[j] > 0.f ? 1u : 0u) << j;
                bp[k * bs] = static_cast<unsigned char>(m);
            }
        }
    }
}

// red[0..c) = sum dz', red[c..2c) = sum dz' * xhat   with dz' = dy * (y > 0 if relu)
template <int K, bool kMask>
__global__ void __launch_bounds__(kRedThreads, 4) bn_bwd_reduce_kernel(
    const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off, const __nv_bfloat16* __restrict__ x, int x_ct, int x_off,
    int c, int64_t pixels, int64_t ppb, int gpb, int rows, const float* __restrict__ scale, const float* __restrict__ shift,
    const float* __restrict__ mean, const float* __restrict__ invstd, int relu, float* __restrict__ red,
    const __nv_bfloat16* __restrict__ mask, int m_ct, int m_off) {
    pdl_prologue();
    extern __shared__ __align__(16) float bn_sp[];      // [2][gpb * 8]: scale, shift (the ReLU mask is recomputed from x)
    const int cw = gpb * 8;
    const int groups = c >> 3;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    for (int lc = threadIdx.x; lc < cw; lc += blockDim.x) {
        const int ch = blockIdx.y * cw + lc;
        if (ch >= c) break;
        bn_sp[lc] = __ldg(scale + ch);
        bn_sp[cw + lc] = __ldg(shift + ch);
    }
    __syncthreads();
    float acc[2][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
    if (row < rows && group < groups) {
        const float* psc = bn_sp + g_local * 8;
        // strided sweep: all CTAs move through the tensor together (measured faster than one contiguous chunk per CTA)
        const int64_t p0 = static_cast<int64_t>(blockIdx.x) * rows;
        const int64_t p1 = pixels;
        const int64_t pstep = static_cast<int64_t>(gridDim.x) * rows;
        (void)ppb;
        const __nv_bfloat16* xb = x + x_off + group * 8;
        const __nv_bfloat16* db = dy + dy_off + group * 8;
        const __nv_bfloat16* mb = kMask ? mask + m_off + group * 8 : nullptr;
        const int64_t first = p0 + row;
        const int cnt = first < p1 ? static_cast<int>((p1 - first + pstep - 1) / pstep) : 0;
        const int xs = static_cast<int>(pstep * x_ct), ds = static_cast<int>(pstep * dy_ct), ms = static_cast<int>(pstep * m_ct);
        const __nv_bfloat16* xp = xb + first * x_ct;
        const __nv_bfloat16* dp = db + first * dy_ct;
        const __nv_bfloat16* mp = kMask ? mb + first * m_ct : nullptr;
#pragma unroll 1
        for (int i = 0; i < cnt; i += K, xp += K * xs, dp += K * ds) {
            if (kMask) mp += (i ? K * ms : 0);
            uint4 xv[K], dv[K], mv[kMask ? K : 1];
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (i + k < cnt) {
                    xv[k] = ldg128(xp + k * xs);
                    dv[k] = ldg128(dp + k * ds);
                    // ReLU mask from the stored block output (BN output + residual), not recomputable from x alone
                    if (kMask) mv[k] = ldg128(mp + k * ms);
                }
            }
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (i + k >= cnt) break;
                float f[8], d[8], m[8];
                unpack8(xv[k], f);
                unpack8(dv[k], d);
                if (kMask) {
                    unpack8(mv[k], m);
                } else if (relu) {
                    float sc[8], sh[8];
                    lds8(psc, sc);
                    lds8(psc + cw, sh);
#pragma unroll
                    for (int j = 0; j < 8; ++j) m[j] = fmaf(f[j], sc[j], sh[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float dz = (relu && m[j] <= 0.f) ? 0.f : d[j];
                    acc[0][j] += dz;
                    acc[1][j] = fmaf(dz, f[j], acc[1][j]);      // sum dz * x; centred and scaled once below
                }
            }
        }
        // sum dz * xhat = invstd * (sum dz*x - mean * sum dz), per thread (a few hundred terms: no cancellation issue)
        float mu[8], is[8];
        load8f(mean + group * 8, mu);
        load8f(invstd + group * 8, is);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[1][j] = is[j] * (acc[1][j] - mu[j] * acc[0][j]);
    }
    block_col_reduce<2>(acc, g_local, row, gpb, rows, group, groups, c, red);
}

// dx = scale * (dz' - red0/M - xhat * red1/M)   (training)   or   scale * dz'   (eval: statistics are constants)
// rewritten per channel as dx = A * dz' + B * x + C with A = scale, B = -scale*invstd*red1/M (xhat = (x-mean)*invstd),
// C = -scale*red0/M - B*mean (per-channel rows in shared memory).
template <int K, bool kMask>
__global__ void __launch_bounds__(kRedThreads, 4) bn_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off, const __nv_bfloat16* __restrict__ x, int x_ct, int x_off,
    __nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off, int groups, int64_t pixels, int64_t ppb, int gpb, int rows,
    float inv_count, const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
    const float* __restrict__ invstd, int relu, int training, const float* __restrict__ red, int c,
    const __nv_bfloat16* __restrict__ mask, int m_ct, int m_off) {
    pdl_prologue();
    extern __shared__ __align__(16) float bn_sp[];      // [4][gpb * 8]: scale, shift, B, C
    const int cw = gpb * 8;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    for (int lc = threadIdx.x; lc < cw; lc += blockDim.x) {       // one channel per thread, coalesced (see bn_apply_kernel)
        const int ch = blockIdx.y * cw + lc;
        if (ch >= c) break;
        const float scj = __ldg(scale + ch);
        float cb = 0.f, cc = 0.f;
        if (training) {
            cb = -scj * __ldg(invstd + ch) * __ldg(red + c + ch) * inv_count;
            cc = -scj * __ldg(red + ch) * inv_count - cb * __ldg(mean + ch);
        }
        bn_sp[lc] = scj;
        bn_sp[cw + lc] = __ldg(shift + ch);
        bn_sp[2 * cw + lc] = cb;
        bn_sp[3 * cw + lc] = cc;
    }
    __syncthreads();
    if (row >= rows || group >= groups) return;
    const float* psc = bn_sp + g_local * 8;
    const int64_t p0 = static_cast<int64_t>(blockIdx.x) * rows;
    const int64_t p1 = pixels;
    const int64_t pstep = static_cast<int64_t>(gridDim.x) * rows;
    (void)ppb;
    const __nv_bfloat16* xb = x + x_off + group * 8;
    const __nv_bfloat16* db = dy + dy_off + group * 8;
    const __nv_bfloat16* mb = kMask ? mask + m_off + group * 8 : nullptr;
    __nv_bfloat16* ob = dx + dx_off + group * 8;
    const int64_t first = p0 + row;
    const int cnt = first < p1 ? static_cast<int>((p1 - first + pstep - 1) / pstep) : 0;
    const int xs = static_cast<int>(pstep * x_ct), ds = static_cast<int>(pstep * dy_ct), ms = static_cast<int>(pstep * m_ct),
              os = static_cast<int>(pstep * dx_ct);
    const __nv_bfloat16* xp = xb + first * x_ct;
    const __nv_bfloat16* dp = db + first * dy_ct;
    const __nv_bfloat16* mp = kMask ? mb + first * m_ct : nullptr;
    __nv_bfloat16* op = ob + first * dx_ct;
#pragma unroll 1
    for (int i = 0; i < cnt; i += K, xp += K * xs, dp += K * ds, op += K * os) {
        if (kMask) mp += (i ? K * ms : 0);
        uint4 xv[K], dv[K], mv[kMask ? K : 1];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (i + k < cnt) {
                xv[k] = ldcs128(xp + k * xs);
                dv[k] = ldcs128(dp + k * ds);
                if (kMask) mv[k] = ldcs128(mp + k * ms);
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (i + k >= cnt) break;
            float f[8], d[8], m[8], sc[8];
            unpack8(xv[k], f);
            unpack8(dv[k], d);
            lds8(psc, sc);
            if (kMask) {
                unpack8(mv[k], m);
            } else if (relu) {
                float sh[8];
                lds8(psc + cw, sh);
#pragma unroll
                for (int j = 0; j < 8; ++j) m[j] = fmaf(f[j], sc[j], sh[j]);
            }
            float cb[8], cc[8];
            lds8(psc + 2 * cw, cb);
            lds8(psc + 3 * cw, cc);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float dz = (relu && m[j] <= 0.f) ? 0.f : d[j];
                f[j] = fmaf(sc[j], dz, fmaf(cb[j], f[j], cc[j]));
            }
            st128(op + k * os, pack8(f));
        }
    }
}

// dx += dy * (y > 0): the identity branch of a residual block joins the conv branch's data gradient
__global__ void __launch_bounds__(256) relu_bwd_add_kernel(__nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off,
                                                            const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                            const __nv_bfloat16* __restrict__ y, int y_ct, int y_off,
                                                            int groups, int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int g;
        const int64_t p = divmod_small(i, groups, total, g);
        float a[8], d[8], m[8];
        __nv_bfloat16* dst = dx + p * dx_ct + dx_off + g * 8;
        unpack8(*reinterpret_cast<const uint4*>(dst), a);
        unpack8(ldg128(dy + p * dy_ct + dy_off + g * 8), d);
        unpack8(ldg128(y + p * y_ct + y_off + g * 8), m);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += m[j] > 0.f ? d[j] : 0.f;
        st128(dst, pack8(a));
    }
}

// dx = (dx + dy * [y > 0]) * [xin > 0]; dy, y and xin are each optional.  The residual-block gradient join with the
// ReLU mask of the block INPUT applied on the way out, so the producer block's BatchNorm backward gets a pre-masked
// gradient and never reads its own output again.
__global__ void __launch_bounds__(256) grad_add_masked_kernel(__nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off,
                                                               const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                               const __nv_bfloat16* __restrict__ y, int y_ct, int y_off,
                                                               const __nv_bfloat16* __restrict__ xin, int xin_ct,
                                                               int xin_off, const unsigned char* __restrict__ xin_bits,
                                                               int groups, int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int g;
        const int64_t p = divmod_small(i, groups, total, g);
        float a[8];
        __nv_bfloat16* dst = dx + p * dx_ct + dx_off + g * 8;
        unpack8(*reinterpret_cast<const uint4*>(dst), a);
        if (dy) {
            float d[8];
            unpack8(ldcs128(dy + p * dy_ct + dy_off + g * 8), d);
            if (y) {
                float m[8];
                unpack8(ldcs128(y + p * y_ct + y_off + g * 8), m);
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] += m[j] > 0.f ? d[j] : 0.f;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] += d[j];
            }
        }
        if (xin_bits) {
            const unsigned int m = __ldg(xin_bits + i);          // dense [pixels][groups]: element i
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = (m >> j) & 1u ? a[j] : 0.f;
        } else if (xin) {
            float m[8];
            unpack8(ldcs128(xin + p * xin_ct + xin_off + g * 8), m);
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = m[j] > 0.f ? a[j] : 0.f;
        }
        st128(dst, pack8(a));
    }
}

// The residual-block gradient join FUSED with the first pass of the next BatchNorm backward in line.  In a ResNet
// backward the joined gradient d = (dx + dy) * [xin > 0] of a block input IS the (pre-masked) output gradient of the block
// before it, whose bn3 backward — and its downsample BN's, if it has one — starts by reducing sum d and sum d * xhat over
// all pixels.  Doing that here, while d is in registers, saves that pass one read of d per block and a launch:
//   dx <- d ;  red3[0..c) += sum d,  red3[c..2c) += invstd3 * (sum d*x3 - mean3 * sum d) ;  same for (xd, redd) if kDs
// Column geometry of the reductions (a thread owns 8 channels for its pixel rows); xin's mask comes as bits.
template <bool kDs>
__global__ void __launch_bounds__(kRedThreads, 4) grad_join_reduce_kernel(
    __nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off, const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
    const unsigned char* __restrict__ bits, const __nv_bfloat16* __restrict__ x3, int x3_ct, int x3_off,
    const float* __restrict__ mean3, const float* __restrict__ invstd3, float* __restrict__ red3,
    const __nv_bfloat16* __restrict__ xd, int xd_ct, int xd_off, const float* __restrict__ meand,
    const float* __restrict__ invstdd, float* __restrict__ redd, int c, int64_t pixels, int gpb, int rows) {
    pdl_prologue();
    const int groups = c >> 3;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    float a0[8], a3[8], ad[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a0[j] = a3[j] = ad[j] = 0.f;
    if (row < rows && group < groups) {
        const int64_t pstep = static_cast<int64_t>(gridDim.x) * rows;
        __nv_bfloat16* ob = dx + dx_off + group * 8;
        const __nv_bfloat16* db = dy ? dy + dy_off + group * 8 : nullptr;
        const __nv_bfloat16* x3b = x3 + x3_off + group * 8;
        const __nv_bfloat16* xdb = kDs ? xd + xd_off + group * 8 : nullptr;
#pragma unroll 1
        for (int64_t p = static_cast<int64_t>(blockIdx.x) * rows + row; p < pixels; p += pstep) {
            const uint4 v = *reinterpret_cast<const uint4*>(ob + p * dx_ct);
            const uint4 x3v = ldcs128(x3b + p * x3_ct);
            uint4 dv = make_uint4(0u, 0u, 0u, 0u), xdv = dv;
            if (db) dv = ldcs128(db + p * dy_ct);
            if (kDs) xdv = ldcs128(xdb + p * xd_ct);
            const unsigned int m = __ldg(bits + p * groups + group);
            float a[8], d[8], f[8];
            unpack8(v, a);
            unpack8(dv, d);
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = (m >> j) & 1u ? a[j] + d[j] : 0.f;
            const uint4 packed = pack8(a);
            st128(ob + p * dx_ct, packed);
            // the sums are those of the gradient AS STORED (bf16), which is what the apply pass will read
            unpack8(packed, a);
            unpack8(x3v, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                a0[j] += a[j];
                a3[j] = fmaf(a[j], f[j], a3[j]);
            }
            if (kDs) {
                unpack8(xdv, f);
#pragma unroll
                for (int j = 0; j < 8; ++j) ad[j] = fmaf(a[j], f[j], ad[j]);
            }
        }
    }
    float acc[2][8];
    {
        float mu[8], is[8];
        if (row < rows && group < groups) {
            load8f(mean3 + group * 8, mu);
            load8f(invstd3 + group * 8, is);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc[0][j] = a0[j];
            acc[1][j] = (row < rows && group < groups) ? is[j] * (a3[j] - mu[j] * a0[j]) : 0.f;
        }
    }
    block_col_reduce<2>(acc, g_local, row, gpb, rows, group, groups, c, red3);
    if (kDs) {
        float mu[8], is[8];
        if (row < rows && group < groups) {
            load8f(meand + group * 8, mu);
            load8f(invstdd + group * 8, is);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[1][j] = (row < rows && group < groups) ? is[j] * (ad[j] - mu[j] * a0[j]) : 0.f;
        block_col_reduce<2>(acc, g_local, row, gpb, rows, group, groups, c, redd);
    }
}

// y = relu?(x0 + x1 + ... ) over up to 4 bf16 NHWC slices — the multi-resolution fuse of an HRModule
// (models/hrnet.py:240-249: y = y + fuse(x[j]) ...; relu(y))
struct AddN {
    const __nv_bfloat16* x[4];
    int ct[4], off[4];
    int n;
};
__global__ void __launch_bounds__(256) add_n_relu_kernel(const AddN a, __nv_bfloat16* __restrict__ y, int y_ct, int y_off,
                                                          int groups, int64_t total, int relu) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int g;
        const int64_t p = divmod_small(i, groups, total, g);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (k < a.n) {
                float f[8];
                unpack8(ldcs128(a.x[k] + p * a.ct[k] + a.off[k] + g * 8), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += f[j];
            }
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = fmaxf(acc[j], 0.f);
        }
        st128(y + p * y_ct + y_off + g * 8, pack8(acc));
    }
}

// dx = dy * [y > 0]
__global__ void __launch_bounds__(256) relu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                        const __nv_bfloat16* __restrict__ y, int y_ct, int y_off,
                                                        __nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off, int groups,
                                                        int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int g;
        const int64_t p = divmod_small(i, groups, total, g);
        float d[8], m[8];
        unpack8(ldcs128(dy + p * dy_ct + dy_off + g * 8), d);
        unpack8(ldcs128(y + p * y_ct + y_off + g * 8), m);
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = m[j] > 0.f ? d[j] : 0.f;
        st128(dx + p * dx_ct + dx_off + g * 8, pack8(d));
    }
}

// dgamma = red1, dbeta = red0 (optionally accumulated into existing .grad)
__global__ void bn_param_grad_kernel(const float* __restrict__ red, int c, float* dgamma, float* dbeta, int accumulate) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    if (dbeta) dbeta[i] = accumulate ? dbeta[i] + red[i] : red[i];
    if (dgamma) dgamma[i] = accumulate ? dgamma[i] + red[c + i] : red[c + i];
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_bn_stats(const void* x, int x_c_total, int x_c_offset, int c, int64_t pixels, float* stats,
                     b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_slice("bn_stats", x, x_c_total, x_c_offset, c)) return 1;
    B200SEG_CHECK(stats && pixels > 0, "bn_stats: bad arguments");
    B200SEG_CUDA(cudaMemsetAsync(stats, 0, sizeof(float) * 2 * c, st));
    const ColGeom g = col_geom(c, pixels);
    B200SEG_CUDA(launch_pdl(bn_stats_kernel, dim3(g.gx, g.gy), kRedThreads, 0, st, static_cast<const __nv_bfloat16*>(x), x_c_total, x_c_offset, c,
                                                             pixels, g.pixels_per_block, g.gpb, g.rows, stats));
    return 0;
}

int b200seg_channel_sum(const void* x, int x_c_total, int x_c_offset, int c, int64_t pixels, float* out, int accumulate,
                        b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_slice("channel_sum", x, x_c_total, x_c_offset, c)) return 1;
    B200SEG_CHECK(out && pixels > 0, "channel_sum: bad arguments");
    if (!accumulate) B200SEG_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * c, st));
    const ColGeom g = col_geom(c, pixels);
    B200SEG_CUDA(launch_pdl(channel_sum_kernel, dim3(g.gx, g.gy), kRedThreads, 0, st, static_cast<const __nv_bfloat16*>(x), x_c_total,
                                                                x_c_offset, c, pixels, g.pixels_per_block, g.gpb, g.rows,
                                                                out));
    return 0;
}

int b200seg_bn_finalize(const float* stats, int c, int64_t count, const float* gamma, const float* beta, float eps,
                        float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                        float* save_mean, float* save_invstd, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(stats && scale && shift && c > 0 && count > 0, "bn_finalize: bad arguments");
    const float unbias = count > 1 ? static_cast<float>(static_cast<double>(count) / static_cast<double>(count - 1)) : 1.f;
    B200SEG_CUDA(launch_pdl(bn_finalize_kernel, ceil_div(c, 128), 128, 0, static_cast<cudaStream_t>(stream), 
        stats, c, static_cast<float>(1.0 / static_cast<double>(count)), unbias, gamma, beta, eps, momentum, running_mean,
        running_var, scale, shift, save_mean, save_invstd));
    return 0;
}

int b200seg_bn_eval_affine(int c, const float* gamma, const float* beta, const float* running_mean,
                           const float* running_var, float eps, float* scale, float* shift, float* save_mean,
                           float* save_invstd, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(running_mean && running_var && scale && shift && c > 0, "bn_eval_affine: bad arguments");
    B200SEG_CUDA(launch_pdl(bn_eval_affine_kernel, ceil_div(c, 128), 128, 0, static_cast<cudaStream_t>(stream), 
        c, gamma, beta, running_mean, running_var, eps, scale, shift, save_mean, save_invstd));
    return 0;
}

int b200seg_bn_apply_add(const void* x, int x_c_total, int x_c_offset, void* y, int y_c_total, int y_c_offset, int c,
                         int64_t pixels, const float* scale, const float* shift, int relu, const void* res,
                         int res_c_total, int res_c_offset, const float* res_scale, const float* res_shift,
                         b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("bn_apply(x)", x, x_c_total, x_c_offset, c)) return 1;
    if (check_slice("bn_apply(y)", y, y_c_total, y_c_offset, c)) return 1;
    if (res && check_slice("bn_apply(res)", res, res_c_total, res_c_offset, c)) return 1;
    B200SEG_CHECK(scale && shift && pixels > 0, "bn_apply: bad arguments");
    B200SEG_CHECK((res_scale == nullptr) == (res_shift == nullptr), "bn_apply: residual scale and shift go together");
    const ColGeom g = col_geom(c, pixels, apply_ctas(), apply_gpb());
    static const int kk = bn_env_k("B200SEG_BN_K_APPLY", 4);
#define B200SEG_BN_APPLY_K(RA, BITS, K)     B200SEG_CUDA(launch_pdl(bn_apply_kernel<RA, BITS, K>, dim3(g.gx, g.gy), kRedThreads, sizeof(float) * 8 * g.gpb * ((RA) == 2 ? 4 : 2), static_cast<cudaStream_t>(stream),  \
        static_cast<const __nv_bfloat16*>(x), x_c_total, x_c_offset, static_cast<__nv_bfloat16*>(y), y_c_total, y_c_offset, \
        c / 8, pixels, g.pixels_per_block, g.gpb, g.rows, scale, shift, relu, static_cast<const __nv_bfloat16*>(res), \
        res_c_total, res_c_offset, res_scale, res_shift, BnTrain{}));
    /* loads in flight per thread = K x streams: 4 without a residual, 2 x 2 with one, 1 x 2 with an affine residual (its */ \
    /* two extra parameter rows leave no registers for more at four resident CTAs per SM) */
#define B200SEG_BN_APPLY(RA, BITS)                                                                         \
    {                                                                                                      \
        const int ke = (RA) == 0 ? kk : ((RA) == 1 ? (kk >= 2 ? 2 : 1) : 1);                               \
        if (ke == 4) B200SEG_BN_APPLY_K(RA, BITS, 4) else if (ke == 2) B200SEG_BN_APPLY_K(RA, BITS, 2)     \
        else B200SEG_BN_APPLY_K(RA, BITS, 1)                                                               \
    }
    if (res && res_scale) B200SEG_BN_APPLY(2, false) else if (res) B200SEG_BN_APPLY(1, false) else B200SEG_BN_APPLY(0, false)
#undef B200SEG_BN_APPLY
#undef B200SEG_BN_APPLY_K
    return 0;
}

int b200seg_bn_apply_train(const void* x, int x_c_total, int x_c_offset, void* y, int y_c_total, int y_c_offset, int c,
                           int64_t pixels, const float* stats, const float* gamma, const float* beta, float eps,
                           float momentum, float* running_mean, float* running_var, float* vec, int relu, const void* res,
                           int res_c_total, int res_c_offset, const float* res_scale, const float* res_shift,
                           uint8_t* mask_bits, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("bn_apply_train(x)", x, x_c_total, x_c_offset, c)) return 1;
    if (check_slice("bn_apply_train(y)", y, y_c_total, y_c_offset, c)) return 1;
    if (res && check_slice("bn_apply_train(res)", res, res_c_total, res_c_offset, c)) return 1;
    B200SEG_CHECK(stats && gamma && beta && vec && pixels > 0, "bn_apply_train: bad arguments");
    B200SEG_CHECK((res_scale == nullptr) == (res_shift == nullptr), "bn_apply_train: residual scale and shift go together");
    B200SEG_CHECK(((reinterpret_cast<uintptr_t>(stats) | reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15) == 0,
                  "bn_apply_train: statistics / affine parameters must be 16-byte aligned");
    BnTrain tr;
    tr.stats = stats;
    tr.gamma = gamma;
    tr.beta = beta;
    tr.inv_count = static_cast<float>(1.0 / static_cast<double>(pixels));
    tr.unbias = pixels > 1 ? static_cast<float>(static_cast<double>(pixels) / static_cast<double>(pixels - 1)) : 1.f;
    tr.eps = eps;
    tr.momentum = momentum;
    tr.running_mean = running_mean;
    tr.running_var = running_var;
    tr.vec = vec;
    tr.bits = mask_bits;
    const ColGeom g = col_geom(c, pixels, apply_ctas(), apply_gpb());
    static const int kk = bn_env_k("B200SEG_BN_K_APPLY", 4);
#define B200SEG_BN_APPLY_K(RA, BITS, K)     B200SEG_CUDA(launch_pdl(bn_apply_kernel<RA, BITS, K>, dim3(g.gx, g.gy), kRedThreads, sizeof(float) * 8 * g.gpb * ((RA) == 2 ? 4 : 2), static_cast<cudaStream_t>(stream), \
        static_cast<const __nv_bfloat16*>(x), x_c_total, x_c_offset, static_cast<__nv_bfloat16*>(y), y_c_total, y_c_offset, \
        c / 8, pixels, g.pixels_per_block, g.gpb, g.rows, static_cast<const float*>(nullptr), \
        static_cast<const float*>(nullptr), relu, static_cast<const __nv_bfloat16*>(res), res_c_total, res_c_offset, \
        res_scale, res_shift, tr));
    /* loads in flight per thread = K x streams: 4 without a residual, 2 x 2 with one, 1 x 2 with an affine residual (its */ \
    /* two extra parameter rows leave no registers for more at four resident CTAs per SM) */
#define B200SEG_BN_APPLY(RA, BITS)                                                                         \
    {                                                                                                      \
        const int ke = (RA) == 0 ? kk : ((RA) == 1 ? (kk >= 2 ? 2 : 1) : 1);                               \
        if (ke == 4) B200SEG_BN_APPLY_K(RA, BITS, 4) else if (ke == 2) B200SEG_BN_APPLY_K(RA, BITS, 2)     \
        else B200SEG_BN_APPLY_K(RA, BITS, 1)                                                               \
    }
    const int ra = res ? (res_scale ? 2 : 1) : 0;
    if (ra == 2 && mask_bits) B200SEG_BN_APPLY(2, true) else if (ra == 2) B200SEG_BN_APPLY(2, false)
    else if (ra == 1 && mask_bits) B200SEG_BN_APPLY(1, true) else if (ra == 1) B200SEG_BN_APPLY(1, false)
    else if (mask_bits) B200SEG_BN_APPLY(0, true) else B200SEG_BN_APPLY(0, false)
#undef B200SEG_BN_APPLY
#undef B200SEG_BN_APPLY_K
    return 0;
}

int b200seg_bn_apply(const void* x, int x_c_total, int x_c_offset, void* y, int y_c_total, int y_c_offset, int c,
                     int64_t pixels, const float* scale, const float* shift, int relu, b200seg_stream_t stream) {
    return b200seg_bn_apply_add(x, x_c_total, x_c_offset, y, y_c_total, y_c_offset, c, pixels, scale, shift, relu, nullptr,
                                0, 0, nullptr, nullptr, stream);
}

int b200seg_relu_bwd_add(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                         const void* y, int y_c_total, int y_c_offset, int c, int64_t pixels, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("relu_bwd_add(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    if (check_slice("relu_bwd_add(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("relu_bwd_add(y)", y, y_c_total, y_c_offset, c)) return 1;
    B200SEG_CHECK(pixels > 0, "relu_bwd_add: bad arguments");
    const int64_t total = pixels * (c / 8);
    B200SEG_CUDA(launch_pdl(relu_bwd_add_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        static_cast<__nv_bfloat16*>(dx), dx_c_total, dx_c_offset, static_cast<const __nv_bfloat16*>(dy), dy_c_total,
        dy_c_offset, static_cast<const __nv_bfloat16*>(y), y_c_total, y_c_offset, c / 8, total));
    return 0;
}

int b200seg_add_n_relu(int n_in, const void* const* x, const int* x_c_total, const int* x_c_offset, void* y, int y_c_total,
                       int y_c_offset, int c, int64_t pixels, int relu, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(n_in >= 1 && n_in <= 4 && x && x_c_total && x_c_offset && pixels > 0, "add_n_relu: 1..4 inputs");
    if (check_slice("add_n_relu(y)", y, y_c_total, y_c_offset, c)) return 1;
    AddN a;
    a.n = n_in;
    for (int i = 0; i < 4; ++i) {
        a.x[i] = nullptr;
        a.ct[i] = a.off[i] = 0;
    }
    for (int i = 0; i < n_in; ++i) {
        if (check_slice("add_n_relu(x)", x[i], x_c_total[i], x_c_offset[i], c)) return 1;
        a.x[i] = static_cast<const __nv_bfloat16*>(x[i]);
        a.ct[i] = x_c_total[i];
        a.off[i] = x_c_offset[i];
    }
    const int64_t total = pixels * (c / 8);
    B200SEG_CUDA(launch_pdl(add_n_relu_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), a,
                            static_cast<__nv_bfloat16*>(y), y_c_total, y_c_offset, c / 8, total, relu));
    return 0;
}

int b200seg_relu_bwd(const void* dy, int dy_c_total, int dy_c_offset, const void* y, int y_c_total, int y_c_offset,
                     void* dx, int dx_c_total, int dx_c_offset, int c, int64_t pixels, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("relu_bwd(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("relu_bwd(y)", y, y_c_total, y_c_offset, c)) return 1;
    if (check_slice("relu_bwd(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    B200SEG_CHECK(pixels > 0, "relu_bwd: bad arguments");
    const int64_t total = pixels * (c / 8);
    B200SEG_CUDA(launch_pdl(relu_bwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
                            static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset,
                            static_cast<const __nv_bfloat16*>(y), y_c_total, y_c_offset, static_cast<__nv_bfloat16*>(dx),
                            dx_c_total, dx_c_offset, c / 8, total));
    return 0;
}

int b200seg_grad_add_masked(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                            const void* y, int y_c_total, int y_c_offset, const void* xin, int xin_c_total,
                            int xin_c_offset, int c, int64_t pixels, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("grad_add_masked(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    if (dy && check_slice("grad_add_masked(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (y && check_slice("grad_add_masked(y)", y, y_c_total, y_c_offset, c)) return 1;
    if (xin && check_slice("grad_add_masked(xin)", xin, xin_c_total, xin_c_offset, c)) return 1;
    B200SEG_CHECK(pixels > 0 && (dy || xin) && (dy || !y), "grad_add_masked: bad arguments");
    const int64_t total = pixels * (c / 8);
    B200SEG_CUDA(launch_pdl(grad_add_masked_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        static_cast<__nv_bfloat16*>(dx), dx_c_total, dx_c_offset, static_cast<const __nv_bfloat16*>(dy), dy_c_total,
        dy_c_offset, static_cast<const __nv_bfloat16*>(y), y_c_total, y_c_offset, static_cast<const __nv_bfloat16*>(xin),
        xin_c_total, xin_c_offset, static_cast<const unsigned char*>(nullptr), c / 8, total));
    return 0;
}

int b200seg_grad_add_masked_bits(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                                 const void* y, int y_c_total, int y_c_offset, const uint8_t* xin_bits, int c, int64_t pixels,
                                 b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("grad_add_masked_bits(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    if (dy && check_slice("grad_add_masked_bits(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (y && check_slice("grad_add_masked_bits(y)", y, y_c_total, y_c_offset, c)) return 1;
    B200SEG_CHECK(pixels > 0 && xin_bits && (dy || !y), "grad_add_masked_bits: bad arguments");
    const int64_t total = pixels * (c / 8);
    B200SEG_CUDA(launch_pdl(grad_add_masked_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
        static_cast<__nv_bfloat16*>(dx), dx_c_total, dx_c_offset, static_cast<const __nv_bfloat16*>(dy), dy_c_total,
        dy_c_offset, static_cast<const __nv_bfloat16*>(y), y_c_total, y_c_offset, static_cast<const __nv_bfloat16*>(nullptr),
        0, 0, xin_bits, c / 8, total));
    return 0;
}

int b200seg_grad_join_reduce(void* dx, int dx_c_total, int dx_c_offset, const void* dy, int dy_c_total, int dy_c_offset,
                             const uint8_t* xin_bits, int c, int64_t pixels, const void* x3, int x3_c_total, int x3_c_offset,
                             const float* mean3, const float* invstd3, float* red3, const void* xd, int xd_c_total,
                             int xd_c_offset, const float* meand, const float* invstdd, float* redd,
                             b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("grad_join_reduce(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    if (dy && check_slice("grad_join_reduce(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("grad_join_reduce(x3)", x3, x3_c_total, x3_c_offset, c)) return 1;
    if (xd && check_slice("grad_join_reduce(xd)", xd, xd_c_total, xd_c_offset, c)) return 1;
    B200SEG_CHECK(pixels > 0 && xin_bits && mean3 && invstd3 && red3, "grad_join_reduce: bad arguments");
    B200SEG_CHECK(!xd || (meand && invstdd && redd), "grad_join_reduce: downsample statistics missing");
    const ColGeom g = col_geom(c, pixels, 4, 32);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
#define B200SEG_JOIN(DS)                                                                                                  \
    B200SEG_CUDA(launch_pdl(grad_join_reduce_kernel<DS>, dim3(g.gx, g.gy), kRedThreads, 0, st,                            \
        static_cast<__nv_bfloat16*>(dx), dx_c_total, dx_c_offset, static_cast<const __nv_bfloat16*>(dy), dy_c_total,      \
        dy_c_offset, xin_bits, static_cast<const __nv_bfloat16*>(x3), x3_c_total, x3_c_offset, mean3, invstd3, red3,      \
        static_cast<const __nv_bfloat16*>(xd), xd_c_total, xd_c_offset, meand, invstdd, redd, c, pixels, g.gpb, g.rows));
    if (xd) B200SEG_JOIN(true) else B200SEG_JOIN(false)
#undef B200SEG_JOIN
    return 0;
}

int b200seg_bn_bwd_reduce(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total, int x_c_offset,
                          int c, int64_t pixels, const float* scale, const float* shift, const float* save_mean,
                          const float* save_invstd, int relu, float* red, b200seg_stream_t stream) {
    return b200seg_bn_bwd_reduce_masked(dy, dy_c_total, dy_c_offset, x, x_c_total, x_c_offset, c, pixels, scale, shift,
                                        save_mean, save_invstd, relu, red, nullptr, 0, 0, stream);
}

static int bn_bwd_reduce_impl(bool zero_red, const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                                 int x_c_offset, int c, int64_t pixels, const float* scale, const float* shift,
                                 const float* save_mean, const float* save_invstd, int relu, float* red, const void* mask,
                                 int mask_c_total, int mask_c_offset, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_slice("bn_bwd_reduce(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("bn_bwd_reduce(x)", x, x_c_total, x_c_offset, c)) return 1;
    if (mask && check_slice("bn_bwd_reduce(mask)", mask, mask_c_total, mask_c_offset, c)) return 1;
    B200SEG_CHECK(scale && shift && save_mean && save_invstd && red && pixels > 0, "bn_bwd_reduce: bad arguments");
    if (zero_red) B200SEG_CUDA(cudaMemsetAsync(red, 0, sizeof(float) * 2 * c, st));
    // one wave of 4 resident CTAs per SM (the kernel is held to 64 registers for that: at 99 registers only two fit and
    // the pass ran 1.59 ms per step instead of 1.28); more CTAs than that only add atomics on the same 2c addresses
    static const int ctas = [] {
        const char* e = getenv("B200SEG_BNRED_CTAS");
        const int v = e ? atoi(e) : 4;
        return v > 0 ? v : 4;
    }();
    static const int red_gpb = [] {
        const char* e = getenv("B200SEG_BNRED_GPB");
        const int v = e ? atoi(e) : 32;
        return v >= 8 && v <= 256 ? v : 32;
    }();
    const ColGeom g = col_geom(c, pixels, ctas, red_gpb);
    static const int kk = bn_env_k("B200SEG_BN_K_RED", 2);
#define B200SEG_BN_RED_(K, M)                                                                                                  \
    B200SEG_CUDA(launch_pdl(bn_bwd_reduce_kernel<K, M>, dim3(g.gx, g.gy), kRedThreads, sizeof(float) * 16 * g.gpb, st,                                  \
        static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset, static_cast<const __nv_bfloat16*>(x), x_c_total,     \
        x_c_offset, c, pixels, g.pixels_per_block, g.gpb, g.rows, scale, shift, save_mean, save_invstd, relu, red,           \
        static_cast<const __nv_bfloat16*>(mask), mask_c_total, mask_c_offset));
#define B200SEG_BN_RED(K) { if (mask) B200SEG_BN_RED_(K, true) else B200SEG_BN_RED_(K, false) }
    if (kk == 4) B200SEG_BN_RED(4) else if (kk == 2) B200SEG_BN_RED(2) else B200SEG_BN_RED(1)
#undef B200SEG_BN_RED
#undef B200SEG_BN_RED_
    return 0;
}

int b200seg_bn_bwd_reduce_masked(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                                 int x_c_offset, int c, int64_t pixels, const float* scale, const float* shift,
                                 const float* save_mean, const float* save_invstd, int relu, float* red, const void* mask,
                                 int mask_c_total, int mask_c_offset, b200seg_stream_t stream) {
    return bn_bwd_reduce_impl(true, dy, dy_c_total, dy_c_offset, x, x_c_total, x_c_offset, c, pixels, scale, shift, save_mean,
                              save_invstd, relu, red, mask, mask_c_total, mask_c_offset, stream);
}

int b200seg_bn_bwd_reduce_acc(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total, int x_c_offset,
                              int c, int64_t pixels, const float* scale, const float* shift, const float* save_mean,
                              const float* save_invstd, int relu, float* red, const void* mask, int mask_c_total,
                              int mask_c_offset, b200seg_stream_t stream) {
    return bn_bwd_reduce_impl(false, dy, dy_c_total, dy_c_offset, x, x_c_total, x_c_offset, c, pixels, scale, shift, save_mean,
                              save_invstd, relu, red, mask, mask_c_total, mask_c_offset, stream);
}

int b200seg_bn_bwd_apply(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total, int x_c_offset,
                         void* dx, int dx_c_total, int dx_c_offset, int c, int64_t pixels, const float* scale,
                         const float* shift, const float* save_mean, const float* save_invstd, int relu, int training,
                         const float* red, float* dgamma, float* dbeta, int accumulate_param_grads,
                         b200seg_stream_t stream) {
    return b200seg_bn_bwd_apply_masked(dy, dy_c_total, dy_c_offset, x, x_c_total, x_c_offset, dx, dx_c_total, dx_c_offset, c,
                                       pixels, scale, shift, save_mean, save_invstd, relu, training, red, dgamma, dbeta,
                                       accumulate_param_grads, nullptr, 0, 0, stream);
}

int b200seg_bn_bwd_apply_masked(const void* dy, int dy_c_total, int dy_c_offset, const void* x, int x_c_total,
                                int x_c_offset, void* dx, int dx_c_total, int dx_c_offset, int c, int64_t pixels,
                                const float* scale, const float* shift, const float* save_mean, const float* save_invstd,
                                int relu, int training, const float* red, float* dgamma, float* dbeta,
                                int accumulate_param_grads, const void* mask, int mask_c_total, int mask_c_offset,
                                b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_slice("bn_bwd_apply(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("bn_bwd_apply(x)", x, x_c_total, x_c_offset, c)) return 1;
    if (check_slice("bn_bwd_apply(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    if (mask && check_slice("bn_bwd_apply(mask)", mask, mask_c_total, mask_c_offset, c)) return 1;
    B200SEG_CHECK(scale && shift && save_mean && save_invstd && red && pixels > 0, "bn_bwd_apply: bad arguments");
    const ColGeom g = col_geom(c, pixels, apply_ctas(), apply_gpb());
    static const int kk = bn_env_k("B200SEG_BN_K_BWD", 1);
#define B200SEG_BN_BWD_(K, M)                                                                                                  \
    B200SEG_CUDA(launch_pdl(bn_bwd_apply_kernel<K, M>, dim3(g.gx, g.gy), kRedThreads, sizeof(float) * 32 * g.gpb, st,                                  \
        static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset, static_cast<const __nv_bfloat16*>(x), x_c_total,     \
        x_c_offset, static_cast<__nv_bfloat16*>(dx), dx_c_total, dx_c_offset, c / 8, pixels, g.pixels_per_block, g.gpb,      \
        g.rows, static_cast<float>(1.0 / static_cast<double>(pixels)), scale, shift, save_mean, save_invstd, relu, training, \
        red, c, static_cast<const __nv_bfloat16*>(mask), mask_c_total, mask_c_offset));
#define B200SEG_BN_BWD(K) { if (mask) B200SEG_BN_BWD_(K, true) else B200SEG_BN_BWD_(K, false) }
    if (kk == 4) B200SEG_BN_BWD(4) else if (kk == 2) B200SEG_BN_BWD(2) else B200SEG_BN_BWD(1)
#undef B200SEG_BN_BWD
#undef B200SEG_BN_BWD_
    if (dgamma || dbeta) {
        B200SEG_CUDA(launch_pdl(bn_param_grad_kernel, ceil_div(c, 128), 128, 0, st, red, c, dgamma, dbeta, accumulate_param_grads));
    }
    return 0;
}

}  // extern "C"
