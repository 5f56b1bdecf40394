// Stride-1 "same" convolution as an implicit GEMM on the 5th-gen tensor cores (tcgen05), NHWC bf16.
//
//   D[pixel][k] = sum over sources, taps (r,s), input channels c of
//                 X_src[n, y + (r - R/2)*dil, x + (s - S/2)*dil, c] * W_src[tap][k][c]
//
// One CTA tile = 128 output pixels (a bw x bh x bi patch of the NHW volume, bw*bh*bi = 128) x bn output
// channels.  For every (tap, 64-channel block) the producer issues ONE 4-D TMA box load of the activation patch
// shifted by the tap offset — dilation and padding live in the TMA coordinates; out-of-bounds rows/columns are
// zero-filled by the TMA unit — and one 3-D box load of the weight slab.  Both land in shared memory in the
// 128-byte-swizzled K-major layout tcgen05.mma consumes.  Taps whose shifted patch lies entirely in the padding
// are skipped (their contribution is exactly zero).  Accumulators live in TMEM (2 x 256 columns, double
// buffered) so the epilogue of tile i overlaps the main loop of tile i+1 in this persistent kernel.
//
// Warp roles (256 threads): warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one lane), warp 2 = TMEM
// allocator, warps 4-7 = epilogue (TMEM -> registers -> scale/bias/ReLU -> global, 16-byte stores).
//
// The same kernel is the data-gradient (x = dy, weights packed transposed with mirrored taps) and sums
// several sources in TMEM (the ASPP branches' gradients w.r.t. their shared input).
#include "common.cuh"
#include "ptx.cuh"

#include <string.h>

#include <array>
#include <map>
#include <mutex>
#include <vector>

namespace b200seg {

namespace {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kMaxBN = 256;
constexpr int kStages = 4;
constexpr int kABytes = kBM * kBK * 2;
constexpr int kBBytesMax = kMaxBN * kBK * 2;
constexpr int kStageBytes = kABytes + kBBytesMax;
constexpr int kThreads = 384;          // warps 0-3: TMA / MMA / TMEM alloc / idle, warps 4-11: epilogue
constexpr int kEpiThreads = 256;
constexpr int kBarrierBytes = 256;
constexpr int kOutSlab = 64;                       // output channels per TMA store box (128 bytes of bf16)
constexpr int kOutBufBytes = kBM * kOutSlab * 2;   // 16 KB: one 128-pixel x 64-channel slab, 128-byte swizzled
constexpr int kOutBufs = 2;
constexpr int kSmemBytes = kStages * kStageBytes + kOutBufs * kOutBufBytes + kBarrierBytes + 1024;   // +1024: alignment
constexpr int kTmemCols = 512;

struct IgemmSeg {
    int r, s, dil, cblks;
};

struct __align__(64) IgemmParams {
    CUtensorMap a_map[B200SEG_MAX_CONV_SRC];
    CUtensorMap b_map[B200SEG_MAX_CONV_SRC];
    CUtensorMap y_map;       // bf16 output slice (channels, W, H, N), box = 64 channels x pixel patch
    IgemmSeg seg[B200SEG_MAX_CONV_SRC];
    int nseg;
    int N, H, W;
    int bw, bh, bi;          // patch dims (powers of two)
    int lbw, lbh;            // log2
    int tiles_x, tiles_y, tiles_i, tiles_n;
    int bn;
    int num_tiles;
    void* y;
    int y_c_total, y_c_offset, k_store, out_fp32;
    const float* scale;
    const float* bias;
    int relu;
    float* stats;
    long long* dbg;
    int dbg_slots;
};

long long* g_dbg = nullptr;
int g_dbg_slots = 0;

__device__ __forceinline__ long long gtime() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

struct TileCoord {
    int n_tile, x0, y0, n0;
};

__device__ __forceinline__ TileCoord decode_tile(const IgemmParams& p, int tile) {
    TileCoord t;
    t.n_tile = tile % p.tiles_n;
    int m = tile / p.tiles_n;
    t.x0 = (m % p.tiles_x) * p.bw;
    m /= p.tiles_x;
    t.y0 = (m % p.tiles_y) * p.bh;
    t.n0 = (m / p.tiles_y) * p.bi;
    return t;
}

__device__ __forceinline__ bool tap_in_bounds(const IgemmParams& p, const TileCoord& t, int dy, int dx) {
    return (t.y0 + dy < p.H) && (t.y0 + p.bh - 1 + dy >= 0) && (t.x0 + dx < p.W) && (t.x0 + p.bw - 1 + dx >= 0);
}

__global__ void __launch_bounds__(kThreads, 1) igemm_kernel(const __grid_constant__ IgemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* out_buf = smem + kStages * kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(out_buf + kOutBufs * kOutBufBytes);
    uint64_t* full = bars;                    // [kStages] TMA -> MMA
    uint64_t* empty = bars + kStages;         // [kStages] MMA -> TMA
    uint64_t* tfull = bars + 2 * kStages;     // [2] MMA -> epilogue
    uint64_t* tempty = bars + 2 * kStages + 2;   // [2] epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < p.nseg; ++i) {
            tma_prefetch_desc(&p.a_map[i]);
            tma_prefetch_desc(&p.b_map[i]);
        }
        if (!p.out_fp32) tma_prefetch_desc(&p.y_map);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], p.out_fp32 ? 4 : 8);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ------------------------------------------------------------ TMA producer
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx_bytes = kABytes + p.bn * kBK * 2;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const TileCoord t = decode_tile(p, tile);
                for (int sg = 0; sg < p.nseg; ++sg) {
                    const IgemmSeg seg = p.seg[sg];
                    for (int r = 0; r < seg.r; ++r) {
                        const int dy = (r - seg.r / 2) * seg.dil;
                        for (int s = 0; s < seg.s; ++s) {
                            const int dx = (s - seg.s / 2) * seg.dil;
                            if (!tap_in_bounds(p, t, dy, dx)) continue;
                            const int tap = r * seg.s + s;
                            for (int cb = 0; cb < seg.cblks; ++cb) {
                                mbar_wait(&empty[stage], phase ^ 1);
                                uint8_t* sa = smem + stage * kStageBytes;
                                mbar_arrive_expect_tx(&full[stage], tx_bytes);
                                tma_load_4d(sa, &p.a_map[sg], &full[stage], cb * kBK, t.x0 + dx, t.y0 + dy, t.n0);
                                tma_load_3d(sa + kABytes, &p.b_map[sg], &full[stage], cb * kBK, t.n_tile * p.bn, tap);
                                if (++stage == kStages) {
                                    stage = 0;
                                    phase ^= 1;
                                }
                            }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ------------------------------------------------------------ MMA issuer
            const uint32_t idesc = umma_idesc_bf16(kBM, p.bn, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            int lt = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++lt) {
                const TileCoord t = decode_tile(p, tile);
                int kiters = 0;
                for (int sg = 0; sg < p.nseg; ++sg) {
                    const IgemmSeg seg = p.seg[sg];
                    for (int r = 0; r < seg.r; ++r)
                        for (int s = 0; s < seg.s; ++s)
                            if (tap_in_bounds(p, t, (r - seg.r / 2) * seg.dil, (s - seg.s / 2) * seg.dil))
                                kiters += seg.cblks;
                }
                const int acc = lt & 1;
                const uint32_t acc_phase = (lt >> 1) & 1;
                mbar_wait(&tempty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * kMaxBN;
                for (int it = 0; it < kiters; ++it) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                    const uint64_t a_desc = umma_smem_desc(sa, 16, 1024);
                    const uint64_t b_desc = umma_smem_desc(sa + kABytes, 16, 1024);
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k) {
                        // +32 bytes (16 bf16) along K inside the 128-byte swizzle row: +2 in the address field
                        tc_mma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (it | k) != 0);
                    }
                    tc_commit(&empty[stage]);
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                tc_commit(&tfull[acc]);
            }
        }
    } else if (warp >= 4) {
        // ---------------------------------------------------------------- epilogue
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int wl = row & (p.bw - 1);
        const int hl = (row >> p.lbw) & (p.bh - 1);
        const int il = row >> (p.lbw + p.lbh);
        if (!p.out_fp32) {
            // bf16 output: TMEM -> registers -> (scale/bias/ReLU) -> 128-byte-swizzled shared-memory slab (64 channels
            // x 128 pixels) -> one TMA store per slab (the TMA unit clips pixels outside the image and channels
            // beyond the slice).  Eight warps: warp w reads TMEM lanes 32*(w%4).., warps 4-7 take the low 32 columns
            // of the slab, warps 8-11 the high 32 (two warps per scheduler hide each other's latencies).
            // BN statistics are column sums of the staged bf16 slab: thread (row group rg, channel pair cp) keeps
            // running sums in registers over all tiles of this CTA and flushes once.
            const int et = threadIdx.x - 128;            // 0..255
            const int half = et >> 7;
            const int cp = et & 31, rg = et >> 5;        // rg 0..7: 16 rows each
            const bool affine = p.scale != nullptr || p.bias != nullptr || p.relu != 0;
            float st1[4][2], st2[4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i) st1[i][0] = st1[i][1] = st2[i][0] = st2[i][1] = 0.f;
            int cur_n_tile = -1;
            int slab_seq = 0;
            // Flush: the row groups' partial sums meet in shared memory (the staging buffers are idle then), so each
            // channel costs one global atomic per CTA, issued by consecutive lanes on consecutive channels.
            auto flush_stats = [&](int n_tile) {
                float* red = reinterpret_cast<float*>(out_buf);        // [rg 8][slab 4][sum/sumsq 2][64]
                if (et == 0) tma_store_wait_read<0>();
                named_bar_sync(1, kEpiThreads);
#pragma unroll
                for (int sl = 0; sl < 4; ++sl) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        red[((rg * 4 + sl) * 2 + 0) * 64 + cp * 2 + e] = st1[sl][e];
                        red[((rg * 4 + sl) * 2 + 1) * 64 + cp * 2 + e] = st2[sl][e];
                        st1[sl][e] = st2[sl][e] = 0.f;
                    }
                }
                named_bar_sync(2, kEpiThreads);
                for (int i = et; i < 4 * 2 * 64; i += kEpiThreads) {
                    const int c = i & 63, which = (i >> 6) & 1, sl = i >> 7;
                    const int ch = n_tile * p.bn + sl * kOutSlab + c;
                    if (sl * kOutSlab < p.bn && ch < p.k_store) {
                        float v = 0.f;
#pragma unroll
                        for (int r = 0; r < 8; ++r) v += red[((r * 4 + sl) * 2 + which) * 64 + c];
                        atomicAdd(p.stats + which * p.k_store + ch, v);
                    }
                }
                named_bar_sync(1, kEpiThreads);      // red[] is read before the next slab is staged over it
            };
            int lt = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++lt) {
                const TileCoord t = decode_tile(p, tile);
                const int acc = lt & 1;
                const uint32_t acc_phase = (lt >> 1) & 1;
                // rows outside the image exist only in border tiles (uniform test); they are clipped by the store
                // and must be zero for the statistics
                const bool border = (t.x0 + p.bw > p.W) || (t.y0 + p.bh > p.H) || (t.n0 + p.bi > p.N);
                const bool valid = !border || ((t.x0 + wl < p.W) && (t.y0 + hl < p.H) && (t.n0 + il < p.N));
                if (p.stats && t.n_tile != cur_n_tile) {
                    if (cur_n_tile >= 0) flush_stats(cur_n_tile);
                    cur_n_tile = t.n_tile;
                }
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + acc * kMaxBN + (static_cast<uint32_t>(q * 32) << 16) + half * 32;
#pragma unroll
                for (int sl = 0; sl < 4; ++sl) {
                    if (sl * kOutSlab < p.bn) {
                        uint8_t* buf = out_buf + (slab_seq & 1) * kOutBufBytes;
                        long long* dbg = (p.dbg && blockIdx.x == 0 && et == 0 && slab_seq < p.dbg_slots) ? p.dbg + slab_seq * 8
                                                                                                   : nullptr;
                        ++slab_seq;
                        if (dbg) dbg[0] = gtime();
                        const int k0 = t.n_tile * p.bn + sl * kOutSlab;
                        const int kh = k0 + half * 32;
                        uint32_t v[32];
                        const bool have = sl * kOutSlab + half * 32 < p.bn;
                        if (have) {
                            tmem_ld_32x32(taddr + sl * kOutSlab, v);
                            tmem_ld_wait();
                        }
                        if (!have || !valid) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = 0u;
                        }
                        if (dbg) dbg[1] = gtime();
                        // the buffer is free once the store issued two slabs ago has read it (thread 128 waited for
                        // that before arriving here)
                        named_bar_sync(1, kEpiThreads);
                        if (dbg) dbg[2] = gtime();
                        uint8_t* rowp = buf + row * 128;
                        const int sw = row & 7;
                        if (affine) {
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                const int kg = kh + g * 8;
                                float f[8];
#pragma unroll
                                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g * 8 + j]);
                                if (kg < p.k_store) {
                                    if (p.scale) {
                                        const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.scale + kg));
                                        const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.scale + kg) + 1);
                                        f[0] *= s0.x; f[1] *= s0.y; f[2] *= s0.z; f[3] *= s0.w;
                                        f[4] *= s1.x; f[5] *= s1.y; f[6] *= s1.z; f[7] *= s1.w;
                                    }
                                    if (p.bias) {
                                        const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + kg));
                                        const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + kg) + 1);
                                        f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
                                        f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
                                    }
                                    if (p.relu) {
#pragma unroll
                                        for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
                                    }
                                }
                                uint4 o;
                                o.x = pack_bf16x2(f[0], f[1]);
                                o.y = pack_bf16x2(f[2], f[3]);
                                o.z = pack_bf16x2(f[4], f[5]);
                                o.w = pack_bf16x2(f[6], f[7]);
                                *reinterpret_cast<uint4*>(rowp + (((half * 4 + g) ^ sw) << 4)) = o;
                            }
                        } else {
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                uint4 o;
                                o.x = pack_bf16x2(__uint_as_float(v[g * 8 + 0]), __uint_as_float(v[g * 8 + 1]));
                                o.y = pack_bf16x2(__uint_as_float(v[g * 8 + 2]), __uint_as_float(v[g * 8 + 3]));
                                o.z = pack_bf16x2(__uint_as_float(v[g * 8 + 4]), __uint_as_float(v[g * 8 + 5]));
                                o.w = pack_bf16x2(__uint_as_float(v[g * 8 + 6]), __uint_as_float(v[g * 8 + 7]));
                                *reinterpret_cast<uint4*>(rowp + (((half * 4 + g) ^ sw) << 4)) = o;
                            }
                        }
                        if (dbg) dbg[3] = gtime();
                        fence_proxy_async();
                        if (dbg) dbg[4] = gtime();
                        named_bar_sync(2, kEpiThreads);
                        if (dbg) dbg[5] = gtime();
                        if (et == 0) {
                            tma_store_4d(&p.y_map, buf, k0, t.x0, t.y0, t.n0);
                            tma_store_commit();
                            tma_store_wait_read<1>();      // the previous slab's buffer may be overwritten
                        }
                        if (dbg) dbg[6] = gtime();
                        if (p.stats) {
                            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
                            const uint8_t* colp = buf + rg * 16 * 128 + (cp & 3) * 4;
                            const int cq = cp >> 2;
#pragma unroll
                            for (int r = 0; r < 16; ++r) {
                                // rows rg*16 + r: (row & 7) == (r & 7) because rg*16 is a multiple of 8
                                const uint32_t wv = *reinterpret_cast<const uint32_t*>(colp + r * 128 + ((cq ^ (r & 7)) << 4));
                                const float x0 = __uint_as_float(wv << 16), x1 = __uint_as_float(wv & 0xffff0000u);
                                a0 += x0;
                                a1 += x1;
                                b0 = fmaf(x0, x0, b0);
                                b1 = fmaf(x1, x1, b1);
                            }
                            st1[sl][0] += a0;
                            st1[sl][1] += a1;
                            st2[sl][0] += b0;
                            st2[sl][1] += b1;
                        }
                        if (dbg) dbg[7] = gtime();
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[acc]);
            }
            if (p.stats && cur_n_tile >= 0) flush_stats(cur_n_tile);
            if (et == 0) tma_store_wait_all<0>();
        } else if (warp < 8) {
            // fp32 class map (cls_conv): direct stores, four warps
            int lt = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++lt) {
                const TileCoord t = decode_tile(p, tile);
                const int acc = lt & 1;
                const uint32_t acc_phase = (lt >> 1) & 1;
                const int px = t.x0 + wl, py = t.y0 + hl, pn = t.n0 + il;
                const bool valid = (px < p.W) && (py < p.H) && (pn < p.N);
                const size_t pix = (static_cast<size_t>(pn) * p.H + py) * p.W + px;
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + acc * kMaxBN + (static_cast<uint32_t>(q * 32) << 16);
                for (int ch = 0; ch < p.bn; ch += 32) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + ch, v);
                    tmem_ld_wait();
                    const int k0 = t.n_tile * p.bn + ch;
                    if (valid) {
    #pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const int kg = k0 + g * 8;
                            if (kg < p.k_store) {
                                float f[8];
    #pragma unroll
                                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g * 8 + j]);
                                if (p.scale) {
    #pragma unroll
                                    for (int j = 0; j < 8; ++j) f[j] *= __ldg(p.scale + kg + j);
                                }
                                if (p.bias) {
    #pragma unroll
                                    for (int j = 0; j < 8; ++j) f[j] += __ldg(p.bias + kg + j);
                                }
                                if (p.relu) {
    #pragma unroll
                                    for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
                                }
                                const size_t off = pix * p.y_c_total + p.y_c_offset + kg;
                                if (p.out_fp32) {
                                    float4* dst = reinterpret_cast<float4*>(static_cast<float*>(p.y) + off);
                                    dst[0] = make_float4(f[0], f[1], f[2], f[3]);
                                    dst[1] = make_float4(f[4], f[5], f[6], f[7]);
                                } else {
                                    uint4 o;
                                    o.x = pack_bf16x2(f[0], f[1]);
                                    o.y = pack_bf16x2(f[2], f[3]);
                                    o.z = pack_bf16x2(f[4], f[5]);
                                    o.w = pack_bf16x2(f[6], f[7]);
                                    *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.y) + off) = o;
                                }
                            }
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[acc]);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

int ilog2(int v) {
    int l = 0;
    while ((1 << l) < v) ++l;
    return l;
}

// Total pipeline iterations (TMA box pairs) the whole problem costs with a bw x bh x bi patch.
long long patch_cost(int N, int H, int W, int bw, int bh, int bi, int nseg, const IgemmSeg* seg) {
    const int tx = ceil_div(W, bw), ty = ceil_div(H, bh), ti = ceil_div(N, bi);
    long long per_image = 0;
    for (int y = 0; y < ty; ++y)
        for (int x = 0; x < tx; ++x) {
            const int y0 = y * bh, x0 = x * bw;
            for (int g = 0; g < nseg; ++g)
                for (int r = 0; r < seg[g].r; ++r)
                    for (int s = 0; s < seg[g].s; ++s) {
                        const int dy = (r - seg[g].r / 2) * seg[g].dil, dx = (s - seg[g].s / 2) * seg[g].dil;
                        if ((y0 + dy < H) && (y0 + bh - 1 + dy >= 0) && (x0 + dx < W) && (x0 + bw - 1 + dx >= 0))
                            per_image += seg[g].cblks;
                    }
        }
    return per_image * ti;
}

int launch_igemm(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                 cudaStream_t st) {
    clear_error();
    B200SEG_CHECK(n > 0 && h > 0 && w > 0, "conv: bad activation shape %dx%dx%d", n, h, w);
    B200SEG_CHECK(n_src >= 1 && n_src <= B200SEG_MAX_CONV_SRC, "conv: n_src %d out of range", n_src);
    B200SEG_CHECK(out && out->y, "conv: null output");
    B200SEG_CHECK(out->k > 0 && out->k % 8 == 0 && out->y_c_offset % 8 == 0 && out->y_c_total % 8 == 0,
                  "conv: output channels (k=%d, offset=%d, total=%d) must be multiples of 8", out->k,
                  out->y_c_offset, out->y_c_total);
    B200SEG_CHECK(out->y_c_offset + out->k <= out->y_c_total, "conv: output slice exceeds buffer channels");

    IgemmParams p;
    memset(&p, 0, sizeof(p));
    p.nseg = n_src;
    p.N = n;
    p.H = h;
    p.W = w;
    for (int i = 0; i < n_src; ++i) {
        const b200seg_conv_src& s = src[i];
        B200SEG_CHECK(s.x && s.w, "conv: source %d has null pointers", i);
        B200SEG_CHECK((s.r == 1 && s.s == 1) || (s.r == 3 && s.s == 3), "conv: only 1x1 and 3x3 filters (got %dx%d)",
                      s.r, s.s);
        B200SEG_CHECK(s.dil >= 1, "conv: dilation %d", s.dil);
        B200SEG_CHECK(s.c > 0 && s.c % 8 == 0 && s.x_c_offset % 8 == 0 && s.x_c_total % 8 == 0,
                      "conv: input channels (c=%d, offset=%d, total=%d) must be multiples of 8", s.c, s.x_c_offset,
                      s.x_c_total);
        B200SEG_CHECK(s.x_c_offset + s.c <= s.x_c_total, "conv: input slice exceeds buffer channels");
        B200SEG_CHECK(s.w_rows >= out->k, "conv: packed weight has %d rows < %d output channels", s.w_rows, out->k);
        p.seg[i].r = s.r;
        p.seg[i].s = s.s;
        p.seg[i].dil = s.dil;
        p.seg[i].cblks = ceil_div(s.c, kBK);
    }

    // output-channel tiling
    p.tiles_n = ceil_div(out->k, kMaxBN);
    // whole 64-channel store slabs per n-tile (a narrow single tile may be 32 wide: the store clips at k)
    p.bn = out->k <= 32 ? 32 : round_up(ceil_div(out->k, p.tiles_n), kOutSlab);
    B200SEG_CHECK(p.bn <= kMaxBN, "conv: internal bn %d", p.bn);
    B200SEG_CHECK(!(out->stats && (out->scale || out->bias || out->relu || out->out_fp32)),
                  "conv: fused BN statistics are those of the stored bf16 conv result; scale/bias/relu/fp32 output "
                  "cannot be combined with them");

    // pixel patch: the power-of-two bw x bh x bi = 128 shape that needs the fewest pipeline iterations.  The search
    // walks every tile of the map, so its result is memoised per geometry (a training step repeats the same few).
    long long best = -1;
    {
        static std::mutex mu;
        static std::map<std::vector<int>, std::array<int, 3>> memo;
        std::vector<int> key = {n, h, w, n_src};
        for (int i = 0; i < n_src; ++i) {
            key.push_back(p.seg[i].r);
            key.push_back(p.seg[i].dil);
            key.push_back(p.seg[i].cblks);
        }
        std::lock_guard<std::mutex> lock(mu);
        auto it = memo.find(key);
        if (it != memo.end()) {
            p.bw = it->second[0];
            p.bh = it->second[1];
            p.bi = it->second[2];
            best = 1;
        } else {
            for (int bi = 1; bi <= 128; bi *= 2) {
                for (int bw = 1; bw * bi <= 128; bw *= 2) {
                    const int bh = 128 / (bw * bi);
                    const long long c = patch_cost(n, h, w, bw, bh, bi, n_src, p.seg);
                    if (best < 0 || c < best || (c == best && bw > p.bw)) {
                        best = c;
                        p.bw = bw;
                        p.bh = bh;
                        p.bi = bi;
                    }
                }
            }
            if (best > 0) memo[key] = {p.bw, p.bh, p.bi};
        }
    }
    B200SEG_CHECK(best > 0, "conv: no tile shape for %dx%dx%d", n, h, w);
    p.lbw = ilog2(p.bw);
    p.lbh = ilog2(p.bh);
    p.tiles_x = ceil_div(w, p.bw);
    p.tiles_y = ceil_div(h, p.bh);
    p.tiles_i = ceil_div(n, p.bi);
    p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_i * p.tiles_n;

    for (int i = 0; i < n_src; ++i) {
        const b200seg_conv_src& s = src[i];
        const uint64_t adims[4] = {(uint64_t)s.c, (uint64_t)w, (uint64_t)h, (uint64_t)n};
        const uint64_t astr[3] = {(uint64_t)s.x_c_total * 2, (uint64_t)w * s.x_c_total * 2,
                                  (uint64_t)h * w * s.x_c_total * 2};
        const uint32_t abox[4] = {kBK, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bi};
        const char* abase = static_cast<const char*>(s.x) + (size_t)s.x_c_offset * 2;
        if (encode_tmap_bf16(&p.a_map[i], abase, 4, adims, astr, abox)) return 1;
        const uint64_t bdims[3] = {(uint64_t)s.c, (uint64_t)s.w_rows, (uint64_t)(s.r * s.s)};
        const uint64_t bstr[2] = {(uint64_t)s.c * 2, (uint64_t)s.w_rows * s.c * 2};
        const uint32_t bbox[3] = {kBK, (uint32_t)p.bn, 1};
        if (encode_tmap_bf16(&p.b_map[i], s.w, 3, bdims, bstr, bbox)) return 1;
    }

    if (!out->out_fp32) {
        const uint64_t ydims[4] = {(uint64_t)out->k, (uint64_t)w, (uint64_t)h, (uint64_t)n};
        const uint64_t ystr[3] = {(uint64_t)out->y_c_total * 2, (uint64_t)w * out->y_c_total * 2,
                                  (uint64_t)h * w * out->y_c_total * 2};
        const uint32_t ybox[4] = {kOutSlab, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bi};
        const char* ybase = static_cast<const char*>(out->y) + (size_t)out->y_c_offset * 2;
        if (encode_tmap_bf16(&p.y_map, ybase, 4, ydims, ystr, ybox)) return 1;
    }
    p.y = out->y;
    p.y_c_total = out->y_c_total;
    p.y_c_offset = out->y_c_offset;
    p.k_store = out->k;
    p.out_fp32 = out->out_fp32;
    p.scale = out->scale;
    p.bias = out->bias;
    p.relu = out->relu;
    p.stats = out->stats;
    p.dbg = g_dbg;
    p.dbg_slots = g_dbg_slots;
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(out->y) & 15) == 0, "conv: output pointer not 16-byte aligned");

    static bool attr_set = false;
    if (!attr_set) {
        B200SEG_CUDA(cudaFuncSetAttribute(igemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    const int sms = sm_count();
    B200SEG_CHECK(sms > 0, "conv: cannot query SM count");
    const int grid = p.num_tiles < sms ? p.num_tiles : sms;
    igemm_kernel<<<grid, kThreads, kSmemBytes, st>>>(p);
    B200SEG_LAUNCH_CHECK();
    return 0;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_debug_timing(int64_t* buf, int slots) {
    b200seg::g_dbg = reinterpret_cast<long long*>(buf);
    b200seg::g_dbg_slots = buf ? slots : 0;
    return 0;
}

int b200seg_conv_fprop(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st) {
    return b200seg::launch_igemm(n, h, w, n_src, src, out, static_cast<cudaStream_t>(st));
}

int b200seg_conv_dgrad(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st) {
    return b200seg::launch_igemm(n, h, w, n_src, src, out, static_cast<cudaStream_t>(st));
}

}  // extern "C"
