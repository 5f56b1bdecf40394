// Convolution weight gradient on tcgen05: for every filter tap a GEMM whose reduction runs over pixels.
//
//   dW[tap][k][c] = sum over pixels p of dY[p][k] * X[p + tap_offset][c]
//
// Both operands are NHWC (channels contiguous, pixels strided), i.e. "MN-major" for this GEMM, which
// tcgen05.mma consumes directly through MN-major shared-memory descriptors — no transposes are materialised.
// One pipeline stage = a patch of 64 pixels (bw x bh x bi): TMA loads (64 channels x patch) boxes of the
// M-side tensor (2 boxes = 128 rows of D) and of the N-side tensor (bn/64 boxes), the X boxes shifted by the
// tap offset with out-of-bounds pixels zero-filled (that is the padding).  Patches whose shifted window is
// entirely padding are skipped.  The reduction is split across CTAs (split-K) when taps x tiles alone do not
// fill the GPU.  The split CTAs of one output tile form a thread-block CLUSTER (2, 4 or 8 CTAs): each drains its TMEM
// accumulator into its own shared memory, and after a cluster barrier CTA r sums column slice r of all ranks through
// distributed shared memory in rank order (deterministic) — on chip, no workspace round trip.  With splits <= cluster
// size the sums go straight into the OIHW fp32 gradient from the epilogue; only beyond that do per-cluster partials go
// to a caller-provided fp32 workspace for a (now 2-8x smaller) fixed-order second-stage reduce.
//
// The M side is whichever of (dY, X) has more channels to tile by 128; the other is the N side (<= 256 per
// tile).  For cls_conv (21 classes) that puts X's 384 channels on M and the padded 32 classes on N.
#include "common.cuh"
#include "pdl.cuh"
#include "ptx.cuh"

#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <vector>

namespace b200seg {
namespace {

constexpr int kBM = 128;
constexpr int kBKP = 64;            // pixels per stage
constexpr int kMaxBN = 256;
constexpr int kStages = 4;
constexpr int kBoxBytes = 64 * kBKP * 2;          // 8 KB: 64 channels x 64 pixels
constexpr int kABytes = 2 * kBoxBytes;
constexpr int kBBytesMax = (kMaxBN / 64) * kBoxBytes;
constexpr int kStageBytes = kABytes + kBBytesMax;
constexpr int kThreads = 256;
constexpr int kSmemBytes = kStages * kStageBytes + 256 + 1024;

struct __align__(64) WgradParams {
    CUtensorMap x_map[4];  // X views (channels, W, H, N): one for stride 1, the four parity sub-grids for stride 2
    CUtensorMap y_map;     // dY (channels, W, H, N)
    int N, H, W;           // the dY grid
    int R, S;
    short tap_map[9], tap_dy[9], tap_dx[9];   // per filter tap: which X view it reads and its pixel shift there
    int shift_m;           // 1: the M side is X; 0: the N side is X
    int bw, bh, bi;
    int patches_x, patches_y, patches;
    int tiles_m, tiles_n, bn;
    int splits, patches_per_split;
    int m_pad, n_pad;      // workspace matrix dims per (split, tap)
    float* ws;
    int tap_group;         // filter taps per CTA (>1 only with X on the N side): the taps' shifted X boxes sit side by
                           // side as one N = tap_group * bn operand, so the dY boxes are loaded once per group
    int no_skip;           // anchored window X: the buffer carries its own padding, every (tap, patch) is in bounds
    int cs;                // cluster size along the split dimension (1 = no cluster); splits = splits_hi * cs
    int direct;            // 1: this launch's sums are final — write dw (fp32 OIHW) from the epilogue, no workspace
    float* dw;
    int k_out, c_in, accumulate;
    // 5-D views (64 ch, W, H, N, channel block) of the same tensors, built when the channel count is a multiple of 64: ONE
    // TMA instruction then brings all the 64-channel boxes of an operand (2 on the M side, bn/64 on the N side) — the
    // producer is a single thread and each TMA issue costs it ~200 clk, which at six issues per stage was the whole stage
    CUtensorMap x_map5[4];
    CUtensorMap y_map5;
    int x5, y5;
};

// final element (m, n) of filter tap `tap` -> dw[k][c][tap] (+)= v        (m, n) = (c, k) with X on the M side, else (k, c)
__device__ __forceinline__ void wgrad_store_final(const WgradParams& p, int taps, int tap, int m, int n, float v) {
    const int kk = p.shift_m ? n : m, cc = p.shift_m ? m : n;
    if (kk < p.k_out && cc < p.c_in) {
        float* d = p.dw + (static_cast<size_t>(kk) * p.c_in + cc) * taps + tap;
        *d = p.accumulate ? *d + v : v;
    }
}

// Walks the pixel patches kp = begin, begin+1, ... with running (x, y, image) patch counters: the producer and the
// MMA issuer are single threads, so a division per pipeline stage would sit on the critical path.
struct PatchWalker {
    int px, py, pn;
    __device__ __forceinline__ void init(const WgradParams& p, int kp) {
        px = kp % p.patches_x;
        const int t = kp / p.patches_x;
        py = t % p.patches_y;
        pn = t / p.patches_y;
    }
    __device__ __forceinline__ void next(const WgradParams& p) {
        if (++px == p.patches_x) {
            px = 0;
            if (++py == p.patches_y) {
                py = 0;
                ++pn;
            }
        }
    }
    __device__ __forceinline__ bool in_bounds(const WgradParams& p, int dy, int dx) const {
        if (p.no_skip) return true;
        const int x0 = px * p.bw, y0 = py * p.bh;
        return (y0 + dy < p.H) && (y0 + p.bh - 1 + dy >= 0) && (x0 + dx < p.W) && (x0 + p.bw - 1 + dx >= 0);
    }
};

// kGrouped = false is the one-tap-per-CTA kernel; the grouped code lives in its own instantiation so the single-thread
// producer / MMA loops of the common case stay as small as they were (sharing one body cost the multi-tap launches
// 5-20 %).
template <bool kGrouped>
__global__ void __launch_bounds__(kThreads, 1) wgrad_kernel(const __grid_constant__ WgradParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + kStages;
    uint64_t* tfull = bars + 2 * kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // tile decode: cluster rank (low part of the split index) fastest, then n_tile, m_tile, tap, split (high part)
    const int rank = p.cs > 1 ? static_cast<int>(blockIdx.x % p.cs) : 0;
    int tile = blockIdx.x / p.cs;
    const int n_tile = tile % p.tiles_n;
    tile /= p.tiles_n;
    const int m_tile = tile % p.tiles_m;
    tile /= p.tiles_m;
    const int taps = p.R * p.S;
    const int G = kGrouped ? p.tap_group : 1;
    const int ngroups = (taps + G - 1) / G;
    const int tap = (tile % ngroups) * G;          // first tap of this CTA's group
    const int gt = min(G, taps - tap);             // taps in the group
    const int split_hi = tile / ngroups;
    const int split = split_hi * p.cs + rank;
    const int dy = p.tap_dy[tap];
    const int dx = p.tap_dx[tap];
    const CUtensorMap* xm = &p.x_map[p.tap_map[tap]];
    const CUtensorMap* m_map = p.shift_m ? xm : &p.y_map;
    const CUtensorMap* n_map = p.shift_m ? &p.y_map : xm;
    const CUtensorMap* xm5 = &p.x_map5[p.tap_map[tap]];
    const CUtensorMap* m_map5 = p.shift_m ? xm5 : &p.y_map5;
    const CUtensorMap* n_map5 = p.shift_m ? &p.y_map5 : xm5;
    const bool m5 = (p.shift_m ? p.x5 : p.y5) != 0, n5 = (p.shift_m ? p.y5 : p.x5) != 0;
    // a patch is skipped only when every tap of the group reads nothing but padding there (a tap that does is
    // zero-filled by the TMA unit)
    int gdy[4], gdx[4];                            // the group's shifts in registers: the walkers below are single
#pragma unroll                                     // threads on the critical path
    for (int j = 0; j < 4; ++j) {
        gdy[j] = kGrouped ? p.tap_dy[min(tap + j, taps - 1)] : dy;
        gdx[j] = kGrouped ? p.tap_dx[min(tap + j, taps - 1)] : dx;
    }
    // (a by-reference lambda over the __grid_constant__ parameter block made every p.* read in the single-thread
    // walkers a generic load: 10-20 % on the multi-tap launches; the geometry is copied to registers instead)
    const int gH = p.H, gW = p.W, gbw = p.bw, gbh = p.bh;
    const bool no_skip = p.no_skip != 0;
#define B200SEG_TAP_IN(w, ddy, ddx)                                                                                   \
    (((w).py * gbh + (ddy) < gH) && ((w).py * gbh + gbh - 1 + (ddy) >= 0) && ((w).px * gbw + (ddx) < gW) &&           \
     ((w).px * gbw + gbw - 1 + (ddx) >= 0))
#define group_in_bounds(w)                                                                                            \
    (kGrouped ? (no_skip || B200SEG_TAP_IN(w, dy, dx) || (gt > 1 && B200SEG_TAP_IN(w, gdy[1], gdx[1])) ||             \
                 (gt > 2 && B200SEG_TAP_IN(w, gdy[2], gdx[2])) || (gt > 3 && B200SEG_TAP_IN(w, gdy[3], gdx[3])))      \
              : (w).in_bounds(p, dy, dx))
    const int kp_begin = split * p.patches_per_split;
    const int kp_end = min(kp_begin + p.patches_per_split, p.patches);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(m5 ? m_map5 : m_map);
        tma_prefetch_desc(n5 ? n_map5 : n_map);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        mbar_init(tfull, 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_slot, kMaxBN);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // PDL: everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped the previous kernel's
    // tail; no global memory is touched before this point
    pdl_launch_dependents();
    pdl_wait();
    const int n_boxes = p.bn / 64;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx_bytes = kABytes + gt * n_boxes * kBoxBytes;
            const int mdx = p.shift_m ? dx : 0, mdy = p.shift_m ? dy : 0;
            const int ndx = p.shift_m ? 0 : dx, ndy = p.shift_m ? 0 : dy;
            PatchWalker pw;
            pw.init(p, kp_begin);
            for (int kp = kp_begin; kp < kp_end; ++kp, pw.next(p)) {
                if (!group_in_bounds(pw)) continue;
                const int x0 = pw.px * p.bw, y0 = pw.py * p.bh, n0 = pw.pn * p.bi;
                mbar_wait(&empty[stage], phase ^ 1);
                uint8_t* sa = smem + stage * kStageBytes;
                mbar_arrive_expect_tx(&full[stage], tx_bytes);
                if (m5) {
                    tma_load_5d(sa, m_map5, &full[stage], 0, x0 + mdx, y0 + mdy, n0, m_tile * (kBM / 64));
                } else {
                    for (int b = 0; b < 2; ++b)
                        tma_load_4d(sa + b * kBoxBytes, m_map, &full[stage], m_tile * kBM + b * 64, x0 + mdx, y0 + mdy, n0);
                }
                if (!kGrouped) {
                    if (n5) {
                        tma_load_5d(sa + kABytes, n_map5, &full[stage], 0, x0 + ndx, y0 + ndy, n0, n_tile * n_boxes);
                    } else {
                        for (int b = 0; b < n_boxes; ++b)
                            tma_load_4d(sa + kABytes + b * kBoxBytes, n_map, &full[stage], n_tile * p.bn + b * 64, x0 + ndx,
                                        y0 + ndy, n0);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (j < gt)
                            for (int b = 0; b < n_boxes; ++b)
                                tma_load_4d(sa + kABytes + (j * n_boxes + b) * kBoxBytes, &p.x_map[p.tap_map[tap + j]],
                                            &full[stage], b * 64, x0 + gdx[j], y0 + gdy[j], n0);
                }
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(kBM, gt * p.bn, 1, 1);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            PatchWalker pw;
            pw.init(p, kp_begin);
            for (int kp = kp_begin; kp < kp_end; ++kp, pw.next(p)) {
                if (!group_in_bounds(pw)) continue;
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * kStageBytes);
#pragma unroll
                for (int k = 0; k < kBKP / 16; ++k) {
                    // 16 pixels (K) = 16 rows of 128 B; 64-channel chunks are one 8 KB box apart (LBO),
                    // 8-pixel groups 1 KB apart (SBO)
                    const uint64_t a_desc = umma_smem_desc(sa + k * 2048, kBoxBytes, 1024);
                    const uint64_t b_desc = umma_smem_desc(sa + kABytes + k * 2048, kBoxBytes, 1024);
                    tc_mma_f16(tmem_base, a_desc, b_desc, idesc, (it | k) != 0);
                }
                tc_commit(&empty[stage]);
                ++it;
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            if (it > 0)
                tc_commit(tfull);
            else
                mbar_arrive(tfull);
        }
    } else if (warp >= 4) {
        // any work at all for this (tap, split)?  (uniform across the CTA)
        bool any = false;
        PatchWalker pw;
        pw.init(p, kp_begin);
        for (int kp = kp_begin; kp < kp_end && !any; ++kp, pw.next(p)) any = group_in_bounds(pw);
        const int q = warp & 3;
        const int row = q * 32 + lane;
        mbar_wait(tfull, 0);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        if (p.cs > 1) {
            // stage the accumulator tile in this CTA's shared memory (the pipeline stages are idle: every MMA has
            // completed), column-major [col][128 rows] so lanes write consecutive words; the cluster reduces it below
            float* stage_buf = reinterpret_cast<float*>(smem);
            for (int col = 0; col < gt * p.bn; col += 32) {
                uint32_t v[32];
                if (any) {
                    tmem_ld_32x32(taddr + col, v);
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = 0;
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) stage_buf[(col + j) * kBM + row] = __uint_as_float(v[j]);
            }
        } else if (p.direct) {
            // a single split: the accumulator IS the gradient
            for (int g = 0; g < gt; ++g)
                for (int ch = 0; ch < p.bn; ch += 32) {
                    uint32_t v[32];
                    if (any) {
                        tmem_ld_32x32(taddr + g * p.bn + ch, v);
                        tmem_ld_wait();
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = 0;
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        wgrad_store_final(p, taps, tap + g, m_tile * kBM + row, n_tile * p.bn + ch + j, __uint_as_float(v[j]));
                }
        } else {
            for (int g = 0; g < gt; ++g) {
                // tap g of the group: accumulator columns [g * bn, (g + 1) * bn) -> its own plane of the workspace
                float* dst = p.ws + ((static_cast<size_t>(split) * taps + tap + g) * p.m_pad + (m_tile * kBM + row)) * p.n_pad +
                             n_tile * p.bn;
                for (int ch = 0; ch < p.bn; ch += 32) {
                    uint32_t v[32];
                    if (any) {
                        tmem_ld_32x32(taddr + g * p.bn + ch, v);
                        tmem_ld_wait();
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = 0;
                    }
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<uint4*>(dst + ch + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                }
            }
        }
    }

    if (p.cs > 1) {
        // ---- split-K reduction on chip: CTA `rank` owns accumulator columns [rank * wc, (rank + 1) * wc) and adds them
        // up over the cluster's ranks in rank order through distributed shared memory
        __syncwarp();
        cluster_sync_all();
        if (warp >= 4) {
            const int row = (warp & 3) * 32 + lane;
            const int wtot = gt * p.bn, wc = wtot / p.cs;           // bn is a multiple of 64, cs of {2, 4, 8}
            const uint32_t local = smem_u32(smem);
            uint32_t peer[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) peer[r] = r < p.cs ? dsmem_addr(local, r) : 0u;
            for (int col = rank * wc; col < (rank + 1) * wc; col += 4) {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    if (r < p.cs) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[j] += dsmem_ld_f32(peer[r] + ((col + j) * kBM + row) * 4);
                    }
                }
                const int g = col / p.bn, ch = col - g * p.bn;      // 4 consecutive columns never straddle a tap
                if (p.direct) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        wgrad_store_final(p, taps, tap + g, m_tile * kBM + row, n_tile * p.bn + ch + j, acc[j]);
                } else {
                    float* dst = p.ws + ((static_cast<size_t>(split_hi) * taps + tap + g) * p.m_pad + (m_tile * kBM + row)) * p.n_pad +
                                 n_tile * p.bn + ch;
                    *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
                }
            }
        }
        __syncwarp();
        cluster_sync_all();          // nobody leaves while a peer still reads its shared memory
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kMaxBN);
    }
#undef group_in_bounds
#undef B200SEG_TAP_IN
}

// dw[k][c][r][s] (+)= sum_split ws[split][tap][m][n], (m, n) = (k, c) or (c, k)
__global__ void wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int k_out, int c, int taps,
                                    int splits, int m_pad, int n_pad, int m_is_c, int accumulate) {
    pdl_prologue();
    const long long total = static_cast<long long>(k_out) * c * taps;
    const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (i >= total) return;
    // thread index enumerates the workspace-contiguous order (tap, m, n) so reads coalesce
    const int n_dim = m_is_c ? k_out : c;
    const int m_dim = m_is_c ? c : k_out;
    const int n = static_cast<int>(i % n_dim);
    const int m = static_cast<int>((i / n_dim) % m_dim);
    const int tap = static_cast<int>(i / (static_cast<long long>(n_dim) * m_dim));
    const size_t plane = static_cast<size_t>(m_pad) * n_pad;
    const float* src = ws + (static_cast<size_t>(tap) * m_pad + m) * n_pad + n;
    // up to 148 partials per element: eight loads in flight per thread (the adds keep their fixed order); one load
    // per trip made the small-weight launches pure latency chains of 16-32 us
    float acc = 0.f;
    const size_t sstride = static_cast<size_t>(taps) * plane;
    int s = 0;
    for (; s + 8 <= splits; s += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcs(src + static_cast<size_t>(s + u) * sstride);
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += v[u];
    }
    for (; s < splits; ++s) acc += __ldcs(src + static_cast<size_t>(s) * sstride);
    const int kk = m_is_c ? n : m;
    const int cc = m_is_c ? m : n;
    float* d = dw + (static_cast<size_t>(kk) * c + cc) * taps + tap;
    *d = accumulate ? (*d + acc) : acc;
}

bool cluster_reduce_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200SEG_WGRAD_CLUSTER");
        return !(e && e[0] == '0');
    }();
    return on;
}

bool direct_write_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200SEG_WGRAD_DIRECT");
        return !(e && e[0] == '0');
    }();
    return on;
}

int cluster_capacity(int cs);

bool tap_grouping_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200SEG_WGRAD_TAPGROUP");
        return !(e && e[0] == '0');
    }();
    return on;
}

void set_wgrad_attrs() {
    static std::once_flag once;
    std::call_once(once, [] {
        cudaFuncSetAttribute(wgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        cudaFuncSetAttribute(wgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    });
}

// co-resident clusters of `cs` CTAs of this kernel (one CTA per SM: ~193 KB of shared memory) — asked of the driver once
int cluster_capacity(int cs) {
    static std::mutex mu;
    static std::map<int, int> memo;
    std::lock_guard<std::mutex> lock(mu);
    auto it = memo.find(cs);
    if (it != memo.end()) return it->second;
    set_wgrad_attrs();
    int sms = gemm_sm_count();
    if (sms <= 0) sms = 148;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(sms / cs * cs));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, wgrad_kernel<false>, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = sms / cs * 3 / 4;          // conservative guess: GPCs of ~18 SMs leave a remainder per cluster size
    }
    const int lim = sms / cs;
    if (n > lim) n = lim;
    memo[cs] = n;
    return n;
}

struct WgradPlan {
    int shift_m, bw, bh, bi, patches_x, patches_y, patches_i, tiles_m, tiles_n, bn, splits, pps, m_pad, n_pad, tap_group;
    int cs, splits_hi, direct;             // cluster size along the splits, workspace planes, epilogue writes dw itself
    size_t ws_bytes;
    long long iters_run, iters_dense;      // (tap, patch) pairs with / without the all-padding ones skipped
};

int plan_wgrad_uncached(const b200seg_wgrad_desc* d, WgradPlan* pl) {
    B200SEG_CHECK(d && d->dy && d->x && d->dw, "wgrad: null pointer");
    B200SEG_CHECK(d->n > 0 && d->h > 0 && d->w > 0, "wgrad: bad shape");
    const bool anchored = d->x_rows || d->x_cols;
    if (anchored) {
        B200SEG_CHECK(d->r >= 1 && d->s >= 1 && d->r * d->s <= 9 && d->dil == 1 && d->stride == 1,
                      "wgrad: anchored window input with %dx%d taps, dilation %d, stride %d", d->r, d->s, d->dil, d->stride);
        B200SEG_CHECK(d->x_c_total > 0 && d->c > 0, "wgrad: bad window");
        const int span = ceil_div(d->x_c_offset + d->c, d->x_c_total);
        B200SEG_CHECK(d->h + d->r - 1 <= d->x_rows && d->w + d->s - 1 + span - 1 <= d->x_cols,
                      "wgrad: %dx%d grid with %dx%d taps and %d-pixel windows exceeds the %dx%d buffer", d->h, d->w, d->r,
                      d->s, span, d->x_rows, d->x_cols);
    } else {
        B200SEG_CHECK((d->r == 1 && d->s == 1) || (d->r == 3 && d->s == 3), "wgrad: only 1x1 and 3x3 filters");
        B200SEG_CHECK(d->stride == 1 || (d->stride == 2 && d->dil == 1), "wgrad: stride %d / dilation %d not supported",
                      d->stride, d->dil);
    }
    B200SEG_CHECK(d->k % 8 == 0 && d->dy_c_offset % 8 == 0 && d->dy_c_total % 8 == 0 && d->c % 8 == 0 &&
                      d->x_c_offset % 8 == 0 && d->x_c_total % 8 == 0,
                  "wgrad: channel counts/offsets must be multiples of 8");
    B200SEG_CHECK(d->k_out > 0 && d->k_out <= d->k, "wgrad: k_out %d vs k %d", d->k_out, d->k);
    // M side = the tensor with more channels
    pl->shift_m = d->c > d->k ? 1 : 0;
    const int m_ch = pl->shift_m ? d->c : d->k;
    const int n_ch = pl->shift_m ? d->k : d->c;
    pl->tiles_m = ceil_div(m_ch, kBM);
    pl->tiles_n = ceil_div(n_ch, kMaxBN);
    pl->bn = round_up(ceil_div(n_ch, pl->tiles_n), 64);
    pl->m_pad = pl->tiles_m * kBM;
    pl->n_pad = pl->tiles_n * pl->bn;
    // 64-pixel patch with the fewest non-padding (tap, patch) pairs
    long long best = -1;
    for (int bi = 1; bi <= 64; bi *= 2) {
        for (int bw = 1; bw * bi <= 64; bw *= 2) {
            const int bh = 64 / (bw * bi);
            const int px = ceil_div(d->w, bw), py = ceil_div(d->h, bh), pi = ceil_div(d->n, bi);
            long long cost = 0;
            for (int y = 0; y < py; ++y)
                for (int x = 0; x < px; ++x)
                    for (int r = 0; r < d->r; ++r)
                        for (int s = 0; s < d->s; ++s) {
                            int dy = (r - d->r / 2) * d->dil, dx = (s - d->s / 2) * d->dil;
                            if (d->stride == 2) {
                                dy = dy >= 0 ? dy / 2 : -((-dy + 1) / 2);
                                dx = dx >= 0 ? dx / 2 : -((-dx + 1) / 2);
                            }
                            const int y0 = y * bh, x0 = x * bw;
                            if (anchored || (y0 + dy < d->h) && (y0 + bh - 1 + dy >= 0) && (x0 + dx < d->w) && (x0 + bw - 1 + dx >= 0))
                                ++cost;
                        }
            cost *= pi;
            if (best < 0 || cost < best || (cost == best && bw > pl->bw)) {
                best = cost;
                pl->bw = bw;
                pl->bh = bh;
                pl->bi = bi;
                pl->patches_x = px;
                pl->patches_y = py;
                pl->patches_i = pi;
            }
        }
    }
    B200SEG_CHECK(best > 0, "wgrad: no patch shape");
    const int patches = pl->patches_x * pl->patches_y * pl->patches_i;
    pl->iters_run = best;
    pl->iters_dense = static_cast<long long>(patches) * d->r * d->s;
    // tap groups: with X on the N side and a single narrow n-tile, several taps share one CTA (and one load of dY)
    const int taps = d->r * d->s;
    pl->tap_group = 1;
    if (!pl->shift_m && pl->tiles_n == 1 && taps > 1 && pl->bn <= kMaxBN / 4 && tap_grouping_enabled()) {   // measured: pairs of 128-wide taps do not pay
        const int gmax = kMaxBN / pl->bn;
        const int ngroups = ceil_div(taps, gmax);
        pl->tap_group = ceil_div(taps, ngroups);
    }
    const int base_tiles = ceil_div(taps, pl->tap_group) * pl->tiles_m * pl->tiles_n;
    int sms = gemm_sm_count();
    if (sms <= 0) sms = 148;
    int want = sms / base_tiles;                    // splits that fill the GPU once
    if (want < 1) want = 1;
    const int max_splits = patches / 8 > 0 ? patches / 8 : 1;
    if (want > max_splits) want = max_splits;
    // Cluster size along the splits: clusters are placed inside one GPC, so fewer CTAs are co-resident the larger the
    // cluster (cluster_capacity asks the driver); take the largest cluster that keeps >= 80 % of the CTAs a plain launch
    // would have — the on-chip reduction is worth more than the last few SMs on these HBM-bound launches
    pl->cs = 1;
    pl->splits_hi = want;
    if (want >= 2 && cluster_reduce_enabled()) {
        static const int max_cs = [] {
            const char* e = getenv("B200SEG_WGRAD_MAXCS");
            const int v = e ? atoi(e) : 8;
            return v == 2 || v == 4 || v == 8 ? v : 8;
        }();
        for (int cs = max_cs; cs >= 2; cs /= 2) {
            if (cs > want) continue;
            const int cap = cluster_capacity(cs);                  // co-resident clusters of this size
            int hi = cap / base_tiles;
            if (hi > want / cs) hi = want / cs;
            if (hi < 1) continue;
            if (static_cast<long long>(base_tiles) * hi * cs * 10 >= static_cast<long long>(base_tiles) * want * 8) {
                pl->cs = cs;
                pl->splits_hi = hi;
                break;
            }
        }
    }
    int splits = pl->splits_hi * pl->cs;
    pl->pps = ceil_div(patches, splits);
    if (pl->cs == 1) {
        splits = ceil_div(patches, pl->pps);        // drop empty tail splits
        pl->splits_hi = splits;
    }
    pl->splits = splits;
    pl->direct = pl->splits_hi == 1 && direct_write_enabled() ? 1 : 0;
    pl->ws_bytes = pl->direct ? 256 : static_cast<size_t>(pl->splits_hi) * d->r * d->s * pl->m_pad * pl->n_pad * sizeof(float);
    return 0;
}

// The patch search walks every patch of the map: memoise the plan per geometry.
int plan_wgrad(const b200seg_wgrad_desc* d, WgradPlan* pl) {
    B200SEG_CHECK(d, "wgrad: null descriptor");
    static std::mutex mu;
    static std::map<std::vector<int>, WgradPlan> memo;
    const std::vector<int> key = {d->n, d->h, d->w, d->k, d->c, d->r, d->s, d->dil, d->stride, d->k_out, d->dy_c_offset % 8,
                                  d->dy_c_total % 8, d->x_c_offset % 8, d->x_c_total % 8, d->x_rows, d->x_cols,
                                  d->x_c_total, d->x_c_offset};
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = memo.find(key);
        if (it != memo.end() && d->dy && d->x && d->dw) {
            *pl = it->second;
            return 0;
        }
    }
    if (plan_wgrad_uncached(d, pl)) return 1;
    std::lock_guard<std::mutex> lock(mu);
    memo[key] = *pl;
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant (b200seg_conv_wgrad_noalloc): the split-K workspace lives in this file-scope device array
// instead of caller-allocated memory.  Every GPU that loads the library holds it until the process exits.  All calls
// share it and it keeps the last call's contents — harmless here: wgrad_kernel writes every (plane, tap, tile) of the
// workspace it declares before wgrad_reduce_kernel reads it, nothing has to start at zero — but two calls that overlap
// on the GPU (different streams, or graph branches) race on it.
static __device__ __align__(256) unsigned char g_wgrad_static_ws[B200SEG_NOALLOC_WGRAD_MAX_BYTES];

// The array's name is a host-side stand-in in host code; the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
constexpr int kMaxDevices = 64;
int static_ws_address(void** out) {
    static std::mutex mu;
    static void* addr[kMaxDevices] = {};
    int dev = -1;
    B200SEG_CUDA(cudaGetDevice(&dev));
    B200SEG_CHECK(dev >= 0 && dev < kMaxDevices, "conv_wgrad_noalloc: device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        B200SEG_CUDA(cudaGetSymbolAddress(&p, g_wgrad_static_ws));
        B200SEG_CHECK(p && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "conv_wgrad_noalloc: bad symbol address");
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_conv_wgrad_plan_info(const b200seg_wgrad_desc* d, int64_t* iters_run, int64_t* iters_dense) {
    b200seg::clear_error();
    b200seg::WgradPlan pl;
    memset(&pl, 0, sizeof(pl));
    if (b200seg::plan_wgrad(d, &pl)) return 1;
    if (iters_run) *iters_run = pl.iters_run;
    if (iters_dense) *iters_dense = pl.iters_dense;
    return 0;
}

size_t b200seg_conv_wgrad_workspace(const b200seg_wgrad_desc* d) {
    b200seg::WgradPlan pl;
    memset(&pl, 0, sizeof(pl));
    if (b200seg::plan_wgrad(d, &pl)) return 0;
    return pl.ws_bytes;
}

int b200seg_conv_wgrad(const b200seg_wgrad_desc* d, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    WgradPlan pl;
    memset(&pl, 0, sizeof(pl));
    if (plan_wgrad(d, &pl)) return 1;
    B200SEG_CHECK(d->workspace && d->workspace_bytes >= pl.ws_bytes, "wgrad: workspace of %zu bytes needed, %zu given",
                  pl.ws_bytes, d->workspace_bytes);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(d->workspace) & 15) == 0, "wgrad: workspace not 16-byte aligned");

    WgradParams p;
    memset(&p, 0, sizeof(p));
    p.N = d->n;
    p.H = d->h;
    p.W = d->w;
    p.R = d->r;
    p.S = d->s;
    p.shift_m = pl.shift_m;
    p.bw = pl.bw;
    p.bh = pl.bh;
    p.bi = pl.bi;
    p.patches_x = pl.patches_x;
    p.patches_y = pl.patches_y;
    p.patches = pl.patches_x * pl.patches_y * pl.patches_i;
    p.tiles_m = pl.tiles_m;
    p.tiles_n = pl.tiles_n;
    p.bn = pl.bn;
    p.tap_group = pl.tap_group;
    p.splits = pl.splits;
    p.patches_per_split = pl.pps;
    p.m_pad = pl.m_pad;
    p.n_pad = pl.n_pad;
    p.ws = static_cast<float*>(d->workspace);

    // n, h, w describe the dY grid; X is (stride*h) x (stride*w).  Tap (r, s) reads X pixel stride*o + (r - R/2)*dil:
    // for stride 2 its parity picks one of four strided views of X and floor(./2) is the shift inside that view.
    const uint32_t box[4] = {64, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bi};
    const int cs = d->stride;
    const int hx = d->h * cs, wx = d->w * cs;
    const uint64_t ydims[4] = {(uint64_t)d->k, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n};
    const uint64_t ystr[3] = {(uint64_t)d->dy_c_total * 2, (uint64_t)d->w * d->dy_c_total * 2,
                              (uint64_t)d->h * d->w * d->dy_c_total * 2};
    const char* ybase = static_cast<const char*>(d->dy) + (size_t)d->dy_c_offset * 2;
    if (encode_tmap_bf16(&p.y_map, ybase, 4, ydims, ystr, box)) return 1;
    const bool anchored = d->x_rows || d->x_cols;
    p.no_skip = anchored ? 1 : 0;
    if (anchored) {
        // window view: see launch_fprop in conv_igemm.cu
        const int span = ceil_div(d->x_c_offset + d->c, d->x_c_total);
        const uint64_t wdims[4] = {(uint64_t)d->c, (uint64_t)(d->x_cols - span + 1), (uint64_t)d->x_rows, (uint64_t)d->n};
        const uint64_t wstr[3] = {(uint64_t)d->x_c_total * 2, (uint64_t)d->x_cols * d->x_c_total * 2,
                                  (uint64_t)d->x_rows * d->x_cols * d->x_c_total * 2};
        const char* wbase = static_cast<const char*>(d->x) + (size_t)d->x_c_offset * 2;
        if (encode_tmap_bf16(&p.x_map[0], wbase, 4, wdims, wstr, box)) return 1;
    }
    const uint64_t xdims[4] = {(uint64_t)d->c, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n};
    const uint64_t xstr[3] = {(uint64_t)cs * d->x_c_total * 2, (uint64_t)cs * wx * d->x_c_total * 2,
                              (uint64_t)hx * wx * d->x_c_total * 2};
    for (int vy = 0; vy < (anchored ? 0 : cs); ++vy)
        for (int vx = 0; vx < cs; ++vx) {
            const char* xbase = static_cast<const char*>(d->x) + ((size_t)vy * wx + vx) * d->x_c_total * 2 +
                                (size_t)d->x_c_offset * 2;
            if (encode_tmap_bf16(&p.x_map[vy * 2 + vx], xbase, 4, xdims, xstr, box)) return 1;
        }
    // 5-D views: the channel axis split into (64, blocks); the block axis is the outermost box dimension, so one box is
    // the same run of consecutive 8 KB boxes in shared memory that the per-box loads produce
    static const bool tma5 = [] {
        const char* e = getenv("B200SEG_WGRAD_TMA5");
        return !(e && e[0] == '0');
    }();
    const int n_boxes_host = pl.bn / 64;
    p.y5 = tma5 && d->k % 64 == 0 ? 1 : 0;
    p.x5 = tma5 && d->c % 64 == 0 && !anchored ? 1 : 0;
    if (p.y5) {
        const uint32_t yb = pl.shift_m ? n_boxes_host : 2;       // dY is the N side when X is on the M side
        const uint32_t box5[5] = {64, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bi, yb};
        const uint64_t dims5[5] = {64, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n, (uint64_t)d->k / 64};
        const uint64_t str5[4] = {ystr[0], ystr[1], ystr[2], 128};
        if (encode_tmap_bf16(&p.y_map5, ybase, 5, dims5, str5, box5)) return 1;
    }
    if (p.x5) {
        const uint32_t xb = pl.shift_m ? 2 : n_boxes_host;
        const uint32_t box5[5] = {64, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bi, xb};
        const uint64_t dims5[5] = {64, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n, (uint64_t)d->c / 64};
        const uint64_t str5[4] = {xstr[0], xstr[1], xstr[2], 128};
        for (int vy = 0; vy < cs; ++vy)
            for (int vx = 0; vx < cs; ++vx) {
                const char* xbase = static_cast<const char*>(d->x) + ((size_t)vy * wx + vx) * d->x_c_total * 2 +
                                    (size_t)d->x_c_offset * 2;
                if (encode_tmap_bf16(&p.x_map5[vy * 2 + vx], xbase, 5, dims5, str5, box5)) return 1;
            }
    }
    for (int r = 0; r < d->r; ++r)
        for (int s = 0; s < d->s; ++s) {
            const int oy = (r - d->r / 2) * d->dil, ox = (s - d->s / 2) * d->dil;
            const int t = r * d->s + s;
            if (anchored) {
                p.tap_map[t] = 0;
                p.tap_dy[t] = static_cast<short>(r);
                p.tap_dx[t] = static_cast<short>(s);
            } else if (cs == 1) {
                p.tap_map[t] = 0;
                p.tap_dy[t] = static_cast<short>(oy);
                p.tap_dx[t] = static_cast<short>(ox);
            } else {
                p.tap_map[t] = static_cast<short>((oy & 1) * 2 + (ox & 1));
                p.tap_dy[t] = static_cast<short>(oy >= 0 ? oy / 2 : -((-oy + 1) / 2));
                p.tap_dx[t] = static_cast<short>(ox >= 0 ? ox / 2 : -((-ox + 1) / 2));
            }
        }

    set_wgrad_attrs();
    p.cs = pl.cs;
    p.direct = pl.direct;
    p.dw = static_cast<float*>(d->dw);
    p.k_out = d->k_out;
    p.c_in = d->c;
    p.accumulate = d->accumulate;
    p.splits = pl.splits;
    const int taps = d->r * d->s;
    const int grid = pl.splits * ceil_div(taps, pl.tap_group) * pl.tiles_m * pl.tiles_n;
    if (pl.tap_group > 1)
        B200SEG_CUDA(launch_pdl_cluster(wgrad_kernel<true>, grid, kThreads, kSmemBytes, st, pl.cs, p));
    else
        B200SEG_CUDA(launch_pdl_cluster(wgrad_kernel<false>, grid, kThreads, kSmemBytes, st, pl.cs, p));
    if (pl.direct) return 0;

    const long long total = static_cast<long long>(d->k_out) * d->c * taps;
    const int rb = 256;
    B200SEG_CUDA(launch_pdl(wgrad_reduce_kernel, static_cast<unsigned>((total + rb - 1) / rb), rb, 0, st,
        p.ws, d->dw, d->k_out, d->c, taps, pl.splits_hi, pl.m_pad, pl.n_pad, pl.shift_m, d->accumulate));
    return 0;
}

int b200seg_conv_wgrad_noalloc(const b200seg_wgrad_desc* d, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(d, "conv_wgrad_noalloc: null descriptor");
    WgradPlan pl;
    memset(&pl, 0, sizeof(pl));
    if (plan_wgrad(d, &pl)) return 1;
    B200SEG_CHECK(pl.ws_bytes <= static_cast<size_t>(B200SEG_NOALLOC_WGRAD_MAX_BYTES),
                  "conv_wgrad_noalloc: this weight gradient needs %zu bytes of split-K scratch, the static scratch holds "
                  "%zu (B200SEG_NOALLOC_WGRAD_MAX_BYTES); use b200seg_conv_wgrad with a caller-owned workspace",
                  pl.ws_bytes, static_cast<size_t>(B200SEG_NOALLOC_WGRAD_MAX_BYTES));
    void* ws = nullptr;
    if (static_ws_address(&ws)) return 1;
    b200seg_wgrad_desc dd = *d;                 // the caller's workspace fields are ignored
    dd.workspace = ws;
    dd.workspace_bytes = static_cast<size_t>(B200SEG_NOALLOC_WGRAD_MAX_BYTES);
    return b200seg_conv_wgrad(&dd, stream);
}

}  // extern "C"
