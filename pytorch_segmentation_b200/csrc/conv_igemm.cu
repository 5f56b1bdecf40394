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
#include "pdl.cuh"
#include "ptx.cuh"

#include <string.h>

#include <stdlib.h>

#include <array>
#include <atomic>
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

constexpr int kMaxSeg = 8;
constexpr int kMaxTaps = 9;

// Division of a tile index (< 2^31) by a launch constant as multiply-high + shift (Granlund-Montgomery): the three
// single-thread / per-warp control loops each decode every tile, and four run-time divisions (~100 clk apiece when
// dependent) were a quarter of the per-tile bookkeeping on the many-small-tile launches.
struct FastDiv {
    unsigned mul, shift;
};
inline FastDiv make_fastdiv(int d) {
    FastDiv f;
    unsigned l = 0;
    while ((1u << l) < static_cast<unsigned>(d)) ++l;
    f.shift = l;
    f.mul = static_cast<unsigned>(((static_cast<unsigned long long>(1) << 32) * ((1ull << l) - d)) / d + 1);
    return f;
}
__device__ __forceinline__ int fast_div(int n, const FastDiv& f) {
    return static_cast<int>((__umulhi(f.mul, static_cast<unsigned>(n)) + static_cast<unsigned>(n)) >> f.shift);
}

// One reduction segment: an activation view (tensor map), a packed weight, and the filter taps that read this view
// — each tap is a pixel shift (dy, dx) in the view's grid plus the index of its weight slab.  A stride-1 conv is one
// segment with all R*S taps (shift = (r - R/2) * dilation); a stride-2 conv is one segment per input parity
// sub-grid (a strided view of the input) with the taps that land on it.
struct IgemmSeg {
    int ntaps, cblks;
    short dy[kMaxTaps], dx[kMaxTaps], wtap[kMaxTaps];
    int packed[kMaxTaps];        // dy (low 12 bits, signed) | dx << 12 (12 bits, signed) | wtap << 24: one load per tap in the producer loop
};

struct __align__(64) IgemmParams {
    CUtensorMap a_map[kMaxSeg];
    CUtensorMap b_map[kMaxSeg];
    CUtensorMap y_map;       // bf16 output view (channels, W, H, N), box = 64 channels x pixel patch
    IgemmSeg seg[kMaxSeg];
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
    int no_skip;             // anchored window sources: the buffer carries its own padding, every tap is in bounds
    int* sched;              // dynamic tile scheduler: {next tile - gridDim.x, CTAs done}; nullptr = static striding
    int n_fast;              // A/B switch (B200SEG_IGEMM_NFAST=1, static striding only): n_tile as the fastest tile index
    int dy_min, dy_max, dx_min, dx_max, iters_all;   // extremes of the tap shifts; pipeline iterations of a tile none of whose taps is all padding
    // Operand-role swap for launches with <= 64 output channels (layer1, the stem): D^T[k][pixel] = W[k][c] . X^T[c][pixel] —
    // the weights are the M operand (64 real rows of 128, the rest zero-filled by the TMA unit), a 256-pixel patch the N
    // operand, so one issued MMA does twice the useful work of the pixel-major form and the single-thread issue loops stop
    // being the bound.  The accumulator is channel-per-lane, pixel-per-column; the epilogue transposes it through the
    // staging slabs (one 128-pixel slab per half of the patch).  sub_dim / sub_len: the patch dimension (0 x, 1 y, 2 n)
    // along which the patch splits into its two 128-pixel halves, and the half's extent there.
    int swap, sub_dim, sub_len;
    FastDiv d_mt, d_tx, d_ty, d_tn;   // divisions by m_tiles, tiles_x, tiles_y, tiles_n in decode_tile
    int kpack;               // (tap, 64-channel block) operand pairs per pipeline stage: 2 when two fit a stage buffer
                             // (bn <= 64), else 1 — see the producer loop
};

long long* g_dbg = nullptr;
int g_dbg_slots = 0;

// Pipeline iterations ((tap, 64-channel block, tile) triples) of the launches made by the last conv_fprop / conv_dgrad
// call of this thread: those actually run, and what the same tiling would run with no all-padding tap skipped.  Their
// ratio turns nominal FLOPs (padding taps counted) into executed FLOPs (bench.py, roofline.frac_executed).
thread_local long long g_iters_run = 0, g_iters_dense = 0;

__device__ __forceinline__ long long gtime() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

struct TileCoord {
    int n_tile, x0, y0, n0;
};

__device__ __forceinline__ TileCoord decode_tile(const IgemmParams& p, int tile) {
    // n_tile is the SLOWEST index: a CTA's consecutive tiles (static stride or dynamic) then share their output-channel
    // block, so the per-channel BN statistics it carries in registers are flushed once per block, not once per tile
    TileCoord t;
    const int m_tiles = p.tiles_x * p.tiles_y * p.tiles_i;
    int m;
    if (p.n_fast) {
        m = fast_div(tile, p.d_tn);
        t.n_tile = tile - m * p.tiles_n;
    } else {
        t.n_tile = fast_div(tile, p.d_mt);
        m = tile - t.n_tile * m_tiles;
    }
    const int my = fast_div(m, p.d_tx);
    t.x0 = (m - my * p.tiles_x) * p.bw;
    const int mn = fast_div(my, p.d_ty);
    t.y0 = (my - mn * p.tiles_y) * p.bh;
    t.n0 = mn * p.bi;
    return t;
}

__device__ __forceinline__ bool tap_in_bounds(const IgemmParams& p, const TileCoord& t, int dy, int dx) {
    if (p.no_skip) return true;
    return (t.y0 + dy < p.H) && (t.y0 + p.bh - 1 + dy >= 0) && (t.x0 + dx < p.W) && (t.x0 + p.bw - 1 + dx >= 0);
}

// every tap of the tile reads at least one pixel inside the image (the common case away from large dilations): the
// control loops then skip the per-tap test
__device__ __forceinline__ bool all_taps_in_bounds(const IgemmParams& p, const TileCoord& t) {
    if (p.no_skip) return true;
    return (t.y0 + p.dy_max < p.H) && (t.y0 + p.bh - 1 + p.dy_min >= 0) && (t.x0 + p.dx_max < p.W) &&
           (t.x0 + p.bw - 1 + p.dx_min >= 0);
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
    uint64_t* sfull = bars + 2 * kStages + 4;    // [2] scheduler (the producer thread) -> MMA + epilogue: a tile id is posted
    uint64_t* sempty = bars + 2 * kStages + 6;   // [2] MMA + epilogue warps -> scheduler: the slot was read
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 8);
    volatile int* sched_tile = reinterpret_cast<volatile int*>(tmem_slot + 1);      // [2] posted tile ids (-1 = no more)
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
            mbar_init(&sfull[i], 1);
            mbar_init(&sempty[i], p.out_fp32 ? 5 : 9);      // the MMA thread + one lane per epilogue warp
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
    // PDL: everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped the previous kernel's
    // tail; no global memory is touched before this point
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            // ------------------------------------------------------------ TMA producer
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx_bytes = kABytes + (p.swap ? kMaxBN : p.bn) * kBK * 2;
            // This thread is also the CTA's TILE SCHEDULER.  The first tile is blockIdx.x; every later one comes from a
            // global counter (atomicAdd) instead of a static stride, so a CTA that starts late or runs on a shared SM
            // (weight-gradient kernels on the side stream, NCCL under data parallelism) simply takes fewer tiles instead
            // of holding the launch's tail.  Tile ids reach the MMA and epilogue warps through a two-slot mailbox.
            int slot = 0;
            uint32_t sphase = 0;
            int tile = blockIdx.x;
            for (;;) {
                mbar_wait(&sempty[slot], sphase ^ 1);
                sched_tile[slot] = tile < p.num_tiles ? tile : -1;
                mbar_arrive(&sfull[slot]);
                if (++slot == 2) {
                    slot = 0;
                    sphase ^= 1;
                }
                if (tile >= p.num_tiles) break;
                // the next id is requested before this tile's loads are issued: the atomic's latency hides under them
                const int next = p.sched ? atomicAdd(p.sched, 1) + static_cast<int>(gridDim.x) : tile + static_cast<int>(gridDim.x);
                const TileCoord t = decode_tile(p, tile);
                // With a narrow N tile the tensor work of one (tap, 64-channel block) pair is 128 clk while the two
                // single-thread control loops (this one and the MMA issuer) need 300-750 clk of issue latency per
                // pipeline stage (try_wait ~90 clk even when complete, arrive, TMA / MMA issues, tap bookkeeping:
                // profiles/r02_igemm_ablation.md): a stage therefore carries p.kpack pairs when they fit its buffer —
                // one wait, one expect_tx and one commit for both.  expect_tx is posted after the loads it covers
                // (the phase cannot complete before this thread's arrive, the tx count may run negative meanwhile).
                int sub = 0;
                const bool all_in = all_taps_in_bounds(p, t);
                for (int sg = 0; sg < p.nseg; ++sg) {
                    const IgemmSeg& seg = p.seg[sg];
                    for (int tp = 0; tp < seg.ntaps; ++tp) {
                        const int pk = seg.packed[tp];
                        const int dy = (pk << 20) >> 20, dx = (pk << 8) >> 20;
                        if (!all_in && !tap_in_bounds(p, t, dy, dx)) continue;
                        const int tap = pk >> 24;
                        for (int cb = 0; cb < seg.cblks; ++cb) {
                            if (sub == 0) mbar_wait(&empty[stage], phase ^ 1);
                            uint8_t* sa = smem + stage * kStageBytes + sub * tx_bytes;
                            if (p.swap) {       // M operand = 128 weight rows, N operand = the 256-pixel patch
                                tma_load_3d(sa, &p.b_map[sg], &full[stage], cb * kBK, 0, tap);
                                tma_load_4d(sa + kABytes, &p.a_map[sg], &full[stage], cb * kBK, t.x0 + dx, t.y0 + dy, t.n0);
                            } else {
                                tma_load_4d(sa, &p.a_map[sg], &full[stage], cb * kBK, t.x0 + dx, t.y0 + dy, t.n0);
                                tma_load_3d(sa + kABytes, &p.b_map[sg], &full[stage], cb * kBK, t.n_tile * p.bn, tap);
                            }
                            if (++sub == p.kpack) {
                                mbar_arrive_expect_tx(&full[stage], sub * tx_bytes);
                                sub = 0;
                                if (++stage == kStages) {
                                    stage = 0;
                                    phase ^= 1;
                                }
                            }
                        }
                    }
                }
                if (sub) {                                   // odd tail of the tile: a stage with one pair
                    mbar_arrive_expect_tx(&full[stage], sub * tx_bytes);
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                tile = next;
            }
            if (p.sched) {
                // the last CTA to run out of tiles leaves the counters zeroed for the next launch that uses this slot
                __threadfence();
                if (atomicAdd(p.sched + 1, 1) == static_cast<int>(gridDim.x) - 1) {
                    p.sched[0] = 0;
                    p.sched[1] = 0;
                    __threadfence();
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ------------------------------------------------------------ MMA issuer
            const uint32_t idesc = umma_idesc_bf16(kBM, p.swap ? kMaxBN : p.bn, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            int lt = 0;
            int slot = 0;
            uint32_t sphase = 0;
            for (;; ++lt) {
                mbar_wait(&sfull[slot], sphase);
                const int tile = sched_tile[slot];
                mbar_arrive(&sempty[slot]);
                if (++slot == 2) {
                    slot = 0;
                    sphase ^= 1;
                }
                if (tile < 0) break;
                const TileCoord t = decode_tile(p, tile);
                int kiters = p.iters_all;
                if (!all_taps_in_bounds(p, t)) {
                    kiters = 0;
                    for (int sg = 0; sg < p.nseg; ++sg) {
                        const IgemmSeg& seg = p.seg[sg];
                        for (int tp = 0; tp < seg.ntaps; ++tp)
                            if (tap_in_bounds(p, t, seg.dy[tp], seg.dx[tp])) kiters += seg.cblks;
                    }
                }
                const int acc = lt & 1;
                const uint32_t acc_phase = (lt >> 1) & 1;
                mbar_wait(&tempty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * kMaxBN;
                const uint32_t pair_bytes = kABytes + (p.swap ? kMaxBN : p.bn) * kBK * 2;
                for (int it = 0; it < kiters; it += p.kpack) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const int pairs = min(p.kpack, kiters - it);      // operand pairs this stage holds
                    for (int j = 0; j < pairs; ++j) {
                        const uint32_t sa = smem_u32(smem + stage * kStageBytes) + j * pair_bytes;
                        const uint64_t a_desc = umma_smem_desc(sa, 16, 1024);
                        const uint64_t b_desc = umma_smem_desc(sa + kABytes, 16, 1024);
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k) {
                            // +32 bytes (16 bf16) along K inside the 128-byte swizzle row: +2 in the address field
                            tc_mma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (it | j | k) != 0);
                        }
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
        if (!p.out_fp32 && p.swap) {
            // ---- operand-role swap: TMEM lane = output channel, column = pixel of the 256-pixel patch.  Warps with
            // q = 0 / 1 own channels 0-31 / 32-63; the warp pair (q = 0, 1) of each half h = 0 / 1 (warps 4, 5 / 8, 9)
            // transposes pixels [128 h, 128 h + 128) into staging slab h and one of its lanes stores it; warps with
            // q = 2 / 3 hold zero rows and only take part in the handshakes.
            const int h = (warp - 4) >> 2;
            const bool active = q < 2;
            const int ch = q * 32 + lane;                       // this thread's output channel
            const bool issuer = active && q == 0 && lane == 0;
            float sc = 1.f, bi = 0.f;
            if (active && ch < p.k_store) {
                if (p.scale) sc = __ldg(p.scale + ch);
                if (p.bias) bi = __ldg(p.bias + ch);
            }
            const bool affine = p.scale != nullptr || p.bias != nullptr || p.relu != 0;
            float s1 = 0.f, s2 = 0.f;
            uint8_t* buf = out_buf + h * kOutBufBytes;
            // byte offset of this channel inside a 128-byte pixel row, for each of the eight swizzle phases (row & 7)
            const int chunk = ch >> 3, within = (ch & 7) * 2;
            int sw_off[8];
#pragma unroll
            for (int m = 0; m < 8; ++m) sw_off[m] = ((chunk ^ m) << 4) | within;
            const bool want_stats = p.stats != nullptr;
            int lt = 0;
            int slot = 0;
            uint32_t sphase = 0;
            for (;; ++lt) {
                mbar_wait(&sfull[slot], sphase);
                const int tile = sched_tile[slot];
                __syncwarp();
                if (lane == 0) mbar_arrive(&sempty[slot]);
                if (++slot == 2) {
                    slot = 0;
                    sphase ^= 1;
                }
                if (tile < 0) break;
                const TileCoord t = decode_tile(p, tile);
                const int acc = lt & 1;
                const uint32_t acc_phase = (lt >> 1) & 1;
                const bool border = (t.x0 + p.bw > p.W) || (t.y0 + p.bh > p.H) || (t.n0 + p.bi > p.N);
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                if (active) {
                    // the slab is free once the store issued for the previous tile has read it
                    if (issuer) tma_store_wait_read<0>();
                    named_bar_sync(3 + h, 64);
                    const uint32_t taddr = tmem_base + acc * kMaxBN + (static_cast<uint32_t>(q * 32) << 16) + h * 128;
#pragma unroll 1
                    for (int j = 0; j < 4; ++j) {
                        uint32_t v[32];
                        tmem_ld_32x32(taddr + j * 32, v);
                        tmem_ld_wait();
                        uint8_t* rows = buf + j * 32 * 128;
                        if (affine || border) {                                // the uncommon forms, element by element
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                float f = __uint_as_float(v[i]);
                                if (affine) {
                                    f = fmaf(f, sc, bi);
                                    if (p.relu) f = fmaxf(f, 0.f);
                                }
                                if (border) {
                                    const int pr = h * 128 + j * 32 + i;       // pixel inside the patch
                                    const int wl2 = pr & (p.bw - 1), hl2 = (pr >> p.lbw) & (p.bh - 1), il2 = pr >> (p.lbw + p.lbh);
                                    if (!((t.x0 + wl2 < p.W) && (t.y0 + hl2 < p.H) && (t.n0 + il2 < p.N))) f = 0.f;
                                }
                                v[i] = __float_as_uint(f);
                            }
                        }
                        // two pixels per conversion; 16-bit stores at precomputed swizzled offsets (i & 7 is static)
#pragma unroll
                        for (int i = 0; i < 32; i += 2) {
                            const uint32_t pk = pack_bf16x2(__uint_as_float(v[i]), __uint_as_float(v[i + 1]));
                            *reinterpret_cast<uint16_t*>(rows + i * 128 + sw_off[i & 7]) = static_cast<uint16_t>(pk & 0xffffu);
                            *reinterpret_cast<uint16_t*>(rows + (i + 1) * 128 + sw_off[(i + 1) & 7]) = static_cast<uint16_t>(pk >> 16);
                            if (want_stats) {                                  // of the values as stored
                                const float f0 = __uint_as_float(pk << 16), f1 = __uint_as_float(pk & 0xffff0000u);
                                s1 += f0 + f1;
                                s2 = fmaf(f0, f0, fmaf(f1, f1, s2));
                            }
                        }
                    }
                    fence_proxy_async();
                    named_bar_sync(3 + h, 64);
                    if (issuer) {
                        const int off = h * p.sub_len;
                        tma_store_4d(&p.y_map, buf, 0, t.x0 + (p.sub_dim == 0 ? off : 0), t.y0 + (p.sub_dim == 1 ? off : 0),
                                     t.n0 + (p.sub_dim == 2 ? off : 0));
                        tma_store_commit();
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[acc]);
            }
            if (p.stats && active && ch < p.k_store) {
                atomicAdd(p.stats + ch, s1);
                atomicAdd(p.stats + p.k_store + ch, s2);
            }
            if (issuer) tma_store_wait_all<0>();
        } else if (!p.out_fp32) {
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
            int slot = 0;
            uint32_t sphase = 0;
            for (;; ++lt) {
                mbar_wait(&sfull[slot], sphase);
                const int tile = sched_tile[slot];
                __syncwarp();
                if (lane == 0) mbar_arrive(&sempty[slot]);
                if (++slot == 2) {
                    slot = 0;
                    sphase ^= 1;
                }
                if (tile < 0) break;
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
            int slot = 0;
            uint32_t sphase = 0;
            for (;; ++lt) {
                mbar_wait(&sfull[slot], sphase);
                const int tile = sched_tile[slot];
                __syncwarp();
                if (lane == 0) mbar_arrive(&sempty[slot]);
                if (++slot == 2) {
                    slot = 0;
                    sphase ^= 1;
                }
                if (tile < 0) break;
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

// Dynamic tile scheduler state: a ring of {tile counter, done counter} pairs, one pair per launch (a launch leaves its
// pair zeroed; consecutive launches, which programmatic dependent launch lets overlap, never share one).  Allocated on
// first use — which must not be inside a stream capture (trainer.TrainStep runs eager steps before it captures).
constexpr int kSchedSlots = 2048;
int* sched_slot() {
    static int* ring = nullptr;
    static std::once_flag once;
    static std::atomic<unsigned> seq{0};
    static const bool on = [] {
        const char* e = getenv("B200SEG_IGEMM_DYNAMIC");
        return !(e && e[0] == '0');
    }();
    if (!on) return nullptr;
    std::call_once(once, [] {
        int* p = nullptr;
        if (cudaMalloc(&p, sizeof(int) * 2 * kSchedSlots) == cudaSuccess && cudaMemset(p, 0, sizeof(int) * 2 * kSchedSlots) == cudaSuccess) {
            cudaDeviceSynchronize();
            ring = p;
        } else {
            cudaGetLastError();
        }
    });
    if (!ring) return nullptr;
    return ring + 2 * (seq.fetch_add(1) % kSchedSlots);
}

int ilog2(int v) {
    int l = 0;
    while ((1 << l) < v) ++l;
    return l;
}

// Host-side description of one segment before its tensor maps are encoded.
struct SegPlan {
    const char* a_base;        // first element of the activation view
    uint64_t adims[4];         // (channels, W, H, N) of the view
    uint64_t astr[3];          // byte strides of W, H, N
    const void* w;             // packed weights [taps][w_rows][c]
    int w_rows, c, w_taps;
    IgemmSeg seg;
};

// Output view of one launch: (channels, W, H, N) with byte strides (a parity sub-grid of dX is a strided view).
struct OutPlan {
    char* base;
    uint64_t dims[4];
    uint64_t str[3];
};

// Total pipeline iterations (TMA box pairs) the whole problem costs with a bw x bh x bi patch.
long long patch_cost(int N, int H, int W, int bw, int bh, int bi, int nseg, const SegPlan* sp, bool no_skip) {
    const int tx = ceil_div(W, bw), ty = ceil_div(H, bh), ti = ceil_div(N, bi);
    long long per_image = 0;
    for (int g = 0; g < nseg; ++g) {
        const IgemmSeg& seg = sp[g].seg;
        for (int tp = 0; tp < seg.ntaps; ++tp) {
            if (no_skip) {
                per_image += static_cast<long long>(ty) * tx * seg.cblks;
                continue;
            }
            const int dy = seg.dy[tp], dx = seg.dx[tp];
            // the in-bounds test is separable: count tile rows and tile columns that touch the image
            int ny = 0, nx = 0;
            for (int y = 0; y < ty; ++y) ny += (y * bh + dy < H) && (y * bh + bh - 1 + dy >= 0);
            for (int x = 0; x < tx; ++x) nx += (x * bw + dx < W) && (x * bw + bw - 1 + dx >= 0);
            per_image += static_cast<long long>(ny) * nx * seg.cblks;
        }
    }
    return per_image * ti;
}

// Encodes the maps and launches the persistent kernel over an n x h x w output grid.
int launch_planned(int n, int h, int w, int nseg, const SegPlan* sp, const OutPlan& op, const b200seg_conv_out* out,
                   cudaStream_t st, bool no_skip = false) {
    IgemmParams p;
    memset(&p, 0, sizeof(p));
    p.no_skip = no_skip ? 1 : 0;
    p.nseg = nseg;
    p.N = n;
    p.H = h;
    p.W = w;
    bool has_center = false;
    for (int i = 0; i < nseg; ++i) {
        p.seg[i] = sp[i].seg;
        for (int tp = 0; tp < sp[i].seg.ntaps; ++tp) has_center |= sp[i].seg.dy[tp] == 0 && sp[i].seg.dx[tp] == 0;
    }
    B200SEG_CHECK(has_center, "conv: internal: launch without an unshifted tap");

    // output-channel tiling: whole 64-channel store slabs per n-tile (a narrow single tile may be 32 wide: the
    // store clips at k)
    p.tiles_n = ceil_div(out->k, kMaxBN);
    p.bn = out->k <= 32 ? 32 : round_up(ceil_div(out->k, p.tiles_n), kOutSlab);
    B200SEG_CHECK(p.bn <= kMaxBN, "conv: internal bn %d", p.bn);

    // operand-role swap (see IgemmParams::swap): narrow outputs over enough pixels to fill the GPU with 256-pixel patches
    static const int swap_mode = [] {               // B200SEG_IGEMM_SWAP: 0 off, 1 (default) where it pays, 2 wherever legal (tests)
        const char* e = getenv("B200SEG_IGEMM_SWAP");
        return e ? atoi(e) : 1;
    }();
    {
        int sms = gemm_sm_count();
        if (sms <= 0) sms = 148;
        // "where it pays": the transposing epilogue (four warps, 16-bit shared-memory stores) has to hide under the
        // reduction of the patch — at least four (tap, 64-channel block) pairs (measured: 3x3 64->64 66.6 -> 50.2 us,
        // its dgrad 64.5 -> 46.3 us, 1x1 256->64 37.9 -> 35.8 us; B200SEG_IGEMM_SWAP_DEPTH)
        static const int min_depth = [] {
            const char* e = getenv("B200SEG_IGEMM_SWAP_DEPTH");
            return e ? atoi(e) : 4;
        }();
        int depth = 0;
        for (int i = 0; i < nseg; ++i) depth += p.seg[i].ntaps * p.seg[i].cblks;
        p.swap = (swap_mode > 0 && out->k <= 64 && !out->out_fp32 && p.tiles_n == 1 &&
                  (swap_mode > 1 || (depth >= min_depth && static_cast<long long>(n) * h * w >= 256LL * sms))) ? 1 : 0;
    }
    const int patch_px = p.swap ? 256 : 128;

    // pixel patch: the power-of-two bw x bh x bi = 128 shape that needs the fewest pipeline iterations (memoised
    // per geometry: a training step repeats the same few)
    long long best = -1;
    {
        static std::mutex mu;
        static std::map<std::vector<int>, std::array<long long, 5>> memo;
        std::vector<int> key = {n, h, w, nseg, p.no_skip, patch_px};
        for (int i = 0; i < nseg; ++i) {
            key.push_back(p.seg[i].ntaps);
            key.push_back(p.seg[i].cblks);
            for (int tp = 0; tp < p.seg[i].ntaps; ++tp) key.push_back(p.seg[i].dy[tp] * 65536 + p.seg[i].dx[tp]);
        }
        std::lock_guard<std::mutex> lock(mu);
        auto it = memo.find(key);
        if (it != memo.end()) {
            p.bw = static_cast<int>(it->second[0]);
            p.bh = static_cast<int>(it->second[1]);
            p.bi = static_cast<int>(it->second[2]);
            best = 1;
            g_iters_run += it->second[3];
            g_iters_dense += it->second[4];
        } else {
            for (int bi = 1; bi <= patch_px; bi *= 2) {
                for (int bw = 1; bw * bi <= patch_px; bw *= 2) {
                    const int bh = patch_px / (bw * bi);
                    const long long c = patch_cost(n, h, w, bw, bh, bi, nseg, sp, no_skip);
                    if (best < 0 || c < best || (c == best && bw > p.bw)) {
                        best = c;
                        p.bw = bw;
                        p.bh = bh;
                        p.bi = bi;
                    }
                }
            }
            if (best > 0) {
                const long long dense = patch_cost(n, h, w, p.bw, p.bh, p.bi, nseg, sp, true);
                memo[key] = {p.bw, p.bh, p.bi, best, dense};
                g_iters_run += best;
                g_iters_dense += dense;
            }
        }
    }
    B200SEG_CHECK(best > 0, "conv: no tile shape for %dx%dx%d", n, h, w);
    p.lbw = ilog2(p.bw);
    p.lbh = ilog2(p.bh);
    p.tiles_x = ceil_div(w, p.bw);
    p.tiles_y = ceil_div(h, p.bh);
    p.tiles_i = ceil_div(n, p.bi);
    p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_i * p.tiles_n;
    p.dy_min = p.dx_min = 1 << 20;
    p.dy_max = p.dx_max = -(1 << 20);
    p.iters_all = 0;
    for (int i = 0; i < nseg; ++i) {
        IgemmSeg& sgm = p.seg[i];
        p.iters_all += sgm.ntaps * sgm.cblks;
        for (int tp = 0; tp < sgm.ntaps; ++tp) {
            const int dy = sgm.dy[tp], dx = sgm.dx[tp];
            B200SEG_CHECK(dy > -2048 && dy < 2048 && dx > -2048 && dx < 2048 && sgm.wtap[tp] >= 0 && sgm.wtap[tp] < 128,
                          "conv: tap shift (%d, %d) out of the packed range", dy, dx);
            sgm.packed[tp] = (dy & 0xfff) | ((dx & 0xfff) << 12) | (static_cast<int>(sgm.wtap[tp]) << 24);
            p.dy_min = dy < p.dy_min ? dy : p.dy_min;
            p.dy_max = dy > p.dy_max ? dy : p.dy_max;
            p.dx_min = dx < p.dx_min ? dx : p.dx_min;
            p.dx_max = dx > p.dx_max ? dx : p.dx_max;
        }
    }
    p.d_mt = make_fastdiv(p.tiles_x * p.tiles_y * p.tiles_i);
    p.d_tx = make_fastdiv(p.tiles_x);
    p.d_ty = make_fastdiv(p.tiles_y);
    p.d_tn = make_fastdiv(p.tiles_n);

    for (int i = 0; i < nseg; ++i) {
        const uint32_t abox[4] = {kBK, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bi};
        if (encode_tmap_bf16(&p.a_map[i], sp[i].a_base, 4, sp[i].adims, sp[i].astr, abox)) return 1;
        const uint64_t bdims[3] = {(uint64_t)sp[i].c, (uint64_t)sp[i].w_rows, (uint64_t)sp[i].w_taps};
        const uint64_t bstr[2] = {(uint64_t)sp[i].c * 2, (uint64_t)sp[i].w_rows * sp[i].c * 2};
        const uint32_t bbox[3] = {kBK, (uint32_t)(p.swap ? kBM : p.bn), 1};      // swap: the weights are the M operand
        if (encode_tmap_bf16(&p.b_map[i], sp[i].w, 3, bdims, bstr, bbox)) return 1;
    }
    if (p.swap) {
        // the two 128-pixel halves of the patch: split along its slowest dimension that has more than one element
        uint32_t sb[3] = {(uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bi};
        p.sub_dim = p.bi >= 2 ? 2 : (p.bh >= 2 ? 1 : 0);
        sb[p.sub_dim] /= 2;
        p.sub_len = static_cast<int>(sb[p.sub_dim]);
        const uint32_t ybox[4] = {kOutSlab, sb[0], sb[1], sb[2]};
        if (encode_tmap_bf16(&p.y_map, op.base, 4, op.dims, op.str, ybox)) return 1;
    } else if (!out->out_fp32) {
        const uint32_t ybox[4] = {kOutSlab, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bi};
        if (encode_tmap_bf16(&p.y_map, op.base, 4, op.dims, op.str, ybox)) return 1;
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

    static bool attr_set = false;
    if (!attr_set) {
        B200SEG_CUDA(cudaFuncSetAttribute(igemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    const int sms = gemm_sm_count();
    B200SEG_CHECK(sms > 0, "conv: cannot query SM count");
    const int grid = p.num_tiles < sms ? p.num_tiles : sms;
    p.sched = p.num_tiles > grid ? sched_slot() : nullptr;       // a single wave has nothing to schedule
    static const bool n_fast = [] {
        const char* e = getenv("B200SEG_IGEMM_NFAST");
        return e && e[0] == '1';
    }();
    p.n_fast = n_fast && !p.sched ? 1 : 0;
    static const bool kpack_on = [] {
        const char* e = getenv("B200SEG_IGEMM_KPACK");
        return !(e && e[0] == '0');
    }();
    p.kpack = (kpack_on && !p.swap && 2 * (kABytes + p.bn * kBK * 2) <= kStageBytes) ? 2 : 1;
    B200SEG_CUDA(launch_pdl(igemm_kernel, grid, kThreads, kSmemBytes, st, p));
    return 0;
}

int check_common(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out) {
    B200SEG_CHECK(n > 0 && h > 0 && w > 0, "conv: bad activation shape %dx%dx%d", n, h, w);
    B200SEG_CHECK(n_src >= 1 && n_src <= B200SEG_MAX_CONV_SRC, "conv: n_src %d out of range", n_src);
    B200SEG_CHECK(out && out->y, "conv: null output");
    B200SEG_CHECK(out->k > 0 && out->k % 8 == 0 && out->y_c_offset % 8 == 0 && out->y_c_total % 8 == 0,
                  "conv: output channels (k=%d, offset=%d, total=%d) must be multiples of 8", out->k,
                  out->y_c_offset, out->y_c_total);
    B200SEG_CHECK(out->y_c_offset + out->k <= out->y_c_total, "conv: output slice exceeds buffer channels");
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(out->y) & 15) == 0, "conv: output pointer not 16-byte aligned");
    B200SEG_CHECK(!(out->stats && (out->scale || out->bias || out->relu || out->out_fp32)),
                  "conv: fused BN statistics are those of the stored bf16 conv result; scale/bias/relu/fp32 output "
                  "cannot be combined with them");
    for (int i = 0; i < n_src; ++i) {
        const b200seg_conv_src& s = src[i];
        B200SEG_CHECK(s.x && s.w, "conv: source %d has null pointers", i);
        if (s.x_rows || s.x_cols) {
            // anchored window source (include/b200seg.h): geometry checked here, the regular rules do not apply
            B200SEG_CHECK(n_src == 1, "conv: an anchored window source must be the only source");
            B200SEG_CHECK(s.r >= 1 && s.s >= 1 && s.r * s.s <= kMaxTaps && s.dil == 1 && s.stride == 1,
                          "conv: anchored window source with %dx%d taps, dilation %d, stride %d", s.r, s.s, s.dil, s.stride);
            B200SEG_CHECK(s.c > 0 && s.c % 8 == 0 && s.x_c_offset % 8 == 0 && s.x_c_total % 8 == 0,
                          "conv: input channels (c=%d, offset=%d, total=%d) must be multiples of 8", s.c, s.x_c_offset,
                          s.x_c_total);
            const int span = ceil_div(s.x_c_offset + s.c, s.x_c_total);      // buffer pixels one window touches
            B200SEG_CHECK(h + s.r - 1 <= s.x_rows && w + s.s - 1 + span - 1 <= s.x_cols,
                          "conv: %dx%d outputs with %dx%d taps and %d-pixel windows exceed the %dx%d buffer", h, w, s.r,
                          s.s, span, s.x_rows, s.x_cols);
            B200SEG_CHECK(s.w_rows >= out->k, "conv: packed weight has %d rows < %d output channels", s.w_rows, out->k);
            continue;
        }
        B200SEG_CHECK((s.r == 1 && s.s == 1) || (s.r == 3 && s.s == 3), "conv: only 1x1 and 3x3 filters (got %dx%d)",
                      s.r, s.s);
        B200SEG_CHECK(s.dil >= 1 && s.dil <= 4096, "conv: dilation %d", s.dil);
        B200SEG_CHECK(s.stride == 1 || s.stride == 2, "conv: stride %d (only 1 and 2)", s.stride);
        B200SEG_CHECK(s.stride == 1 || s.dil == 1, "conv: stride 2 with dilation %d is not supported", s.dil);
        B200SEG_CHECK(s.c > 0 && s.c % 8 == 0 && s.x_c_offset % 8 == 0 && s.x_c_total % 8 == 0,
                      "conv: input channels (c=%d, offset=%d, total=%d) must be multiples of 8", s.c, s.x_c_offset,
                      s.x_c_total);
        B200SEG_CHECK(s.x_c_offset + s.c <= s.x_c_total, "conv: input slice exceeds buffer channels");
        B200SEG_CHECK(s.w_rows >= out->k, "conv: packed weight has %d rows < %d output channels", s.w_rows, out->k);
    }
    return 0;
}

// Activation view: the (hv x wv) sub-grid of an (hf x wf) NHWC tensor starting at pixel (oy, ox) with pixel step
// `step` in both directions.
void set_view(SegPlan* sp, const void* x, int c_total, int c_off, int c, int n, int hf, int wf, int oy, int ox, int step,
              int hv, int wv) {
    sp->a_base = static_cast<const char*>(x) + (static_cast<size_t>(oy) * wf + ox) * c_total * 2 + static_cast<size_t>(c_off) * 2;
    sp->adims[0] = c;
    sp->adims[1] = wv;
    sp->adims[2] = hv;
    sp->adims[3] = n;
    sp->astr[0] = static_cast<uint64_t>(step) * c_total * 2;
    sp->astr[1] = static_cast<uint64_t>(step) * wf * c_total * 2;
    sp->astr[2] = static_cast<uint64_t>(hf) * wf * c_total * 2;
}

void plain_out(OutPlan* op, const b200seg_conv_out* out, int n, int h, int w) {
    op->base = static_cast<char*>(out->y) + static_cast<size_t>(out->y_c_offset) * 2;
    op->dims[0] = out->k;
    op->dims[1] = w;
    op->dims[2] = h;
    op->dims[3] = n;
    op->str[0] = static_cast<uint64_t>(out->y_c_total) * 2;
    op->str[1] = static_cast<uint64_t>(w) * out->y_c_total * 2;
    op->str[2] = static_cast<uint64_t>(h) * w * out->y_c_total * 2;
}

inline int floor_div2(int v) { return v >= 0 ? v / 2 : -((-v + 1) / 2); }

// Forward: n x h x w is the OUTPUT grid; a stride-2 source reads a (2h x 2w) input through its four parity views.
int launch_fprop(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                 cudaStream_t st) {
    clear_error();
    if (check_common(n, h, w, n_src, src, out)) return 1;
    SegPlan sp[kMaxSeg];
    memset(sp, 0, sizeof(sp));
    int nseg = 0;
    if (src[0].x_rows || src[0].x_cols) {
        // anchored window source: tap (r, q) of output pixel (y, x) reads the c contiguous elements that start at
        // buffer pixel (y + r, x + q) — with c > x_c_total the view's innermost dimension runs on into the following
        // pixels, i.e. neighbouring windows overlap in memory (the TMA unit only sees strides)
        const b200seg_conv_src& s = src[0];
        SegPlan& g = sp[nseg++];
        const int span = ceil_div(s.x_c_offset + s.c, s.x_c_total);
        g.a_base = static_cast<const char*>(s.x) + static_cast<size_t>(s.x_c_offset) * 2;
        g.adims[0] = s.c;
        g.adims[1] = s.x_cols - span + 1;
        g.adims[2] = s.x_rows;
        g.adims[3] = n;
        g.astr[0] = static_cast<uint64_t>(s.x_c_total) * 2;
        g.astr[1] = static_cast<uint64_t>(s.x_cols) * s.x_c_total * 2;
        g.astr[2] = static_cast<uint64_t>(s.x_rows) * s.x_cols * s.x_c_total * 2;
        g.w = s.w;
        g.w_rows = s.w_rows;
        g.c = s.c;
        g.w_taps = s.r * s.s;
        g.seg.cblks = ceil_div(s.c, kBK);
        for (int r = 0; r < s.r; ++r)
            for (int q = 0; q < s.s; ++q) {
                const int t = g.seg.ntaps++;
                g.seg.dy[t] = static_cast<short>(r);
                g.seg.dx[t] = static_cast<short>(q);
                g.seg.wtap[t] = static_cast<short>(r * s.s + q);
            }
        OutPlan wop;
        plain_out(&wop, out, n, h, w);
        return launch_planned(n, h, w, nseg, sp, wop, out, st, true);
    }
    for (int i = 0; i < n_src; ++i) {
        const b200seg_conv_src& s = src[i];
        if (s.stride == 1) {
            B200SEG_CHECK(nseg < kMaxSeg, "conv: too many segments");
            SegPlan& g = sp[nseg++];
            set_view(&g, s.x, s.x_c_total, s.x_c_offset, s.c, n, h, w, 0, 0, 1, h, w);
            g.w = s.w;
            g.w_rows = s.w_rows;
            g.c = s.c;
            g.w_taps = s.r * s.s;
            g.seg.cblks = ceil_div(s.c, kBK);
            for (int r = 0; r < s.r; ++r)
                for (int q = 0; q < s.s; ++q) {
                    const int t = g.seg.ntaps++;
                    g.seg.dy[t] = static_cast<short>((r - s.r / 2) * s.dil);
                    g.seg.dx[t] = static_cast<short>((q - s.s / 2) * s.dil);
                    g.seg.wtap[t] = static_cast<short>(r * s.s + q);
                }
        } else {
            // input pixel = 2*o + (r - R/2): parity picks the view, floor((r - R/2)/2) is the shift inside it
            for (int py = 0; py < 2; ++py)
                for (int px = 0; px < 2; ++px) {
                    SegPlan g;
                    memset(&g, 0, sizeof(g));
                    for (int r = 0; r < s.r; ++r)
                        for (int q = 0; q < s.s; ++q) {
                            const int oy = r - s.r / 2, ox = q - s.s / 2;
                            if ((oy & 1) != py || (ox & 1) != px) continue;
                            const int t = g.seg.ntaps++;
                            g.seg.dy[t] = static_cast<short>(floor_div2(oy));
                            g.seg.dx[t] = static_cast<short>(floor_div2(ox));
                            g.seg.wtap[t] = static_cast<short>(r * s.s + q);
                        }
                    if (g.seg.ntaps == 0) continue;
                    B200SEG_CHECK(nseg < kMaxSeg, "conv: too many segments");
                    set_view(&g, s.x, s.x_c_total, s.x_c_offset, s.c, n, 2 * h, 2 * w, py, px, 2, h, w);
                    g.w = s.w;
                    g.w_rows = s.w_rows;
                    g.c = s.c;
                    g.w_taps = s.r * s.s;
                    g.seg.cblks = ceil_div(s.c, kBK);
                    sp[nseg++] = g;
                }
        }
    }
    OutPlan op;
    plain_out(&op, out, n, h, w);
    return launch_planned(n, h, w, nseg, sp, op, out, st);
}

// Backward-data: n x h x w is the dX grid.  A source that was a stride-1 conv brings a dY of the same size; a
// stride-2 source brings a dY of (h/2 x w/2).  Weights are the dgrad packing (taps mirrored), so
//   dX[i] = sum_t' dY[(i + (t' - R/2) * dil) / stride] * Wd[t']      over the t' for which the division is exact.
// With any stride-2 source the launch is split into the four parity sub-grids of dX (strided output views); each
// source contributes, per parity, the taps that land on it.
int launch_dgrad(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                 cudaStream_t st) {
    clear_error();
    if (check_common(n, h, w, n_src, src, out)) return 1;
    bool any_s2 = false;
    for (int i = 0; i < n_src; ++i) {
        any_s2 |= src[i].stride == 2;
        B200SEG_CHECK(!src[i].x_rows && !src[i].x_cols, "conv dgrad: anchored window sources are forward-only");
    }
    if (!any_s2) return launch_fprop(n, h, w, n_src, src, out, st);     // same GEMM, mirrored weights

    B200SEG_CHECK(h % 2 == 0 && w % 2 == 0, "conv dgrad: stride 2 needs even dX size (got %dx%d)", h, w);
    B200SEG_CHECK(!out->out_fp32 && !out->stats, "conv dgrad: stride 2 writes bf16 without statistics");
    const int hs = h / 2, ws = w / 2;
    for (int py = 0; py < 2; ++py)
        for (int px = 0; px < 2; ++px) {
            SegPlan sp[kMaxSeg];
            memset(sp, 0, sizeof(sp));
            int nseg = 0;
            for (int i = 0; i < n_src; ++i) {
                const b200seg_conv_src& s = src[i];
                if (s.stride == 2) {
                    SegPlan g;
                    memset(&g, 0, sizeof(g));
                    for (int r = 0; r < s.r; ++r)
                        for (int q = 0; q < s.s; ++q) {
                            const int ey = py + r - s.r / 2, ex = px + q - s.s / 2;
                            if ((ey & 1) || (ex & 1)) continue;
                            const int t = g.seg.ntaps++;
                            g.seg.dy[t] = static_cast<short>(ey / 2);
                            g.seg.dx[t] = static_cast<short>(ex / 2);
                            g.seg.wtap[t] = static_cast<short>(r * s.s + q);
                        }
                    if (g.seg.ntaps == 0) continue;
                    B200SEG_CHECK(nseg < kMaxSeg, "conv dgrad: too many segments");
                    set_view(&g, s.x, s.x_c_total, s.x_c_offset, s.c, n, hs, ws, 0, 0, 1, hs, ws);
                    g.w = s.w;
                    g.w_rows = s.w_rows;
                    g.c = s.c;
                    g.w_taps = s.r * s.s;
                    g.seg.cblks = ceil_div(s.c, kBK);
                    sp[nseg++] = g;
                } else {
                    // full-resolution dY seen from the parity sub-grid: pixel 2i' + p + o -> view (p+o)&1, shift floor
                    for (int vy = 0; vy < 2; ++vy)
                        for (int vx = 0; vx < 2; ++vx) {
                            SegPlan g;
                            memset(&g, 0, sizeof(g));
                            for (int r = 0; r < s.r; ++r)
                                for (int q = 0; q < s.s; ++q) {
                                    const int oy = py + (r - s.r / 2) * s.dil, ox = px + (q - s.s / 2) * s.dil;
                                    if ((oy & 1) != vy || (ox & 1) != vx) continue;
                                    const int t = g.seg.ntaps++;
                                    g.seg.dy[t] = static_cast<short>(floor_div2(oy));
                                    g.seg.dx[t] = static_cast<short>(floor_div2(ox));
                                    g.seg.wtap[t] = static_cast<short>(r * s.s + q);
                                }
                            if (g.seg.ntaps == 0) continue;
                            B200SEG_CHECK(nseg < kMaxSeg, "conv dgrad: too many segments");
                            set_view(&g, s.x, s.x_c_total, s.x_c_offset, s.c, n, h, w, vy, vx, 2, hs, ws);
                            g.w = s.w;
                            g.w_rows = s.w_rows;
                            g.c = s.c;
                            g.w_taps = s.r * s.s;
                            g.seg.cblks = ceil_div(s.c, kBK);
                            sp[nseg++] = g;
                        }
                }
            }
            B200SEG_CHECK(nseg > 0, "conv dgrad: parity (%d,%d) of dX receives no gradient from any source; zero-fill "
                                    "is not implemented (combine the stride-2 1x1 with a stride-1 source)", py, px);
            OutPlan op;
            op.base = static_cast<char*>(out->y) + (static_cast<size_t>(py) * w + px) * out->y_c_total * 2 +
                      static_cast<size_t>(out->y_c_offset) * 2;
            op.dims[0] = out->k;
            op.dims[1] = ws;
            op.dims[2] = hs;
            op.dims[3] = n;
            op.str[0] = 2ull * out->y_c_total * 2;
            op.str[1] = 2ull * w * out->y_c_total * 2;
            op.str[2] = static_cast<uint64_t>(h) * w * out->y_c_total * 2;
            if (launch_planned(n, hs, ws, nseg, sp, op, out, st)) return 1;
        }
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

int b200seg_conv_last_plan(int64_t* iters_run, int64_t* iters_dense) {
    if (iters_run) *iters_run = b200seg::g_iters_run;
    if (iters_dense) *iters_dense = b200seg::g_iters_dense;
    return 0;
}

int b200seg_conv_fprop(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st) {
    b200seg::g_iters_run = b200seg::g_iters_dense = 0;
    return b200seg::launch_fprop(n, h, w, n_src, src, out, static_cast<cudaStream_t>(st));
}

int b200seg_conv_dgrad(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st) {
    b200seg::g_iters_run = b200seg::g_iters_dense = 0;
    return b200seg::launch_dgrad(n, h, w, n_src, src, out, static_cast<cudaStream_t>(st));
}

}  // extern "C"
