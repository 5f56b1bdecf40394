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

namespace b200seg {

namespace {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kMaxBN = 256;
constexpr int kStages = 4;
constexpr int kABytes = kBM * kBK * 2;
constexpr int kBBytesMax = kMaxBN * kBK * 2;
constexpr int kStageBytes = kABytes + kBBytesMax;
constexpr int kThreads = 256;
constexpr int kBarrierBytes = 256;
constexpr int kSmemBytes = kStages * kStageBytes + kBarrierBytes + 1024;   // +1024: manual 1 KB alignment
constexpr int kTmemCols = 512;

struct IgemmSeg {
    int r, s, dil, cblks;
};

struct __align__(64) IgemmParams {
    CUtensorMap a_map[B200SEG_MAX_CONV_SRC];
    CUtensorMap b_map[B200SEG_MAX_CONV_SRC];
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
};

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
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
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
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], 4);
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
                if (p.stats) {
                    // column sums over this warp's 32 rows by recursive halving: 31 shuffles leave lane j
                    // holding the total of channel k0 + j (rows outside the image contribute zero)
                    float s1[32], s2[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        // statistics of the values as stored (bf16-rounded), so the saved mean / inv-std are
                        // exactly those of the tensor the backward pass re-reads
                        float a = valid ? __uint_as_float(v[j]) : 0.f;
                        if (!p.out_fp32) a = __bfloat162float(__float2bfloat16(a));
                        s1[j] = a;
                        s2[j] = a * a;
                    }
#pragma unroll
                    for (int off = 16; off >= 1; off >>= 1) {
                        const bool upper = (lane & off) != 0;
#pragma unroll
                        for (int j = 0; j < off; ++j) {
                            const float send1 = upper ? s1[j] : s1[j + off];
                            const float keep1 = upper ? s1[j + off] : s1[j];
                            s1[j] = keep1 + __shfl_xor_sync(0xffffffffu, send1, off);
                            const float send2 = upper ? s2[j] : s2[j + off];
                            const float keep2 = upper ? s2[j + off] : s2[j];
                            s2[j] = keep2 + __shfl_xor_sync(0xffffffffu, send2, off);
                        }
                    }
                    if (k0 + lane < p.k_store) {
                        atomicAdd(p.stats + k0 + lane, s1[0]);
                        atomicAdd(p.stats + p.k_store + k0 + lane, s2[0]);
                    }
                }
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
    p.bn = round_up(ceil_div(out->k, p.tiles_n), 32);
    B200SEG_CHECK(p.bn <= kMaxBN, "conv: internal bn %d", p.bn);

    // pixel patch: the power-of-two bw x bh x bi = 128 shape that needs the fewest pipeline iterations
    long long best = -1;
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

    p.y = out->y;
    p.y_c_total = out->y_c_total;
    p.y_c_offset = out->y_c_offset;
    p.k_store = out->k;
    p.out_fp32 = out->out_fp32;
    p.scale = out->scale;
    p.bias = out->bias;
    p.relu = out->relu;
    p.stats = out->stats;
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

int b200seg_conv_fprop(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st) {
    return b200seg::launch_igemm(n, h, w, n_src, src, out, static_cast<cudaStream_t>(st));
}

int b200seg_conv_dgrad(int n, int h, int w, int n_src, const b200seg_conv_src* src, const b200seg_conv_out* out,
                       b200seg_stream_t st) {
    return b200seg::launch_igemm(n, h, w, n_src, src, out, static_cast<cudaStream_t>(st));
}

}  // extern "C"
