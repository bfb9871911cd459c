// 3x3 stride-1 pad-1 convolution with halo reuse and resident filters (tcgen05 / TMEM / TMA, sm_100a).
// Reference layers: every `nn.Conv2d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)` of BasicBlock
// (pytorch_code/model_ops/resnet.py:19-22, resnet_split.py:28,34) and the 3x3 convs of the VGG configuration; forward,
// data gradient (same kernel on mirrored filters) and weight gradient (conv_halo_wgrad_kernel below).
//
// The generic implicit-GEMM path (gemm_tcgen05.cu, OP_CONVK) fetches one shifted activation box per filter tap, i.e.
// every input element travels L2 -> shared memory nine times, and re-fetches the filter tile for every output tile.
// For the early ResNet / VGG stages (C = 64 / 128, 32x32 and 16x16 maps) that traffic, not the tensor pipe, is the
// limit (ncu: 19% tensor active, see profiles/). Here:
//   * ONE TMA box per (tile, 64-channel block) brings a zero-padded halo image [rows][W+2][64ch] into shared memory
//     (out-of-bounds columns / rows are zero-filled by TMA: that IS the padding);
//   * in that padded-flat space the input of tap (r, s) for output position p is position p + (r-1)(W+2) + (s-1):
//     a pure shift, so all nine A operands are views of the same buffer -- the UMMA descriptor start address simply
//     moves by multiples of 128 bytes (measured on B200: the 128B-swizzle XOR is taken from the absolute shared-memory
//     address, exactly as TMA wrote it, so the descriptor's base-offset field must stay 0; setting it to
//     (addr >> 7) & 7 gives wrong data -- bench/kernel_check.py conv_halo_diag);
//   * the filter bank of this CTA's 64 output channels (9 taps x C/64 blocks x 8 KiB) is loaded once and stays
//     resident while the persistent CTA walks its tiles.
// Output positions that fall on the two padding columns of the padded-flat space are computed and discarded.
#include "common.cuh"
#include "../include/pdnn_b200.h"
#include <stdlib.h>

namespace pdnn {

struct HaloParams {
    int batch, H, W, C, Kout;
    int Wp;              // W + 2
    int Rb;              // halo rows per box
    int tiles_per_img;   // ceil(H * Wp / 128)
    int KB;              // C / 64
    int n_cblocks;       // Kout / 64
    int stage_bytes;     // Rb * Wp * 128 rounded up to 1024
    int stages;
    void* out;
    long long ldo;
    int relu;
    const float* bias;
    const __nv_bfloat16* residual;
    const uint32_t* guard_closed;
    const uint32_t* guard_step;
    float* bn_stats;     // optional [2][Kout]: per-channel sum and sum of squares of the (bf16-rounded) output
};

constexpr int HALO_THREADS = 256;
constexpr int HALO_BN = 64;
// With N = 64 one tcgen05.mma occupies the tensor pipe for ~32 cycles, but a single thread needs ~90 cycles to issue
// one (descriptor moves into uniform registers; measured with ncu: tensor pipe 35% active, issuing warp 100% busy).
// Three warps therefore issue for three different tiles at once, each into its own TMEM accumulator.
// Every issuer owns a private slice of the stage ring (SI stages) and two private accumulators, so each mbarrier has
// exactly one producer sequence and one consumer sequence in program order. (Sharing one ring / one accumulator pool
// between issuers lets a fast issuer get two barrier phases ahead of a slow one; a parity wait cannot tell phase n
// from phase n+2 -- found as a hang at batch 1024 with a 6-stage ring.)
constexpr int HALO_ISSUERS = 3;
constexpr int HALO_NACC = 6;

__global__ void __launch_bounds__(HALO_THREADS, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const HaloParams p) {
    if (p.guard_closed != nullptr) {
        __shared__ int s_abort;
        if (threadIdx.x == 0) s_abort = ld_relaxed_sys(p.guard_closed) >= ld_relaxed_sys(p.guard_step);
        __syncthreads();
        if (s_abort) return;
    }
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = smem_u32(smem);
    const int w_bytes = 9 * p.KB * 8192;
    const uint32_t w_base = smem_base;
    const uint32_t a_base = smem_base + w_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + w_bytes + p.stages * p.stage_bytes);
    const uint32_t bar_base = smem_u32(bars);
    // barrier slots: [0..S) full, [S..2S) empty, then NACC tmem-full, NACC tmem-empty, 1 weights
    const int S = p.stages;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (S + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * S + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * S + HALO_NACC + a); };
    const uint32_t wbar = bar_base + 8u * (2 * S + 2 * HALO_NACC);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 2 * HALO_NACC + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmW);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < HALO_NACC; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }
        mbar_init(wbar, 1);
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc(smem_u32(tmem_slot), 512);      // 6 accumulators x 64 columns = 384 -> next power of two
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                 // everything above may overlap the tail of the previous kernel (programmatic dependent launch)
    pdl_launch_dependents();

    // work split: CTA -> (output-channel block, strided subset of tiles)
    const int cb = blockIdx.x % p.n_cblocks;
    const int cta_in_cb = blockIdx.x / p.n_cblocks;
    const int ctas_per_cb = (gridDim.x - cb + p.n_cblocks - 1) / p.n_cblocks;
    const int total_tiles = p.batch * p.tiles_per_img;

    if (warp == 0) {
        if (lane == 0) {
            // resident filter bank of this CTA's 64 output channels
            mbar_expect_tx(wbar, (uint32_t)w_bytes);
            for (int tap = 0; tap < 9; ++tap)
                for (int kb = 0; kb < p.KB; ++kb)
                    tma_load_2d(w_base + (tap * p.KB + kb) * 8192, &tmW, wbar, tap * p.C + kb * 64, cb * HALO_BN);
            const int SI = S / HALO_ISSUERS;           // stages per issuer
            int seq = 0;
            for (int tile = cta_in_cb; tile < total_tiles; tile += ctas_per_cb, ++seq) {
                const int n = tile / p.tiles_per_img;
                const int p0 = (tile - n * p.tiles_per_img) * 128;
                int num = p0 - p.Wp - 1;
                const int h_lo = num >= 0 ? num / p.Wp : -((-num + p.Wp - 1) / p.Wp);   // floor: first image row of the halo box
                const int issuer = seq % HALO_ISSUERS;
                const int t_i = seq / HALO_ISSUERS;
                for (int kb = 0; kb < p.KB; ++kb) {
                    const int m = t_i * p.KB + kb;                 // this issuer's unit counter
                    const int stage = issuer * SI + (m % SI);
                    const uint32_t phase = (uint32_t)(m / SI) & 1u;
                    mbar_wait(empty_bar(stage), phase ^ 1);
                    const uint32_t fb = full_bar(stage);
                    mbar_expect_tx(fb, (uint32_t)(p.Rb * p.Wp * 128));
                    tma_load_4d(a_base + stage * p.stage_bytes, &tmX, fb, kb * 64, -1, h_lo, n);
                }
            }
        }
    } else if (warp <= HALO_ISSUERS) {
        // ===================== MMA issuers (warps 1..3): issuer i owns the tiles with sequence number == i mod 3 =====
        const int issuer = warp - 1;
        constexpr uint32_t idesc = make_idesc_bf16(128, HALO_BN, 0, 0);
        // descriptor arithmetic: the upper word is constant, the lower word is (addr >> 4) | LBO, so moving the start
        // address by `b` bytes is `lo += b >> 4`.
        const uint32_t desc_hi = (uint32_t)(make_smem_desc_sw128(0, 16, 1024) >> 32);
        int tap_off[9];
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) tap_off[tap] = ((tap / 3 - 1) * p.Wp + (tap % 3 - 1)) * 8;   // 128 B / 16
        mbar_wait(wbar, 0);
        tc_fence_after();
        const uint32_t w_lo = (uint32_t)make_smem_desc_sw128(w_base, 16, 1024);
        const uint32_t issue = lane == 0 ? 1u : 0u;
        int seq = 0;
        for (int tile = cta_in_cb; tile < total_tiles; tile += ctas_per_cb, ++seq) {
            if (seq % HALO_ISSUERS != issuer) continue;
            const int n = tile / p.tiles_per_img;
            const int p0 = (tile - n * p.tiles_per_img) * 128;
            int num = p0 - p.Wp - 1;
            const int fl = num >= 0 ? num / p.Wp : -((-num + p.Wp - 1) / p.Wp);
            // Output position q = row * Wp + (w + 1) reads, for tap (r, s), input pixel (row + r - 1, w + s - 1), which sits
            // at index q + (r - 1 - h_lo) * Wp + (s - 1) of a halo box whose first row is image row h_lo = fl and whose
            // first column is w = -1.  origin = p0 - fl * Wp >= Wp + 1 keeps every index non-negative.
            const int origin = p0 - fl * p.Wp;
            const int t_i = seq / HALO_ISSUERS;
            const int acc = issuer * 2 + (t_i & 1);
            const uint32_t acc_phase = (uint32_t)(t_i >> 1) & 1u;
            mbar_wait(tempty_bar(acc), acc_phase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * HALO_BN;
            const int SI = S / HALO_ISSUERS;
            for (int kb = 0; kb < p.KB; ++kb) {
                const int m = t_i * p.KB + kb;
                const int stage = issuer * SI + (m % SI);
                const uint32_t phase = (uint32_t)(m / SI) & 1u;
                mbar_wait(full_bar(stage), phase);
                tc_fence_after();
                const uint32_t a_lo = (uint32_t)make_smem_desc_sw128(a_base + stage * p.stage_bytes + (uint32_t)origin * 128u, 16, 1024);
                const uint32_t b_lo = w_lo + (uint32_t)kb * 512u;           // 8192 B per (tap, kb) tile
#pragma unroll
                for (int tap = 0; tap < 9; ++tap) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint64_t adesc = ((uint64_t)desc_hi << 32) | (uint32_t)(a_lo + tap_off[tap] + j * 2);
                        const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (uint32_t)(b_lo + (uint32_t)(tap * p.KB) * 512u + j * 2);
                        umma_bf16_ss_pred(tmem_d, adesc, bdesc, idesc, (kb > 0 || tap > 0 || j > 0) ? 1u : 0u, issue);
                    }
                }
                umma_commit_pred(empty_bar(stage), issue);
                if (kb == p.KB - 1) umma_commit_pred(tfull_bar(acc), issue);
                __syncwarp();
            }
        }
    } else if (warp >= 4) {
        const int ew = warp & 3;
        int seq = 0;
        // BatchNorm statistics of the output. Every thread keeps private sums for its 64 columns over ALL tiles of this
        // CTA (the CTA's output channels never change; only the row a thread holds does), so the epilogue pays two FMAs per
        // element and the cross-lane reduction (31 shuffles per 32 columns) happens once per CTA, not once per tile.
        float st_sum[HALO_BN / 32][32], st_sq[HALO_BN / 32][32];
#pragma unroll
        for (int c = 0; c < HALO_BN / 32; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) { st_sum[c][i] = 0.f; st_sq[c][i] = 0.f; }
        for (int tile = cta_in_cb; tile < total_tiles; tile += ctas_per_cb, ++seq) {
            const int acc = (seq % HALO_ISSUERS) * 2 + ((seq / HALO_ISSUERS) & 1);
            const uint32_t acc_phase = (uint32_t)((seq / HALO_ISSUERS) >> 1) & 1u;
            const int n = tile / p.tiles_per_img;
            const int q = (tile - n * p.tiles_per_img) * 128 + ew * 32 + lane;
            const int row = q / p.Wp;
            const int colp = q - row * p.Wp;
            const bool valid = row < p.H && colp >= 1 && colp <= p.W;
            const long long pix = ((long long)n * p.H + row) * p.W + (colp - 1);
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < HALO_BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + acc * HALO_BN + c * 32, v);
                tmem_ld_wait();
                const int col0 = cb * HALO_BN + c * 32;
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
                if (valid) {
                    if (p.bias != nullptr) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) f[i] += __ldg(p.bias + col0 + i);
                    }
                    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + pix * p.ldo + col0;
                    // 32-byte accesses: one full sector per thread and instruction (ldo % 16 == 0 and 32-byte aligned
                    // bases are checked on the host)
                    if (p.residual != nullptr) {
                        const __nv_bfloat16* rp = p.residual + pix * p.ldo + col0;
                        const u32x8 r0 = ld_global_nc_256(rp), r1 = ld_global_nc_256(rp + 16);
#pragma unroll
                        for (int q4 = 0; q4 < 8; ++q4) {
                            const float2 a = unpack_bf16x2(r0.v[q4]), b = unpack_bf16x2(r1.v[q4]);
                            f[q4 * 2] += a.x; f[q4 * 2 + 1] += a.y; f[16 + q4 * 2] += b.x; f[16 + q4 * 2 + 1] += b.y;
                        }
                    }
                    if (p.relu) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
                    }
                    u32x8 w0, w1;
#pragma unroll
                    for (int q4 = 0; q4 < 8; ++q4) {
                        w0.v[q4] = pack_bf16x2(f[q4 * 2], f[q4 * 2 + 1]);
                        w1.v[q4] = pack_bf16x2(f[16 + q4 * 2], f[16 + q4 * 2 + 1]);
                    }
                    st_global_256(o, w0);
                    st_global_256(o + 16, w1);
                }
                if (p.bn_stats != nullptr && valid) {
                    // statistics of exactly what was stored (bf16-rounded); the padding columns are not counted
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float r = __bfloat162float(__float2bfloat16(f[i]));
                        st_sum[c][i] += r;
                        st_sq[c][i] = fmaf(r, r, st_sq[c][i]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        if (p.bn_stats != nullptr) {
#pragma unroll
            for (int c = 0; c < HALO_BN / 32; ++c) {
                const float s1 = warp_column_sums(st_sum[c], lane);
                const float s2 = warp_column_sums(st_sq[c], lane);
                atomicAdd(p.bn_stats + cb * HALO_BN + c * 32 + lane, s1);
                atomicAdd(p.bn_stats + p.Kout + cb * HALO_BN + c * 32 + lane, s2);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace pdnn

using namespace pdnn;

int pdnn_make_tmap_bf16(CUtensorMap* tm, const void* base, int rank, const long long* dims, const long long* strides_elems,
                        const int* box);   // gemm_tcgen05.cu

extern "C" int pdnn_conv_halo_supported(int batch, int H, int W, int C, int Kout, int R, int S, int pad) {
    if (R != 3 || S != 3 || pad != 1) return 0;
    if (C != 64 && C != 128) return 0;
    if (Kout % 64 != 0) return 0;
    const int Wp = W + 2;
    const int Rb = (3 * Wp + 128) / Wp + 1;
    if (Wp > 256 || Rb > 256) return 0;
    const int stage_bytes = (Rb * Wp * 128 + 1023) / 1024 * 1024;
    const int w_bytes = 9 * (C / 64) * 8192;
    const int stages = (227 * 1024 - 2560 - w_bytes) / stage_bytes;
    if (stages < 3) return 0;          // one private stage per issuer warp at least
    if (W < 8 || H < 4) return 0;     // small maps waste most of a 128-position tile on padding: generic path is better
    return 1;
}

static int conv_halo_fprop_impl(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                                const pdnn_epilogue_t* epi, float* bn_stats, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(pdnn_conv_halo_supported(batch, H, W, C, Kout, 3, 3, 1), "conv_halo: unsupported geometry (C=%d Kout=%d %dx%d)", C, Kout, H, W);
    PDNN_REQUIRE(epi != nullptr && epi->out != nullptr && epi->out_mode == 0, "conv_halo: bf16 output required");
    PDNN_REQUIRE(epi->ldo % 16 == 0 && (reinterpret_cast<uintptr_t>(epi->out) & 31) == 0 &&
                     (reinterpret_cast<uintptr_t>(epi->residual) & 31) == 0,
                 "conv_halo: ldo must be a multiple of 16 and out / residual 32-byte aligned (32-byte epilogue stores)");
    HaloParams p{};
    p.batch = batch; p.H = H; p.W = W; p.C = C; p.Kout = Kout;
    p.Wp = W + 2;
    p.Rb = (3 * p.Wp + 128) / p.Wp + 1;
    p.tiles_per_img = (H * p.Wp + 127) / 128;
    p.KB = C / 64;
    p.n_cblocks = Kout / 64;
    p.stage_bytes = (p.Rb * p.Wp * 128 + 1023) / 1024 * 1024;
    const int w_bytes = 9 * p.KB * 8192;
    p.stages = (227 * 1024 - 2560 - w_bytes) / p.stage_bytes;
    p.stages = p.stages >= 6 ? 6 : 3;   // a multiple of the issuer count: every issuer owns stages / 3 of the ring
    p.out = epi->out; p.ldo = epi->ldo; p.relu = epi->relu; p.bias = epi->bias;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(epi->residual);
    p.guard_closed = epi->guard_closed; p.guard_step = epi->guard_step;
    p.bn_stats = bn_stats;
    CUtensorMap tmX, tmW;
    {
        long long dims[4] = {C, W, H, batch};
        long long str[4] = {1, C, (long long)W * C, (long long)H * W * C};
        int box[4] = {64, p.Wp, p.Rb, 1};
        if (int rc = pdnn_make_tmap_bf16(&tmX, x, 4, dims, str, box)) return rc;
        long long wd[2] = {9ll * C, Kout};
        long long ws[2] = {1, 9ll * C};
        int wb[2] = {64, 64};
        if (int rc = pdnn_make_tmap_bf16(&tmW, wgt, 2, wd, ws, wb)) return rc;
    }
    const int smem = w_bytes + p.stages * p.stage_bytes + 512 + 1024;
    static int attr_smem = 0;
    if (smem > attr_smem) {
        PDNN_CUDA_CHECK(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem = smem;
    }
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    int grid = sms / p.n_cblocks * p.n_cblocks;
    const int total_units = batch * p.tiles_per_img * p.n_cblocks;
    if (grid > total_units) grid = (total_units / p.n_cblocks) * p.n_cblocks;
    if (grid < p.n_cblocks) grid = p.n_cblocks;
    if (epi->max_ctas > 0 && grid > epi->max_ctas) grid = (epi->max_ctas / p.n_cblocks) * p.n_cblocks;
    if (grid < p.n_cblocks) grid = p.n_cblocks;
    PDNN_CUDA_CHECK(pdnn_launch(conv_halo_kernel, dim3(grid), dim3(HALO_THREADS), (size_t)smem, stream, tmX, tmW, p));
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_conv_halo_fprop_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                                         const pdnn_epilogue_t* epi, void* stream) {
    return conv_halo_fprop_impl(x, wgt, batch, H, W, C, Kout, epi, nullptr, stream);
}

// Same convolution; additionally adds the per-output-channel sum and sum of squares of the stored (bf16) output into
// bn_stats[0][k] / bn_stats[1][k] (fp32, caller zero-fills): the statistics pass of the BatchNorm that follows.
extern "C" int pdnn_conv_halo_fprop_stats_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                                               const pdnn_epilogue_t* epi, float* bn_stats, void* stream) {
    PDNN_REQUIRE(bn_stats != nullptr, "conv_halo_fprop_stats: bn_stats is NULL");
    return conv_halo_fprop_impl(x, wgt, batch, H, W, C, Kout, epi, bn_stats, stream);
}

// =================================================================================================================
// Weight gradient with halo reuse
// =================================================================================================================
// dW[k][r][s][c] = sum over (n, row, w) of dY[n, row, w, k] * X[n, row + r - 1, w + s - 1, c].
// The reduction runs over positions, so both operands are MN-major (positions are the K dimension). A chunk of Hc
// image rows is brought in ONCE per CTA step: dY as a padded-flat tile [Hc][W+2][64 ch] (TMA zero-fills the two
// padding columns, so padding positions contribute nothing) and X as a halo [Hc+2][W+2][64 ch]. In that space tap
// (r, s) is again a pure row shift of the B operand: index q + r (W+2) + (s - 1). A CTA owns one (cin block, cout
// block, tap group) and keeps ALL accumulators of its taps in TMEM for its whole life (taps x 64 columns), walking
// the (image, row chunk) list with split-K across CTAs; one fp32 red.add epilogue at the very end. Three issuer
// warps drive different taps (= different accumulators) of the same chunk concurrently.
namespace pdnn {

struct HaloWgradParams {
    int batch, H, W, C, Kout;
    int Wp, Hc;            // W + 2 ; image rows per chunk (Hc * Wp is a multiple of 16)
    int chunks_per_img;    // H / Hc
    int KS;                // k16 steps per chunk = Hc * Wp / 16
    int m_chunks;          // 1 (Kout tile of 64, rows 64..127 of the MMA alias rows 0..63) or 2 (Kout tile of 128)
    int n_cin_blocks, n_cout_blocks;
    int dy_bytes, halo_off, stage_bytes, stages;
    int ctas_group_a;      // CTAs working on taps 0..5 ; the rest work on taps 6..8
    int wide;              // 1: one N = 192 MMA per filter row (three taps through LBO = one pixel); 0: one N = 64 MMA per tap
    float* out;            // [Kout][9 * C] fp32, red.add
    const uint32_t* guard_closed;
    const uint32_t* guard_step;
};

__global__ void __launch_bounds__(HALO_THREADS, 1)
conv_halo_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX, const HaloWgradParams p) {
    if (p.guard_closed != nullptr) {
        __shared__ int s_abort;
        if (threadIdx.x == 0) s_abort = ld_relaxed_sys(p.guard_closed) >= ld_relaxed_sys(p.guard_step);
        __syncthreads();
        if (s_abort) return;
    }
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = smem_u32(smem);
    const int S = p.stages;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * p.stage_bytes);
    const uint32_t bar_base = smem_u32(bars);
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (S + s); };
    const uint32_t done_bar = bar_base + 8u * (2 * S);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // which (tap group, cin block, cout block) does this CTA accumulate, and which share of the chunks does it take?
    const int combos = p.n_cin_blocks * p.n_cout_blocks;
    const bool group_a = (int)blockIdx.x < p.ctas_group_a;
    const int gidx = group_a ? blockIdx.x : blockIdx.x - p.ctas_group_a;
    const int gsize = group_a ? p.ctas_group_a : (int)gridDim.x - p.ctas_group_a;
    const int combo = gidx % combos;
    const int part = gidx / combos;
    const int parts = (gsize - combo + combos - 1) / combos;
    const int cin_blk = combo % p.n_cin_blocks;
    const int cout_blk = combo / p.n_cin_blocks;
    const int tap0 = group_a ? 0 : 6;
    const int ntaps = group_a ? 6 : 3;
    const int total_chunks = p.batch * p.chunks_per_img;

    // zero the guard rows around every halo buffer once: positions on the padding columns multiply a zero dY row with
    // whatever sits one row before / after the halo, and that must not be a NaN bit pattern
    {
        const int halo_end = p.halo_off + (p.Hc + 2) * p.Wp * 128;
        for (int s = 0; s < S; ++s) {
            uint8_t* st = smem + s * p.stage_bytes;
            for (int i = threadIdx.x; i < 1024 / 16; i += blockDim.x)
                reinterpret_cast<uint4*>(st + p.halo_off - 1024)[i] = make_uint4(0, 0, 0, 0);
            for (int i = threadIdx.x; i < (p.stage_bytes - halo_end) / 16; i += blockDim.x)
                reinterpret_cast<uint4*>(st + halo_end)[i] = make_uint4(0, 0, 0, 0);
        }
    }
    fence_proxy_async_smem();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmDY);
        tma_prefetch_desc(&tmX);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), HALO_ISSUERS); }
        mbar_init(done_bar, HALO_ISSUERS);
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc(smem_u32(tmem_slot), 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                 // everything above may overlap the tail of the previous kernel (programmatic dependent launch)
    pdl_launch_dependents();
    const int halo_bytes = (p.Hc + 2) * p.Wp * 128;
    const int dy_tile_bytes = p.Hc * p.Wp * 128;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int ch = part; ch < total_chunks; ch += parts) {
                const int n = ch / p.chunks_per_img;
                const int h0 = (ch - n * p.chunks_per_img) * p.Hc;
                mbar_wait(empty_bar(stage), phase ^ 1);
                const uint32_t fb = full_bar(stage);
                const uint32_t sb = smem_base + stage * p.stage_bytes;
                mbar_expect_tx(fb, (uint32_t)(p.m_chunks * dy_tile_bytes + halo_bytes));
                for (int j = 0; j < p.m_chunks; ++j)
                    tma_load_4d(sb + j * dy_tile_bytes, &tmDY, fb, cout_blk * (64 * p.m_chunks) + j * 64, -1, h0, n);
                tma_load_4d(sb + p.halo_off, &tmX, fb, cin_blk * 64, -1, h0 - 1, n);
                if (++stage == S) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp <= HALO_ISSUERS) {
        // issuer i drives taps tap0 + i, tap0 + i + 3 (group A) or tap0 + i (group B); accumulator column = local tap * 64
        const int issuer = warp - 1;
        constexpr uint32_t idesc = make_idesc_bf16(128, HALO_BN, 1, 1);     // both operands MN-major
        const uint32_t issue = lane == 0 ? 1u : 0u;
        // MN-major SW128 descriptors: SBO = 1024 (8 K-rows of 128 B); LBO = stride between 64-wide MN chunks. With a
        // 64-wide Kout tile LBO = 0 makes MMA rows 64..127 alias rows 0..63 (ignored by the epilogue) instead of costing
        // a second, empty dY tile.
        const uint64_t a_hi_lbo = make_smem_desc_sw128(0, p.m_chunks == 2 ? (uint32_t)dy_tile_bytes : 0u, 1024);
        const uint64_t b_hi_lbo = make_smem_desc_sw128(0, 0, 1024);
        int stage = 0;
        uint32_t phase = 0;
        bool first = true;
        for (int ch = part; ch < total_chunks; ch += parts) {
            mbar_wait(full_bar(stage), phase);
            tc_fence_after();
            const uint32_t sb = smem_base + stage * p.stage_bytes;
            const uint64_t a0 = a_hi_lbo + (uint64_t)((sb & 0x3FFFF) >> 4);
            const uint64_t b0 = b_hi_lbo + (uint64_t)(((sb + p.halo_off) & 0x3FFFF) >> 4);
            if (p.wide) {
                // The three taps of a filter row read the halo at shifts -1, 0, +1 pixels = consecutive 128-byte rows of
                // the MN-major B operand. A descriptor whose 64-wide N chunks are LBO = 128 bytes apart therefore covers all
                // three in ONE N = 192 MMA: dY (the A operand) is read from shared memory once per filter row instead of
                // once per tap, which is what bounded the N = 64 form (192 B/clk of operand reads against 128 B/clk).
                constexpr uint32_t idesc3 = make_idesc_bf16(128, 3 * HALO_BN, 1, 1);
                const uint64_t b_lbo128 = make_smem_desc_sw128(0, 128, 1024) - b_hi_lbo;      // LBO field only
                for (int lr = issuer; lr < ntaps / 3; lr += HALO_ISSUERS) {
                    const int r = tap0 / 3 + lr;
                    const int shift = r * p.Wp - 1;
                    const uint32_t tmem_d = tmem_base + lr * 3 * HALO_BN;
                    const uint64_t b_row = b0 + b_lbo128 + (uint64_t)(int64_t)(shift * 8);
#pragma unroll 4
                    for (int k = 0; k < p.KS; ++k) {
                        umma_bf16_ss_pred(tmem_d, a0 + (uint64_t)(k * 128), b_row + (uint64_t)(k * 128), idesc3, (first && k == 0) ? 0u : 1u, issue);
                    }
                }
            } else
            for (int lt = issuer; lt < ntaps; lt += HALO_ISSUERS) {
                const int tap = tap0 + lt;
                const int r = tap / 3, s = tap - r * 3;
                const int shift = r * p.Wp + (s - 1);                       // rows of 128 B; >= -1 (guard row in front)
                const uint32_t tmem_d = tmem_base + lt * HALO_BN;
                const uint64_t b_tap = b0 + (uint64_t)(int64_t)(shift * 8);  // 128 B = 8 x 16 B
#pragma unroll 4
                for (int k = 0; k < p.KS; ++k) {
                    umma_bf16_ss_pred(tmem_d, a0 + (uint64_t)(k * 128), b_tap + (uint64_t)(k * 128), idesc, (first && k == 0) ? 0u : 1u, issue);
                }
            }
            umma_commit_pred(empty_bar(stage), issue);
            __syncwarp();
            first = false;
            if (++stage == S) { stage = 0; phase ^= 1; }
        }
        umma_commit_pred(done_bar, issue);
    } else if (warp >= 4) {
        // ===================== epilogue: once, after every chunk of this CTA has been accumulated =====================
        const int ew = warp & 3;
        mbar_wait(done_bar, 0);
        tc_fence_after();
        const int krow = cout_blk * (64 * p.m_chunks) + ew * 32 + lane;
        const bool row_ok = (ew * 32 + lane) < 64 * p.m_chunks && krow < p.Kout && part < total_chunks;
        for (int lt = 0; lt < ntaps; ++lt) {
            const int tap = tap0 + lt;
#pragma unroll 1
            for (int c = 0; c < HALO_BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + lt * HALO_BN + c * 32, v);
                tmem_ld_wait();
                if (!row_ok) continue;
                float* o = p.out + (long long)krow * (9 * p.C) + tap * p.C + cin_blk * 64 + c * 32;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + q * 4), "f"(__uint_as_float(v[q * 4])),
                                 "f"(__uint_as_float(v[q * 4 + 1])), "f"(__uint_as_float(v[q * 4 + 2])), "f"(__uint_as_float(v[q * 4 + 3]))
                                 : "memory");
            }
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace pdnn

static int halo_wgrad_geometry(int H, int W, int C, int Kout, HaloWgradParams* p) {
    if (C != 64 && C != 128) return 0;
    if (Kout % 64 != 0) return 0;
    const int Wp = W + 2;
    int Hc = 0;
    for (int cand : {8, 16, 4}) {
        if (H % cand == 0 && (cand * Wp) % 16 == 0 && cand * Wp <= 512 && cand + 2 <= 256) { Hc = cand; break; }
    }
    if (Hc == 0 || Wp > 256 || W < 8) return 0;
    p->Wp = Wp; p->Hc = Hc;
    p->chunks_per_img = H / Hc;
    p->KS = Hc * Wp / 16;
    p->m_chunks = (Kout % 128 == 0) ? 2 : 1;
    p->n_cin_blocks = C / 64;
    p->n_cout_blocks = Kout / (64 * p->m_chunks);
    const int dy_tile = Hc * Wp * 128;
    p->dy_bytes = (p->m_chunks * dy_tile + 1023) / 1024 * 1024;
    p->halo_off = p->dy_bytes + 1024;                                 // 1 KiB zero guard in front of the halo
    p->stage_bytes = (p->halo_off + (Hc + 2) * Wp * 128 + 1023) / 1024 * 1024 + 1024;   // + zero guard behind it
    p->stages = (227 * 1024 - 2048) / p->stage_bytes;
    if (p->stages > 4) p->stages = 4;
    return p->stages >= 2;
}

extern "C" int pdnn_conv_halo_wgrad_supported(int batch, int H, int W, int C, int Kout, int R, int S, int pad) {
    if (R != 3 || S != 3 || pad != 1) return 0;
    HaloWgradParams p{};
    return halo_wgrad_geometry(H, W, C, Kout, &p);
}

extern "C" int pdnn_conv_halo_wgrad_bf16(const void* x, const void* dy, int batch, int H, int W, int C, int Kout, float* dw,
                                         const uint32_t* guard_closed, const uint32_t* guard_step, int max_ctas, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    HaloWgradParams p{};
    PDNN_REQUIRE(halo_wgrad_geometry(H, W, C, Kout, &p), "conv_halo_wgrad: unsupported geometry (C=%d Kout=%d %dx%d)", C, Kout, H, W);
    p.batch = batch; p.H = H; p.W = W; p.C = C; p.Kout = Kout;
    p.out = dw;
    p.guard_closed = guard_closed; p.guard_step = guard_step;
    CUtensorMap tmDY, tmX;
    {
        long long dd[4] = {Kout, W, H, batch};
        long long ds[4] = {1, Kout, (long long)W * Kout, (long long)H * W * Kout};
        int db[4] = {64, p.Wp, p.Hc, 1};
        if (int rc = pdnn_make_tmap_bf16(&tmDY, dy, 4, dd, ds, db)) return rc;
        long long xd[4] = {C, W, H, batch};
        long long xs[4] = {1, C, (long long)W * C, (long long)H * W * C};
        int xb[4] = {64, p.Wp, p.Hc + 2, 1};
        if (int rc = pdnn_make_tmap_bf16(&tmX, x, 4, xd, xs, xb)) return rc;
    }
    const int smem = p.stages * p.stage_bytes + 512 + 1024;
    static int attr_smem = 0;
    if (smem > attr_smem) {
        PDNN_CUDA_CHECK(cudaFuncSetAttribute(conv_halo_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem = smem;
    }
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    int grid = sms;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    const int combos = p.n_cin_blocks * p.n_cout_blocks;
    if (grid < 2 * combos) grid = 2 * combos;
    // filter rows 0..1 carry twice the MMA work of row 2 but the same TMA traffic. Measured (ResNet-18 b1024 step): 2 : 1
    // split of the CTAs 6.98 ms, 3 : 2 7.02 ms, 72 % 7.07 ms with the N = 192 form; the N = 64 form preferred 3 : 2.
    static const bool narrow = getenv("PDNN_HALO_WGRAD_N64") != nullptr;
    int a = (narrow ? grid * 3 / 5 : grid * 2 / 3) / combos * combos;
    if (a < combos) a = combos;
    if (grid - a < combos) a = grid - combos;
    p.ctas_group_a = a;
    p.wide = narrow ? 0 : 1;
    if (const char* e = getenv("PDNN_HALO_WGRAD_SPLIT")) {      // percentage of the CTAs on filter rows 0..1 (tuning knob)
        int a2 = grid * atoi(e) / 100 / combos * combos;
        if (a2 >= combos && grid - a2 >= combos) p.ctas_group_a = a2;
    }
    PDNN_CUDA_CHECK(pdnn_launch(conv_halo_wgrad_kernel, dim3(grid), dim3(HALO_THREADS), (size_t)smem, stream, tmDY, tmX, p));
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}
