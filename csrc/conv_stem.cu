// Stem convolution (3x3, stride 1, pad 1, C <= 3 input channels, 64 output channels) for sm_100a.
//
// The first layer of the CIFAR networks (reference pytorch_code/model_ops/resnet.py:71 `conv1 = nn.Conv2d(3, 64, 3, 1, 1)`,
// vgg.py:30-37 first entry of cfg) has K = 27: far too thin for a TMA-fed implicit GEMM (a 54-byte row is not a TMA
// box), and the two-pass form (im2col kernel + GEMM over a 64-byte-row matrix) measured 133 us + 87 us at batch 1024
// where the output write alone is 21 us of HBM time. This kernel builds the im2col rows in registers instead:
//
//   warps 0-3  gather: thread t owns pixel t of the 128-pixel tile, reads its 3x3xC patch (L1-resident: neighbours
//              share 8/9 of it), packs K = 27 (+5 zeros) bf16 into four 16-byte chunks and writes them at the
//              128B-swizzled positions of row t of the A stage (the XOR is a function of the absolute shared-memory
//              address, as for TMA); fence.proxy.async; arrive on the stage's mbarrier (count 128). Optionally the same
//              64 bytes go to the global im2col matrix the weight-gradient GEMM reads.
//   warp 8     issues two tcgen05.mma (128x64x16) per tile against the filter bank (8 KB, loaded once by TMA);
//              accumulators double-buffered in TMEM.
//   warps 4-7  epilogue: tcgen05.ld, bias / ReLU, bf16 pack, 32-byte stores; per-thread column sums over all tiles of
//              the CTA give the BatchNorm statistics of the output (one cross-lane reduction per CTA at the end).
#include "common.cuh"
#include "../include/pdnn_b200.h"

namespace pdnn {

constexpr int STEM_THREADS = 288;
constexpr int STEM_STAGES = 4;
constexpr int STEM_A_BYTES = 128 * 128;     // 128 pixels x 128-byte swizzle rows (only the first 64 bytes carry K)
constexpr int STEM_B_BYTES = 64 * 128;      // 64 filters x 128-byte rows
constexpr int STEM_N = 64;
constexpr int STEM_KPAD = 32;

struct StemParams {
    const __nv_bfloat16* x;     // [batch, H, W, C]
    __nv_bfloat16* col;         // optional [M, 32] im2col matrix for the weight gradient
    void* out;                  // bf16 [M, ldo]
    const float* bias;
    float* bn_stats;            // optional [2][64]
    int batch, H, W, C;
    int ldo, relu;
    int total_tiles;
    const uint32_t* guard_closed;
    const uint32_t* guard_step;
};

__global__ void __launch_bounds__(STEM_THREADS, 1)
conv_stem_kernel(const __grid_constant__ CUtensorMap tmW, const StemParams p) {
    if (p.guard_closed != nullptr) {
        __shared__ int s_abort;
        if (threadIdx.x == 0) s_abort = ld_relaxed_sys(p.guard_closed) >= ld_relaxed_sys(p.guard_step);
        __syncthreads();
        if (s_abort) return;
    }
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t sB = smem_base + STEM_STAGES * STEM_A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STEM_STAGES * STEM_A_BYTES + STEM_B_BYTES);
    const uint32_t bar_base = smem_u32(bars);
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STEM_STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STEM_STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STEM_STAGES + 2 + a); };
    const uint32_t wfull_bar = bar_base + 8u * (2 * STEM_STAGES + 4);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STEM_STAGES + 5);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 8) {
        if (lane == 0) {
            tma_prefetch_desc(&tmW);
            for (int s = 0; s < STEM_STAGES; ++s) {
                mbar_init(full_bar(s), 128);
                mbar_init(empty_bar(s), 1);
            }
            for (int a = 0; a < 2; ++a) {
                mbar_init(tfull_bar(a), 1);
                mbar_init(tempty_bar(a), 4);
            }
            mbar_init(wfull_bar, 1);
            mbar_fence_init();
        }
        __syncwarp();
        tmem_alloc(smem_u32(tmem_slot), 128);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                 // everything above may overlap the tail of the previous kernel (programmatic dependent launch)
    pdl_launch_dependents();

    if (warp < 4) {
        // ===================== gather producers =====================
        const int t = threadIdx.x;                  // row of the tile
        const int hw = p.H * p.W;
        const int C = p.C;
        int stage = 0;
        uint32_t phase = 0;
        // the patch of the NEXT tile is fetched before the current one is packed and stored: the L2 latency of the
        // three image rows a tile touches overlaps with the stage hand-over instead of adding to it
        auto fetch = [&](int tile, unsigned short (&vals)[STEM_KPAD]) {
#pragma unroll
            for (int i = 0; i < STEM_KPAD; ++i) vals[i] = 0;
            if (tile >= p.total_tiles) return;
            const long long m = (long long)tile * 128 + t;
            const int n = (int)(m / hw);
            const int rem = (int)(m - (long long)n * hw);
            const int oh = rem / p.W, ow = rem - oh * p.W;
            // k = (r * 3 + s) * C + c, the KRSC order of the filters; 27 values + zeros up to 32
            const __nv_bfloat16* img = p.x + (long long)n * hw * C;
            if (C == 3) {
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const int h = oh + r - 1;
                    const bool hok = h >= 0 && h < p.H;
#pragma unroll
                    for (int s = 0; s < 3; ++s) {
                        const int w = ow + s - 1;
                        if (hok && w >= 0 && w < p.W) {
                            const unsigned short* px = reinterpret_cast<const unsigned short*>(img + ((long long)h * p.W + w) * 3);
                            vals[(r * 3 + s) * 3 + 0] = __ldg(px);
                            vals[(r * 3 + s) * 3 + 1] = __ldg(px + 1);
                            vals[(r * 3 + s) * 3 + 2] = __ldg(px + 2);
                        }
                    }
                }
            } else {
                for (int rs = 0; rs < 9; ++rs) {
                    const int h = oh + rs / 3 - 1, w = ow + rs % 3 - 1;
                    if (h >= 0 && h < p.H && w >= 0 && w < p.W) {
                        const unsigned short* px = reinterpret_cast<const unsigned short*>(img + ((long long)h * p.W + w) * C);
                        for (int c = 0; c < C; ++c) {
                            const unsigned short v = __ldg(px + c);
                            const int k = rs * C + c;
#pragma unroll
                            for (int i = 0; i < STEM_KPAD; ++i)
                                if (i == k) vals[i] = v;        // keeps vals[] in registers
                        }
                    }
                }
            }
        };
        unsigned short nxt[STEM_KPAD];
        fetch(blockIdx.x, nxt);
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
            const long long m = (long long)tile * 128 + t;
            uint32_t packed[16];
            unsigned short vals[STEM_KPAD];
#pragma unroll
            for (int i = 0; i < STEM_KPAD; ++i) vals[i] = nxt[i];
            fetch(tile + (int)gridDim.x, nxt);
#pragma unroll
            for (int i = 0; i < 16; ++i) packed[i] = (uint32_t)vals[2 * i] | ((uint32_t)vals[2 * i + 1] << 16);
            mbar_wait(empty_bar(stage), phase ^ 1);
            const uint32_t row = smem_base + stage * STEM_A_BYTES + t * 128;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t dst = row + (uint32_t)((j ^ (t & 7)) << 4);
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(packed[4 * j]), "r"(packed[4 * j + 1]),
                             "r"(packed[4 * j + 2]), "r"(packed[4 * j + 3])
                             : "memory");
            }
            fence_proxy_async_smem();
            mbar_arrive(full_bar(stage));
            if (p.col != nullptr) {
                u32x8 c0, c1;
#pragma unroll
                for (int i = 0; i < 8; ++i) { c0.v[i] = packed[i]; c1.v[i] = packed[8 + i]; }
                __nv_bfloat16* cp = p.col + m * STEM_KPAD;
                st_global_256(cp, c0);
                st_global_256(cp + 16, c1);
            }
            if (++stage == STEM_STAGES) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 8) {
        // ===================== filter load + MMA issuer =====================
        if (lane == 0) {
            mbar_expect_tx(wfull_bar, STEM_B_BYTES);
            tma_load_2d(sB, &tmW, wfull_bar, 0, 0);
        }
        mbar_wait(wfull_bar, 0);
        constexpr uint32_t idesc = make_idesc_bf16(128, STEM_N, 0u, 0u);
        const uint64_t b0 = make_smem_desc_sw128(sB, 16, 1024);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
            mbar_wait(tempty_bar(acc), acc_phase ^ 1);
            mbar_wait(full_bar(stage), phase);
            tc_fence_after();
            if (lane == 0) {
                const uint64_t a0 = make_smem_desc_sw128(smem_base + stage * STEM_A_BYTES, 16, 1024);
                const uint32_t tmem_d = tmem_base + acc * STEM_N;
                umma_bf16_ss(tmem_d, a0, b0, idesc, 0u);
                umma_bf16_ss(tmem_d, a0 + 2, b0 + 2, idesc, 1u);
                umma_commit(empty_bar(stage));
                umma_commit(tfull_bar(acc));
            }
            __syncwarp();
            if (++stage == STEM_STAGES) { stage = 0; phase ^= 1; }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int ew = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        float st_sum[2][32], st_sq[2][32];
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) { st_sum[c][i] = 0.f; st_sq[c][i] = 0.f; }
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
            const long long row = (long long)tile * 128 + ew * 32 + lane;
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + acc * STEM_N + c * 32, v);
                tmem_ld_wait();
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
                if (p.bias != nullptr) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) f[i] += __ldg(p.bias + c * 32 + i);
                }
                if (p.relu) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
                }
                u32x8 w0, w1;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    w0.v[q] = pack_bf16x2(f[q * 2], f[q * 2 + 1]);
                    w1.v[q] = pack_bf16x2(f[16 + q * 2], f[16 + q * 2 + 1]);
                }
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + row * p.ldo + c * 32;
                st_global_256(o, w0);
                st_global_256(o + 16, w1);
                if (p.bn_stats != nullptr) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float r = __bfloat162float(__float2bfloat16(f[i]));     // statistics of what was stored
                        st_sum[c][i] += r;
                        st_sq[c][i] = fmaf(r, r, st_sq[c][i]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (p.bn_stats != nullptr) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const float s1 = warp_column_sums(st_sum[c], lane);
                const float s2 = warp_column_sums(st_sq[c], lane);
                atomicAdd(p.bn_stats + c * 32 + lane, s1);
                atomicAdd(p.bn_stats + STEM_N + c * 32 + lane, s2);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

}  // namespace pdnn

using namespace pdnn;

int pdnn_make_tmap_bf16(CUtensorMap* tm, const void* base, int rank, const long long* dims, const long long* strides_elems,
                        const int* box);   // gemm_tcgen05.cu

extern "C" int pdnn_conv_stem_supported(int batch, int H, int W, int C, int Kout, int R, int S, int stride, int pad) {
    if (R != 3 || S != 3 || stride != 1 || pad != 1) return 0;
    if (C < 1 || 9 * C > STEM_KPAD || Kout != STEM_N) return 0;
    const long long m = (long long)batch * H * W;
    return (m % 128 == 0 && m / 128 < (1ll << 31)) ? 1 : 0;
}

// y = conv3x3(x, w) (+ bias, ReLU); w_pad: [64][32] bf16 = KRSC filters, rows zero-padded from 9*C to 32.
// col (optional): receives the [M][32] im2col matrix. bn_stats (optional, [2][64] fp32, zeroed by the caller) receives
// the per-channel sum / sum of squares of the stored output.
extern "C" int pdnn_conv_stem_bf16(const void* x, const void* w_pad, int batch, int H, int W, int C, int Kout,
                                   const pdnn_epilogue_t* epi, void* col, float* bn_stats, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(pdnn_conv_stem_supported(batch, H, W, C, Kout, 3, 3, 1, 1), "conv_stem: unsupported geometry (C=%d Kout=%d %dx%d batch %d)",
                 C, Kout, H, W, batch);
    PDNN_REQUIRE(epi != nullptr && epi->out != nullptr && epi->out_mode == 0 && epi->residual == nullptr,
                 "conv_stem: bf16 output without residual required");
    PDNN_REQUIRE(epi->ldo % 16 == 0 && (reinterpret_cast<uintptr_t>(epi->out) & 31) == 0 && (reinterpret_cast<uintptr_t>(col) & 31) == 0,
                 "conv_stem: ldo must be a multiple of 16 and out / col 32-byte aligned");
    StemParams p{};
    p.x = reinterpret_cast<const __nv_bfloat16*>(x);
    p.col = reinterpret_cast<__nv_bfloat16*>(col);
    p.out = epi->out; p.ldo = epi->ldo; p.relu = epi->relu; p.bias = epi->bias;
    p.bn_stats = bn_stats;
    p.batch = batch; p.H = H; p.W = W; p.C = C;
    p.total_tiles = (int)((long long)batch * H * W / 128);
    p.guard_closed = epi->guard_closed; p.guard_step = epi->guard_step;
    CUtensorMap tmW;
    {
        long long wd[2] = {STEM_KPAD, Kout};
        long long ws[2] = {1, STEM_KPAD};
        int wb[2] = {64, 64};                 // the 32 columns past the tensor are zero-filled by TMA
        if (int rc = pdnn_make_tmap_bf16(&tmW, w_pad, 2, wd, ws, wb)) return rc;
    }
    const int smem = STEM_STAGES * STEM_A_BYTES + STEM_B_BYTES + 256 + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        PDNN_CUDA_CHECK(cudaFuncSetAttribute(conv_stem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int grid = p.total_tiles < sms ? p.total_tiles : sms;
    if (epi->max_ctas > 0 && grid > epi->max_ctas) grid = epi->max_ctas;
    PDNN_CUDA_CHECK(pdnn_launch(conv_stem_kernel, dim3(grid), dim3(STEM_THREADS), (size_t)smem, stream, tmW, p));
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}
