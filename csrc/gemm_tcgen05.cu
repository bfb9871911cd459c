// tcgen05 / TMEM / TMA GEMM and implicit-GEMM convolution for sm_100a.
//
// One persistent, warp-specialised kernel family covers every GEMM-shaped op of the framework:
//   * Linear fwd / dgrad / wgrad        (replaces nn.Linear + BLAS sgemm: reference
//                                        pytorch_code/model_ops/lenet.py:45-46, resnet.py:78;
//                                        MPI_code/src/util/util.h:35-81 cblas_dgemm NN/NT/TN)
//   * 3x3 stride-1 conv fprop / dgrad   (implicit GEMM, activation tiles fetched by *tiled* 4-D TMA
//                                        with out-of-bounds zero fill acting as the padding;
//                                        reference resnet.py:19-21, resnet_split.py:28,34)
//   * 3x3 stride-1 conv wgrad           (split-K over pixels, MN-major operands, fp32 red.add
//                                        epilogue straight into the flat gradient bucket)
//   * im2col-lowered convs (stride 2, 1x1, stems) as plain GEMMs.
//
// Roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane issues
// tcgen05.mma), warp 2 = TMEM allocator, warps 4-7 = epilogue (tcgen05.ld -> registers -> global).
// Pipelines: smem ring full/empty (TMA <-> MMA), TMEM double buffer full/empty (MMA <-> epilogue),
// static persistent tile scheduler.
#include "common.cuh"
#include "../include/pdnn_b200.h"

#include <mutex>
#include <string.h>

namespace pdnn {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;   // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 256;

// OP_CONVMN_G3: wgrad B operand = the three s-shifts of one filter row side by side (N = 3 x 64 channels), so the
// dY tile is fetched once per three shifts (used when Cin <= 128, where a single-shift N tile starves the MMA).
enum OperMode : int { OP_K2D = 0, OP_MN2D = 1, OP_CONVK = 2, OP_CONVMN = 3, OP_CONVMN_G3 = 4 };

struct GemmParams {
    int M, N, K;
    int m_tiles, n_tiles, k_splits;
    int k_iters_total, k_iters_per_split;
    // geometry of the activation operand in the conv modes
    int H, W, C, R, S, pad;   // H, W: OUTPUT map (tile decode); the activation tensor map carries the input geometry
    int stride;               // conv stride (1 or 2): the TMA box walks the input with this element stride
    // tap-list mode (stride-2 dgrad by output parity): k-iteration = (tap t, 64-channel block); A box shifted by
    // (tap_dh[t], tap_dw[t]) pixels, B columns start at tap_koff[t]
    int ntaps;
    int tap_dh[4], tap_dw[4], tap_koff[4];
    // output row remap of the same mode: tile row (n, i, j) of the dy map -> dx pixel (n, 2i + par_a, 2j + par_b)
    int out_remap, par_a, par_b;
    int box_h, box_n;        // pixel box (box_w == W)
    int cblks;               // C / 64 (OP_CONVK) or C / BN (OP_CONVMN)
    // epilogue
    void* out;
    long long ldo;           // row stride of out, in elements
    int out_mode;            // 0: bf16 store, 1: fp32 store, 2: fp32 red.add (split-K / accumulate)
    int relu;
    float alpha;
    const float* bias;       // per output column, may be null
    const __nv_bfloat16* residual;  // optional bf16 [M][ldo] added before relu (bf16 out only)
    int tma_store;           // bf16 output leaves through shared memory + TMA bulk tensor stores (set by launch_gemm)
    // straggler abort guard: if *guard_closed >= *guard_step the kernel returns immediately
    const uint32_t* guard_closed;
    const uint32_t* guard_step;
};

// HALF_B: CTA-pair mode, a stage holds the A tile and only this CTA's half of the B tile -> more stages fit
template <int BN, bool HALF_B = false, bool NO_EPI = false>
struct SmemLayout {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;   // 16 KiB
    static constexpr int B_BYTES = BN * BLOCK_K * 2;        // the whole N tile (the pair MMA sees it as two halves)
    static constexpr int B_LOCAL = HALF_B ? B_BYTES / 2 : B_BYTES;
    static constexpr int STAGE_BYTES = A_BYTES + B_LOCAL;
    static constexpr int STAGES = HALF_B ? (BN <= 128 ? 8 : 6) : ((BN <= 64) ? 8 : (BN <= 128 ? 6 : 4));
    // epilogue staging for TMA stores: per epilogue warp two slabs of 32 rows x 128 bytes (64 bf16 columns, 128B-swizzled)
    static constexpr int EPI_OFFSET = STAGES * STAGE_BYTES;
    // (none for the pair schedule and for the MN-major-B instantiations = weight gradients, whose fp32 output is accumulated
    // with red.add: their smaller footprint lets a BatchNorm CTA of another stream share the SM)
    static constexpr int EPI_BYTES = (HALF_B || NO_EPI) ? 0 : 4 * 2 * 4096;
    static constexpr int BAR_OFFSET = EPI_OFFSET + EPI_BYTES;
    static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;   // barriers + alignment slack
};

// CL == 3: clusters of two CTAs that work on two M-tiles of the same N-tile as a CTA PAIR (cta_group::2). Each CTA loads its own A tile and only ITS half of the B tile;
// the leader (cluster rank 0) issues M = 256 MMAs that read both CTAs' shared memory and write each CTA's own TMEM. Every
// SM then ingests 16 KB + B/2 per k-block instead of 16 KB + B: ncu shows these kernels bound by that ingest (tensor pipe
// 51 % active on the stage-3 conv), which multicast cannot change because it still delivers all of B to every SM.
template <int AMODE, int BMODE, int BN, int CL = 1>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmO, const GemmParams p) {
    using L = SmemLayout<BN, CL == 3, BMODE != OP_K2D>;
    constexpr int STAGES = L::STAGES;
    constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));

    constexpr bool CLUSTER = (CL >= 2);       // two CTAs per cluster, tiles (t, t + 1) = two M-tiles of one N-tile
    constexpr bool PAIR = (CL == 3);          // cta_group::2
    uint32_t cta_rank = 0;
    if (CLUSTER) cta_rank = cluster_ctarank();
    if (p.guard_closed != nullptr) {
        // one thread samples the guard so that the whole CTA (the whole cluster) takes the same branch
        __shared__ uint32_t s_abort;
        if (CLUSTER) {
            if (threadIdx.x == 0 && cta_rank == 0) {
                const uint32_t v = ld_relaxed_sys(p.guard_closed) >= ld_relaxed_sys(p.guard_step) ? 1u : 0u;
                s_abort = v;
                st_shared_remote_u32(smem_u32(&s_abort), 1u, v);
            }
            cluster_sync_all();
        } else {
            if (threadIdx.x == 0) s_abort = ld_relaxed_sys(p.guard_closed) >= ld_relaxed_sys(p.guard_step) ? 1u : 0u;
            __syncthreads();
        }
        if (s_abort) return;
    }

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = smem_u32(smem);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFFSET);
    const uint32_t bar_base = smem_u32(bars);
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if (CL == 1 && p.tma_store) tma_prefetch_desc(&tmO);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), PAIR ? 8 : 4);     // pair: the leader's barrier collects both CTAs' epilogue warps
        }
        mbar_fence_init();
    }
    if (warp == 2) {
        if (PAIR) {
            tmem_alloc_2cta(smem_u32(tmem_slot), TMEM_COLS);
            tmem_relinquish_2cta();
        } else {
            tmem_alloc(smem_u32(tmem_slot), TMEM_COLS);
            tmem_relinquish();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CLUSTER) cluster_sync_all();          // the peer's barriers exist before anything is sent to them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                               // the prologue above may overlap the tail of the previous kernel (programmatic dependent launch)
    pdl_launch_dependents();

    const int total_tiles = p.m_tiles * p.n_tiles * p.k_splits;
    const int hw = p.H * p.W;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int ks = tile % p.k_splits;
                const int rest = tile / p.k_splits;
                // cluster: M-tile fastest, so that the two CTAs of a cluster (tiles t, t + 1) share the N-tile
                const int nt = CLUSTER ? rest / p.m_tiles : rest % p.n_tiles;
                const int mt = CLUSTER ? rest % p.m_tiles : rest / p.n_tiles;
                const int k_begin = ks * p.k_iters_per_split;
                const int k_end = min(k_begin + p.k_iters_per_split, p.k_iters_total);
                // per-tile constants of the conv modes
                int a_n0 = 0, a_h0 = 0;
                if (AMODE == OP_CONVK) {
                    const int p0 = mt * BLOCK_M;
                    a_n0 = p0 / hw;
                    a_h0 = (p0 - a_n0 * hw) / p.W;
                }
                int b_r = 0, b_s = 0, b_c0 = 0;
                if (BMODE == OP_CONVMN) {
                    const int rs = nt / p.cblks;
                    b_r = rs / p.S;
                    b_s = rs - b_r * p.S;
                    b_c0 = (nt - rs * p.cblks) * BN;
                }
                for (int kit = k_begin; kit < k_end; ++kit) {
                    mbar_wait(empty_bar(stage), phase ^ 1);
                    const uint32_t sa = smem_base + stage * L::STAGE_BYTES;
                    const uint32_t sb = sa + L::A_BYTES;
                    // pair: ONE barrier (the leader's) collects the bytes of both CTAs' halves of the stage; every load lands in
                    // the issuing CTA's shared memory and signals that barrier (cta_group::2 form of the TMA load)
                    // (both CTAs use the cta_group::2 form: with the leader on the plain form the convs ran 12 % slower)
                    const bool pair_ld = PAIR;
                    const uint32_t fb = PAIR ? mapa_rank(full_bar(stage), 0u) : full_bar(stage);
                    if (!PAIR || cta_rank == 0) mbar_expect_tx(full_bar(stage), PAIR ? (uint32_t)(2 * L::STAGE_BYTES) : (uint32_t)L::STAGE_BYTES);
                    // ---- A ----
                    if (AMODE == OP_K2D) {
                        if (pair_ld) tma_load_2d_2cta(sa, &tmA, fb, kit * BLOCK_K, mt * BLOCK_M);
                        else tma_load_2d(sa, &tmA, fb, kit * BLOCK_K, mt * BLOCK_M);
                    } else if (AMODE == OP_MN2D) {
                        tma_load_2d(sa, &tmA, fb, mt * BLOCK_M, kit * BLOCK_K);
                        tma_load_2d(sa + 8192, &tmA, fb, mt * BLOCK_M + 64, kit * BLOCK_K);
                    } else if (AMODE == OP_CONVK) {
                        const int rs = kit / p.cblks;
                        const int cb = kit - rs * p.cblks;
                        if (p.ntaps > 0) {
                            tma_load_4d(sa, &tmA, fb, cb * BLOCK_K, p.tap_dw[rs], a_h0 + p.tap_dh[rs], a_n0);
                        } else {
                            const int r = rs / p.S;
                            const int s = rs - r * p.S;
                            if (pair_ld) tma_load_4d_2cta(sa, &tmA, fb, cb * BLOCK_K, s - p.pad, a_h0 * p.stride + r - p.pad, a_n0);
                            else tma_load_4d(sa, &tmA, fb, cb * BLOCK_K, s - p.pad, a_h0 * p.stride + r - p.pad, a_n0);
                        }
                    }
                    // ---- B ----
                    if (BMODE == OP_K2D) {
                        int kcol = kit * BLOCK_K;
                        if (AMODE == OP_CONVK && p.ntaps > 0) {
                            const int t = kit / p.cblks;
                            kcol = p.tap_koff[t] + (kit - t * p.cblks) * BLOCK_K;
                        }
                        if (pair_ld)     // my half of the rows, into my own stage (the pair MMA reads both halves in place)
                            tma_load_2d_2cta(sb, &tmB, fb, kcol, nt * BN + (int)cta_rank * (BN / 2));
                        else
                            tma_load_2d(sb, &tmB, fb, kcol, nt * BN);
                    } else if (BMODE == OP_MN2D) {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(sb + j * 8192, &tmB, fb, nt * BN + j * 64, kit * BLOCK_K);
                    } else if (BMODE == OP_CONVMN) {
                        const int p0 = kit * BLOCK_K;   // 64 pixels per k-iteration
                        const int n0 = p0 / hw;
                        const int h0 = (p0 - n0 * hw) / p.W;
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_4d(sb + j * 8192, &tmB, fb, b_c0 + j * 64, b_s - p.pad,
                                        h0 * p.stride + b_r - p.pad, n0);
                    } else if (BMODE == OP_CONVMN_G3) {
                        const int p0 = kit * BLOCK_K;
                        const int n0 = p0 / hw;
                        const int h0 = (p0 - n0 * hw) / p.W;
                        const int r = nt / p.cblks;
                        const int cb = nt - r * p.cblks;
#pragma unroll
                        for (int j = 0; j < 3; ++j)
                            tma_load_4d(sb + j * 8192, &tmB, fb, cb * 64, j - p.pad, h0 * p.stride + r - p.pad, n0);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t A_MN = (AMODE == OP_MN2D) ? 1u : 0u;
        constexpr uint32_t B_MN = (BMODE == OP_MN2D || BMODE == OP_CONVMN || BMODE == OP_CONVMN_G3) ? 1u : 0u;
        constexpr uint32_t idesc = make_idesc_bf16(PAIR ? 2 * BLOCK_M : BLOCK_M, BN, A_MN, B_MN);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        if (PAIR && cta_rank != 0) {
            // the peer issues no MMA: the leader's M = 256 instructions read this CTA's shared memory and write its TMEM
        } else
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int ks = tile % p.k_splits;
            const int k_begin = ks * p.k_iters_per_split;
            const int k_end = min(k_begin + p.k_iters_per_split, p.k_iters_total);
            if (PAIR) mbar_wait_cluster(tempty_bar(acc), acc_phase ^ 1);
            else mbar_wait(tempty_bar(acc), acc_phase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * BN;
            for (int kit = k_begin; kit < k_end; ++kit) {
                if (PAIR) mbar_wait_cluster(full_bar(stage), phase);     // completed by both CTAs' loads
                else mbar_wait(full_bar(stage), phase);
                tc_fence_after();
                {
                    const uint32_t issue = lane == 0 ? 1u : 0u;
                    const uint32_t sa = smem_base + stage * L::STAGE_BYTES;
                    const uint32_t sb = sa + L::A_BYTES;
                    // K-major: advance 32 bytes inside the 128B swizzle row; SBO = 8 rows * 128 B.
                    // MN-major: advance 16 K-rows (2048 B); SBO = 8 K-rows (1024 B); LBO = next 64-wide MN chunk (8192 B).
                    // Only the low descriptor word changes with the start address (units of 16 bytes): one add per
                    // operand per MMA instead of rebuilding the descriptor.
                    const uint64_t a0 = A_MN ? make_smem_desc_sw128(sa, 8192, 1024) : make_smem_desc_sw128(sa, 16, 1024);
                    const uint64_t b0 = B_MN ? make_smem_desc_sw128(sb, 8192, 1024) : make_smem_desc_sw128(sb, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t adesc = a0 + (uint64_t)(k * (A_MN ? 128 : 2));
                        const uint64_t bdesc = b0 + (uint64_t)(k * (B_MN ? 128 : 2));
                        if (PAIR) umma_bf16_ss_pred_2cta(tmem_d, adesc, bdesc, idesc, (kit > k_begin || k > 0) ? 1u : 0u, issue);
                        else umma_bf16_ss_pred(tmem_d, adesc, bdesc, idesc, (kit > k_begin || k > 0) ? 1u : 0u, issue);
                    }
                    if (PAIR) {          // both CTAs' stages / accumulators are released or published by the one commit
                        umma_commit_2cta_multicast_pred(empty_bar(stage), (uint16_t)0x3, issue);
                        if (kit == k_end - 1) umma_commit_2cta_multicast_pred(tfull_bar(acc), (uint16_t)0x3, issue);
                    } else {
                        umma_commit_pred(empty_bar(stage), issue);
                        if (kit == k_end - 1) umma_commit_pred(tfull_bar(acc), issue);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int ew = warp & 3;           // TMEM lane quarter this warp may touch
        int acc = 0;
        uint32_t acc_phase = 0;
        // TMA-store path: this warp's two staging slabs; a slab is rewritten only after the bulk store that read it is done
        const bool tma_st = (CL == 1) && (BMODE == OP_K2D) && p.tma_store != 0;
        const uint32_t slab_base = smem_base + L::EPI_OFFSET + (uint32_t)ew * 8192u;
        uint32_t slab = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int rest = tile / p.k_splits;
            const int nt = CLUSTER ? rest / p.m_tiles : rest % p.n_tiles;
            const int mt = CLUSTER ? rest % p.m_tiles : rest / p.n_tiles;
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            long long row = (long long)mt * BLOCK_M + ew * 32 + lane;
            const bool row_ok = row < p.M;
            if (p.out_remap) {   // (n, i, j) on the dy map -> dx pixel (n, 2i + a, 2j + b)
                const int hw_o = p.H * p.W;
                const int n_ = (int)(row / hw_o);
                const int rem = (int)(row - (long long)n_ * hw_o);
                const int i_ = rem / p.W, j_ = rem - i_ * p.W;
                row = ((long long)n_ * (2 * p.H) + 2 * i_ + p.par_a) * (2 * p.W) + 2 * j_ + p.par_b;
            }
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + acc * BN + c * 32, v);
                tmem_ld_wait();
                int col0 = nt * BN + c * 32;
                if (BMODE == OP_CONVMN_G3) {   // tile columns = (shift s, 64 channels); output columns are KRSC
                    const int r = nt / p.cblks;
                    const int cb = nt - r * p.cblks;
                    const int j = c * 32;
                    col0 = (r * p.S + (j >> 6)) * p.C + cb * 64 + (j & 63);
                }
                if (tma_st) {
                    if (col0 >= p.N) continue;          // warp-uniform (N % 64 == 0): both halves of a 64-column store or neither
                    if ((c & 1) == 0) {                 // slab `slab` was last read by the store issued two stores ago
                        if (lane == 0) bulk_wait_group_read<1>();
                        __syncwarp();
                    }
                } else if (!row_ok || col0 >= p.N) continue;
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
                const bool full = (col0 + 32 <= p.N);
                if (p.bias != nullptr) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (full || col0 + i < p.N) f[i] += __ldg(p.bias + col0 + i);
                }
                if (p.out_mode == 0) {
                    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + row * p.ldo + col0;
                    const bool vec = full && ((p.ldo & 7) == 0);
                    // 32-byte form: rows and both base pointers on 32-byte boundaries (CTA-uniform)
                    const bool vec32 = vec && ((p.ldo & 15) == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 31) == 0) &&
                                       ((reinterpret_cast<uintptr_t>(p.residual) & 31) == 0);
                    if (p.residual != nullptr && row_ok) {
                        const __nv_bfloat16* rp = p.residual + row * p.ldo + col0;
                        if (vec32) {   // two 32-byte loads, both issued before the first use
                            const u32x8 r0 = ld_global_nc_256(rp), r1 = ld_global_nc_256(rp + 16);
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                const float2 a2 = unpack_bf16x2(r0.v[q]), b2 = unpack_bf16x2(r1.v[q]);
                                f[q * 2] += a2.x; f[q * 2 + 1] += a2.y; f[16 + q * 2] += b2.x; f[16 + q * 2 + 1] += b2.y;
                            }
                        } else if (vec) {
                            uint4 rq[4];
#pragma unroll
                            for (int q = 0; q < 4; ++q) rq[q] = __ldg(reinterpret_cast<const uint4*>(rp) + q);
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float2 r0 = unpack_bf16x2(rq[q].x), r1 = unpack_bf16x2(rq[q].y);
                                const float2 r2 = unpack_bf16x2(rq[q].z), r3 = unpack_bf16x2(rq[q].w);
                                f[q * 8 + 0] += r0.x; f[q * 8 + 1] += r0.y; f[q * 8 + 2] += r1.x; f[q * 8 + 3] += r1.y;
                                f[q * 8 + 4] += r2.x; f[q * 8 + 5] += r2.y; f[q * 8 + 6] += r3.x; f[q * 8 + 7] += r3.y;
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (full || col0 + i < p.N) f[i] += __bfloat162float(rp[i]);
                        }
                    }
                    if (p.relu) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
                    }
                    if (tma_st) {
                        // row `lane` of the slab, 16-byte chunk j of its 128 bytes lives at chunk j ^ (row & 7) (128B swizzle)
                        const uint32_t row_addr = slab_base + slab * 4096u + (uint32_t)lane * 128u;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint32_t j = (uint32_t)((c & 1) * 4 + q);
                            st_shared_v4(row_addr + ((j ^ ((uint32_t)lane & 7u)) << 4),
                                         pack_bf16x2(f[q * 8 + 0], f[q * 8 + 1]), pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]),
                                         pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]), pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]));
                        }
                        if (c & 1) {                    // 64 columns complete: hand the slab to the TMA engine
                            fence_proxy_async_shared();
                            __syncwarp();
                            if (lane == 0) {
                                tma_store_2d(&tmO, slab_base + slab * 4096u, col0 - 32, mt * BLOCK_M + ew * 32);
                                bulk_commit_group();
                            }
                            slab ^= 1u;
                        }
                    } else if (vec32) {       // one full 32-byte sector per store
                        u32x8 w0, w1;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            w0.v[q] = pack_bf16x2(f[q * 2], f[q * 2 + 1]);
                            w1.v[q] = pack_bf16x2(f[16 + q * 2], f[16 + q * 2 + 1]);
                        }
                        st_global_256(o, w0);
                        st_global_256(o + 16, w1);
                    } else if (vec) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            uint4 w;
                            w.x = pack_bf16x2(f[q * 8 + 0], f[q * 8 + 1]);
                            w.y = pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]);
                            w.z = pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]);
                            w.w = pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]);
                            *reinterpret_cast<uint4*>(o + q * 8) = w;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (col0 + i < p.N) o[i] = __float2bfloat16(f[i]);
                    }
                } else {
                    float* o = reinterpret_cast<float*>(p.out) + row * p.ldo + col0;
                    if (p.relu) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
                    }
                    const bool vec = full && ((p.ldo & 3) == 0);
                    if (p.out_mode == 1) {
                        if (vec && ((p.ldo & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 31) == 0)) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                u32x8 w;
#pragma unroll
                                for (int i = 0; i < 8; ++i) w.v[i] = __float_as_uint(f[q * 8 + i]);
                                st_global_256(o + q * 8, w);
                            }
                        } else if (vec) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                *reinterpret_cast<float4*>(o + q * 4) =
                                    make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (col0 + i < p.N) o[i] = f[i];
                        }
                    } else {
                        if (vec) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + q * 4),
                                             "f"(f[q * 4]), "f"(f[q * 4 + 1]), "f"(f[q * 4 + 2]),
                                             "f"(f[q * 4 + 3])
                                             : "memory");
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (col0 + i < p.N) atomicAdd(o + i, f[i]);
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR && cta_rank != 0) mbar_arrive_remote(tempty_bar(acc), 0u);     // the leader's barrier counts all eight warps
                else mbar_arrive(tempty_bar(acc));
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (tma_st && lane == 0) bulk_wait_group<0>();       // every store of this warp has left shared memory and landed
    }

    tc_fence_before();
    __syncthreads();
    if (CLUSTER) cluster_sync_all();          // nobody leaves while the peer can still write into it or signal it
    if (warp == 2) {
        tc_fence_after();
        if (PAIR) tmem_dealloc_2cta(tmem_base, TMEM_COLS);
        else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    });
    return fn;
}

// bf16 tensor map, up to 4 dims, dims[0] innermost (contiguous). strides in elements for dims 1..rank-1.
static int make_tmap_bf16(CUtensorMap* tm, const void* base, int rank, const long long* dims,
                          const long long* strides_elems, const int* box, const int* elem_strides = nullptr) {
    EncodeTiledFn enc = get_encode_fn();
    PDNN_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t gdim[5];
    cuuint64_t gstr[4];
    cuuint32_t bdim[5];
    cuuint32_t estr[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = (cuuint64_t)dims[i];
        bdim[i] = (cuuint32_t)box[i];
        estr[i] = elem_strides != nullptr ? (cuuint32_t)elem_strides[i] : 1;
        if (i > 0) gstr[i - 1] = (cuuint64_t)strides_elems[i] * 2ull;
    }
    PDNN_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base pointer must be 16-byte aligned");
    for (int i = 1; i < rank; ++i)
        PDNN_REQUIRE((gstr[i - 1] & 15) == 0, "TMA stride of dim %d (%llu bytes) is not a multiple of 16", i,
                     (unsigned long long)gstr[i - 1]);
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr,
                     bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    PDNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d)", (int)r, rank);
    return 0;
}

static int g_num_sms = 0;
static int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

template <int AMODE, int BMODE, int BN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int max_ctas,
                       cudaStream_t stream) {
    using L = SmemLayout<BN, false, BMODE != OP_K2D>;
    auto kern = gemm_tcgen05_kernel<AMODE, BMODE, BN>;
    static bool attr_set = false;
    if (!attr_set) {
        PDNN_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
        attr_set = true;
    }
    const int total = p.m_tiles * p.n_tiles * p.k_splits;
    int grid = total < num_sms() ? total : num_sms();
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    // bf16 outputs whose rows are whole 128-byte column groups leave as TMA bulk tensor stores of [32 rows][64 columns]
    // boxes (full lines, coalesced by the copy engine) instead of one 32-byte sector per thread and instruction
    GemmParams q = p;
    CUtensorMap tmO;
    memset(&tmO, 0, sizeof(tmO));
    static const bool no_tma_store = getenv("PDNN_NO_TMA_STORE") != nullptr && getenv("PDNN_NO_TMA_STORE")[0] == '1';
    q.tma_store = 0;
    if (!no_tma_store && p.out_mode == 0 && p.out_remap == 0 && BMODE == OP_K2D && p.N % 64 == 0 &&
        (p.ldo * 2) % 16 == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0) {
        long long dims[2] = {p.N, p.M}, str[2] = {1, p.ldo};
        int box[2] = {64, 32};
        if (int rc = make_tmap_bf16(&tmO, p.out, 2, dims, str, box)) return rc;
        q.tma_store = 1;
    }
    PDNN_CUDA_CHECK(pdnn_launch(kern, dim3(grid), dim3(NUM_THREADS), L::TOTAL, stream, tmA, tmB, tmO, q));
    PDNN_COUNT_LAUNCH();
    return 0;
}

// Cluster-of-two launch (see the kernel comment). The grid is two CTAs per cluster that can be co-resident, so that
// the persistent loop still runs as one wave.
template <int AMODE, int BMODE, int BN, int CLV>
static int launch_gemm_cluster(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int max_ctas,
                               cudaStream_t stream) {
    using L = SmemLayout<BN, CLV == 3>;
    auto kern = gemm_tcgen05_kernel<AMODE, BMODE, BN, CLV>;
    static int max_clusters = -1;
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = L::TOTAL;
    cfg.stream = stream;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (max_clusters < 0) {
        PDNN_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
        cfg.gridDim = dim3(num_sms() / 2 * 2);
        int n = 0;
        PDNN_CUDA_CHECK(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
        max_clusters = n > 0 ? n : 1;
    }
    const int total = p.m_tiles * p.n_tiles;
    int grid = 2 * max_clusters;
    if (grid > total) grid = total / 2 * 2;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas / 2 * 2;
    if (grid < 2) grid = 2;
    cfg.gridDim = dim3(grid);
    CUtensorMap tmO;                           // the pair schedule keeps the per-thread stores
    memset(&tmO, 0, sizeof(tmO));
    PDNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmO, p));
    PDNN_COUNT_LAUNCH();
    return 0;
}
// Opt-in (PDNN_CLUSTER=2): the CTA-pair schedule. A two-CTA schedule that TMA-multicast the B tile into both CTAs was
// measured exactly as fast as the single-CTA one (conv 3x3 n1024 8x8 c256 k256: 72.6 us vs 72.3 us) and has been removed:
// the operand pointers of this entry point are the caller's, and TMA multicast loads here read only this GPU's memory,
// since on B200 a TMA multicast load from another GPU's memory hung the GPU. Every TMA load of this library is unicast.
static int g_cluster_mode = -1;   // -1: not read yet; 0 off, 2 CTA pair / cta_group::2 (CL = 3)
static int want_cluster(int bn, const GemmParams& p) {
    if (g_cluster_mode < 0) {
        const char* e = getenv("PDNN_CLUSTER");
        g_cluster_mode = (e != nullptr && e[0] == '2') ? 2 : 0;
    }
    const bool ok = bn >= 128 && p.k_splits == 1 && (p.m_tiles % 2) == 0 && p.m_tiles * p.n_tiles >= 4 && p.out_remap == 0;
    return ok ? g_cluster_mode : 0;
}
template <int AMODE, int BMODE>
static int dispatch_cluster(int mode, int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int max_ctas,
                            cudaStream_t stream) {
    (void)mode;      // the CTA pair is the only cluster schedule
    if (bn == 128) return launch_gemm_cluster<AMODE, BMODE, 128, 3>(tmA, tmB, p, max_ctas, stream);
    return launch_gemm_cluster<AMODE, BMODE, 256, 3>(tmA, tmB, p, max_ctas, stream);
}

template <int AMODE, int BMODE>
static int dispatch_bn(int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int max_ctas,
                       cudaStream_t stream) {
    switch (bn) {
        case 64: return launch_gemm<AMODE, BMODE, 64>(tmA, tmB, p, max_ctas, stream);
        case 128: return launch_gemm<AMODE, BMODE, 128>(tmA, tmB, p, max_ctas, stream);
        case 256: return launch_gemm<AMODE, BMODE, 256>(tmA, tmB, p, max_ctas, stream);
        default: pdnn_set_error("unsupported BLOCK_N %d", bn); return -2;
    }
}

static int pick_bn(int N) {
    if (N <= 64) return 64;
    if (N <= 128) return 128;
    if (N % 256 == 0) return 256;
    if (N % 128 == 0) return 128;
    return N >= 512 ? 256 : 128;
}

static void fill_epilogue(GemmParams& p, const pdnn_epilogue_t* epi) {
    p.out = epi->out;
    p.ldo = epi->ldo;
    p.out_mode = epi->out_mode;
    p.relu = epi->relu;
    p.alpha = epi->alpha;
    p.bias = epi->bias;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(epi->residual);
    p.guard_closed = epi->guard_closed;
    p.guard_step = epi->guard_step;
}

static int choose_splits(int tiles_mn, int k_iters, int requested) {
    if (requested > 0) return requested > k_iters ? k_iters : requested;
    const int sms = num_sms();
    if (tiles_mn >= sms || k_iters < 8) return 1;
    int s = (sms + tiles_mn - 1) / tiles_mn;
    if (s > k_iters / 4) s = k_iters / 4;
    return s < 1 ? 1 : s;
}

static void finish_splits(GemmParams& p) {
    p.k_iters_per_split = (p.k_iters_total + p.k_splits - 1) / p.k_splits;
    p.k_splits = (p.k_iters_total + p.k_iters_per_split - 1) / p.k_iters_per_split;   // no empty split
}

}  // namespace pdnn

using namespace pdnn;

extern "C" int pdnn_set_cluster_mode(int on) {
    const int prev = pdnn::g_cluster_mode;
    pdnn::g_cluster_mode = on == 2 ? 2 : 0;
    return prev < 0 ? 0 : prev;
}

// shared with conv_halo.cu
int pdnn_make_tmap_bf16(CUtensorMap* tm, const void* base, int rank, const long long* dims, const long long* strides_elems,
                        const int* box) {
    return pdnn::make_tmap_bf16(tm, base, rank, dims, strides_elems, box);
}

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" int pdnn_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major,
                              long long ldb, int M, int N, int K, const pdnn_epilogue_t* epi, int k_splits,
                              void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
    PDNN_REQUIRE(epi != nullptr && epi->out != nullptr, "gemm: missing epilogue / output");
    const int bn = pick_bn(N);
    CUtensorMap tmA, tmB;
    {
        long long dims[2], str[2];
        int box[2];
        if (!a_mn_major) { dims[0] = K; dims[1] = M; str[1] = lda; box[0] = 64; box[1] = BLOCK_M; }
        else             { dims[0] = M; dims[1] = K; str[1] = lda; box[0] = 64; box[1] = 64; }
        str[0] = 1;
        if (int rc = make_tmap_bf16(&tmA, A, 2, dims, str, box)) return rc;
    }
    GemmParams p{};
    p.stride = 1;
    p.M = M; p.N = N; p.K = K;
    p.m_tiles = (M + BLOCK_M - 1) / BLOCK_M;
    p.n_tiles = (N + bn - 1) / bn;
    p.k_iters_total = (K + BLOCK_K - 1) / BLOCK_K;
    fill_epilogue(p, epi);
    p.k_splits = (p.out_mode == 2) ? choose_splits(p.m_tiles * p.n_tiles, p.k_iters_total, k_splits) : 1;
    finish_splits(p);
    p.H = p.W = 1;
    const int cluster = (!a_mn_major && !b_mn_major) ? want_cluster(bn, p) : 0;
    {
        long long dims[2], str[2];
        int box[2];
        str[0] = 1;
        if (!b_mn_major) { dims[0] = K; dims[1] = N; str[1] = ldb; box[0] = 64; box[1] = cluster ? bn / 2 : bn; }
        else             { dims[0] = N; dims[1] = K; str[1] = ldb; box[0] = 64; box[1] = 64; }
        if (int rc = make_tmap_bf16(&tmB, B, 2, dims, str, box)) return rc;
    }
    if (cluster) return dispatch_cluster<OP_K2D, OP_K2D>(cluster, bn, tmA, tmB, p, epi->max_ctas, stream);
    if (!a_mn_major && !b_mn_major) return dispatch_bn<OP_K2D, OP_K2D>(bn, tmA, tmB, p, epi->max_ctas, stream);
    if (!a_mn_major && b_mn_major) return dispatch_bn<OP_K2D, OP_MN2D>(bn, tmA, tmB, p, epi->max_ctas, stream);
    if (a_mn_major && b_mn_major) return dispatch_bn<OP_MN2D, OP_MN2D>(bn, tmA, tmB, p, epi->max_ctas, stream);
    pdnn_set_error("gemm: A MN-major with B K-major is not instantiated");
    return -2;
}

// Pixel box for a tile of `pixels` consecutive NHWC pixels (box_w == W).
static int pixel_box(int H, int W, int batch, int pixels, int* box_h, int* box_n) {
    const int hw = H * W;
    if (W > pixels || (pixels % W) != 0) return -1;
    if (hw >= pixels) {
        if (hw % pixels != 0) return -1;
        *box_h = pixels / W;
        *box_n = 1;
    } else {
        if (pixels % hw != 0) return -1;
        *box_h = H;
        *box_n = pixels / hw;
        if (batch % *box_n != 0) return -1;
    }
    return 0;
}

// y[n,h,w,:] = sum_{r,s,c} x[n,h+r-pad,w+s-pad,c] * wgt[k,r,s,c]   (stride 1, "same" size output)
extern "C" int pdnn_conv_fprop_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                                    int R, int S, int pad, const pdnn_epilogue_t* epi, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(C % 64 == 0, "conv_fprop: C=%d must be a multiple of 64 (use the im2col path)", C);
    PDNN_REQUIRE(2 * pad == R - 1 && 2 * pad == S - 1, "conv_fprop: only 'same' stride-1 geometry (R=%d S=%d pad=%d)", R, S, pad);
    GemmParams p{};
    p.stride = 1;
    PDNN_REQUIRE(pixel_box(H, W, batch, BLOCK_M, &p.box_h, &p.box_n) == 0,
                 "conv_fprop: %dx%d images (batch %d) do not tile into 128-pixel boxes", H, W, batch);
    const long long Mll = (long long)batch * H * W;
    PDNN_REQUIRE(Mll % BLOCK_M == 0 && Mll < (1ll << 31), "conv_fprop: pixel count must be a multiple of 128");
    const int bn = pick_bn(Kout);
    CUtensorMap tmA, tmB;
    {
        long long dims[4] = {C, W, H, batch};
        long long str[4] = {1, C, (long long)W * C, (long long)H * W * C};
        int box[4] = {64, W, p.box_h, p.box_n};
        if (int rc = make_tmap_bf16(&tmA, x, 4, dims, str, box)) return rc;
    }
    p.M = (int)Mll; p.N = Kout; p.K = R * S * C;
    p.m_tiles = p.M / BLOCK_M;
    p.n_tiles = (Kout + bn - 1) / bn;
    p.k_iters_total = R * S * (C / 64);
    p.k_splits = 1;
    finish_splits(p);
    p.H = H; p.W = W; p.C = C; p.R = R; p.S = S; p.pad = pad;
    p.cblks = C / 64;
    fill_epilogue(p, epi);
    const int cluster = want_cluster(bn, p);
    {
        long long wd[2] = {(long long)R * S * C, Kout};
        long long ws[2] = {1, (long long)R * S * C};
        int wb[2] = {64, cluster ? bn / 2 : bn};
        if (int rc = make_tmap_bf16(&tmB, wgt, 2, wd, ws, wb)) return rc;
    }
    if (cluster) return dispatch_cluster<OP_CONVK, OP_K2D>(cluster, bn, tmA, tmB, p, epi->max_ctas, stream);
    return dispatch_bn<OP_CONVK, OP_K2D>(bn, tmA, tmB, p, epi->max_ctas, stream);
}

// dW[k,r,s,c] (+)= sum_{n,h,w} dy[n,h,w,k] * x[n,h+r-pad,w+s-pad,c]; fp32 red.add into dw (KRSC).
extern "C" int pdnn_conv_wgrad_bf16(const void* x, const void* dy, int batch, int H, int W, int C, int Kout,
                                    int R, int S, int pad, const pdnn_epilogue_t* epi, int k_splits,
                                    void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(C % 64 == 0, "conv_wgrad: C=%d must be a multiple of 64", C);
    PDNN_REQUIRE(Kout % 8 == 0, "conv_wgrad: Kout=%d must be a multiple of 8", Kout);
    PDNN_REQUIRE(epi->out_mode == 2, "conv_wgrad accumulates with fp32 red.add (out_mode must be 2)");
    GemmParams p{};
    p.stride = 1;
    PDNN_REQUIRE(pixel_box(H, W, batch, BLOCK_K, &p.box_h, &p.box_n) == 0,
                 "conv_wgrad: %dx%d images (batch %d) do not tile into 64-pixel boxes", H, W, batch);
    const long long pix = (long long)batch * H * W;
    PDNN_REQUIRE(pix % BLOCK_K == 0 && pix < (1ll << 31), "conv_wgrad: pixel count must be a multiple of 64");
    const bool g3 = (S == 3 && C <= 128);
    const int bn = g3 ? 192 : ((C % 256 == 0) ? 256 : (C % 128 == 0 ? 128 : 64));
    CUtensorMap tmA, tmB;
    {
        long long ad[2] = {Kout, pix};
        long long as[2] = {1, Kout};
        int ab[2] = {64, 64};
        if (int rc = make_tmap_bf16(&tmA, dy, 2, ad, as, ab)) return rc;
        long long dims[4] = {C, W, H, batch};
        long long str[4] = {1, C, (long long)W * C, (long long)H * W * C};
        int box[4] = {64, W, p.box_h, p.box_n};
        if (int rc = make_tmap_bf16(&tmB, x, 4, dims, str, box)) return rc;
    }
    p.M = Kout; p.N = R * S * C; p.K = (int)pix;
    p.m_tiles = (Kout + BLOCK_M - 1) / BLOCK_M;
    if (g3) {
        p.cblks = C / 64;
        p.n_tiles = R * p.cblks;
    } else {
        p.cblks = C / bn;
        p.n_tiles = R * S * p.cblks;
    }
    p.k_iters_total = (int)(pix / BLOCK_K);
    p.H = H; p.W = W; p.C = C; p.R = R; p.S = S; p.pad = pad;
    fill_epilogue(p, epi);
    p.k_splits = choose_splits(p.m_tiles * p.n_tiles, p.k_iters_total, k_splits);
    finish_splits(p);
    if (g3) return launch_gemm<OP_MN2D, OP_CONVMN_G3, 192>(tmA, tmB, p, epi->max_ctas, stream);
    return dispatch_bn<OP_MN2D, OP_CONVMN>(bn, tmA, tmB, p, epi->max_ctas, stream);
}


// Output-geometry helper of the strided forms: the TMA box spans `stride * count` input elements along W / H and
// picks every stride-th one (elementStrides), so shared memory receives exactly the `count` pixels an output tile needs.
static int strided_box(int count, int stride) { return stride == 1 ? count : stride * count; }

// Strided convolution forward as implicit GEMM (no im2col): y[n,oh,ow,:] = sum x[n, oh*stride+r-pad, ow*stride+s-pad, :] w[k,r,s,:]
extern "C" int pdnn_conv_fprop_strided_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout, int R, int S,
                                            int pad, int stride, const pdnn_epilogue_t* epi, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(C % 64 == 0, "conv_fprop_strided: C=%d must be a multiple of 64", C);
    PDNN_REQUIRE(stride >= 1 && stride <= 2, "conv_fprop_strided: stride %d not supported", stride);
    const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
    GemmParams p{};
    p.stride = stride;
    PDNN_REQUIRE(pixel_box(OH, OW, batch, BLOCK_M, &p.box_h, &p.box_n) == 0,
                 "conv_fprop_strided: %dx%d outputs (batch %d) do not tile into 128-pixel boxes", OH, OW, batch);
    const long long Mll = (long long)batch * OH * OW;
    PDNN_REQUIRE(Mll % BLOCK_M == 0 && Mll < (1ll << 31), "conv_fprop_strided: pixel count must be a multiple of 128");
    PDNN_REQUIRE(strided_box(OW, stride) <= 256 && strided_box(p.box_h, stride) <= 256, "conv_fprop_strided: box too large");
    const int bn = pick_bn(Kout);
    CUtensorMap tmA, tmB;
    {
        long long dims[4] = {C, W, H, batch};
        long long str[4] = {1, C, (long long)W * C, (long long)H * W * C};
        int box[4] = {64, strided_box(OW, stride), strided_box(p.box_h, stride), p.box_n};
        int es[4] = {1, stride, stride, 1};
        if (int rc = make_tmap_bf16(&tmA, x, 4, dims, str, box, es)) return rc;
        long long wd[2] = {(long long)R * S * C, Kout};
        long long ws[2] = {1, (long long)R * S * C};
        int wb[2] = {64, bn};
        if (int rc = make_tmap_bf16(&tmB, wgt, 2, wd, ws, wb)) return rc;
    }
    p.M = (int)Mll; p.N = Kout; p.K = R * S * C;
    p.m_tiles = p.M / BLOCK_M;
    p.n_tiles = (Kout + bn - 1) / bn;
    p.k_iters_total = R * S * (C / 64);
    p.k_splits = 1;
    finish_splits(p);
    p.H = OH; p.W = OW; p.C = C; p.R = R; p.S = S; p.pad = pad;
    p.cblks = C / 64;
    fill_epilogue(p, epi);
    return dispatch_bn<OP_CONVK, OP_K2D>(bn, tmA, tmB, p, epi->max_ctas, stream);
}

// Weight gradient of the strided conv: dw[k,r,s,c] += sum dy[n,oh,ow,k] x[n, oh*stride+r-pad, ow*stride+s-pad, c]
extern "C" int pdnn_conv_wgrad_strided_bf16(const void* x, const void* dy, int batch, int H, int W, int C, int Kout, int R, int S,
                                            int pad, int stride, const pdnn_epilogue_t* epi, int k_splits, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(C % 64 == 0 && Kout % 8 == 0, "conv_wgrad_strided: C=%d / Kout=%d not supported", C, Kout);
    PDNN_REQUIRE(epi->out_mode == 2, "conv_wgrad_strided accumulates with fp32 red.add (out_mode must be 2)");
    PDNN_REQUIRE(stride >= 1 && stride <= 2, "conv_wgrad_strided: stride %d not supported", stride);
    const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
    GemmParams p{};
    p.stride = stride;
    PDNN_REQUIRE(pixel_box(OH, OW, batch, BLOCK_K, &p.box_h, &p.box_n) == 0,
                 "conv_wgrad_strided: %dx%d outputs (batch %d) do not tile into 64-pixel boxes", OH, OW, batch);
    const long long pix = (long long)batch * OH * OW;
    PDNN_REQUIRE(pix % BLOCK_K == 0 && pix < (1ll << 31), "conv_wgrad_strided: pixel count must be a multiple of 64");
    const bool g3 = (S == 3 && C <= 128);
    const int bn = g3 ? 192 : ((C % 256 == 0) ? 256 : (C % 128 == 0 ? 128 : 64));
    CUtensorMap tmA, tmB;
    {
        long long ad[2] = {Kout, pix};
        long long as[2] = {1, Kout};
        int ab[2] = {64, 64};
        if (int rc = make_tmap_bf16(&tmA, dy, 2, ad, as, ab)) return rc;
        long long dims[4] = {C, W, H, batch};
        long long str[4] = {1, C, (long long)W * C, (long long)H * W * C};
        int box[4] = {64, strided_box(OW, stride), strided_box(p.box_h, stride), p.box_n};
        int es[4] = {1, stride, stride, 1};
        if (int rc = make_tmap_bf16(&tmB, x, 4, dims, str, box, es)) return rc;
    }
    p.M = Kout; p.N = R * S * C; p.K = (int)pix;
    p.m_tiles = (Kout + BLOCK_M - 1) / BLOCK_M;
    if (g3) { p.cblks = C / 64; p.n_tiles = R * p.cblks; }
    else { p.cblks = C / bn; p.n_tiles = R * S * p.cblks; }
    p.k_iters_total = (int)(pix / BLOCK_K);
    p.H = OH; p.W = OW; p.C = C; p.R = R; p.S = S; p.pad = pad;
    fill_epilogue(p, epi);
    p.k_splits = choose_splits(p.m_tiles * p.n_tiles, p.k_iters_total, k_splits);
    finish_splits(p);
    if (g3) return launch_gemm<OP_MN2D, OP_CONVMN_G3, 192>(tmA, tmB, p, epi->max_ctas, stream);
    return dispatch_bn<OP_MN2D, OP_CONVMN>(bn, tmA, tmB, p, epi->max_ctas, stream);
}


// Data gradient of a stride-2 convolution without a col2im pass. For stride 2 every input pixel (2i+a, 2j+b) receives
// contributions only from the filter taps whose parity matches (a, b): four small stride-1 convolutions over dy, one per
// output parity class, each with 1, 2, 2 or 4 taps, written to the interleaved positions of dx. wt is the
// pdnn_wgt_flip_transpose()d filter bank [C][R][S][Kout] (tap (r, s) sits at ((R-1-r)*S + (S-1-s)) * Kout).
// dy: [batch][OH][OW][Kout], dx: [batch][2*OH][2*OW][C]. R = S = 3 (pad 1) or R = S = 1 (pad 0).
extern "C" int pdnn_conv_dgrad_s2_bf16(const void* dy, const void* wt, int batch, int OH, int OW, int Kout, int C, int R, int S,
                                       int pad, const pdnn_epilogue_t* epi, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    PDNN_REQUIRE(Kout % 64 == 0, "conv_dgrad_s2: Kout=%d must be a multiple of 64", Kout);
    PDNN_REQUIRE((R == 3 && S == 3 && pad == 1) || (R == 1 && S == 1 && pad == 0), "conv_dgrad_s2: 3x3/pad1 or 1x1/pad0 only");
    PDNN_REQUIRE(epi != nullptr && epi->out != nullptr && epi->out_mode == 0, "conv_dgrad_s2: bf16 output required");
    GemmParams base{};
    base.stride = 1;
    PDNN_REQUIRE(pixel_box(OH, OW, batch, BLOCK_M, &base.box_h, &base.box_n) == 0,
                 "conv_dgrad_s2: %dx%d maps (batch %d) do not tile into 128-pixel boxes", OH, OW, batch);
    const long long Mll = (long long)batch * OH * OW;
    PDNN_REQUIRE(Mll % BLOCK_M == 0 && Mll < (1ll << 31), "conv_dgrad_s2: pixel count must be a multiple of 128");
    const int bn = pick_bn(C);
    CUtensorMap tmA, tmB;
    {
        long long dims[4] = {Kout, OW, OH, batch};
        long long str[4] = {1, Kout, (long long)OW * Kout, (long long)OH * OW * Kout};
        int box[4] = {64, OW, base.box_h, base.box_n};
        if (int rc = make_tmap_bf16(&tmA, dy, 4, dims, str, box)) return rc;
        long long wd[2] = {(long long)R * S * Kout, C};
        long long ws[2] = {1, (long long)R * S * Kout};
        int wb[2] = {64, bn};
        if (int rc = make_tmap_bf16(&tmB, wt, 2, wd, ws, wb)) return rc;
    }
    if (R == 1) {
        PDNN_REQUIRE(epi->residual == nullptr, "conv_dgrad_s2: a residual is only supported for the 3x3 form (every pixel is written)");
        // only the (even, even) pixels receive a gradient: everything else is zero
        PDNN_CUDA_CHECK(cudaMemsetAsync(epi->out, 0, sizeof(__nv_bfloat16) * (size_t)batch * 2 * OH * 2 * OW * (size_t)epi->ldo, stream));
    }
    for (int a = 0; a < 2; ++a) {
        for (int b = 0; b < 2; ++b) {
            GemmParams p = base;
            // taps of parity (a, b): h = 2*oh + r - pad  =>  oh = i + (a + pad - r) / 2 for h = 2i + a
            int nt = 0;
            for (int r = 0; r < R; ++r) {
                if (((a + pad - r) & 1) != 0) continue;
                for (int s2 = 0; s2 < S; ++s2) {
                    if (((b + pad - s2) & 1) != 0) continue;
                    p.tap_dh[nt] = (a + pad - r) / 2;
                    p.tap_dw[nt] = (b + pad - s2) / 2;
                    p.tap_koff[nt] = ((R - 1 - r) * S + (S - 1 - s2)) * Kout;
                    ++nt;
                }
            }
            if (nt == 0) continue;
            p.ntaps = nt;
            p.M = (int)Mll; p.N = C; p.K = nt * Kout;
            p.m_tiles = p.M / BLOCK_M;
            p.n_tiles = (C + bn - 1) / bn;
            p.cblks = Kout / 64;
            p.k_iters_total = nt * p.cblks;
            p.k_splits = 1;
            finish_splits(p);
            p.H = OH; p.W = OW; p.C = Kout; p.R = R; p.S = S; p.pad = pad;
            fill_epilogue(p, epi);
            p.out_remap = 1; p.par_a = a; p.par_b = b;
            if (int rc = dispatch_bn<OP_CONVK, OP_K2D>(bn, tmA, tmB, p, epi->max_ctas, stream)) return rc;
        }
    }
    return 0;
}
