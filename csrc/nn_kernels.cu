// Memory-bound layer kernels for sm_100a: casts/layout, im2col, BatchNorm, ReLU, pooling,
// softmax-cross-entropy, top-k, momentum SGD, sigmoid family and a generic SIMT fp32 GEMM.
// Activations are NHWC; `T` is __nv_bfloat16 on the tensor-core path and float on the small-shape
// (LeNet / FC-MLP) path. All kernels use 16-byte accesses where the layout allows it and grid-stride
// loops sized to the 148 SMs.
//
// What each family stands in for in the reference (which runs them through torch 0.2 CPU autograd, numpy or BLAS):
//   BatchNorm fwd/bwd      nn.BatchNorm2d in pytorch_code/model_ops/resnet.py:20,22,27-29 (training statistics, running buffers)
//   ReLU / max / avg pool  F.relu, F.max_pool2d in model_ops/lenet.py:27-33; F.avg_pool2d(out, 4) in resnet.py:94
//   softmax cross-entropy  nn.CrossEntropyLoss in nn_ops/__init__.py:43; Softmax + LogDot in MPI_code/src/util/util.h:125-144;
//                          SoftmaxOutputLayer in pure_py_code/nn/nn_layer.py:128-145
//   top-k                  accuracy() in nn_ops/__init__.py:13-26
//   momentum SGD           optim.SGD in nn_ops/__init__.py:42 and update_params_dist_version, sync_replicas_master_nn.py:22-28
//   sigmoid family         SigmoidLayer in pure_py_code/nn/nn_layer.py:113-125; Sigmoid / SigmoidDeriv in util.h:101-123
//   SIMT fp32 GEMM         cblas_dgemm wrappers MatrixMultiply / TransA / TransB in util.h:35-81; numpy dot in nn_layer.py:79-100
//   im2col / col2im        the lowering torch's CPU conv uses internally for lenet.py:17-18 and the ResNet stems / shortcuts
#include "common.cuh"
#include "../include/pdnn_b200.h"

namespace pdnn {

static inline int grid_for(long long work_items, int threads, int max_blocks = 148 * 8) {
    long long b = (work_items + threads - 1) / threads;
    if (b < 1) b = 1;
    if (b > max_blocks) b = max_blocks;
    return (int)b;
}

// Straggler abort guard of the layer kernels (BatchNorm, pooling, element-wise, loss): registered per host thread by
// pdnn_set_abort_guard, passed to every launch; a CTA returns at once when *closed >= *step (the master closed the step
// this worker is still computing). The decision is taken by one thread per CTA so that no __syncthreads() can be split.
struct NnGuard { const uint32_t* closed; const uint32_t* step; };
static thread_local NnGuard t_nn_guard = {nullptr, nullptr};
static inline NnGuard nn_guard() { return t_nn_guard; }
#define NN_GUARD_RETURN(g)                                                                         \
    do {                                                                                           \
        if ((g).closed != nullptr) {                                                               \
            __shared__ int s_nn_abort;                                                             \
            if (threadIdx.x == 0) s_nn_abort = ld_relaxed_sys((g).closed) >= ld_relaxed_sys((g).step); \
            __syncthreads();                                                                       \
            if (s_nn_abort) return;                                                                \
        }                                                                                          \
    } while (0)

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

// 8 consecutive bf16 <-> 8 floats
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
// raw 16-byte forms: the BatchNorm passes issue a batch of these before touching any of them, so that every
// thread keeps several loads in flight (one load per thread leaves HBM3e half idle at 1-2 K threads per SM)
// streamed once: keep it out of L1 (measured: statistics pass 4.9 -> 5.7 TB/s on a 537 MB tensor)
__device__ __forceinline__ uint4 ldg16(const __nv_bfloat16* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
#ifndef PDNN_BN_UNROLL
#define PDNN_BN_UNROLL 4
#endif
// Row schedule of the BatchNorm passes. Default: grid-stride (all blocks together sweep one contiguous window per
// iteration). PDNN_BN_SPAN: every block streams through its own contiguous span of rows.
#ifdef PDNN_BN_SPAN
#define BN_ROW_RANGE(U)                                                                                                   \
    const long long span_ = (((M + gridDim.x - 1) / gridDim.x) + (long long)rpi * (U) - 1) / ((long long)rpi * (U)) * ((long long)rpi * (U)); \
    const long long row0 = (long long)blockIdx.x * span_ + rgrp;                                                           \
    const long long rend = (M < ((long long)blockIdx.x + 1) * span_) ? M : ((long long)blockIdx.x + 1) * span_;            \
    const long long rstride = rpi;
#else
#define BN_ROW_RANGE(U)                                           \
    const long long row0 = (long long)blockIdx.x * rpi + rgrp;   \
    const long long rend = M;                                    \
    const long long rstride = (long long)gridDim.x * rpi;
#endif
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
    uint4 u;
    u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
    u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
    *reinterpret_cast<uint4*>(p) = u;
}

// ------------------------------------------------------------------------------------------------
// casts and layout
// ------------------------------------------------------------------------------------------------
__global__ void cast_f32_bf16_kernel(const float* __restrict__ s, __nv_bfloat16* __restrict__ d, long long n) {
    long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    const long long stride = (long long)gridDim.x * blockDim.x * 8;
    for (; i + 8 <= n; i += stride) {
        float4 a = *reinterpret_cast<const float4*>(s + i);
        float4 b = *reinterpret_cast<const float4*>(s + i + 4);
        uint4 u;
        u.x = pack_bf16x2(a.x, a.y); u.y = pack_bf16x2(a.z, a.w);
        u.z = pack_bf16x2(b.x, b.y); u.w = pack_bf16x2(b.z, b.w);
        *reinterpret_cast<uint4*>(d + i) = u;
    }
    if (i < n) for (long long j = i; j < n && j < i + 8; ++j) d[j] = __float2bfloat16(s[j]);
}
__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ s, float* __restrict__ d, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) d[i] = __bfloat162float(s[i]);
}

template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, T* __restrict__ y, int N, int C, int H, int W, int Cpad) {
    const long long total = (long long)N * H * W * Cpad;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % Cpad);
        const long long pix = i / Cpad;
        const int w = (int)(pix % W);
        const int h = (int)((pix / W) % H);
        const int n = (int)(pix / ((long long)W * H));
        float v = 0.f;
        if (c < C) v = x[(((long long)n * C + c) * H + h) * W + w];
        y[i] = from_f<T>(v);
    }
}

// image staging (C <= 4, no padding): one thread per pixel -> every channel plane is read with full 128-byte warps and the
// pixel's channels leave as one contiguous group
template <typename T>
__global__ void nchw_to_nhwc_pixel_kernel(const float* __restrict__ x, T* __restrict__ y, int N, int C, long long HW, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const long long total = (long long)N * HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / HW;
        const long long p = i - n * HW;
        const float* src = x + n * C * HW + p;
        T* dst = y + i * C;
        for (int c = 0; c < C; ++c) dst[c] = from_f<T>(__ldg(src + (long long)c * HW));
    }
}

// wt[c][R-1-r][S-1-s][k] = w[k][r][s][c]
__global__ void wgt_flip_transpose_kernel(const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ wt, int K, int R, int S, int C) {
    const long long total = (long long)K * R * S * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        // i indexes the destination so that writes are coalesced
        const int k = (int)(i % K);
        const int s2 = (int)((i / K) % S);
        const int r2 = (int)((i / ((long long)K * S)) % R);
        const int c = (int)(i / ((long long)K * S * R));
        wt[i] = w[(((long long)k * R + (R - 1 - r2)) * S + (S - 1 - s2)) * C + c];
    }
}

// All convolutions of a network in one launch: blockIdx.y picks the layer from a device table of
// {source offset in the flat weight buffer, destination address, K, R, S, C} (six 64-bit words per layer).
// wt[c][R-1-r][S-1-s][k] = w[k][r][s][c] is a (k, c) transpose per filter tap: 32x32 tiles go through shared memory so
// that both the read (along c) and the write (along k) are coalesced (the element-wise form read one 2-byte value per
// 128-byte line and took 244 us for ResNet-18's 11 M parameters; this one is bound by 45 MB of traffic).
__global__ void __launch_bounds__(256) wgt_flip_transpose_multi_kernel(const __nv_bfloat16* __restrict__ w_flat, const long long* __restrict__ table) {
    __shared__ __nv_bfloat16 tile[32][34];
    const long long* e = table + 6 * blockIdx.y;
    const __nv_bfloat16* w = w_flat + e[0];
    __nv_bfloat16* wt = reinterpret_cast<__nv_bfloat16*>(e[1]);
    const int K = (int)e[2], R = (int)e[3], S = (int)e[4], C = (int)e[5];
    const int RS = R * S;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
    if ((K & 31) == 0 && (C & 31) == 0) {
        const int kt = K / 32, ct = C / 32;
        const int tiles = RS * kt * ct;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
            const int c0 = (t % ct) * 32;
            const int k0 = ((t / ct) % kt) * 32;
            const int rs = t / (ct * kt);
            const int rs_m = RS - 1 - rs;                         // (R-1-r) * S + (S-1-s)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int kk = ty + j * 8;
                tile[kk][tx] = w[((long long)(k0 + kk) * RS + rs) * C + c0 + tx];
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int cc = ty + j * 8;
                wt[((long long)(c0 + cc) * RS + rs_m) * K + k0 + tx] = tile[tx][cc];
            }
            __syncthreads();
        }
        return;
    }
    const long long total = (long long)K * RS * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % K);
        const int s2 = (int)((i / K) % S);
        const int r2 = (int)((i / ((long long)K * S)) % R);
        const int c = (int)(i / ((long long)K * S * R));
        wt[i] = w[(((long long)k * R + (R - 1 - r2)) * S + (S - 1 - s2)) * C + c];
    }
}

__global__ void oihw_krsc_kernel(const float* __restrict__ src, float* __restrict__ dst, int K, int C, int R, int S, int to_krsc) {
    const long long total = (long long)K * C * R * S;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        // i indexes KRSC
        const int c = (int)(i % C);
        const int s = (int)((i / C) % S);
        const int r = (int)((i / ((long long)C * S)) % R);
        const int k = (int)(i / ((long long)C * S * R));
        const long long j = (((long long)k * C + c) * R + r) * S + s;   // OIHW
        if (to_krsc) dst[i] = src[j]; else dst[j] = src[i];
    }
}

// ------------------------------------------------------------------------------------------------
// im2col / col2im (lowered convolutions)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void im2col_kernel(const T* __restrict__ x, T* __restrict__ col, int N, int H, int W, int C, int R, int S,
                              int stride, int pad, int OH, int OW, int Kpad) {
    const long long total = (long long)N * OH * OW * Kpad;
    const int K = R * S * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % Kpad);
        const long long m = i / Kpad;
        T v = from_f<T>(0.f);
        if (k < K) {
            const int c = k % C;
            const int rs = k / C;
            const int s = rs % S, r = rs / S;
            const int ow = (int)(m % OW);
            const int oh = (int)((m / OW) % OH);
            const int n = (int)(m / ((long long)OW * OH));
            const int h = oh * stride - pad + r, w = ow * stride - pad + s;
            if (h >= 0 && h < H && w >= 0 && w < W) v = x[(((long long)n * H + h) * W + w) * C + c];
        }
        col[i] = v;
    }
}
// vectorised variant: C % 8 == 0, Kpad == R*S*C, bf16
__global__ void im2col_vec8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, int N, int H, int W,
                                   int C, int R, int S, int stride, int pad, int OH, int OW, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const int C8 = C / 8;
    const long long total = (long long)N * OH * OW * R * S * C8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % C8);
        long long t = i / C8;
        const int rs = (int)(t % (R * S));
        const long long m = t / (R * S);
        const int s = rs % S, r = rs / S;
        const int ow = (int)(m % OW);
        const int oh = (int)((m / OW) % OH);
        const int n = (int)(m / ((long long)OW * OH));
        const int h = oh * stride - pad + r, w = ow * stride - pad + s;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (h >= 0 && h < H && w >= 0 && w < W)
            v = *reinterpret_cast<const uint4*>(x + (((long long)n * H + h) * W + w) * C + c8 * 8);
        *reinterpret_cast<uint4*>(col + (m * (R * S) + rs) * C + c8 * 8) = v;
    }
}

// stem-style im2col (C < 8): one thread produces 8 consecutive k of one row with a single 16-byte store.
__global__ void im2col_rows8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, int N, int H, int W, int C,
                                    int R, int S, int stride, int pad, int OH, int OW, int Kpad, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const int K = R * S * C;
    const int k8n = Kpad / 8;
    const long long total = (long long)N * OH * OW * k8n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int k8 = (int)(i % k8n);
        const long long m = i / k8n;
        const int ow = (int)(m % OW);
        const int oh = (int)((m / OW) % OH);
        const int n = (int)(m / ((long long)OW * OH));
        __nv_bfloat16 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int k = k8 * 8 + j;
            __nv_bfloat16 e = __float2bfloat16(0.f);
            if (k < K) {
                const int c = k % C;
                const int rs = k / C;
                const int s = rs % S, r = rs / S;
                const int h = oh * stride - pad + r, w = ow * stride - pad + s;
                if (h >= 0 && h < H && w >= 0 && w < W) e = x[(((long long)n * H + h) * W + w) * C + c];
            }
            v[j] = e;
        }
        *reinterpret_cast<uint4*>(col + m * Kpad + k8 * 8) = *reinterpret_cast<const uint4*>(v);
    }
}

// vectorised col2im: C % 8 == 0, one thread gathers 8 channels of one input pixel (16-byte loads).
__global__ void col2im_vec8_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C,
                                   int R, int S, int stride, int pad, int OH, int OW, int Kpad, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const int C8 = C / 8;
    const long long total = (long long)N * H * W * C8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % C8);
        const long long pix = i / C8;
        const int w = (int)(pix % W);
        const int h = (int)((pix / W) % H);
        const int n = (int)(pix / ((long long)W * H));
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int r = 0; r < R; ++r) {
            const int hh = h + pad - r;
            if (hh < 0 || hh % stride != 0) continue;
            const int oh = hh / stride;
            if (oh >= OH) continue;
            for (int s = 0; s < S; ++s) {
                const int ww = w + pad - s;
                if (ww < 0 || ww % stride != 0) continue;
                const int ow = ww / stride;
                if (ow >= OW) continue;
                float v[8];
                load8(dcol + (((long long)n * OH + oh) * OW + ow) * Kpad + (r * S + s) * C + c8 * 8, v);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += v[j];
            }
        }
        store8(dx + pix * C + c8 * 8, acc);
    }
}

template <typename T>
__global__ void col2im_kernel(const T* __restrict__ dcol, T* __restrict__ dx, int N, int H, int W, int C, int R, int S,
                              int stride, int pad, int OH, int OW, int Kpad) {
    const long long total = (long long)N * H * W * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long pix = i / C;
        const int w = (int)(pix % W);
        const int h = (int)((pix / W) % H);
        const int n = (int)(pix / ((long long)W * H));
        float acc = 0.f;
        for (int r = 0; r < R; ++r) {
            const int hh = h + pad - r;
            if (hh < 0 || hh % stride != 0) continue;
            const int oh = hh / stride;
            if (oh >= OH) continue;
            for (int s = 0; s < S; ++s) {
                const int ww = w + pad - s;
                if (ww < 0 || ww % stride != 0) continue;
                const int ow = ww / stride;
                if (ow >= OW) continue;
                acc += to_f<T>(dcol[(((long long)n * OH + oh) * OW + ow) * Kpad + (r * S + s) * C + c]);
            }
        }
        dx[i] = from_f<T>(acc);
    }
}

// ------------------------------------------------------------------------------------------------
// BatchNorm (training statistics), NHWC bf16, C % 8 == 0
// ------------------------------------------------------------------------------------------------
// Pass 1: per-channel sum and sum of squares. Each thread owns 8 channels of a strided set of rows,
// the block folds its row groups through shared memory, one atomicAdd per channel per block.
// __launch_bounds__(256, 2): without the "2" ptxas squeezed MODE 2 into 84 registers (3 blocks per SM) by interleaving
// load / compute pairs -- two loads in flight per thread, 3.5 TB/s; with it all eight are issued first (6+ TB/s).
template <int MODE>   // 0: stats of x ; 1: backward sums (dz, dz*xhat), ReLU mask from y ; 2: same, mask rebuilt from x
__global__ void __launch_bounds__(256, 2) bn_reduce_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                                        const __nv_bfloat16* __restrict__ dy, const float* __restrict__ mean_invstd,
                                                        float* __restrict__ out, long long M, int C, int relu,
                                                        const float* __restrict__ gamma, const float* __restrict__ beta, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    pdl_wait();
    pdl_launch_dependents();
    extern __shared__ float sh[];   // [2][rows_per_iter][C]
    const int tpr = C / 8;                        // threads per row (<= 256)
    const int rpi = max(1, (int)blockDim.x / tpr); // rows per iteration
    const int tid = threadIdx.x;
    const int ct = tid % tpr;
    const int rgrp = tid / tpr;
    if (rgrp < rpi) {
        float a0[8], a1[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { a0[j] = 0.f; a1[j] = 0.f; }
        float mean[8], inv[8], sc[8], sf[8];
        // without a residual the ReLU mask is a function of x alone (y > 0 <=> gamma*xhat + beta > 0): y is not read
        constexpr bool mask_from_x = (MODE == 2);
        if (MODE >= 1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                mean[j] = mean_invstd[ct * 8 + j]; inv[j] = mean_invstd[C + ct * 8 + j];
                sc[j] = 0.f; sf[j] = 0.f;
                if (mask_from_x) { sc[j] = gamma[ct * 8 + j] * inv[j]; sf[j] = beta[ct * 8 + j] - mean[j] * sc[j]; }
                (void)sc[j]; (void)sf[j];
            }
        }
        constexpr int U = PDNN_BN_UNROLL;
        BN_ROW_RANGE(U)
        const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
        for (long long row = row0; row < rend; row += rstride * U) {
            uint4 rx[U], rd[U], ry[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                // rows past the end re-read the last row (an unconditional load keeps the whole batch in one basic
                // block, so all of them are issued before the first use) and are zeroed afterwards
                const long long r = row + u * rstride;
                const long long off = (r < rend ? r : rend - 1) * C + ct * 8;
                rx[u] = ldg16(x + off);
                if (MODE >= 1) rd[u] = ldg16(dy + off);
                if (MODE == 1 && relu) ry[u] = ldg16(y + off);
            }
#pragma unroll
            for (int u = 1; u < U; ++u)
                if (row + u * rstride >= rend) { if (MODE == 0) rx[u] = zero4; else rd[u] = zero4; }   // a zero x / dy row adds nothing
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float xv[8];
                unpack8(rx[u], xv);
                if (MODE == 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) { a0[j] += xv[j]; a1[j] += xv[j] * xv[j]; }
                } else {
                    float dv[8];
                    unpack8(rd[u], dv);
                    if (mask_from_x) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) if (!(xv[j] * sc[j] + sf[j] > 0.f)) dv[j] = 0.f;
                    } else if (relu) {
                        float yv[8];
                        unpack8(ry[u], yv);
#pragma unroll
                        for (int j = 0; j < 8; ++j) if (!(yv[j] > 0.f)) dv[j] = 0.f;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) { a0[j] += dv[j]; a1[j] += dv[j] * (xv[j] - mean[j]) * inv[j]; }
                }
            }
        }
        float* s0 = sh + (long long)rgrp * C + ct * 8;
        float* s1 = sh + (long long)rpi * C + (long long)rgrp * C + ct * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) { s0[j] = a0[j]; s1[j] = a1[j]; }
    }
    __syncthreads();
    for (int c = tid; c < C; c += blockDim.x) {
        float t0 = 0.f, t1 = 0.f;
        for (int g = 0; g < rpi; ++g) { t0 += sh[(long long)g * C + c]; t1 += sh[(long long)rpi * C + (long long)g * C + c]; }
        atomicAdd(out + c, t0);
        atomicAdd(out + C + c, t1);
    }
}

// Apply passes: every thread keeps ONE 8-channel group for its whole life (scale/shift are computed once per
// thread, not once per element) and walks rows with a stride of (rows per block) x (grid size).
__global__ void __launch_bounds__(256) bn_fwd_apply_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ stats,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ y,
                                                           float* __restrict__ mean_invstd, float* running_mean, float* running_var,
                                                           long long M, int C, float eps, float momentum, int relu, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    pdl_wait();
    pdl_launch_dependents();
    const int tpr = C / 8;
    const int rpi = max(1, (int)blockDim.x / tpr);
    const int ct = threadIdx.x % tpr;
    const int rgrp = threadIdx.x / tpr;
    if (rgrp >= rpi) return;
    const float invM = 1.f / (float)M;
    float sc[8], sf[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = ct * 8 + j;
        const float mean = stats[c] * invM;
        const float var = fmaxf(stats[C + c] * invM - mean * mean, 0.f);
        const float inv = rsqrtf(var + eps);
        sc[j] = gamma[c] * inv;
        sf[j] = beta[c] - mean * sc[j];
        if (blockIdx.x == 0 && rgrp == 0) {
            mean_invstd[c] = mean;
            mean_invstd[C + c] = inv;
            if (running_mean != nullptr) {
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                const float unbiased = M > 1 ? var * (float)M / (float)(M - 1) : var;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
            }
        }
    }
    constexpr int U = PDNN_BN_UNROLL;
    BN_ROW_RANGE(U)
    for (long long row = row0; row < rend; row += rstride * U) {
        uint4 rx[U], rr[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long r = row + u * rstride;
            const long long off = (r < rend ? r : rend - 1) * C + ct * 8;     // clamped, unconditional: see bn_reduce_kernel
            rx[u] = ldg16(x + off);
            if (residual != nullptr) rr[u] = ldg16(residual + off);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long r = row + u * rstride;
            if (r >= rend) break;
            float v[8];
            unpack8(rx[u], v);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = v[j] * sc[j] + sf[j];
            if (residual != nullptr) {
                float q[8];
                unpack8(rr[u], q);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] += q[j];
            }
            if (relu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
            }
            store8(y + r * C + ct * 8, v);
        }
    }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                                           const __nv_bfloat16* __restrict__ dy, const float* __restrict__ mean_invstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ sums,
                                                           __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres,
                                                           float* __restrict__ dgamma, float* __restrict__ dbeta, long long M, int C, int relu,
                                                           const float* __restrict__ beta, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    pdl_wait();
    pdl_launch_dependents();
    const int tpr = C / 8;
    const int rpi = max(1, (int)blockDim.x / tpr);
    const int ct = threadIdx.x % tpr;
    const int rgrp = threadIdx.x / tpr;
    if (rgrp >= rpi) return;
    const float invM = 1.f / (float)M;
    // dx = a*dz + b*x + c0 with a = gamma*inv, b = -a*inv*s1/M, c0 = -a*s0/M - b*mean
    float ka[8], kb[8], kc[8], sf[8];
    const bool mask_from_x = relu && (y == nullptr);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = ct * 8 + j;
        const float mean = mean_invstd[c], inv = mean_invstd[C + c];
        const float s0 = sums[c], s1 = sums[C + c];
        ka[j] = gamma[c] * inv;
        sf[j] = mask_from_x ? beta[c] - mean * ka[j] : 0.f;
        kb[j] = -ka[j] * inv * s1 * invM;
        kc[j] = -ka[j] * s0 * invM - kb[j] * mean;
        if (blockIdx.x == 0 && rgrp == 0) { dgamma[c] = s1; dbeta[c] = s0; }
    }
    constexpr int U = PDNN_BN_UNROLL;
    BN_ROW_RANGE(U)
    const bool need_y = relu && !mask_from_x;
    for (long long row = row0; row < rend; row += rstride * U) {
        uint4 rx[U], rd[U], ry[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long r = row + u * rstride;
            const long long off = (r < rend ? r : rend - 1) * C + ct * 8;     // clamped, unconditional: see bn_reduce_kernel
            rx[u] = ldg16(x + off);
            rd[u] = ldg16(dy + off);
            if (need_y) ry[u] = ldg16(y + off);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long r = row + u * rstride;
            if (r >= rend) break;
            const long long off = r * C + ct * 8;
            float xv[8], dv[8];
            unpack8(rx[u], xv);
            unpack8(rd[u], dv);
            if (mask_from_x) {
#pragma unroll
                for (int j = 0; j < 8; ++j) if (!(xv[j] * ka[j] + sf[j] > 0.f)) dv[j] = 0.f;
            } else if (relu) {
                float yv[8];
                unpack8(ry[u], yv);
#pragma unroll
                for (int j = 0; j < 8; ++j) if (!(yv[j] > 0.f)) dv[j] = 0.f;
            }
            if (dres != nullptr) store8(dres + off, dv);
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = ka[j] * dv[j] + kb[j] * xv[j] + kc[j];
            store8(dx + off, o);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// ReLU / add / pooling
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void relu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        y[i] = from_f<T>(fmaxf(to_f<T>(x[i]), 0.f));
}
template <typename T>
__global__ void relu_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long long n, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dx[i] = to_f<T>(y[i]) > 0.f ? dy[i] : from_f<T>(0.f);
}
__global__ void relu_fwd_vec8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n8, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        float v[8];
        load8(x + i * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
        store8(y + i * 8, v);
    }
}
__global__ void relu_bwd_vec8_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, long long n8, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        float yv[8], dv[8];
        load8(y + i * 8, yv);
        load8(dy + i * 8, dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) if (!(yv[j] > 0.f)) dv[j] = 0.f;
        store8(dx + i * 8, dv);
    }
}
__global__ void add_vec8_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b, __nv_bfloat16* __restrict__ o, long long n8, long long n, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        float av[8], bv[8];
        load8(a + i * 8, av);
        load8(b + i * 8, bv);
#pragma unroll
        for (int j = 0; j < 8; ++j) av[j] += bv[j];
        store8(o + i * 8, av);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (long long j = n8 * 8; j < n; ++j) o[j] = __float2bfloat16(__bfloat162float(a[j]) + __bfloat162float(b[j]));
}

template <typename T>
__global__ void maxpool2_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int H, int W, int C, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const int OH = H / 2, OW = W / 2;
    const long long total = (long long)N * OH * OW * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long p = i / C;
        const int ow = (int)(p % OW), oh = (int)((p / OW) % OH), n = (int)(p / ((long long)OW * OH));
        const T* b = x + (((long long)n * H + oh * 2) * W + ow * 2) * C + c;
        float m = fmaxf(fmaxf(to_f<T>(b[0]), to_f<T>(b[C])), fmaxf(to_f<T>(b[(long long)W * C]), to_f<T>(b[(long long)W * C + C])));
        y[i] = from_f<T>(m);
    }
}
// gradient goes to the first max of each window (torch semantics)
template <typename T>
__global__ void maxpool2_bwd_kernel(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, int N, int H, int W, int C, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const int OH = H / 2, OW = W / 2;
    const long long total = (long long)N * OH * OW * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long p = i / C;
        const int ow = (int)(p % OW), oh = (int)((p / OW) % OH), n = (int)(p / ((long long)OW * OH));
        const long long base = (((long long)n * H + oh * 2) * W + ow * 2) * C + c;
        const long long offs[4] = {0, C, (long long)W * C, (long long)W * C + C};
        const float m = to_f<T>(y[i]);
        bool given = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const bool hit = !given && to_f<T>(x[base + offs[q]]) == m;
            dx[base + offs[q]] = hit ? dy[i] : from_f<T>(0.f);
            given = given || hit;
        }
    }
}
__global__ void avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int HW, int C) {
    const long long total = (long long)N * C;
    const float inv = 1.f / (float)HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long n = i / C;
        float acc = 0.f;
        for (int p = 0; p < HW; ++p) acc += __bfloat162float(x[(n * HW + p) * C + c]);
        y[i] = __float2bfloat16(acc * inv);
    }
}
// 8 channels per thread (C % 8 == 0): 16-byte accesses, one scale per vector instead of one division chain per element
__global__ void avgpool_fwd_vec8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int HW, int C8, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const long long total = (long long)N * C8;
    const float inv = 1.f / (float)HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % C8);
        const long long n = i / C8;
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int p = 0; p < HW; ++p) {
            float v[8];
            load8(x + ((n * HW + p) * C8 + c8) * 8, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += v[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] *= inv;
        store8(y + i * 8, acc);
    }
}
__global__ void avgpool_bwd_vec8_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int HW, int C8, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const long long total = (long long)N * HW * C8;
    const float inv = 1.f / (float)HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % C8);
        const long long n = i / ((long long)HW * C8);
        float v[8];
        load8(dy + (n * C8 + c8) * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= inv;
        store8(dx + i * 8, v);
    }
}
__global__ void avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int HW, int C) {
    const long long total = (long long)N * HW * C;
    const float inv = 1.f / (float)HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long n = i / ((long long)HW * C);
        dx[i] = __float2bfloat16(__bfloat162float(dy[n * C + c]) * inv);
    }
}

// column sums over rows (bias gradients). One block per 32-column strip, warps stride over rows.
template <typename T>
__global__ void __launch_bounds__(256) col_sum_kernel(const T* __restrict__ dy, long long ld, float* __restrict__ out, long long M, int C, int accumulate, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    __shared__ float sh[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int w = threadIdx.x >> 5;
    float acc = 0.f;
    if (c < C)
        for (long long r = (long long)blockIdx.y * 8 + w; r < M; r += (long long)gridDim.y * 8) acc += to_f<T>(dy[r * ld + c]);
    sh[w][threadIdx.x & 31] = acc;
    __syncthreads();
    if (w == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int g = 0; g < 8; ++g) t += sh[g][threadIdx.x];
        if (gridDim.y == 1 && !accumulate) out[c] = t; else atomicAdd(out + c, t);
    }
}

// ------------------------------------------------------------------------------------------------
// softmax cross-entropy (mean) forward + backward, one warp per row
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) cross_entropy_kernel(const T* __restrict__ logits, const long long* __restrict__ labels,
                                                            float* __restrict__ loss, T* __restrict__ dlogits, int* __restrict__ correct,
                                                            int B, int classes, int ld, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const float invB = 1.f / (float)B;
    float loss_acc = 0.f;
    int hit_acc = 0;
    for (int row = warp; row < B; row += nwarps) {
        const T* lp = logits + (long long)row * ld;
        const int label = (int)labels[row];
        float mx = -INFINITY;
        int arg = 0x7fffffff;
        for (int c = lane; c < classes; c += 32) {
            const float v = to_f<T>(lp[c]);
            if (v > mx) { mx = v; arg = c; }
        }
        // warp argmax (lowest index wins ties)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float om = __shfl_xor_sync(0xffffffffu, mx, o);
            const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
        }
        float se = 0.f;
        for (int c = lane; c < classes; c += 32) se += __expf(to_f<T>(lp[c]) - mx);
        se = warp_sum(se);
        const float lse = mx + __logf(se);
        if (dlogits != nullptr) {
            T* dp = dlogits + (long long)row * ld;
            for (int c = lane; c < ld; c += 32) {
                float g = 0.f;
                if (c < classes) g = (__expf(to_f<T>(lp[c]) - lse) - (c == label ? 1.f : 0.f)) * invB;
                dp[c] = from_f<T>(g);
            }
        }
        if (lane == 0) {
            loss_acc += (lse - to_f<T>(lp[label])) * invB;
            hit_acc += (arg == label);
        }
    }
    if (lane == 0) {
        if (loss != nullptr) atomicAdd(loss, loss_acc);
        if (correct != nullptr && hit_acc) atomicAdd(correct, hit_acc);
    }
}

__global__ void topk_correct_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, int B, int classes, int ld,
                                    int k0, int k1, int k2, int k3, int nk, int* __restrict__ correct) {
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < B; row += gridDim.x * blockDim.x) {
        const float* lp = logits + (long long)row * ld;
        const int label = (int)labels[row];
        const float lv = lp[label];
        int rank = 0;   // number of classes strictly better (or equal with a lower index)
        for (int c = 0; c < classes; ++c) {
            const float v = lp[c];
            rank += (v > lv) || (v == lv && c < label);
        }
        const int ks[4] = {k0, k1, k2, k3};
        for (int i = 0; i < nk; ++i) if (rank < ks[i]) atomicAdd(correct + i, 1);
    }
}

// ------------------------------------------------------------------------------------------------
// momentum SGD on flat buffers
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sgd_momentum_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m,
                                                           __nv_bfloat16* __restrict__ wb, long long n, float lr, float mu, float scale, int first) {
    const long long n4 = n / 4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 gv = reinterpret_cast<const float4*>(g)[i];
        float4 wv = reinterpret_cast<float4*>(w)[i];
        float4 mv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m != nullptr && !first) mv = reinterpret_cast<float4*>(m)[i];
        mv.x = mu * mv.x + scale * gv.x; mv.y = mu * mv.y + scale * gv.y;
        mv.z = mu * mv.z + scale * gv.z; mv.w = mu * mv.w + scale * gv.w;
        wv.x -= lr * mv.x; wv.y -= lr * mv.y; wv.z -= lr * mv.z; wv.w -= lr * mv.w;
        if (m != nullptr) reinterpret_cast<float4*>(m)[i] = mv;
        reinterpret_cast<float4*>(w)[i] = wv;
        if (wb != nullptr) {
            uint2 u;
            u.x = pack_bf16x2(wv.x, wv.y); u.y = pack_bf16x2(wv.z, wv.w);
            reinterpret_cast<uint2*>(wb)[i] = u;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (long long j = n4 * 4; j < n; ++j) {
            float mv = (m != nullptr && !first) ? m[j] : 0.f;
            mv = mu * mv + scale * g[j];
            w[j] -= lr * mv;
            if (m != nullptr) m[j] = mv;
            if (wb != nullptr) wb[j] = __float2bfloat16(w[j]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// sigmoid family
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void sigmoid_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        y[i] = from_f<T>(1.f / (1.f + __expf(-to_f<T>(x[i]))));
}
template <typename T>
__global__ void sigmoid_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long long n, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float s = to_f<T>(y[i]);
        dx[i] = from_f<T>(to_f<T>(dy[i]) * s * (1.f - s));
    }
}

// ------------------------------------------------------------------------------------------------
// generic SIMT fp32 GEMM: 64x64 tile, 16-deep k step, 256 threads, 4x4 micro tile
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gemm_simt_kernel(const float* __restrict__ A, long long sam, long long sak,
                                                        const float* __restrict__ B, long long sbk, long long sbn,
                                                        float* __restrict__ C, long long ldc, int M, int N, int K, float alpha,
                                                        const float* __restrict__ bias, int relu, int accumulate, int k_per_split, NnGuard guard) {
    NN_GUARD_RETURN(guard);
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int kb = blockIdx.z * k_per_split;
    const int ke = min(K, kb + k_per_split);
    float acc[4][4] = {};
    for (int k0 = kb; k0 < ke; k0 += 16) {
        for (int i = threadIdx.x; i < 64 * 16; i += 256) {
            // consecutive threads walk the contiguous dimension of each operand
            int kk, mm;
            if (sak == 1) { kk = i % 16; mm = i / 16; } else { mm = i % 64; kk = i / 64; }
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < M && gk < ke) ? A[gm * sam + gk * sak] : 0.f;
            if (sbk == 1) { kk = i % 16; mm = i / 16; } else { mm = i % 64; kk = i / 64; }
            const int gn = n0 + mm, gk2 = k0 + kk;
            Bs[kk][mm] = (gn < N && gk2 < ke) ? B[gk2 * sbk + gn * sbn] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; b[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float v = alpha * acc[i][j];
            if (gridDim.z > 1) { atomicAdd(C + gm * ldc + gn, v); continue; }   // split-K (accumulate only)
            if (bias != nullptr) v += bias[gn];
            if (accumulate) v += C[gm * ldc + gn];
            if (relu) v = fmaxf(v, 0.f);
            C[gm * ldc + gn] = v;
        }
    }
}

}  // namespace pdnn

using namespace pdnn;
#define ST(s) reinterpret_cast<cudaStream_t>(s)
#define BF(p) reinterpret_cast<__nv_bfloat16*>(p)
#define CBF(p) reinterpret_cast<const __nv_bfloat16*>(p)
#define LAUNCH_OK()                        \
    do {                                   \
        PDNN_COUNT_LAUNCH();               \
        PDNN_CUDA_CHECK(cudaGetLastError()); \
        return 0;                          \
    } while (0)

extern "C" int pdnn_set_abort_guard(const uint32_t* closed, const uint32_t* step) {
    t_nn_guard.closed = closed;
    t_nn_guard.step = step;
    return 0;
}

extern "C" int pdnn_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
    if (n <= 0) return 0;
    cast_f32_bf16_kernel<<<grid_for((n + 7) / 8, 256), 256, 0, ST(stream)>>>(src, BF(dst), n);
    LAUNCH_OK();
}
extern "C" int pdnn_cast_bf16_to_f32(const void* src, float* dst, long long n, void* stream) {
    if (n <= 0) return 0;
    cast_bf16_f32_kernel<<<grid_for(n, 256), 256, 0, ST(stream)>>>(CBF(src), dst, n);
    LAUNCH_OK();
}
extern "C" int pdnn_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W, int Cpad, void* stream) {
    if (C <= 4 && Cpad == C)
        nchw_to_nhwc_pixel_kernel<__nv_bfloat16><<<grid_for((long long)N * H * W, 256), 256, 0, ST(stream)>>>(x, BF(y), N, C, (long long)H * W, nn_guard());
    else
        nchw_to_nhwc_kernel<__nv_bfloat16><<<grid_for((long long)N * H * W * Cpad, 256), 256, 0, ST(stream)>>>(x, BF(y), N, C, H, W, Cpad);
    LAUNCH_OK();
}
extern "C" int pdnn_nchw_f32_to_nhwc_f32(const float* x, float* y, int N, int C, int H, int W, int Cpad, void* stream) {
    nchw_to_nhwc_kernel<float><<<grid_for((long long)N * H * W * Cpad, 256), 256, 0, ST(stream)>>>(x, y, N, C, H, W, Cpad);
    LAUNCH_OK();
}
extern "C" int pdnn_wgt_flip_transpose(const void* w, void* wt, int K, int R, int S, int C, void* stream) {
    wgt_flip_transpose_kernel<<<grid_for((long long)K * R * S * C, 256), 256, 0, ST(stream)>>>(CBF(w), BF(wt), K, R, S, C);
    LAUNCH_OK();
}
extern "C" int pdnn_wgt_flip_transpose_multi(const void* w_flat, const long long* table_dev, int n_layers, int blocks_per_layer, void* stream) {
    if (n_layers <= 0) return 0;
    PDNN_REQUIRE(n_layers <= 65535 && blocks_per_layer >= 1, "wgt_flip_transpose_multi: %d layers / %d blocks per layer", n_layers, blocks_per_layer);
    wgt_flip_transpose_multi_kernel<<<dim3(blocks_per_layer, n_layers), 256, 0, ST(stream)>>>(CBF(w_flat), table_dev);
    LAUNCH_OK();
}
extern "C" int pdnn_oihw_to_krsc_f32(const float* src, float* dst, int K, int C, int R, int S, int to_krsc, void* stream) {
    oihw_krsc_kernel<<<grid_for((long long)K * R * S * C, 256), 256, 0, ST(stream)>>>(src, dst, K, C, R, S, to_krsc);
    LAUNCH_OK();
}
extern "C" int pdnn_im2col_bf16(const void* x, void* col, int N, int H, int W, int C, int R, int S, int stride, int pad, int OH,
                                int OW, int Kpad, void* stream) {
    if (C % 8 == 0 && Kpad == R * S * C) {
        const long long total = (long long)N * OH * OW * R * S * (C / 8);
        im2col_vec8_kernel<<<grid_for(total, 256, 148 * 16), 256, 0, ST(stream)>>>(CBF(x), BF(col), N, H, W, C, R, S, stride, pad, OH, OW, nn_guard());
    } else if (Kpad % 8 == 0) {
        const long long total = (long long)N * OH * OW * (Kpad / 8);
        im2col_rows8_kernel<<<grid_for(total, 256, 148 * 16), 256, 0, ST(stream)>>>(CBF(x), BF(col), N, H, W, C, R, S, stride, pad, OH, OW, Kpad, nn_guard());
    } else {
        const long long total = (long long)N * OH * OW * Kpad;
        im2col_kernel<__nv_bfloat16><<<grid_for(total, 256, 148 * 16), 256, 0, ST(stream)>>>(CBF(x), BF(col), N, H, W, C, R, S, stride, pad, OH, OW, Kpad);
    }
    LAUNCH_OK();
}
extern "C" int pdnn_im2col_f32(const float* x, float* col, int N, int H, int W, int C, int R, int S, int stride, int pad, int OH,
                               int OW, int Kpad, void* stream) {
    const long long total = (long long)N * OH * OW * Kpad;
    im2col_kernel<float><<<grid_for(total, 256, 148 * 16), 256, 0, ST(stream)>>>(x, col, N, H, W, C, R, S, stride, pad, OH, OW, Kpad);
    LAUNCH_OK();
}
extern "C" int pdnn_col2im_bf16(const void* dcol, void* dx, int N, int H, int W, int C, int R, int S, int stride, int pad, int OH,
                                int OW, int Kpad, void* stream) {
    if (C % 8 == 0 && Kpad % 8 == 0)
        col2im_vec8_kernel<<<grid_for((long long)N * H * W * (C / 8), 256, 148 * 16), 256, 0, ST(stream)>>>(CBF(dcol), BF(dx), N, H, W, C, R, S, stride, pad, OH, OW, Kpad, nn_guard());
    else
        col2im_kernel<__nv_bfloat16><<<grid_for((long long)N * H * W * C, 256, 148 * 16), 256, 0, ST(stream)>>>(CBF(dcol), BF(dx), N, H, W, C, R, S, stride, pad, OH, OW, Kpad);
    LAUNCH_OK();
}
extern "C" int pdnn_col2im_f32(const float* dcol, float* dx, int N, int H, int W, int C, int R, int S, int stride, int pad, int OH,
                               int OW, int Kpad, void* stream) {
    col2im_kernel<float><<<grid_for((long long)N * H * W * C, 256, 148 * 16), 256, 0, ST(stream)>>>(dcol, dx, N, H, W, C, R, S, stride, pad, OH, OW, Kpad);
    LAUNCH_OK();
}

// One resident wave: the BatchNorm passes are grid-stride loops, so a grid larger than what fits on the chip
// runs as full-length extra waves at partial occupancy (592 blocks at 3 per SM = a second wave on a third of the SMs).
template <typename K>
static int resident_blocks(K kernel, int threads, size_t smem) {
    int per_sm = 0, dev = 0, sms = 148;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // one SM's worth of slots stays free: a small kernel of another stream that is resident for the whole step (the
    // parameter-server master's flag-polling warp) then costs nothing, instead of pushing one CTA into a second wave
    return per_sm * sms - per_sm;
}
static int bn_reduce_cfg(int C, int* rpi, size_t* smem) {
    const int tpr = C / 8;
    *rpi = tpr >= 256 ? 1 : 256 / tpr;
    *smem = (size_t)2 * (*rpi) * C * sizeof(float);
    return 0;
}
extern "C" int pdnn_bn_fwd_stats(const void* x, float* stats, long long M, int C, void* stream) {
    PDNN_REQUIRE(C % 8 == 0 && C <= 2048, "bn: C=%d must be a multiple of 8 and <= 2048", C);
    int rpi; size_t smem;
    bn_reduce_cfg(C, &rpi, &smem);
    static const int wave = resident_blocks(bn_reduce_kernel<0>, 256, 16384);
    const int grid = grid_for((M + rpi - 1) / rpi, 1, wave);
    PDNN_CUDA_CHECK(pdnn_launch(bn_reduce_kernel<0>, dim3(grid), dim3(256), (size_t)smem, ST(stream), CBF(x), (const __nv_bfloat16*)nullptr, (const __nv_bfloat16*)nullptr, (const float*)nullptr, stats, M, C, 0, (const float*)nullptr, (const float*)nullptr, nn_guard()));
    LAUNCH_OK();
}
extern "C" int pdnn_bn_fwd_apply(const void* x, const float* stats, const float* gamma, const float* beta, const void* residual,
                                 void* y, float* mean_invstd, float* running_mean, float* running_var, long long M, int C,
                                 float eps, float momentum, int relu, void* stream) {
    PDNN_REQUIRE(C % 8 == 0 && C <= 2048, "bn: C=%d must be a multiple of 8 and <= 2048", C);
    int rpi_a; size_t smem_a;
    bn_reduce_cfg(C, &rpi_a, &smem_a);
    static const int wave = resident_blocks(bn_fwd_apply_kernel, 256, 0);
    PDNN_CUDA_CHECK(pdnn_launch(bn_fwd_apply_kernel, dim3(grid_for((M + rpi_a - 1) / rpi_a, 1, wave)), dim3(256), (size_t)0, ST(stream), CBF(x), stats, gamma, beta,
                                CBF(residual), BF(y), mean_invstd, running_mean, running_var, M, C, eps, momentum, relu, nn_guard()));
    LAUNCH_OK();
}
extern "C" int pdnn_bn_bwd_reduce(const void* x, const void* y, const void* dy, const float* mean_invstd, float* sums, long long M,
                                  int C, int relu, const float* gamma, const float* beta, void* stream) {
    PDNN_REQUIRE(!(relu && y == nullptr) || (gamma != nullptr && beta != nullptr), "bn_bwd_reduce: y == NULL needs gamma and beta to rebuild the ReLU mask");
    PDNN_REQUIRE(C % 8 == 0 && C <= 2048, "bn: C=%d must be a multiple of 8 and <= 2048", C);
    int rpi; size_t smem;
    bn_reduce_cfg(C, &rpi, &smem);
    static const int wave1 = resident_blocks(bn_reduce_kernel<1>, 256, 16384), wave2 = resident_blocks(bn_reduce_kernel<2>, 256, 16384);
    const int grid = grid_for((M + rpi - 1) / rpi, 1, (relu && y == nullptr) ? wave2 : wave1);
    if (relu && y == nullptr)
        PDNN_CUDA_CHECK(pdnn_launch(bn_reduce_kernel<2>, dim3(grid), dim3(256), (size_t)smem, ST(stream), CBF(x), (const __nv_bfloat16*)nullptr, CBF(dy), mean_invstd, sums, M, C, relu, gamma, beta, nn_guard()));
    else
        PDNN_CUDA_CHECK(pdnn_launch(bn_reduce_kernel<1>, dim3(grid), dim3(256), (size_t)smem, ST(stream), CBF(x), CBF(y), CBF(dy), mean_invstd, sums, M, C, relu, gamma, beta, nn_guard()));
    LAUNCH_OK();
}
extern "C" int pdnn_bn_bwd_apply(const void* x, const void* y, const void* dy, const float* mean_invstd, const float* gamma,
                                 const float* sums, void* dx, void* dresidual, float* dgamma, float* dbeta, long long M, int C,
                                 int relu, const float* beta, void* stream) {
    PDNN_REQUIRE(!(relu && y == nullptr) || beta != nullptr, "bn_bwd_apply: y == NULL needs beta to rebuild the ReLU mask");
    PDNN_REQUIRE(C % 8 == 0 && C <= 2048, "bn: C=%d must be a multiple of 8 and <= 2048", C);
    int rpi_a; size_t smem_a;
    bn_reduce_cfg(C, &rpi_a, &smem_a);
    static const int wave = resident_blocks(bn_bwd_apply_kernel, 256, 0);
    PDNN_CUDA_CHECK(pdnn_launch(bn_bwd_apply_kernel, dim3(grid_for((M + rpi_a - 1) / rpi_a, 1, wave)), dim3(256), (size_t)0, ST(stream), CBF(x), CBF(y), CBF(dy),
                                mean_invstd, gamma, sums, BF(dx), BF(dresidual), dgamma, dbeta, M, C, relu, beta, nn_guard()));
    LAUNCH_OK();
}
extern "C" int pdnn_relu_fwd(const void* x, void* y, long long n, void* stream) {
    if (n % 8 == 0) relu_fwd_vec8_kernel<<<grid_for(n / 8, 256), 256, 0, ST(stream)>>>(CBF(x), BF(y), n / 8, nn_guard());
    else relu_fwd_kernel<__nv_bfloat16><<<grid_for(n, 256), 256, 0, ST(stream)>>>(CBF(x), BF(y), n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_relu_bwd(const void* y, const void* dy, void* dx, long long n, void* stream) {
    if (n % 8 == 0) relu_bwd_vec8_kernel<<<grid_for(n / 8, 256), 256, 0, ST(stream)>>>(CBF(y), CBF(dy), BF(dx), n / 8, nn_guard());
    else relu_bwd_kernel<__nv_bfloat16><<<grid_for(n, 256), 256, 0, ST(stream)>>>(CBF(y), CBF(dy), BF(dx), n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_relu_fwd_f32(const float* x, float* y, long long n, void* stream) {
    relu_fwd_kernel<float><<<grid_for(n, 256), 256, 0, ST(stream)>>>(x, y, n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_relu_bwd_f32(const float* y, const float* dy, float* dx, long long n, void* stream) {
    relu_bwd_kernel<float><<<grid_for(n, 256), 256, 0, ST(stream)>>>(y, dy, dx, n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_add_bf16(const void* a, const void* b, void* out, long long n, void* stream) {
    add_vec8_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, ST(stream)>>>(CBF(a), CBF(b), BF(out), n / 8, n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_maxpool2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream) {
    maxpool2_fwd_kernel<__nv_bfloat16><<<grid_for((long long)N * (H / 2) * (W / 2) * C, 256), 256, 0, ST(stream)>>>(CBF(x), BF(y), N, H, W, C, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_maxpool2_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W, int C, void* stream) {
    maxpool2_bwd_kernel<__nv_bfloat16><<<grid_for((long long)N * (H / 2) * (W / 2) * C, 256), 256, 0, ST(stream)>>>(CBF(x), CBF(y), CBF(dy), BF(dx), N, H, W, C, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_maxpool2_fwd_f32(const float* x, float* y, int N, int H, int W, int C, void* stream) {
    maxpool2_fwd_kernel<float><<<grid_for((long long)N * (H / 2) * (W / 2) * C, 256), 256, 0, ST(stream)>>>(x, y, N, H, W, C, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_maxpool2_bwd_f32(const float* x, const float* y, const float* dy, float* dx, int N, int H, int W, int C, void* stream) {
    maxpool2_bwd_kernel<float><<<grid_for((long long)N * (H / 2) * (W / 2) * C, 256), 256, 0, ST(stream)>>>(x, y, dy, dx, N, H, W, C, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
    if (C % 8 == 0) avgpool_fwd_vec8_kernel<<<grid_for((long long)N * (C / 8), 256), 256, 0, ST(stream)>>>(CBF(x), BF(y), N, HW, C / 8, nn_guard());
    else avgpool_fwd_kernel<<<grid_for((long long)N * C, 256), 256, 0, ST(stream)>>>(CBF(x), BF(y), N, HW, C);
    LAUNCH_OK();
}
extern "C" int pdnn_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
    if (C % 8 == 0) avgpool_bwd_vec8_kernel<<<grid_for((long long)N * HW * (C / 8), 256), 256, 0, ST(stream)>>>(CBF(dy), BF(dx), N, HW, C / 8, nn_guard());
    else avgpool_bwd_kernel<<<grid_for((long long)N * HW * C, 256), 256, 0, ST(stream)>>>(CBF(dy), BF(dx), N, HW, C);
    LAUNCH_OK();
}
extern "C" int pdnn_col_sum(const void* dy, long long ld, float* out, long long M, int C, int accumulate, void* stream) {
    int gy = (int)((M + 255) / 256);
    if (gy > 64) gy = 64;
    if (gy < 1) gy = 1;
    if (gy > 1 && !accumulate) PDNN_CUDA_CHECK(cudaMemsetAsync(out, 0, sizeof(float) * C, ST(stream)));
    col_sum_kernel<__nv_bfloat16><<<dim3((C + 31) / 32, gy), 256, 0, ST(stream)>>>(CBF(dy), ld, out, M, C, accumulate, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_col_sum_f32(const float* dy, long long ld, float* out, long long M, int C, int accumulate, void* stream) {
    int gy = (int)((M + 255) / 256);
    if (gy > 64) gy = 64;
    if (gy < 1) gy = 1;
    if (gy > 1 && !accumulate) PDNN_CUDA_CHECK(cudaMemsetAsync(out, 0, sizeof(float) * C, ST(stream)));
    col_sum_kernel<float><<<dim3((C + 31) / 32, gy), 256, 0, ST(stream)>>>(dy, ld, out, M, C, accumulate, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_cross_entropy_fwd_bwd(const void* logits, const long long* labels, float* loss, void* dlogits, int* correct,
                                          int B, int classes, int ld, void* stream) {
    cross_entropy_kernel<__nv_bfloat16><<<grid_for((long long)B * 32, 256, 148), 256, 0, ST(stream)>>>(CBF(logits), labels, loss, BF(dlogits), correct, B, classes, ld, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_cross_entropy_fwd_bwd_f32(const float* logits, const long long* labels, float* loss, float* dlogits, int* correct,
                                              int B, int classes, int ld, void* stream) {
    cross_entropy_kernel<float><<<grid_for((long long)B * 32, 256, 148), 256, 0, ST(stream)>>>(logits, labels, loss, dlogits, correct, B, classes, ld, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_topk_correct(const float* logits, const long long* labels, int B, int classes, int ld, const int* ks_host, int nk,
                                 int* correct, void* stream) {
    PDNN_REQUIRE(nk >= 1 && nk <= 4, "topk: 1..4 k values supported");
    int ks[4] = {0, 0, 0, 0};
    for (int i = 0; i < nk; ++i) ks[i] = ks_host[i];
    topk_correct_kernel<<<grid_for(B, 128), 128, 0, ST(stream)>>>(logits, labels, B, classes, ld, ks[0], ks[1], ks[2], ks[3], nk, correct);
    LAUNCH_OK();
}
extern "C" int pdnn_sgd_momentum(float* w, const float* g, float* m, void* w_bf16, long long n, float lr, float mu, float scale,
                                 int first_step, void* stream) {
    if (n <= 0) return 0;
    sgd_momentum_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, ST(stream)>>>(w, g, m, BF(w_bf16), n, lr, mu, scale, first_step);
    LAUNCH_OK();
}
extern "C" int pdnn_sigmoid_fwd(const void* x, void* y, long long n, void* stream) {
    sigmoid_fwd_kernel<__nv_bfloat16><<<grid_for(n, 256), 256, 0, ST(stream)>>>(CBF(x), BF(y), n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_sigmoid_bwd(const void* y, const void* dy, void* dx, long long n, void* stream) {
    sigmoid_bwd_kernel<__nv_bfloat16><<<grid_for(n, 256), 256, 0, ST(stream)>>>(CBF(y), CBF(dy), BF(dx), n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_sigmoid_fwd_f32(const float* x, float* y, long long n, void* stream) {
    sigmoid_fwd_kernel<float><<<grid_for(n, 256), 256, 0, ST(stream)>>>(x, y, n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_sigmoid_bwd_f32(const float* y, const float* dy, float* dx, long long n, void* stream) {
    sigmoid_bwd_kernel<float><<<grid_for(n, 256), 256, 0, ST(stream)>>>(y, dy, dx, n, nn_guard());
    LAUNCH_OK();
}
extern "C" int pdnn_gemm_simt_f32(const float* A, int transA, long long lda, const float* B, int transB, long long ldb, float* C,
                                  long long ldc, int M, int N, int K, float alpha, const float* bias, int relu, int accumulate,
                                  void* stream) {
    // op(A)[m][k]: transA ? A[k*lda+m] : A[m*lda+k];  op(B)[k][n]: transB ? B[n*ldb+k] : B[k*ldb+n]
    const long long sam = transA ? 1 : lda, sak = transA ? lda : 1;
    const long long sbk = transB ? 1 : ldb, sbn = transB ? ldb : 1;
    dim3 grid((N + 63) / 64, (M + 63) / 64, 1);
    int k_per_split = K;
    if (accumulate && bias == nullptr && !relu && K >= 1024) {   // split-K keeps the SMs busy on wgrad-shaped problems
        const int blocks = grid.x * grid.y;
        int splits = (148 * 2 + blocks - 1) / blocks;
        if (splits > K / 256) splits = K / 256;
        if (splits > 1) {
            k_per_split = ((K + splits - 1) / splits + 15) / 16 * 16;
            grid.z = (K + k_per_split - 1) / k_per_split;
        }
    }
    gemm_simt_kernel<<<grid, 256, 0, ST(stream)>>>(A, sam, sak, B, sbk, sbn, C, ldc, M, N, K, alpha, bias, relu, accumulate, k_per_split, nn_guard());
    LAUNCH_OK();
}

// ------------------------------------------------------------------------------------------------
// Allocation-free BatchNorm forward
// ------------------------------------------------------------------------------------------------
// pdnn_bn_fwd_stats + pdnn_bn_fwd_apply need a [2][C] fp32 statistics buffer that somebody has to allocate
// and zero. Some users of this C ABI have no way to do that at the point where they call us: a plain-C host
// that links the library without any device allocator of its own, a process in which another runtime has
// already reserved all of HBM for its private pool (a raw cudaMalloc then fails), or a step loop that must
// not stall on the device-wide synchronisation an allocation can imply. For them the statistics live in the
// fixed device array below. It is sized for PDNN_BN_NOALLOC_MAX_C channels (2 x 2048 floats = 16 KiB) and
// every GPU that loads this library keeps it until the process exits.
//
// The array name is only a host-side handle: kernels must be given the device address, which
// cudaGetSymbolAddress returns for the device that is current at the time. The address is looked up once per
// device, on that device's first call (so make that first call outside any stream capture), and cached.
// All calls on a device share the one array and it keeps whatever the previous call left in it, so the part
// that is used is cleared at the start of every call; two calls that overlap on the same GPU (different
// streams) would race on it and must be serialised by the caller.
#define PDNN_BN_NOALLOC_MAX_C 2048
static __device__ __align__(256) float g_bn_noalloc_stats[2 * PDNN_BN_NOALLOC_MAX_C];

#include <mutex>
static float* g_bn_noalloc_ptr[64] = {nullptr};
static std::mutex g_bn_noalloc_mutex;

static int bn_noalloc_scratch(float** out) {
    int dev = -1;
    PDNN_CUDA_CHECK(cudaGetDevice(&dev));
    PDNN_REQUIRE(dev >= 0 && dev < 64, "bn_fwd_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_bn_noalloc_mutex);
    if (g_bn_noalloc_ptr[dev] == nullptr) {
        float* p = nullptr;
        PDNN_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_bn_noalloc_stats));   // address on the CURRENT device
        PDNN_REQUIRE(p != nullptr, "bn_fwd_noalloc: cudaGetSymbolAddress returned a null address");
        g_bn_noalloc_ptr[dev] = p;
    }
    *out = g_bn_noalloc_ptr[dev];
    return 0;
}

extern "C" int pdnn_bn_fwd_noalloc(const void* x, const float* gamma, const float* beta, const void* residual, void* y,
                                   float* mean_invstd, float* running_mean, float* running_var, long long M, int C,
                                   float eps, float momentum, int relu, void* stream) {
    PDNN_REQUIRE(C > 0 && C % 8 == 0, "bn_fwd_noalloc: C=%d must be a positive multiple of 8", C);
    PDNN_REQUIRE(C <= PDNN_BN_NOALLOC_MAX_C,
                 "bn_fwd_noalloc: C=%d exceeds the built-in scratch (max %d channels); use pdnn_bn_fwd_stats/apply with "
                 "your own [2][C] buffer", C, PDNN_BN_NOALLOC_MAX_C);
    PDNN_REQUIRE(M > 0 && x != nullptr && y != nullptr && mean_invstd != nullptr, "bn_fwd_noalloc: null or empty arguments");
    float* stats = nullptr;
    if (int rc = bn_noalloc_scratch(&stats)) return rc;
    // the array still holds the sums of the previous call: zero the [2][C] prefix this call uses
    PDNN_CUDA_CHECK(cudaMemsetAsync(stats, 0, sizeof(float) * 2 * (size_t)C, ST(stream)));
    int rpi; size_t smem;
    bn_reduce_cfg(C, &rpi, &smem);
    static const int wave = resident_blocks(bn_reduce_kernel<0>, 256, 16384);
    const int grid = grid_for((M + rpi - 1) / rpi, 1, wave);
    bn_reduce_kernel<0><<<grid, 256, smem, ST(stream)>>>(CBF(x), nullptr, nullptr, nullptr, stats, M, C, 0, nullptr, nullptr, NnGuard{nullptr, nullptr});
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    bn_fwd_apply_kernel<<<grid_for((M + rpi - 1) / rpi, 1, 148 * 8), 256, 0, ST(stream)>>>(
        CBF(x), stats, gamma, beta, CBF(residual), BF(y), mean_invstd, running_mean, running_var, M, C, eps, momentum, relu, NnGuard{nullptr, nullptr});
    LAUNCH_OK();
}
