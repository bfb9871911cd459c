/* C ABI of libpdnn_b200.so — the sm_100a kernel library of pytorch_distributed_nn_b200.
 *
 * Conventions
 *   - every function returns 0 on success; on failure it returns non-zero and pdnn_last_error()
 *     gives the message (thread-local).
 *   - pointers are device pointers unless the comment says otherwise; `stream` is a cudaStream_t.
 *   - activations are NHWC bf16, parameters / gradients / optimizer state are fp32 unless stated.
 *   - "ref:" comments cite the component of chao1224/pytorch_distributed_nn each entry point covers
 *     (paths relative to the reference root).
 */
#ifndef PDNN_B200_H
#define PDNN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* pdnn_last_error(void);
int pdnn_version(void);
/* Number of kernel launches issued through this library by the calling process (all threads). */
long long pdnn_launch_count(void);
void pdnn_reset_launch_count(void);
/* Adds n to the counter: called once per CUDA-graph replay with the number of library kernels captured in that graph. */
void pdnn_add_launch_count(long long n);

/* ---------------------------------------------------------------------------------------------
 * GEMM / convolution on tcgen05 + TMEM + TMA (csrc/gemm_tcgen05.cu)
 * ------------------------------------------------------------------------------------------- */
typedef struct pdnn_epilogue {
    void* out;                 /* [M][ldo] */
    long long ldo;             /* row stride in elements */
    int out_mode;              /* 0 = bf16 store, 1 = fp32 store, 2 = fp32 red.add (caller zero-fills) */
    int relu;                  /* clamp at 0 after bias/residual */
    float alpha;               /* scale applied to the accumulator */
    const float* bias;         /* [N] or NULL */
    const void* residual;      /* bf16 [M][ldo] or NULL (bf16 out only) */
    const uint32_t* guard_closed; /* straggler abort guard (both NULL = disabled): the kernel returns */
    const uint32_t* guard_step;   /*   at once when *guard_closed >= *guard_step                      */
    int max_ctas;              /* 0 = one CTA per SM */
} pdnn_epilogue_t;

/* D[M][N] = alpha * A * B (+bias)(+residual)(relu). A is [M][K] (a_mn_major=0, K contiguous, row
 * stride lda) or stored as [K][M] (a_mn_major=1). B is [N][K] (b_mn_major=0) or stored as [K][N]
 * (b_mn_major=1). bf16 inputs, fp32 accumulation in TMEM. K-strides must be multiples of 8 elements.
 * ref: nn.Linear fwd/bwd (pytorch_code/model_ops/lenet.py:45-46, resnet.py:78), cblas_dgemm NN/NT/TN
 * (MPI_code/src/util/util.h:35-81), numpy dot (pure_py_code/nn/nn_layer.py:79-92). */
int pdnn_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major,
                   long long ldb, int M, int N, int K, const pdnn_epilogue_t* epi, int k_splits,
                   void* stream);

/* Stride-1 "same" convolution forward as implicit GEMM. x: NHWC bf16 [batch][H][W][C], C % 64 == 0;
 * wgt: bf16 [Kout][R][S][C]; out rows are output pixels in NHWC order, ldo >= Kout.
 * Data-gradient of the same conv = this entry point on dy with pdnn_wgt_flip_transpose()'d weights.
 * ref: nn.Conv2d 3x3 fwd / bwd-data (pytorch_code/model_ops/resnet.py:19-21,72; resnet_split.py:28,34). */
/* Stem convolution: 3x3 / stride 1 / pad 1, C <= 3 input channels, 64 output channels (the first layer of the CIFAR
 * ResNets and VGGs). The im2col rows are built in registers and written straight into the swizzled MMA operand stage;
 * no im2col pass, no 64-byte-row GEMM. w_pad: [64][32] bf16 (KRSC rows zero-padded from 9*C to 32). col (optional):
 * receives the [M][32] im2col matrix for the weight-gradient GEMM. bn_stats (optional, [2][64] fp32, zeroed by the
 * caller): per-channel sum / sum of squares of the stored output. */
int pdnn_conv_stem_supported(int batch, int H, int W, int C, int Kout, int R, int S, int stride, int pad);
int pdnn_conv_stem_bf16(const void* x, const void* w_pad, int batch, int H, int W, int C, int Kout,
                        const pdnn_epilogue_t* epi, void* col, float* bn_stats, void* stream);

/* Two-CTA cluster schedule of the K-major GEMM / conv fprop kernels. 0 = off (default), 2 = CTA pair (cta_group::2: the
 * leader issues M = 256 MMAs over both CTAs' shared memory, every CTA loads only its half of B with unicast TMA loads).
 * Any other value means off. Env PDNN_CLUSTER=2 sets it at first use. Returns the previous setting. */
int pdnn_set_cluster_mode(int on);

int pdnn_conv_fprop_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                         int R, int S, int pad, const pdnn_epilogue_t* epi, void* stream);

/* 3x3 / stride 1 / pad 1 convolution with halo reuse and resident filters (csrc/conv_halo.cu): one TMA halo box per
 * tile instead of nine shifted boxes, filter bank resident in shared memory. C in {64, 128}, Kout % 64 == 0.
 * pdnn_conv_halo_supported() tells whether a geometry qualifies; same epilogue options as pdnn_conv_fprop_bf16
 * (bf16 output only). ref: the same nn.Conv2d call sites as pdnn_conv_fprop_bf16. */
int pdnn_conv_halo_supported(int batch, int H, int W, int C, int Kout, int R, int S, int pad);
int pdnn_conv_halo_fprop_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                              const pdnn_epilogue_t* epi, void* stream);
/* Same conv fused with the statistics pass of the BatchNorm that follows: bn_stats [2][Kout] fp32 (zero on entry)
 * receives sum and sum of squares of the stored bf16 output (then call pdnn_bn_fwd_apply directly). */
int pdnn_conv_halo_fprop_stats_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout,
                                    const pdnn_epilogue_t* epi, float* bn_stats, void* stream);

/* Weight gradient of a 3x3 / stride 1 / pad 1 conv with halo reuse (csrc/conv_halo.cu): dw[Kout][3][3][C] fp32 +=
 * (red.add, caller zero-fills). C in {64, 128}, Kout % 64 == 0, see pdnn_conv_halo_wgrad_supported(). */
int pdnn_conv_halo_wgrad_supported(int batch, int H, int W, int C, int Kout, int R, int S, int pad);
int pdnn_conv_halo_wgrad_bf16(const void* x, const void* dy, int batch, int H, int W, int C, int Kout, float* dw,
                              const uint32_t* guard_closed, const uint32_t* guard_step, int max_ctas, void* stream);

/* Weight gradient of the same conv: dw[Kout][R][S][C] (fp32) += dy^T * im2col(x). out_mode must be 2.
 * ref: nn.Conv2d bwd-filter (same call sites). */
int pdnn_conv_wgrad_bf16(const void* x, const void* dy, int batch, int H, int W, int C, int Kout,
                         int R, int S, int pad, const pdnn_epilogue_t* epi, int k_splits, void* stream);

/* Strided (stride 1 or 2) convolution forward / weight gradient as implicit GEMM without an im2col buffer: the
 * activation tensor map walks the input with TMA element strides. H, W are the INPUT map; outputs have
 * OH = (H + 2 pad - R) / stride + 1 rows. ref: the stride-2 3x3 convs and 1x1 shortcuts of resnet.py:19,27. */
int pdnn_conv_fprop_strided_bf16(const void* x, const void* wgt, int batch, int H, int W, int C, int Kout, int R, int S,
                                 int pad, int stride, const pdnn_epilogue_t* epi, void* stream);
int pdnn_conv_wgrad_strided_bf16(const void* x, const void* dy, int batch, int H, int W, int C, int Kout, int R, int S,
                                 int pad, int stride, const pdnn_epilogue_t* epi, int k_splits, void* stream);

/* Data gradient of a stride-2 conv (3x3 pad 1 or 1x1 pad 0) without a col2im pass: four parity-class implicit GEMMs
 * over dy [batch][OH][OW][Kout] write the interleaved pixels of dx [batch][2 OH][2 OW][C]; wt is the
 * pdnn_wgt_flip_transpose()d filter bank. epi: bf16 out = dx, ldo = C, optional residual (added at dx's pixels). */
int pdnn_conv_dgrad_s2_bf16(const void* dy, const void* wt, int batch, int OH, int OW, int Kout, int C, int R, int S,
                            int pad, const pdnn_epilogue_t* epi, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Layer kernels (csrc/nn_kernels.cu)
 * ------------------------------------------------------------------------------------------- */
int pdnn_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
int pdnn_cast_bf16_to_f32(const void* src, float* dst, long long n, void* stream);
/* NCHW fp32 images -> NHWC bf16 with the channel dim zero-padded to Cpad (input staging). */
int pdnn_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W, int Cpad, void* stream);
/* KRSC bf16 -> [C][R][S][K] with (r,s) mirrored: the filter bank that turns dgrad into fprop. */
int pdnn_wgt_flip_transpose(const void* w, void* wt, int K, int R, int S, int C, void* stream);
/* Every convolution of a network in one launch. table_dev: n_layers x 6 int64 on the device =
 * {element offset of the KRSC weights in w_flat, destination address, K, R, S, C}. */
int pdnn_wgt_flip_transpose_multi(const void* w_flat, const long long* table_dev, int n_layers, int blocks_per_layer, void* stream);
/* fp32 OIHW (torch) <-> KRSC (ours). */
int pdnn_oihw_to_krsc_f32(const float* src, float* dst, int K, int C, int R, int S, int to_krsc, void* stream);

/* im2col for the lowered convs (stride 2, 1x1 shortcuts, stems, LeNet 5x5).
 * col[(n,oh,ow)][Kpad], k = (r*S+s)*C + c, columns >= R*S*C are zero. */
int pdnn_im2col_bf16(const void* x, void* col, int N, int H, int W, int C, int R, int S, int stride, int pad,
                     int OH, int OW, int Kpad, void* stream);
/* dx[n,h,w,c] = sum over the (oh,ow,r,s) that touch it of dcol[(n,oh,ow)][(r*S+s)*C+c]  (gather form). */
int pdnn_col2im_bf16(const void* dcol, void* dx, int N, int H, int W, int C, int R, int S, int stride, int pad,
                     int OH, int OW, int Kpad, void* stream);

/* BatchNorm2d, training mode (batch statistics), NHWC bf16, M = N*H*W rows, C channels.
 * ref: nn.BatchNorm2d (pytorch_code/model_ops/resnet.py:20,22,28; resnet_split.py:31,37,46,128).
 * stats: [2][C] fp32 scratch that must be zero on entry to bn_fwd_stats (sum, sum of squares);
 * mean_invstd: [2][C] fp32 saved for backward. running_mean/var updated in place (may be NULL). */
int pdnn_bn_fwd_stats(const void* x, float* stats, long long M, int C, void* stream);
int pdnn_bn_fwd_apply(const void* x, const float* stats, const float* gamma, const float* beta,
                      const void* residual, void* y, float* mean_invstd, float* running_mean,
                      float* running_var, long long M, int C, float eps, float momentum, int relu,
                      void* stream);
/* BatchNorm forward for callers that cannot hand us scratch memory (no counterpart in the reference, which
 * never runs on a device). Same result as pdnn_bn_fwd_stats followed by pdnn_bn_fwd_apply, but the [2][C]
 * statistics live in a fixed device array inside the library (C <= 2048, 16 KiB per GPU) and are cleared on
 * every call, so the call performs no allocation of any kind. Calls on one GPU share that array: do not let two
 * of them overlap. The first call on a device resolves the array's address; issue it before any stream capture.
 * C above the limit is rejected with an error. Inputs and outputs are passed exactly as for the two-call form. */
int pdnn_bn_fwd_noalloc(const void* x, const float* gamma, const float* beta, const void* residual, void* y,
                        float* mean_invstd, float* running_mean, float* running_var, long long M, int C,
                        float eps, float momentum, int relu, void* stream);
/* sums: [2][C] fp32, zero on entry: sum(dz), sum(dz * xhat), dz = dy * (y > 0 if relu).
 * y may be NULL when the forward had no residual: the mask is then rebuilt from x (gamma*xhat + beta > 0), which
 * saves reading y in both backward passes; gamma / beta are only read in that case. */
int pdnn_bn_bwd_reduce(const void* x, const void* y, const void* dy, const float* mean_invstd, float* sums,
                       long long M, int C, int relu, const float* gamma, const float* beta, void* stream);
/* dx = gamma*invstd*(dz - sum_dz/M - xhat*sum_dzxhat/M); dresidual = dz (optional); dgamma/dbeta fp32. */
int pdnn_bn_bwd_apply(const void* x, const void* y, const void* dy, const float* mean_invstd,
                      const float* gamma, const float* sums, void* dx, void* dresidual, float* dgamma,
                      float* dbeta, long long M, int C, int relu, const float* beta, void* stream);

int pdnn_relu_fwd(const void* x, void* y, long long n, void* stream);
int pdnn_relu_bwd(const void* y, const void* dy, void* dx, long long n, void* stream);
int pdnn_add_bf16(const void* a, const void* b, void* out, long long n, void* stream);
/* 2x2/2 max pool, NHWC bf16. ref: nn.MaxPool2d (pytorch_code/model_ops/lenet.py:48). */
int pdnn_maxpool2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream);
int pdnn_maxpool2_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W, int C, void* stream);
/* Average pool over the whole HxW map. ref: F.avg_pool2d(out, 4) (pytorch_code/model_ops/resnet.py:94). */
int pdnn_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int pdnn_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
/* dbias[c] = sum_rows dy[row][c]  (ref: distributed_worker.py:172-173, sum(dL/dlogits, axis=0)). */
int pdnn_col_sum(const void* dy, long long ld, float* out, long long M, int C, int accumulate, void* stream);

/* Fused softmax cross-entropy, mean reduction: loss (1 fp32, zero on entry) and dlogits = (p - onehot)/B.
 * logits bf16 [B][ld]; dlogits bf16 [B][ld] (padding columns written as 0); correct[0] += #argmax hits.
 * ref: nn.CrossEntropyLoss (pytorch_code/distributed_worker.py:86,161; nn_ops/__init__.py:44) and
 * accuracy() (nn_ops/__init__.py:13-26). */
int pdnn_cross_entropy_fwd_bwd(const void* logits, const long long* labels, float* loss, void* dlogits,
                               int* correct, int B, int classes, int ld, void* stream);
/* prec@k counts: correct[i] = #rows whose label is within the top ks[i] logits (fp32 logits). */
int pdnn_topk_correct(const float* logits, const long long* labels, int B, int classes, int ld, const int* ks_host,
                      int nk, int* correct, void* stream);

/* Momentum SGD on flat fp32 buffers: g' = scale*g; m = mu*m + g' (m = g' when first_step); w -= lr*m.
 * Optionally also writes the bf16 copy of w. ref: torch.optim.SGD (nn_ops/__init__.py:43,73),
 * update_params_dist_version (sync_replicas_master_nn.py:22-28). */
int pdnn_sgd_momentum(float* w, const float* g, float* m, void* w_bf16, long long n, float lr, float mu,
                      float scale, int first_step, void* stream);

/* FC/sigmoid family (native counterpart of MPI_code/src/util/util.h:84-177 and pure_py_code/nn). */
int pdnn_sigmoid_fwd(const void* x, void* y, long long n, void* stream);               /* bf16 */
int pdnn_sigmoid_bwd(const void* y, const void* dy, void* dx, long long n, void* stream); /* dx = dy*y*(1-y) */

/* Generic fp32 SIMT GEMM for shapes the tensor-core path cannot take (LeNet: N = 10/20/50, K = 25/500).
 * C[M][N] = alpha * op(A) * op(B) (+bias)(relu), (+= C when accumulate). A/B/C row-major fp32. */
int pdnn_gemm_simt_f32(const float* A, int transA, long long lda, const float* B, int transB, long long ldb,
                       float* C, long long ldc, int M, int N, int K, float alpha, const float* bias, int relu,
                       int accumulate, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Parameter-server kernels over peer memory (csrc/ps_kernels.cu)
 * ------------------------------------------------------------------------------------------- */
#define PDNN_MAX_WORKERS 16

/* Worker -> master gradient push (ref: communicator.Isend(grad, dest=0, tag=88+l),
 * pytorch_code/model_ops/lenet.py:119, distributed_worker.py:175). Copies (and optionally casts
 * fp32 -> bf16) `n` gradient elements into the master's slot for this worker (a peer pointer), then
 * publishes `step` in the master-side arrival flag with release.sys semantics. The grid-completion counter behind the
 * flag is kept per DEVICE: pushes WITH a flag must not overlap on one GPU (one worker per GPU, the normal case). Several
 * workers on one GPU push with flag_peer = NULL and publish with pdnn_ps_publish from the same stream. */
int pdnn_ps_push(const float* grad, void* slot_peer, long long n, int slot_is_bf16, uint32_t* flag_peer,
                 uint32_t step, const uint32_t* guard_closed, const uint32_t* guard_step, int max_ctas,
                 void* stream);

/* Zero-copy push: the bucket is already where the master reads it (the worker's fp32 gradient buffer is its slot and the
 * switch sums the workers' buffers), so only the arrival flag is published -- the weight-gradient kernels' own epilogues
 * (red.add into that buffer) were the transfer. `step_ptr` non-NULL: the step tag is read from that device word.
 * ref: communicator.Isend(grad, dest=0, tag=88+l) right after each layer's backward, resnet_split.py:244-250. */
int pdnn_ps_publish(uint32_t* flag_peer, uint32_t step, const uint32_t* step_ptr, const uint32_t* guard_closed,
                    const uint32_t* guard_step, void* stream);

#define PDNN_MAX_RECEIVERS 24   /* workers + passive listeners (evaluator ranks) */
#define PDNN_MAX_BUCKETS 32

typedef struct pdnn_ps_master {
    /* inputs */
    const void* slots[PDNN_MAX_WORKERS];     /* per-worker gradient slot on the master (bf16 or fp32), whole model */
    uint32_t* arrival_flags;                 /* [n_workers][arrival_stride]: word (w, b) = step whose bucket b of worker w landed */
    int arrival_stride;
    void* recv_weights[2][PDNN_MAX_RECEIVERS];/* peer pointers: landing buffers; the weights of step s land in [s & 1] */
    uint32_t* recv_ready[PDNN_MAX_RECEIVERS];/* peer pointers: weights_ready word of each receiver (workers first) */
    uint32_t* worker_closed[PDNN_MAX_WORKERS];/* peer pointers: step_closed word of each worker */
    void* multicast_weights[2];              /* NVLS multicast address of the landing buffers (per parity) or NULL */
    const void* multicast_slots;             /* NVLS multicast address of the workers' LOCAL gradient slots or NULL: when set
                                                (and all n are aggregated) the sum is one multimem.ld_reduce in the switch */
    float* master_w;                         /* fp32 master weights */
    float* momentum;                         /* fp32 momentum buffer */
    uint32_t* control;                       /* >= 256 words of zero-initialised master-local scratch */
    long long n;                             /* number of parameters */
    int n_workers;
    int n_receivers;                         /* >= n_workers: every receiver gets the weights and the ready flag */
    int k;                                   /* aggregate the first k arrivals of every bucket */
    int slot_is_bf16;
    int weights_are_bf16;
    int avg_mode;                            /* 0: 1/k, 1: 1/n (reference python), 2: 1/received */
    float lr, mu;
    uint32_t step;                           /* step being closed; receivers are released with step+1 */
    long long deadline_ns;                   /* >0: interval policy, close with whatever arrived */
    int max_ctas;
    uint32_t* step_ptr;                      /* non-NULL: read `step` from this device word and advance it at the end, so
                                                that the launches do not depend on the step number (CUDA-graph replay) */
    int poison_slots;                        /* debug (race detection): overwrite every gradient slot the master has summed
                                                with NaN, inside the same kernel and before the workers are released. A
                                                slot that is ever summed without having been rewritten for its step then
                                                turns the weights into NaN instead of silently reusing stale data. */
    int n_buckets;                           /* gradient buckets in arrival order (deepest layers first) */
    long long bucket_lo[PDNN_MAX_BUCKETS];   /* element range of bucket b: [bucket_lo[b], bucket_hi[b]), multiples of 4 */
    long long bucket_hi[PDNN_MAX_BUCKETS];
    long long hard_timeout_ns;               /* a bucket that never sees k arrivals gives up after this long (0 = 30 s) */
    int poll_sleep_ns;                       /* pause between two looks at the arrival flags (0 = 400 ns) */
} pdnn_ps_master_t;

/* Master, one step = for every bucket b in arrival order {wait kernel, reduce kernel} enqueued on `stream`:
 * wait for the first k arrivals of bucket b of `step` (one warp polls the per-(worker, bucket) flags), sum the
 * selected slots in rank order (or in the switch), apply 1/k + momentum SGD to the fp32 master copy, cast and store
 * the new weights into every receiver's landing buffer of parity (step+1) & 1 (unicast peer stores or one multimem
 * store). Bucket b is therefore finished while the workers still compute the gradients of bucket b+1. The last
 * bucket additionally publishes the kill mask to the workers that were not among its first k (step_closed = step in
 * their memory; the reference kills on the k-th arrival of parameter 0 the same way) and releases every receiver
 * (weights_ready = step+1).
 * ref: SyncReplicasMaster_NN.train hot loop, pytorch_code/sync_replicas_master_nn.py:117-228
 * (Waitany/aggregate/kill/update/Bcast) and MPI_code/src/distributed/sync_replicas_master_nn.h:46-137.
 * result (device, 80 words): [0]=selected mask of the last bucket, [1]=kill mask, [2]=#received, [3]=step (0xDEAD:
 * the hard timeout expired before k arrivals), [4]=microseconds from the start of the step to the k-th arrival of the
 * last bucket, [5]=microseconds from that arrival to the release (the un-overlapped tail), [6]=gradient sets applied
 * since the run began, [7]=n_buckets, [8+b]=mask of bucket b, [40+b]=microseconds into the step at which bucket b
 * had its k arrivals (device %globaltimer), [56+w]=number of steps in which worker w was among the first k of the last
 * bucket since the run began. */
int pdnn_ps_master_step(const pdnn_ps_master_t* args, uint32_t* result, void* stream);
/* The {wait, reduce} pair of ONE bucket of that step; pdnn_ps_master_step(args) == for b in 0..n_buckets-1 this call.
 * Buckets must be enqueued in order on one stream. */
int pdnn_ps_master_bucket(const pdnn_ps_master_t* args, uint32_t* result, int b, void* stream);

/* Worker: copy this step's landing buffer (land1 when the step is odd) into the private compute copy. The step is
 * read from *step_word when that is non-NULL (CUDA-graph replay), else taken from `step`. */
int pdnn_ps_fetch_weights(const void* land0, const void* land1, void* dst, long long nbytes, const uint32_t* step_word,
                          uint32_t step, void* stream);

/* Worker: block the stream until *ready_flag >= step (ref: distributed_worker.py:202-208 irecv tag 10). */
int pdnn_ps_wait_flag(const uint32_t* flag, uint32_t value, uint32_t* cur_step_word, long long timeout_ns,
                      void* stream);
/* Step-number-free forms for CUDA-graph replay: the push reads its step tag from *step_ptr; wait_next advances
 * *step_word by one and waits until *flag reaches it. */
int pdnn_ps_push_dev(const float* grad, void* slot_peer, long long n, int slot_is_bf16, uint32_t* flag_peer,
                     const uint32_t* step_ptr, const uint32_t* guard_closed, const uint32_t* guard_step, int max_ctas,
                     void* stream);
int pdnn_ps_wait_next(const uint32_t* flag, uint32_t* step_word, long long timeout_ns, void* stream);
/* Device-side straggler injection: busy-wait `ns` nanoseconds on %globaltimer (aborts early when the
 * guard closes). ref: pure_py_code/distributed_worker.py:130-132 (time.sleep(0.5) per layer). */
int pdnn_delay(long long ns, const uint32_t* guard_closed, const uint32_t* guard_step, void* stream);
/* The same injection with the delay SAMPLED on the device: a pure function of (seed, worker, step, layer) through
 * Philox4x32-10, the step being read from the worker's device-side step word, so the launch does not change from step
 * to step and replays inside a CUDA graph. Distributions: the reference's constant subset delay and its (commented-out)
 * uniform 300..1200 ms (distributed_TF/src/distributed_train.py:288-290), plus shifted-exponential / log-normal / Pareto
 * tails. `log_ns` (nullable) receives the drawn nanoseconds at index `layer`. */
typedef struct pdnn_delay_spec {
    int kind;                        /* 0 none, 1 const, 2 uniform, 3 shifted exponential, 4 log-normal, 5 Pareto */
    float a, b;                      /* const: a = ms | uniform: [a, b] ms | exp: shift a, mean b | lognormal: mu a, sigma b | pareto: xm a, alpha b */
    float scale;                     /* multiplies the drawn milliseconds */
    float prob;                      /* probability that a (worker, step, layer) draw delays at all */
    unsigned long long rank_mask;    /* bit (worker + 1) set: that worker is affected (reference ranks 1..n); 0 = everyone */
    unsigned long long seed;
} pdnn_delay_spec_t;
int pdnn_delay_sampled(const pdnn_delay_spec_t* spec, int worker, int layer, const uint32_t* step_word,
                       const uint32_t* guard_closed, uint32_t* log_ns, void* stream);
int pdnn_set_word(uint32_t* p, uint32_t v, void* stream);
/* Straggler abort guard of the layer kernels (BatchNorm passes, pooling, ReLU / add, im2col / col2im, column sums,
 * cross-entropy): every such kernel launched by THIS host thread afterwards returns at once when *closed >= *step.
 * (NULL, NULL) switches it off. GEMM / conv / push entry points take the same pair through their own arguments.
 * ref: the kill check between backward ops, pytorch_code/model_ops/lenet.py:168-178. */
int pdnn_set_abort_guard(const uint32_t* closed, const uint32_t* step);

/* Peer-memory plumbing (CUDA IPC): handle is 64 bytes. */
int pdnn_ipc_get_handle(void* dev_ptr, void* handle_out64);
int pdnn_ipc_open_handle(const void* handle64, void** dev_ptr_out);
int pdnn_ipc_close(void* dev_ptr);
int pdnn_enable_peer_access(int peer_device);

/* Link microbenchmarks: copy n bytes local->dst (peer) with 16-byte stores. */
int pdnn_p2p_copy(const void* src, void* dst, long long nbytes, int max_ctas, void* stream);
/* n bytes local -> every member of a multicast group with multimem.st (one store, the switch replicates). */
int pdnn_mc_store(const void* src, void* mc_dst, long long nbytes, int max_ctas, void* stream);
/* dst[i] = sum over the multicast group's members of src[i] with multimem.ld_reduce (fp32, summed in the switch). */
int pdnn_mc_ld_reduce_f32(const void* mc_src, float* dst, long long n, int max_ctas, void* stream);
/* Flag round trip between two GPUs: `iters` release-store / acquire-load ping-pongs; *out_ns (device) = elapsed ns.
 * Launch with initiator = 1 on one GPU and 0 on the other, `mine` local, `peers` the other GPU's word. */
int pdnn_flag_pingpong(uint32_t* mine, uint32_t* peers, int iters, int initiator, uint32_t base, unsigned long long* out_ns,
                       void* stream);

/* fp32 variants used by the small-shape (LeNet / FC-MLP) path; same semantics as the bf16 entry points above. */
int pdnn_nchw_f32_to_nhwc_f32(const float* x, float* y, int N, int C, int H, int W, int Cpad, void* stream);
int pdnn_im2col_f32(const float* x, float* col, int N, int H, int W, int C, int R, int S, int stride, int pad, int OH, int OW,
                    int Kpad, void* stream);
int pdnn_col2im_f32(const float* dcol, float* dx, int N, int H, int W, int C, int R, int S, int stride, int pad, int OH, int OW,
                    int Kpad, void* stream);
int pdnn_relu_fwd_f32(const float* x, float* y, long long n, void* stream);
int pdnn_relu_bwd_f32(const float* y, const float* dy, float* dx, long long n, void* stream);
int pdnn_maxpool2_fwd_f32(const float* x, float* y, int N, int H, int W, int C, void* stream);
int pdnn_maxpool2_bwd_f32(const float* x, const float* y, const float* dy, float* dx, int N, int H, int W, int C, void* stream);
int pdnn_col_sum_f32(const float* dy, long long ld, float* out, long long M, int C, int accumulate, void* stream);
int pdnn_cross_entropy_fwd_bwd_f32(const float* logits, const long long* labels, float* loss, float* dlogits, int* correct,
                                   int B, int classes, int ld, void* stream);
int pdnn_sigmoid_fwd_f32(const float* x, float* y, long long n, void* stream);
int pdnn_sigmoid_bwd_f32(const float* y, const float* dy, float* dx, long long n, void* stream);

/* raw device allocations for peer-mapped communication arenas (zero-filled) */
int pdnn_malloc(void** out, long long nbytes);
int pdnn_free(void* p);

/* ---------------------------------------------------------------------------------------------
 * Native fully-connected parameter server (csrc/fc_ps_native.cu): master / worker / evaluator roles as host
 * threads of one process, one GPU per role where available. ref: MPI_code/ (the reference's only native code).
 * ------------------------------------------------------------------------------------------- */
typedef struct pdnn_fc_ps_config {
    int n_dims;                 /* number of entries of dims, e.g. 9 for 784-500-500-800-800-200-100-100-10 */
    int dims[64];
    int batch;                  /* reference: 128 */
    float lr;                   /* reference: 1e-3 */
    int n_workers;              /* reference: n_procs - 2 = 6 */
    int n_to_collect;           /* reference: n_procs - 2 - 4 = 2 */
    int n_train_iters;          /* reference N_TRAIN_ITERS: 100 */
    int shortcircuit;           /* reference SHORTCIRCUIT: workers abandon a step when a newer one is published */
    int generate_timeline;      /* reference GENERATE_TIMELINE: outfiles/timeline_out_<scheme> */
    int with_evaluator;         /* run the evaluator role (outfiles/time_loss_out_<scheme>: step time_ms loss err) */
    int eval_batches;
    unsigned long long delay_worker_mask;  /* bit w set: worker w sleeps delay_us_per_layer before each backward layer */
    long long delay_us_per_layer;
    unsigned seed;
    int synthetic_examples;     /* size of the synthetic set used when data_dir has no MNIST IDX files */
    int max_devices;            /* 0 = use every visible GPU */
    char data_dir[256];
    char out_dir[256];
    int device_protocol;        /* 1: roles talk through the parameter-server kernels (per-(worker, layer) device flags, peer
                                   stores, pdnn_ps_master_step per step); 0: host threads with atomics + peer copies */
} pdnn_fc_ps_config_t;

typedef struct pdnn_fc_ps_result {
    int steps;
    double total_ms, mean_step_ms;
    long long gradients_accepted, gradients_dropped_stale, steps_abandoned;
    double avg_gradients_per_step;   /* reference visualize_timeline.py: average #gradients received per step */
    float first_loss, final_loss, first_err, final_err;
    char scheme[128];                /* reference scheme_full_name(): SyncReplicasWithBackup<k>_<n>[_shortcircuit] */
} pdnn_fc_ps_result_t;

int pdnn_fc_ps_run(const pdnn_fc_ps_config_t* cfg, pdnn_fc_ps_result_t* result);

/* Same run with caller-side inputs / outputs (all HOST pointers, any of them may be NULL): the data set, the initial
 * weights, the final weights and the list of gradients the master accumulated, in accumulation order. Which gradients
 * win a step is timing dependent (first n_to_collect per layer, sync_replicas_master_nn.h:71-101); with the log a
 * caller can replay exactly that set through an oracle. Weight layout: layers concatenated, each
 * [(dims[l] + 1) x dims[l+1]] row-major with the bias as last row. */
typedef struct pdnn_fc_ps_io {
    const float* train_images;       /* [n_train][dims[0]] */
    const long long* train_labels;
    int n_train;
    const float* test_images;        /* evaluator set; NULL = the training set */
    const long long* test_labels;
    int n_test;
    const float* init_weights;       /* NULL = N(0, 0.1^2), zero bias */
    float* final_weights;            /* out */
    int* accept_log;                 /* out: rows of (step, worker, layer, first row of the worker's batch) */
    long long accept_log_rows;       /* capacity in rows */
    long long accept_log_used;       /* out: rows written */
} pdnn_fc_ps_io_t;
int pdnn_fc_ps_run_ex(const pdnn_fc_ps_config_t* cfg, pdnn_fc_ps_result_t* result, pdnn_fc_ps_io_t* io);
int pdnn_fc_ps_selftest(int iters, float* first_loss, float* last_loss);

#ifdef __cplusplus
}
#endif
#endif /* PDNN_B200_H */
