// Native (C++/CUDA) fully-connected parameter server: master / worker / evaluator roles without Python in the loop.
//
// Counterpart of the reference's only native component, MPI_code/ (C++11 + MPI + cblas_dgemm):
//   NNParams            nn_params.h:16-76          -> fcps::NNParams (layer list, batch, lr, dimension validation)
//   NNLayer / NN        nn_layer.h, nn.h           -> fcps::Net (bias folded in as the last weight row, sigmoid hidden
//                                                    layers, softmax cross-entropy output, loss / error rate)
//   SyncReplicasMasterNN sync_replicas_master_nn.h -> fcps::master_loop ((layer, step)-addressed slots, arrival-order
//                                                    accumulation, stale contributions dropped by step tag, per-layer
//                                                    mean over what arrived, n_to_collect threshold)
//   WorkerNN            worker_nn.h                -> fcps::worker_loop (just-in-time per-layer weight fetch, per-layer
//                                                    gradient push during backward, SHORTCIRCUIT: abandon the step the
//                                                    moment a newer step is published)
//   EvaluatorNN         evaluator_nn.h             -> fcps::evaluator_loop (outfiles/time_loss_out_<scheme>)
//   mnist.h             IDX reader, /255, shuffle  -> fcps::load_mnist (+ synthetic fallback, loader self-test)
//
// B200 design: roles are host threads of ONE process, each bound to a GPU (all on one GPU when only one exists); weights,
// slots and activations live in device memory and the math runs on this library's kernels. Two transports:
//   device protocol (default, cfg.device_protocol = 1): the roles talk through the parameter-server kernels of
//     ps_kernels.cu -- per-(worker, layer) step-tagged device flags, peer stores (pdnn_ps_push), the per-bucket master
//     (pdnn_ps_master_step), parity landing buffers + ready words, abort guards -- and the threads only enqueue;
//   host protocol (cfg.device_protocol = 0): step tags in host atomics, condition variables, device-to-device copies on
//     the sender's stream.
#include "common.cuh"
#include "../include/pdnn_b200.h"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <fstream>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>
#include <string.h>
#include <sys/stat.h>

namespace fcps {

#define FC_CUDA(expr)                                                                               \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess) {                                                                    \
            fprintf(stderr, "fc_ps: %s:%d %s -> %s\n", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            failed.store(true);                                                                     \
            return;                                                                                 \
        }                                                                                           \
    } while (0)

static double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

__global__ void axpy_kernel(float* __restrict__ y, const float* __restrict__ x, float alpha, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) y[i] += alpha * x[i];
}
__global__ void apply_grad_kernel(const float* __restrict__ w_old, const float* __restrict__ acc, float* __restrict__ w_new, float alpha, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        w_new[i] = w_old[i] - alpha * acc[i];
}
__global__ void error_count_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, int B, int C, int* wrong) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < B; r += gridDim.x * blockDim.x) {
        int best = 0;
        for (int c = 1; c < C; ++c) if (logits[(long long)r * C + c] > logits[(long long)r * C + best]) best = c;
        if (best != (int)labels[r]) atomicAdd(wrong, 1);
    }
}

struct NNParams {
    std::vector<int> dims;
    int batch = 128;
    float lr = 1e-3f;
    bool validate(std::string* why) const {
        if (dims.size() < 2) { *why = "need at least an input and an output dimension"; return false; }
        for (int d : dims) if (d <= 0) { *why = "non-positive layer dimension"; return false; }
        if (batch <= 0) { *why = "non-positive batch size"; return false; }
        return true;
    }
    int n_layers() const { return (int)dims.size() - 1; }
    long long layer_elems(int l) const { return (long long)(dims[l] + 1) * dims[l + 1]; }   // bias folded in as last row
};

struct Dataset {
    std::vector<float> images;       // [n][784] scaled to [0,1]
    std::vector<long long> labels;
    int n = 0, dim = 784;
};

// big-endian IDX files (reference mnist.h:25-86); /255 scaling (:102-109); Fisher-Yates shuffle (:132-138)
static bool read_idx(const std::string& img, const std::string& lab, Dataset* d) {
    std::ifstream fi(img, std::ios::binary), fl(lab, std::ios::binary);
    if (!fi || !fl) return false;
    auto be32 = [](std::ifstream& f) { unsigned char b[4]; f.read((char*)b, 4); return (int)((b[0] << 24) | (b[1] << 16) | (b[2] << 8) | b[3]); };
    if (be32(fi) != 2051) return false;
    int n = be32(fi), r = be32(fi), c = be32(fi);
    if (be32(fl) != 2049 || be32(fl) != n) return false;
    d->n = n; d->dim = r * c;
    std::vector<unsigned char> raw((size_t)n * r * c), lraw(n);
    fi.read((char*)raw.data(), raw.size());
    fl.read((char*)lraw.data(), n);
    d->images.resize(raw.size());
    for (size_t i = 0; i < raw.size(); ++i) d->images[i] = raw[i] / 255.f;
    d->labels.assign(lraw.begin(), lraw.end());
    return true;
}
static void synthetic(Dataset* d, int n, int dim, int classes, unsigned seed) {
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::normal_distribution<float> g(0.f, 1.f);
    std::vector<float> proj((size_t)dim * classes);
    std::mt19937 prng(12345);
    for (auto& p : proj) p = g(prng);
    d->n = n; d->dim = dim;
    d->images.resize((size_t)n * dim);
    d->labels.resize(n);
    for (int i = 0; i < n; ++i) {
        float mean = 0.f;
        for (int j = 0; j < dim; ++j) { d->images[(size_t)i * dim + j] = u(rng); mean += d->images[(size_t)i * dim + j]; }
        mean /= dim;
        int best = 0; float bv = -1e30f;
        for (int c = 0; c < classes; ++c) {
            float s = 0.f;
            for (int j = 0; j < dim; ++j) s += (d->images[(size_t)i * dim + j] - mean) * proj[(size_t)j * classes + c];
            if (s > bv) { bv = s; best = c; }
        }
        d->labels[i] = best;
    }
}
static void shuffle(Dataset* d, unsigned seed) {
    std::mt19937 rng(seed);
    for (int i = d->n - 1; i > 0; --i) {
        int j = (int)(rng() % (unsigned)(i + 1));
        if (i == j) continue;
        for (int k = 0; k < d->dim; ++k) std::swap(d->images[(size_t)i * d->dim + k], d->images[(size_t)j * d->dim + k]);
        std::swap(d->labels[i], d->labels[j]);
    }
}
static bool load_mnist(const char* dir, Dataset* d, int synth_n, int dim, int classes, unsigned seed) {
    if (dir != nullptr && dir[0]) {
        std::string b(dir);
        if (read_idx(b + "/train-images-idx3-ubyte", b + "/train-labels-idx1-ubyte", d)) { shuffle(d, seed); return true; }
    }
    synthetic(d, synth_n, dim, classes, seed);
    return false;
}

// A network instance of one role: device buffers on that role's GPU.
struct Net {
    NNParams p;
    int device = 0;
    cudaStream_t stream = nullptr;
    std::vector<float*> W, G, Z, D;   // per layer: weights, gradient, activation output, delta
    float* X = nullptr;               // input batch
    long long* Y = nullptr;
    float* loss = nullptr;
    int* wrong = nullptr;
    std::atomic<bool>& failed;
    Net(const NNParams& prm, int dev, std::atomic<bool>& f) : p(prm), device(dev), failed(f) {}
    // flat = layer offsets into ONE weight and ONE gradient vector (device protocol: the parameter-server kernels work on
    // flat buffers and address a layer as a bucket = an element range)
    float* Wflat = nullptr;
    float* Gflat = nullptr;
    void init(const std::vector<long long>* flat_off = nullptr, long long flat_total = 0) {
        FC_CUDA(cudaSetDevice(device));
        FC_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        const int L = p.n_layers();
        W.resize(L); G.resize(L); Z.resize(L); D.resize(L);
        if (flat_off != nullptr) {
            FC_CUDA(cudaMalloc(&Wflat, sizeof(float) * flat_total));
            FC_CUDA(cudaMalloc(&Gflat, sizeof(float) * flat_total));
            FC_CUDA(cudaMemset(Wflat, 0, sizeof(float) * flat_total));
            FC_CUDA(cudaMemset(Gflat, 0, sizeof(float) * flat_total));
        }
        for (int l = 0; l < L; ++l) {
            if (flat_off != nullptr) {
                W[l] = Wflat + (*flat_off)[l];
                G[l] = Gflat + (*flat_off)[l];
            } else {
                FC_CUDA(cudaMalloc(&W[l], sizeof(float) * p.layer_elems(l)));
                FC_CUDA(cudaMalloc(&G[l], sizeof(float) * p.layer_elems(l)));
            }
            FC_CUDA(cudaMalloc(&Z[l], sizeof(float) * (size_t)p.batch * p.dims[l + 1]));
            FC_CUDA(cudaMalloc(&D[l], sizeof(float) * (size_t)p.batch * p.dims[l + 1]));
        }
        FC_CUDA(cudaMalloc(&X, sizeof(float) * (size_t)p.batch * p.dims[0]));
        FC_CUDA(cudaMalloc(&Y, sizeof(long long) * p.batch));
        FC_CUDA(cudaMalloc(&loss, sizeof(float)));
        FC_CUDA(cudaMalloc(&wrong, sizeof(int)));
    }
    void destroy() {
        cudaSetDevice(device);
        if (Wflat != nullptr) { cudaFree(Wflat); cudaFree(Gflat); for (auto v : {&Z, &D}) for (float* q : *v) cudaFree(q); }
        else for (auto v : {&W, &G, &Z, &D}) for (float* q : *v) cudaFree(q);
        cudaFree(X); cudaFree(Y); cudaFree(loss); cudaFree(wrong);
        if (stream) cudaStreamDestroy(stream);
    }
    void fill_batch(const Dataset& d, int* cursor, int* used = nullptr) {
        if (*cursor + p.batch > d.n) *cursor = 0;
        if (used != nullptr) *used = *cursor;
        FC_CUDA(cudaMemcpyAsync(X, d.images.data() + (size_t)*cursor * d.dim, sizeof(float) * (size_t)p.batch * d.dim, cudaMemcpyHostToDevice, stream));
        FC_CUDA(cudaMemcpyAsync(Y, d.labels.data() + *cursor, sizeof(long long) * p.batch, cudaMemcpyHostToDevice, stream));
        *cursor += p.batch;
    }
    // S = [Z|1] W ; Z = sigmoid(S) on hidden layers (reference nn_layer.h:97-142)
    bool forward_layer(int l) {
        const int L = p.n_layers();
        const float* in = l == 0 ? X : Z[l - 1];
        const int n_in = p.dims[l], n_out = p.dims[l + 1];
        if (pdnn_gemm_simt_f32(in, 0, n_in, W[l], 0, n_out, Z[l], n_out, p.batch, n_out, n_in, 1.f, W[l] + (size_t)n_in * n_out, 0, 0, stream)) return false;
        if (l < L - 1 && pdnn_sigmoid_fwd_f32(Z[l], Z[l], (long long)p.batch * n_out, stream)) return false;
        return true;
    }
    bool output_delta() {   // softmax + cross-entropy: D = (softmax - onehot) / B, loss = mean CE
        const int L = p.n_layers();
        cudaMemsetAsync(loss, 0, sizeof(float), stream);
        return pdnn_cross_entropy_fwd_bwd_f32(Z[L - 1], Y, loss, D[L - 1], nullptr, p.batch, p.dims[L], p.dims[L], stream) == 0;
    }
    // grad = [Z|1]^T D ; D_prev = (D W^T) o sigmoid'  (reference nn_layer.h:144-178)
    bool backward_layer(int l) {
        const float* in = l == 0 ? X : Z[l - 1];
        const int n_in = p.dims[l], n_out = p.dims[l + 1];
        cudaMemsetAsync(G[l], 0, sizeof(float) * p.layer_elems(l), stream);
        if (pdnn_gemm_simt_f32(in, 1, n_in, D[l], 0, n_out, G[l], n_out, n_in, n_out, p.batch, 1.f, nullptr, 0, 1, stream)) return false;
        if (pdnn_col_sum_f32(D[l], n_out, G[l] + (size_t)n_in * n_out, p.batch, n_out, 1, stream)) return false;
        if (l > 0) {
            if (pdnn_gemm_simt_f32(D[l], 0, n_out, W[l], 1, n_out, D[l - 1], n_in, p.batch, n_in, n_out, 1.f, nullptr, 0, 0, stream)) return false;
            if (pdnn_sigmoid_bwd_f32(Z[l - 1], D[l - 1], D[l - 1], (long long)p.batch * n_in, stream)) return false;
        }
        return true;
    }
};

struct Arrival { int worker, layer, step, batch_start; };

struct Board {
    NNParams p;
    pdnn_fc_ps_config_t cfg;
    int n_workers = 0;
    std::atomic<int> cur_step{0};
    std::atomic<bool> failed{false};
    std::mutex mu;
    std::condition_variable cv_step, cv_arrival;
    std::deque<Arrival> arrivals;
    // master-side device memory (device 0): weights double-buffered by step parity, one slot per (worker, layer)
    std::vector<float*> Wm[2];
    std::vector<std::vector<float*>> slots;
    std::vector<float*> acc;
    int master_device = 0;
    // evaluator hand-over: the master copies the newest weights into We whenever the evaluator is idle, so the evaluator
    // never reads a buffer the master is writing (Wm[par] is rewritten one step later)
    std::vector<float*> We;
    std::mutex mu_eval;
    std::condition_variable cv_eval;
    int eval_step = 0;            // step whose weights sit in We (0 = none yet, -1 = finished)
    bool eval_busy = false;       // evaluator still copying We into its own network
    bool we_writing = false;      // master is filling We
    // optional acceptance log for replay against an oracle: (step, worker, layer, batch_start) in accumulation order
    int* accept_log = nullptr;
    long long accept_cap = 0, accept_n = 0;
    // statistics
    std::vector<double> step_ms;
    long long accepted = 0, dropped_stale = 0, abandoned_steps = 0;
    std::vector<float> eval_loss, eval_err;
    std::vector<double> timeline_ms;
    std::vector<int> timeline_step, timeline_who;
    std::string scheme;
};

static void copy_d2d(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes, cudaStream_t s) {
    if (dst_dev == src_dev) cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s);
    else cudaMemcpyPeerAsync(dst, dst_dev, src, src_dev, bytes, s);
}

static void worker_loop(Board* b, int w, int device, const Dataset* data) {
    std::atomic<bool>& failed = b->failed;
    Net net(b->p, device, failed);
    net.init();
    if (failed.load()) return;
    const int L = b->p.n_layers();
    int my_step = 0, cursor = (w * 997 * b->p.batch) % std::max(1, data->n - b->p.batch);
    const long long delay_us = (b->cfg.delay_worker_mask >> w) & 1 ? b->cfg.delay_us_per_layer : 0;
    auto step_changed = [&]() { return b->cfg.shortcircuit && b->cur_step.load(std::memory_order_acquire) != my_step; };
    while (!failed.load()) {
        {
            std::unique_lock<std::mutex> lk(b->mu);
            b->cv_step.wait(lk, [&] { return b->cur_step.load() != my_step || failed.load(); });
            my_step = b->cur_step.load();
        }
        if (my_step > b->cfg.n_train_iters || my_step < 0) break;
        const int par = my_step & 1;
        int batch_start = 0;
        net.fill_batch(*data, &cursor, &batch_start);
        bool abandoned = false;
        for (int l = 0; l < L && !abandoned; ++l) {
            if (step_changed()) { abandoned = true; break; }
            // just-in-time fetch of layer l's weights, overlapped with the previous layer's compute (worker_nn.h:56-74)
            copy_d2d(net.W[l], device, b->Wm[par][l], b->master_device, sizeof(float) * b->p.layer_elems(l), net.stream);
            if (!net.forward_layer(l)) { failed.store(true); break; }
        }
        if (!abandoned && !net.output_delta()) failed.store(true);
        for (int l = L - 1; l >= 0 && !abandoned && !failed.load(); --l) {
            if (delay_us > 0) std::this_thread::sleep_for(std::chrono::microseconds(delay_us));   // injected straggler (pure_py worker:130-132)
            if (step_changed()) { abandoned = true; break; }
            if (!net.backward_layer(l)) { failed.store(true); break; }
            // push this layer's gradient into my slot on the master and announce it, tagged with the step
            copy_d2d(b->slots[w][l], b->master_device, net.G[l], device, sizeof(float) * b->p.layer_elems(l), net.stream);
            cudaStreamSynchronize(net.stream);
            {
                std::lock_guard<std::mutex> lk(b->mu);
                b->arrivals.push_back({w, l, my_step, batch_start});
            }
            b->cv_arrival.notify_one();
        }
        cudaStreamSynchronize(net.stream);
        if (abandoned) { std::lock_guard<std::mutex> lk(b->mu); b->abandoned_steps++; }
    }
    net.destroy();
}

static void evaluator_loop(Board* b, int device, const Dataset* data) {
    std::atomic<bool>& failed = b->failed;
    Net net(b->p, device, failed);
    net.init();
    if (failed.load()) return;
    const int L = b->p.n_layers();
    int my_step = 0;
    const double t0 = now_ms();
    mkdir("outfiles", 0755);
    std::string path = std::string(b->cfg.out_dir[0] ? b->cfg.out_dir : "outfiles") + "/time_loss_out_" + b->scheme;
    mkdir(b->cfg.out_dir[0] ? b->cfg.out_dir : "outfiles", 0755);
    FILE* f = fopen(path.c_str(), "w");
    while (!failed.load()) {
        {
            std::unique_lock<std::mutex> lk(b->mu_eval);
            while (!((b->eval_step != my_step && !b->we_writing) || failed.load())) b->cv_eval.wait_for(lk, std::chrono::milliseconds(100));
            if (failed.load() || b->eval_step < 0) break;
            my_step = b->eval_step;
            b->eval_busy = true;
        }
        for (int l = 0; l < L; ++l)
            copy_d2d(net.W[l], device, b->We[l], b->master_device, sizeof(float) * b->p.layer_elems(l), net.stream);
        cudaStreamSynchronize(net.stream);
        {
            std::lock_guard<std::mutex> lk(b->mu_eval);
            b->eval_busy = false;            // We may be refreshed from here on
        }
        b->cv_eval.notify_all();
        float loss_sum = 0.f; int wrong_sum = 0, seen = 0, cursor = 0;
        for (int it = 0; it < b->cfg.eval_batches && cursor + b->p.batch <= data->n; ++it) {
            net.fill_batch(*data, &cursor);
            for (int l = 0; l < L; ++l) net.forward_layer(l);
            net.output_delta();
            cudaMemsetAsync(net.wrong, 0, sizeof(int), net.stream);
            error_count_kernel<<<4, 128, 0, net.stream>>>(net.Z[L - 1], net.Y, b->p.batch, b->p.dims[L], net.wrong);
            float hl; int hw;
            cudaMemcpyAsync(&hl, net.loss, sizeof(float), cudaMemcpyDeviceToHost, net.stream);
            cudaMemcpyAsync(&hw, net.wrong, sizeof(int), cudaMemcpyDeviceToHost, net.stream);
            cudaStreamSynchronize(net.stream);
            loss_sum += hl; wrong_sum += hw; seen += b->p.batch;
        }
        const float loss = loss_sum / std::max(1, b->cfg.eval_batches), err = (float)wrong_sum / std::max(1, seen);
        {
            std::lock_guard<std::mutex> lk(b->mu);
            b->eval_loss.push_back(loss);
            b->eval_err.push_back(err);
        }
        if (f) { fprintf(f, "%d %.3f %.6f %.6f\n", my_step, now_ms() - t0, loss, err); fflush(f); }   // step time loss err (evaluator_nn.h:58)
        if (my_step > b->cfg.n_train_iters) break;      // that was the final model
    }
    if (f) fclose(f);
    net.destroy();
}

// Weights of `step` (in Wm[step & 1]) -> We, when the evaluator is idle (or, with `must`, after waiting for it).
static void hand_to_evaluator(Board* b, int step, cudaStream_t stream, bool must) {
    if (!b->cfg.with_evaluator) return;
    const int L = b->p.n_layers();
    {
        std::unique_lock<std::mutex> lk(b->mu_eval);
        if (must) { while (b->eval_busy && !b->failed.load()) b->cv_eval.wait_for(lk, std::chrono::milliseconds(100)); }
        else if (b->eval_busy) return;
        if (b->failed.load()) return;
        b->we_writing = true;
    }
    // the evaluator reads We only while eval_busy is set, and cannot set it while we_writing is
    for (int l = 0; l < L; ++l)
        cudaMemcpyAsync(b->We[l], b->Wm[step & 1][l], sizeof(float) * b->p.layer_elems(l), cudaMemcpyDeviceToDevice, stream);
    cudaStreamSynchronize(stream);
    {
        std::lock_guard<std::mutex> lk(b->mu_eval);
        b->eval_step = step;
        b->we_writing = false;
    }
    b->cv_eval.notify_all();
}

static void master_loop(Board* b) {
    std::atomic<bool>& failed = b->failed;
    const int L = b->p.n_layers();
    FC_CUDA(cudaSetDevice(b->master_device));
    cudaStream_t stream;
    FC_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    const int n_collect = b->cfg.n_to_collect;
    std::vector<int> count(L);
    std::vector<std::vector<char>> seen(b->n_workers, std::vector<char>(L, 0));
    const double t_begin = now_ms();
    for (int step = 1; step <= b->cfg.n_train_iters && !failed.load(); ++step) {
        const double t0 = now_ms();
        const int par = step & 1;
        std::fill(count.begin(), count.end(), 0);
        for (auto& s : seen) std::fill(s.begin(), s.end(), 0);
        for (int l = 0; l < L; ++l) cudaMemsetAsync(b->acc[l], 0, sizeof(float) * b->p.layer_elems(l), stream);
        hand_to_evaluator(b, step, stream, false);
        {
            std::lock_guard<std::mutex> lk(b->mu);
            b->cur_step.store(step, std::memory_order_release);      // the step message + weights "broadcast" (weights are already in Wm[par])
            if (b->cfg.generate_timeline) { b->timeline_ms.push_back(now_ms() - t_begin); b->timeline_step.push_back(step); b->timeline_who.push_back(-1); }
        }
        b->cv_step.notify_all();
        bool enough = false;
        while (!enough && !failed.load()) {
            Arrival a;
            {
                std::unique_lock<std::mutex> lk(b->mu);
                b->cv_arrival.wait(lk, [&] { return !b->arrivals.empty() || failed.load(); });
                if (failed.load()) break;
                a = b->arrivals.front();
                b->arrivals.pop_front();
                if (a.step != step) { b->dropped_stale++; continue; }            // stale: tag != cur_step (master.h:85)
                b->accepted++;
                if (b->cfg.generate_timeline) { b->timeline_ms.push_back(now_ms() - t_begin); b->timeline_step.push_back(step); b->timeline_who.push_back(a.worker); }
            }
            if (seen[a.worker][a.layer]) continue;
            seen[a.worker][a.layer] = 1;
            if (b->accept_log != nullptr && b->accept_n < b->accept_cap) {
                int* row = b->accept_log + 4 * b->accept_n++;
                row[0] = step; row[1] = a.worker; row[2] = a.layer; row[3] = a.batch_start;
            }
            const long long n = b->p.layer_elems(a.layer);
            axpy_kernel<<<(int)std::min<long long>((n + 255) / 256, 592), 256, 0, stream>>>(b->acc[a.layer], b->slots[a.worker][a.layer], 1.f, n);
            count[a.layer]++;
            enough = true;
            for (int l = 0; l < L; ++l) enough = enough && count[l] >= n_collect;
        }
        // apply lr / gradients_accumulated[layer]: mean over what actually arrived, per layer (master.h:124-128)
        for (int l = 0; l < L; ++l) {
            const long long n = b->p.layer_elems(l);
            apply_grad_kernel<<<(int)std::min<long long>((n + 255) / 256, 592), 256, 0, stream>>>(b->Wm[par][l], b->acc[l], b->Wm[par ^ 1][l],
                                                                                                b->p.lr / (float)std::max(1, count[l]), n);
        }
        FC_CUDA(cudaStreamSynchronize(stream));
        b->step_ms.push_back(now_ms() - t0);
    }
    {
        std::lock_guard<std::mutex> lk(b->mu);
        b->cur_step.store(b->cfg.n_train_iters + 1, std::memory_order_release);   // the last step message ends the workers
    }
    b->cv_step.notify_all();
    // the evaluator always gets the final model (step n_train_iters + 1), then the end marker
    if (!failed.load()) hand_to_evaluator(b, b->cfg.n_train_iters + 1, stream, true);
    {
        std::unique_lock<std::mutex> lk(b->mu_eval);
        if (!b->cfg.with_evaluator || failed.load()) b->eval_step = -1;
    }
    b->cv_eval.notify_all();
    cudaStreamDestroy(stream);
}


// =================================================================================================================
// Device protocol (cfg.device_protocol): the same roles driven by the parameter-server kernels of ps_kernels.cu instead of
// host atomics, condition variables and cudaMemcpyPeerAsync. A layer is a bucket of ONE flat parameter vector:
//   worker    pdnn_ps_wait_flag (ready >= step) -> per layer, just in time: pdnn_ps_fetch_weights of that layer's range
//             from the landing buffer of the step's parity -> forward; backward layer by layer, each gradient pushed with
//             pdnn_ps_push (peer stores + step-tagged flag (worker, layer)); with SHORTCIRCUIT the kernels carry the abort
//             guard, so a worker the master has closed the step for falls through the rest of it (worker_nn.h:77-108);
//   master    pdnn_ps_master_step: per layer in arrival order the first n_to_collect flags, mean over them, W -= lr * mean,
//             new weights stored into every receiver's landing buffer, kill word + ready flag (sync_replicas_master_nn.h);
//   evaluator a passive receiver: polls its ready word, snapshots the landing buffer of that parity, scores it.
// The host threads only enqueue; nobody synchronises inside a step except the master's per-step read of the result block
// (selection masks for the acceptance log and the timeline).
// =================================================================================================================
struct DevState {
    int L = 0, n_workers = 0, n_recv = 0, stride = PDNN_MAX_BUCKETS;
    std::vector<long long> off, len;      // per layer: offset into the flat vector (64-float aligned), padded length (x4)
    long long total = 0;
    int master_device = 0;
    float* w_master = nullptr;            // master device: fp32 weights, momentum (unused: mu = 0), slots, control words
    float* mom = nullptr;
    std::vector<float*> slot;             // [worker] flat gradient slot on the master device
    uint32_t* ctl = nullptr;              // control[256] | result[80] | arrival[n_workers][stride]
    uint32_t* result = nullptr;
    uint32_t* arrival = nullptr;
    std::vector<int> recv_device;         // receivers: workers 0..n-1, then the evaluator
    std::vector<float*> land[2];          // [parity][receiver] landing buffers on the receiver's device
    std::vector<uint32_t*> flags;         // [receiver]: word 0 ready, 1 closed, 2 step
    std::vector<int> batch_start;         // [worker * (iters + 2) + step]: first row of the batch the worker used at that step
    std::atomic<int> inited{0};           // roles whose device buffers exist: the master launches nothing before all of them do
                                          // (an allocation can block the host while a kernel that polls a flag is resident)
};

static void dev_worker_loop(Board* b, DevState* d, int w, int device, const Dataset* data) {
    std::atomic<bool>& failed = b->failed;
    Net net(b->p, device, failed);
    net.init(&d->off, d->total);
    const int L = d->L, iters = b->cfg.n_train_iters;
    int cursor = (w * 997 * b->p.batch) % std::max(1, data->n - b->p.batch);
    if (!failed.load()) {
        // Dry run of every kernel this role will launch. With CUDA's lazy module loading the FIRST launch of a kernel loads
        // it, and that load can wait for the whole context to drain -- which never happens while another role's kernel is
        // polling a flag. So every role touches its kernels now, and nobody polls before all roles have done so.
        int c0 = cursor;
        pdnn_ps_wait_flag(d->flags[w] + 3, 0u, d->flags[w] + 4, 0, net.stream);
        net.fill_batch(*data, &c0);
        for (int l = 0; l < L; ++l) {
            pdnn_ps_fetch_weights(d->land[0][w] + d->off[l], d->land[1][w] + d->off[l], net.W[l], sizeof(float) * d->len[l], nullptr, 1u, net.stream);
            net.forward_layer(l);
        }
        net.output_delta();
        for (int l = L - 1; l >= 0; --l) {
            net.backward_layer(l);
            pdnn_ps_push(net.G[l], net.G[l], b->p.layer_elems(l), 0, nullptr, 0u, nullptr, nullptr, 0, net.stream);   // onto itself: the slots stay zero for step 0
        }
        pdnn_delay(1000, nullptr, nullptr, net.stream);
        pdnn_ps_publish(d->flags[w] + 5, 0u, nullptr, nullptr, nullptr, net.stream);
        cudaStreamSynchronize(net.stream);
    }
    cudaDeviceSynchronize();
    d->inited.fetch_add(1);
    while (d->inited.load() < d->n_recv + 1 && !failed.load()) std::this_thread::sleep_for(std::chrono::microseconds(100));
    if (failed.load()) return;
    uint32_t* fl = d->flags[w];
    const uint32_t* closed = b->cfg.shortcircuit ? fl + 1 : nullptr;
    const uint32_t* stepw = b->cfg.shortcircuit ? fl + 2 : nullptr;
    const long long delay_ns = ((b->cfg.delay_worker_mask >> w) & 1) ? b->cfg.delay_us_per_layer * 1000ll : 0;
    pdnn_set_abort_guard(closed, stepw);                 // thread-local: the layer kernels this thread launches carry it
    for (int step = 1; step <= iters && !failed.load(); ++step) {
        if (pdnn_ps_wait_flag(fl, (uint32_t)step, fl + 2, 30000000000ll, net.stream)) { failed.store(true); break; }   // bounded: 30 s
        int used = 0;
        net.fill_batch(*data, &cursor, &used);
        d->batch_start[(size_t)w * (iters + 2) + step] = used;
        bool ok = true;
        for (int l = 0; l < L && ok; ++l) {
            // just-in-time fetch of layer l's weights from this step's landing buffer (worker_nn.h:56-74)
            ok = pdnn_ps_fetch_weights(d->land[0][w] + d->off[l], d->land[1][w] + d->off[l], net.W[l], sizeof(float) * d->len[l], nullptr,
                                       (uint32_t)step, net.stream) == 0 && net.forward_layer(l);
        }
        ok = ok && net.output_delta();
        static const bool dbg = getenv("PDNN_FC_DEBUG") != nullptr;
        if (dbg) { cudaError_t e = cudaStreamSynchronize(net.stream); fprintf(stderr, "fc_ps dbg: worker %d step %d forward done ok=%d cuda=%s\n", w, step, (int)ok, cudaGetErrorString(e)); }
        for (int l = L - 1; l >= 0 && ok; --l) {
            if (delay_ns > 0) pdnn_delay(delay_ns, closed, stepw, net.stream);      // injected straggler, on the device
            // copy, then publish the (worker, layer) flag from a second tiny kernel: several workers may share one GPU here, and
            // the single-kernel form keeps its grid-completion counter per device, not per stream
            ok = net.backward_layer(l) &&
                 pdnn_ps_push(net.G[l], d->slot[w] + d->off[l], b->p.layer_elems(l), 0, nullptr, (uint32_t)step, closed, stepw, 0, net.stream) == 0 &&
                 pdnn_ps_publish(d->arrival + (size_t)w * d->stride + (L - 1 - l), (uint32_t)step, nullptr, closed, stepw, net.stream) == 0;
            if (dbg) {
                cudaError_t e = cudaStreamSynchronize(net.stream);
                uint32_t fv = 0;
                cudaMemcpy(&fv, d->arrival + (size_t)w * d->stride + (L - 1 - l), 4, cudaMemcpyDeviceToHost);
                fprintf(stderr, "fc_ps dbg: worker %d step %d layer %d pushed ok=%d cuda=%s flag=%u\n", w, step, l, (int)ok, cudaGetErrorString(e), fv);
            }
        }
        if (!ok) { fprintf(stderr, "fc_ps: worker %d failed at step %d: %s\n", w, step, pdnn_last_error()); failed.store(true); break; }
    }
    cudaStreamSynchronize(net.stream);
    pdnn_set_abort_guard(nullptr, nullptr);
    net.destroy();
}

static void dev_master_loop(Board* b, DevState* d) {
    std::atomic<bool>& failed = b->failed;
    FC_CUDA(cudaSetDevice(d->master_device));
    cudaStream_t stream;
    FC_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    const int L = d->L, iters = b->cfg.n_train_iters;
    pdnn_ps_master_t a;
    memset(&a, 0, sizeof(a));
    for (int w = 0; w < d->n_workers; ++w) { a.slots[w] = d->slot[w]; a.worker_closed[w] = d->flags[w] + 1; }
    for (int r = 0; r < d->n_recv; ++r) { a.recv_weights[0][r] = d->land[0][r]; a.recv_weights[1][r] = d->land[1][r]; a.recv_ready[r] = d->flags[r]; }
    a.arrival_flags = d->arrival; a.arrival_stride = d->stride;
    a.master_w = d->w_master; a.momentum = d->mom; a.control = d->ctl;
    a.n = d->total; a.n_workers = d->n_workers; a.n_receivers = d->n_recv; a.k = b->cfg.n_to_collect;
    a.slot_is_bf16 = 0; a.weights_are_bf16 = 0; a.avg_mode = 0; a.lr = b->p.lr; a.mu = 0.f;
    a.n_buckets = L;
    a.hard_timeout_ns = 10000000000ll;      // a bucket that never sees n_to_collect arrivals: give up after 10 s
    for (int l = 0; l < L; ++l) { a.bucket_lo[L - 1 - l] = d->off[l]; a.bucket_hi[L - 1 - l] = d->off[l] + d->len[l]; }   // deepest layer arrives first
    std::vector<uint32_t> res(80);
    // Step 0 is the initial broadcast: every arrival word still holds 0, every slot is zero, so the master "aggregates" zero
    // gradients, leaves the weights as they are, stores them into every receiver's landing buffer of parity 1 and publishes
    // ready = 1 (the first step message + weight broadcast, sync_replicas_master_nn.h:46-60). It also loads the master's
    // kernels before anybody polls (see dev_worker_loop).
    a.step = 0u;
    if (pdnn_ps_master_step(&a, d->result, stream)) { failed.store(true); }
    FC_CUDA(cudaStreamSynchronize(stream));
    while (d->inited.load() < d->n_recv && !failed.load()) std::this_thread::sleep_for(std::chrono::microseconds(100));
    d->inited.fetch_add(1);                     // releases the other roles into step 1
    const double t_begin = now_ms();
    for (int step = 1; step <= iters && !failed.load(); ++step) {
        const double t0 = now_ms();
        if (b->cfg.generate_timeline) { b->timeline_ms.push_back(t0 - t_begin); b->timeline_step.push_back(step); b->timeline_who.push_back(-1); }
        a.step = (uint32_t)step;
        if (pdnn_ps_master_step(&a, d->result, stream)) { failed.store(true); break; }
        FC_CUDA(cudaMemcpyAsync(res.data(), d->result, sizeof(uint32_t) * 80, cudaMemcpyDeviceToHost, stream));
        FC_CUDA(cudaStreamSynchronize(stream));
        if (res[3] == 0xDEADu) { fprintf(stderr, "fc_ps: master timed out at step %d\n", step); failed.store(true); break; }
        // acceptance log in the order the master summed: bucket by bucket (layer L-1 first), workers by rank
        for (int bkt = 0; bkt < L; ++bkt) {
            const int l = L - 1 - bkt;
            for (int w = 0; w < d->n_workers; ++w) {
                if (!((res[8 + bkt] >> w) & 1u)) continue;
                b->accepted++;
                if (b->accept_log != nullptr && b->accept_n < b->accept_cap) {
                    int* row = b->accept_log + 4 * b->accept_n++;
                    row[0] = step; row[1] = w; row[2] = l; row[3] = d->batch_start[(size_t)w * (iters + 2) + step];
                }
                if (b->cfg.generate_timeline) { b->timeline_ms.push_back(t0 - t_begin + res[40 + bkt] / 1000.0); b->timeline_step.push_back(step); b->timeline_who.push_back(w); }
            }
        }
        if (b->cfg.shortcircuit) b->abandoned_steps += __builtin_popcount(res[1]);       // workers told to drop the step
        b->dropped_stale += (long long)__builtin_popcount(res[1]) * L;                    // their layer gradients never count
        b->step_ms.push_back(now_ms() - t0);
    }
    cudaStreamDestroy(stream);
}

static void dev_evaluator_loop(Board* b, DevState* d, int device, const Dataset* data) {
    std::atomic<bool>& failed = b->failed;
    Net net(b->p, device, failed);
    net.init(&d->off, d->total);
    const int L = d->L, r_idx = d->n_workers, iters = b->cfg.n_train_iters;
    if (!failed.load()) {                      // dry run: see dev_worker_loop
        int c0 = 0;
        net.fill_batch(*data, &c0);
        for (int l = 0; l < L; ++l) net.forward_layer(l);
        net.output_delta();
        error_count_kernel<<<4, 128, 0, net.stream>>>(net.Z[L - 1], net.Y, b->p.batch, b->p.dims[L], net.wrong);
        cudaStreamSynchronize(net.stream);
    }
    cudaDeviceSynchronize();
    d->inited.fetch_add(1);
    while (d->inited.load() < d->n_recv + 1 && !failed.load()) std::this_thread::sleep_for(std::chrono::microseconds(100));
    if (failed.load()) return;
    const double t0 = now_ms();
    const std::string dir = b->cfg.out_dir[0] ? b->cfg.out_dir : "outfiles";
    mkdir(dir.c_str(), 0755);
    FILE* f = fopen((dir + "/time_loss_out_" + b->scheme).c_str(), "w");
    uint32_t seen = 0;
    while (!failed.load()) {
        uint32_t r = 0;
        cudaMemcpyAsync(&r, d->flags[r_idx], sizeof(r), cudaMemcpyDeviceToHost, net.stream);
        cudaStreamSynchronize(net.stream);
        if (r == seen || r == 0) { std::this_thread::sleep_for(std::chrono::microseconds(200)); continue; }
        // snapshot the landing buffer of step r; the master refills it two steps later, so re-read the flag (seqlock)
        cudaMemcpyAsync(net.Wflat, d->land[r & 1][r_idx], sizeof(float) * d->total, cudaMemcpyDeviceToDevice, net.stream);
        uint32_t r2 = 0;
        cudaMemcpyAsync(&r2, d->flags[r_idx], sizeof(r2), cudaMemcpyDeviceToHost, net.stream);
        cudaStreamSynchronize(net.stream);
        if (r2 != r) continue;                          // a newer step was released meanwhile: this parity may already be refilled
        seen = r;
        float loss_sum = 0.f; int wrong_sum = 0, n_seen = 0, cursor = 0, n_it = 0;
        for (int it = 0; it < b->cfg.eval_batches && cursor + b->p.batch <= data->n; ++it, ++n_it) {
            net.fill_batch(*data, &cursor);
            for (int l = 0; l < L; ++l) net.forward_layer(l);
            net.output_delta();
            cudaMemsetAsync(net.wrong, 0, sizeof(int), net.stream);
            error_count_kernel<<<4, 128, 0, net.stream>>>(net.Z[L - 1], net.Y, b->p.batch, b->p.dims[L], net.wrong);
            float hl; int hw;
            cudaMemcpyAsync(&hl, net.loss, sizeof(float), cudaMemcpyDeviceToHost, net.stream);
            cudaMemcpyAsync(&hw, net.wrong, sizeof(int), cudaMemcpyDeviceToHost, net.stream);
            cudaStreamSynchronize(net.stream);
            loss_sum += hl; wrong_sum += hw; n_seen += b->p.batch;
        }
        const float loss = loss_sum / std::max(1, n_it), err = (float)wrong_sum / std::max(1, n_seen);
        {
            std::lock_guard<std::mutex> lk(b->mu);
            b->eval_loss.push_back(loss);
            b->eval_err.push_back(err);
        }
        if (f) { fprintf(f, "%u %.3f %.6f %.6f\n", r, now_ms() - t0, loss, err); fflush(f); }
        if ((int)r > iters) break;                      // that was the final model
    }
    if (f) fclose(f);
    net.destroy();
}

// Set up the device-protocol state from the host-protocol board (whose Wm[1][l] hold the initial weights), run the
// three roles, leave the final weights in Wm[(iters + 1) & 1][l] where the caller reads them.
static int run_device_protocol(Board* b, const Dataset* train, const Dataset* test, int n_dev) {
    DevState st;
    DevState* d = &st;
    const int L = b->p.n_layers(), nw = b->n_workers, iters = b->cfg.n_train_iters;
    d->L = L; d->n_workers = nw; d->n_recv = nw + (b->cfg.with_evaluator ? 1 : 0);
    d->off.resize(L); d->len.resize(L);
    long long off = 0;
    for (int l = 0; l < L; ++l) {
        d->off[l] = off;
        d->len[l] = (b->p.layer_elems(l) + 3) / 4 * 4;
        off += (d->len[l] + 63) / 64 * 64;
    }
    d->total = off;
    d->batch_start.assign((size_t)nw * (iters + 2), 0);
    auto dev_of_worker = [&](int w) { return n_dev > 1 ? 1 + w % (n_dev - 1) : 0; };
    for (int w = 0; w < nw; ++w) d->recv_device.push_back(dev_of_worker(w));
    if (b->cfg.with_evaluator) d->recv_device.push_back(n_dev > 1 ? n_dev - 1 : 0);
    PDNN_CUDA_CHECK(cudaSetDevice(d->master_device));
    PDNN_CUDA_CHECK(cudaMalloc(&d->w_master, sizeof(float) * d->total));
    PDNN_CUDA_CHECK(cudaMalloc(&d->mom, sizeof(float) * d->total));
    PDNN_CUDA_CHECK(cudaMemset(d->w_master, 0, sizeof(float) * d->total));
    PDNN_CUDA_CHECK(cudaMemset(d->mom, 0, sizeof(float) * d->total));
    const size_t ctl_words = 256 + 80 + (size_t)nw * d->stride;
    PDNN_CUDA_CHECK(cudaMalloc(&d->ctl, sizeof(uint32_t) * ctl_words));
    PDNN_CUDA_CHECK(cudaMemset(d->ctl, 0, sizeof(uint32_t) * ctl_words));
    d->result = d->ctl + 256;
    d->arrival = d->ctl + 256 + 80;
    d->slot.assign(nw, nullptr);
    for (int w = 0; w < nw; ++w) {
        PDNN_CUDA_CHECK(cudaMalloc(&d->slot[w], sizeof(float) * d->total));
        PDNN_CUDA_CHECK(cudaMemset(d->slot[w], 0, sizeof(float) * d->total));
    }
    for (int l = 0; l < L; ++l)
        PDNN_CUDA_CHECK(cudaMemcpy(d->w_master + d->off[l], b->Wm[1][l], sizeof(float) * b->p.layer_elems(l), cudaMemcpyDeviceToDevice));
    for (int par = 0; par < 2; ++par) d->land[par].assign(d->n_recv, nullptr);
    d->flags.assign(d->n_recv, nullptr);
    for (int r = 0; r < d->n_recv; ++r) {
        PDNN_CUDA_CHECK(cudaSetDevice(d->recv_device[r]));
        for (int par = 0; par < 2; ++par) {
            PDNN_CUDA_CHECK(cudaMalloc(&d->land[par][r], sizeof(float) * d->total));
            PDNN_CUDA_CHECK(cudaMemset(d->land[par][r], 0, sizeof(float) * d->total));
        }
        PDNN_CUDA_CHECK(cudaMalloc(&d->flags[r], 256));
        PDNN_CUDA_CHECK(cudaMemset(d->flags[r], 0, 256));
    }
    // batches are copied from these host arrays every step: page-lock them so that the copies are true async DMA
    cudaHostRegister(const_cast<float*>(train->images.data()), train->images.size() * sizeof(float), cudaHostRegisterDefault);
    cudaHostRegister(const_cast<long long*>(train->labels.data()), train->labels.size() * sizeof(long long), cudaHostRegisterDefault);
    if (test->images.data() != train->images.data()) {
        cudaHostRegister(const_cast<float*>(test->images.data()), test->images.size() * sizeof(float), cudaHostRegisterDefault);
        cudaHostRegister(const_cast<long long*>(test->labels.data()), test->labels.size() * sizeof(long long), cudaHostRegisterDefault);
    }
    cudaGetLastError();
    // first-use allocations of the library (the push kernel's per-device counter block) happen now, not while kernels that
    // spin on flags are resident
    for (int r = 0; r < nw; ++r) {
        PDNN_CUDA_CHECK(cudaSetDevice(d->recv_device[r]));
        if (int rc = pdnn_ps_push(d->land[0][r], d->land[0][r], 4, 0, nullptr, 0, nullptr, nullptr, 1, nullptr)) return rc;
    }
    for (int i = 0; i < n_dev; ++i) { cudaSetDevice(i); cudaDeviceSynchronize(); }
    std::vector<std::thread> threads;
    for (int w = 0; w < nw; ++w) threads.emplace_back(dev_worker_loop, b, d, w, dev_of_worker(w), train);
    if (b->cfg.with_evaluator) threads.emplace_back(dev_evaluator_loop, b, d, d->recv_device[nw], test);
    std::thread master(dev_master_loop, b, d);
    master.join();
    if (b->failed.load()) {           // unblock whoever still waits on a flag: publish an impossible-to-miss ready value
        const uint32_t big = 0x7fffffffu;
        for (int r = 0; r < d->n_recv; ++r) { cudaSetDevice(d->recv_device[r]); cudaMemcpy(d->flags[r], &big, sizeof(big), cudaMemcpyHostToDevice); }
    }
    for (auto& t : threads) t.join();
    cudaHostUnregister(const_cast<float*>(train->images.data()));
    cudaHostUnregister(const_cast<long long*>(train->labels.data()));
    if (test->images.data() != train->images.data()) {
        cudaHostUnregister(const_cast<float*>(test->images.data()));
        cudaHostUnregister(const_cast<long long*>(test->labels.data()));
    }
    cudaGetLastError();
    PDNN_CUDA_CHECK(cudaSetDevice(d->master_device));
    for (int l = 0; l < L; ++l)
        PDNN_CUDA_CHECK(cudaMemcpy(b->Wm[(iters + 1) & 1][l], d->w_master + d->off[l], sizeof(float) * b->p.layer_elems(l), cudaMemcpyDeviceToDevice));
    cudaFree(d->w_master); cudaFree(d->mom); cudaFree(d->ctl);
    for (int w = 0; w < nw; ++w) cudaFree(d->slot[w]);
    for (int r = 0; r < d->n_recv; ++r) {
        cudaSetDevice(d->recv_device[r]);
        cudaFree(d->land[0][r]); cudaFree(d->land[1][r]); cudaFree(d->flags[r]);
    }
    cudaSetDevice(d->master_device);
    return 0;
}

}  // namespace fcps

extern "C" int pdnn_fc_ps_run(const pdnn_fc_ps_config_t* cfg, pdnn_fc_ps_result_t* res) {
    return pdnn_fc_ps_run_ex(cfg, res, nullptr);
}

extern "C" int pdnn_fc_ps_run_ex(const pdnn_fc_ps_config_t* cfg, pdnn_fc_ps_result_t* res, pdnn_fc_ps_io_t* io) {
    using namespace fcps;
    PDNN_REQUIRE(cfg != nullptr && res != nullptr, "fc_ps: null arguments");
    Board board;
    Board* b = &board;
    b->cfg = *cfg;
    for (int i = 0; i < cfg->n_dims; ++i) b->p.dims.push_back(cfg->dims[i]);
    b->p.batch = cfg->batch;
    b->p.lr = cfg->lr;
    std::string why;
    PDNN_REQUIRE(b->p.validate(&why), "fc_ps: invalid NNParams: %s", why.c_str());
    PDNN_REQUIRE(cfg->n_workers >= 1 && cfg->n_workers <= 32, "fc_ps: n_workers out of range");
    PDNN_REQUIRE(cfg->n_to_collect >= 1 && cfg->n_to_collect <= cfg->n_workers, "fc_ps: n_to_collect must be in [1, n_workers]");
    b->n_workers = cfg->n_workers;
    int n_dev = 0;
    PDNN_CUDA_CHECK(cudaGetDeviceCount(&n_dev));
    PDNN_REQUIRE(n_dev >= 1, "fc_ps: no CUDA device");
    if (cfg->max_devices > 0 && n_dev > cfg->max_devices) n_dev = cfg->max_devices;
    char scheme[128];
    snprintf(scheme, sizeof(scheme), "SyncReplicasWithBackup%d_%d%s", cfg->n_to_collect, cfg->n_workers, cfg->shortcircuit ? "_shortcircuit" : "");
    b->scheme = scheme;      // reference scheme_full_name(), distributed_defines.h:30-41
    // peer access between the GPUs in use
    for (int i = 0; i < n_dev; ++i)
        for (int j = 0; j < n_dev; ++j)
            if (i != j) { cudaSetDevice(i); int can = 0; cudaDeviceCanAccessPeer(&can, i, j); if (can) { cudaDeviceEnablePeerAccess(j, 0); cudaGetLastError(); } }
    const int L = b->p.n_layers();
    PDNN_CUDA_CHECK(cudaSetDevice(0));
    std::mt19937 rng(cfg->seed);
    b->slots.assign(cfg->n_workers, std::vector<float*>(L, nullptr));
    for (int par = 0; par < 2; ++par) b->Wm[par].assign(L, nullptr);
    b->acc.assign(L, nullptr);
    b->We.assign(L, nullptr);
    if (io != nullptr && io->accept_log != nullptr) { b->accept_log = io->accept_log; b->accept_cap = io->accept_log_rows; }
    long long w_off = 0;
    for (int l = 0; l < L; ++l) {
        const long long n = b->p.layer_elems(l);
        for (int par = 0; par < 2; ++par) PDNN_CUDA_CHECK(cudaMalloc(&b->Wm[par][l], sizeof(float) * n));
        PDNN_CUDA_CHECK(cudaMalloc(&b->acc[l], sizeof(float) * n));
        PDNN_CUDA_CHECK(cudaMalloc(&b->We[l], sizeof(float) * n));
        for (int w = 0; w < cfg->n_workers; ++w) PDNN_CUDA_CHECK(cudaMalloc(&b->slots[w][l], sizeof(float) * n));
        // W ~ N(0, 0.1^2), bias row = 0 (pure_py_code/nn/nn_layer.py:58-110)
        std::vector<float> h(n, 0.f);
        std::normal_distribution<float> g(0.f, 0.1f);
        for (long long i = 0; i < (long long)b->p.dims[l] * b->p.dims[l + 1]; ++i) h[i] = g(rng);
        const float* src = (io != nullptr && io->init_weights != nullptr) ? io->init_weights + w_off : h.data();
        PDNN_CUDA_CHECK(cudaMemcpy(b->Wm[1][l], src, sizeof(float) * n, cudaMemcpyHostToDevice));   // step 1 reads parity 1
        w_off += n;
    }
    Dataset train, test;
    if (io != nullptr && io->train_images != nullptr) {
        PDNN_REQUIRE(io->train_labels != nullptr && io->n_train > 0, "fc_ps: caller-provided training set needs labels and a size");
        train.n = io->n_train; train.dim = b->p.dims[0];
        train.images.assign(io->train_images, io->train_images + (size_t)io->n_train * train.dim);
        train.labels.assign(io->train_labels, io->train_labels + io->n_train);
        if (io->test_images != nullptr && io->n_test > 0) {
            test.n = io->n_test; test.dim = train.dim;
            test.images.assign(io->test_images, io->test_images + (size_t)io->n_test * test.dim);
            test.labels.assign(io->test_labels, io->test_labels + io->n_test);
        } else { test = train; }
    } else {
        const bool real = load_mnist(cfg->data_dir, &train, cfg->synthetic_examples > 0 ? cfg->synthetic_examples : 4096, b->p.dims[0], b->p.dims.back(), cfg->seed);
        if (real) { test = train; test.n = std::min(train.n, 2048); }
        else synthetic(&test, 1024, b->p.dims[0], b->p.dims.back(), cfg->seed + 99);
    }
    PDNN_REQUIRE(train.dim == b->p.dims[0], "fc_ps: dataset dimension %d does not match the input layer %d", train.dim, b->p.dims[0]);
    PDNN_REQUIRE(train.n >= b->p.batch, "fc_ps: dataset smaller than one batch");

    const double t0 = now_ms();
    if (cfg->device_protocol && L <= PDNN_MAX_BUCKETS && cfg->n_workers <= PDNN_MAX_WORKERS) {
        if (int rc = run_device_protocol(b, &train, &test, n_dev)) return rc;
    } else {
        std::vector<std::thread> threads;
        for (int w = 0; w < cfg->n_workers; ++w) threads.emplace_back(worker_loop, b, w, n_dev > 1 ? 1 + w % (n_dev - 1) : 0, &train);
        if (cfg->with_evaluator) threads.emplace_back(evaluator_loop, b, n_dev > 1 ? n_dev - 1 : 0, &test);
        std::thread master(master_loop, b);
        master.join();
        for (auto& t : threads) t.join();
    }
    const double total = now_ms() - t0;

    memset(res, 0, sizeof(*res));
    res->steps = (int)b->step_ms.size();
    res->total_ms = total;
    double sum = 0;
    for (double v : b->step_ms) sum += v;
    res->mean_step_ms = b->step_ms.empty() ? 0 : sum / b->step_ms.size();
    res->gradients_accepted = b->accepted;
    res->gradients_dropped_stale = b->dropped_stale;
    res->steps_abandoned = b->abandoned_steps;
    res->avg_gradients_per_step = b->step_ms.empty() ? 0 : (double)b->accepted / (double)b->step_ms.size() / L;
    if (!b->eval_loss.empty()) {
        res->first_loss = b->eval_loss.front(); res->final_loss = b->eval_loss.back();
        res->first_err = b->eval_err.front(); res->final_err = b->eval_err.back();
    }
    snprintf(res->scheme, sizeof(res->scheme), "%s", scheme);
    if (cfg->generate_timeline) {
        std::string path = std::string(cfg->out_dir[0] ? cfg->out_dir : "outfiles") + "/timeline_out_" + scheme;
        mkdir(cfg->out_dir[0] ? cfg->out_dir : "outfiles", 0755);
        if (FILE* f = fopen(path.c_str(), "w")) {
            for (size_t i = 0; i < b->timeline_ms.size(); ++i) fprintf(f, "%.3f %d %d\n", b->timeline_ms[i], b->timeline_step[i], b->timeline_who[i] < 0 ? 1 : 0);
            fclose(f);
        }
    }
    cudaSetDevice(0);
    if (io != nullptr) {
        io->accept_log_used = b->accept_n;
        if (io->final_weights != nullptr && !b->failed.load()) {      // weights of step n_train_iters + 1
            long long off = 0;
            for (int l = 0; l < L; ++l) {
                const long long n = b->p.layer_elems(l);
                PDNN_CUDA_CHECK(cudaMemcpy(io->final_weights + off, b->Wm[(cfg->n_train_iters + 1) & 1][l], sizeof(float) * n, cudaMemcpyDeviceToHost));
                off += n;
            }
        }
    }
    for (int l = 0; l < L; ++l) {
        for (int par = 0; par < 2; ++par) cudaFree(b->Wm[par][l]);
        cudaFree(b->acc[l]);
        cudaFree(b->We[l]);
        for (int w = 0; w < cfg->n_workers; ++w) cudaFree(b->slots[w][l]);
    }
    PDNN_REQUIRE(!b->failed.load(), "fc_ps: a role failed (see stderr)");
    return 0;
}

// Loader / network self-tests of the reference (mnist.h:155-162 test_load_data, nn.h:158-207 test_nn):
// returns 0 when a 784-100-10 net trained for `iters` single-worker steps reduces its loss.
extern "C" int pdnn_fc_ps_selftest(int iters, float* first_loss, float* last_loss) {
    pdnn_fc_ps_config_t c;
    memset(&c, 0, sizeof(c));
    c.n_dims = 3; c.dims[0] = 784; c.dims[1] = 100; c.dims[2] = 10;
    c.batch = 128; c.lr = 0.5f; c.n_workers = 1; c.n_to_collect = 1; c.n_train_iters = iters; c.with_evaluator = 1; c.eval_batches = 2;
    c.seed = 7; c.synthetic_examples = 2048; c.max_devices = 1; c.device_protocol = 1;
    snprintf(c.out_dir, sizeof(c.out_dir), "/tmp/pdnn_fc_selftest");
    pdnn_fc_ps_result_t r;
    if (int rc = pdnn_fc_ps_run(&c, &r)) return rc;
    if (first_loss) *first_loss = r.first_loss;
    if (last_loss) *last_loss = r.final_loss;
    PDNN_REQUIRE(r.final_loss < r.first_loss, "fc_ps selftest: loss did not decrease (%f -> %f)", r.first_loss, r.final_loss);
    return 0;
}
