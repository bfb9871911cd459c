// Parameter-server data path over NVLink peer memory for sm_100a.
//
// The reference moves every gradient / weight tensor through mpi4py (Isend/Irecv/Bcast of float64
// numpy buffers, pytorch_code/sync_replicas_master_nn.py:235-314, distributed_worker.py:202-264).
// Here all movement happens inside kernels: workers store their gradient buckets straight into the
// master's slots and publish a step-tagged arrival flag; one master kernel selects the first k
// arrivals, signals the stragglers, reduces, applies momentum SGD and stores the new weights into
// every worker's buffer (peer stores or one NVLS multicast store), then releases the workers.
// No host round trip sits between "k-th gradient landed" and "workers see the new weights".
#include "common.cuh"
#include "../include/pdnn_b200.h"
#include <string.h>

namespace pdnn {

// control words of the master (zero-initialised by the caller once)
enum { CTL_MASK = 0, CTL_MASK_STEP = 1, CTL_DONE = 2, CTL_NSEL = 3, CTL_ERR = 4, CTL_PUSH_DONE = 8 };

#ifndef PDNN_PS_HARD_TIMEOUT_NS
#define PDNN_PS_HARD_TIMEOUT_NS 30000000000ull   // a master that never sees k gradients gives up after 30 s
#endif

// ------------------------------------------------------------------------------------------------
// worker -> master push
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) ps_push_kernel(const float* __restrict__ grad, void* __restrict__ slot, long long n,
                                                      int slot_is_bf16, uint32_t* flag_peer, uint32_t step,
                                                      uint32_t* local_counter, const uint32_t* guard_closed,
                                                      const uint32_t* guard_step, const uint32_t* step_ptr) {
    if (step_ptr != nullptr) step = ld_relaxed_sys(step_ptr);   // CUDA-graph mode: the step number lives on the device
    __shared__ int s_abort;
    if (threadIdx.x == 0) s_abort = guard_closed != nullptr && ld_relaxed_sys(guard_closed) >= ld_relaxed_sys(guard_step);
    __syncthreads();
    // a killed worker drops its contribution (CTA-uniform decision). The decision can differ between the CTAs of one
    // grid (the master's `closed` store may land while the grid is running), so an aborting CTA still takes part in the
    // completion handshake below and only vetoes the flag: the counter always returns to zero.
    const bool abort = s_abort != 0;
    if (abort && flag_peer == nullptr) return;
    if (!abort) {
        const long long n8 = n / 8;
        const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const long long nth = (long long)gridDim.x * blockDim.x;
        if (slot_is_bf16) {
            __nv_bfloat16* d = reinterpret_cast<__nv_bfloat16*>(slot);
            for (long long i = tid; i < n8; i += nth) {
                const float4 a = *reinterpret_cast<const float4*>(grad + i * 8);
                const float4 b = *reinterpret_cast<const float4*>(grad + i * 8 + 4);
                uint4 u;
                u.x = pack_bf16x2(a.x, a.y); u.y = pack_bf16x2(a.z, a.w);
                u.z = pack_bf16x2(b.x, b.y); u.w = pack_bf16x2(b.z, b.w);
                st_na_v4(d + i * 8, u);
            }
            if (tid == 0) for (long long j = n8 * 8; j < n; ++j) d[j] = __float2bfloat16(grad[j]);
        } else {
            float* d = reinterpret_cast<float*>(slot);
            for (long long i = tid; i < n8 * 2; i += nth) st_na_v4(d + i * 4, *reinterpret_cast<const uint4*>(grad + i * 4));
            if (tid == 0) for (long long j = n8 * 8; j < n; ++j) d[j] = grad[j];
        }
        if (flag_peer == nullptr) return;
    }
    // grid completion -> one release store of the step-tagged arrival flag (unless any CTA of the grid aborted)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        if (abort) atomicOr(local_counter + 1, 1u);
        const uint32_t done = atom_add_acqrel_gpu(local_counter, 1u) + 1u;
        if (done == gridDim.x) {
            const uint32_t vetoed = atomicExch(local_counter + 1, 0u);
            st_relaxed_sys(local_counter, 0u);
            __threadfence_system();
            if (!vetoed) st_release_sys(flag_peer, step);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// master: select first k, kill the rest, reduce + momentum SGD + broadcast
// ------------------------------------------------------------------------------------------------
struct MasterArgs {
    pdnn_ps_master_t a;
    uint32_t* result;
};

__device__ __forceinline__ void multimem_st_v4(void* mc_addr, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

__global__ void __launch_bounds__(512) ps_master_step_kernel(const __grid_constant__ MasterArgs margs) {
    const pdnn_ps_master_t& a = margs.a;
    // CUDA-graph mode: the step to close is read from device memory (and advanced by the last block below)
    const uint32_t cur_step = a.step_ptr != nullptr ? ld_relaxed_sys(a.step_ptr) : a.step;
    __shared__ uint32_t s_mask;
    __shared__ uint32_t s_nsel;
    uint32_t* ctl = a.control;

    // -------- phase A: decide the contributing set (one thread of block 0), everyone else waits locally
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) {
            const uint64_t t0 = globaltimer_ns();
            uint32_t mask = 0, count = 0, err = 0;
            const uint32_t want = (uint32_t)a.k;
            for (;;) {
                for (int w = 0; w < a.n_workers && count < want; ++w) {
                    if (!(mask & (1u << w)) && ld_acquire_sys(a.arrival_flags + w) == cur_step) {
                        mask |= 1u << w;
                        ++count;
                    }
                }
                if (count >= want) break;
                const uint64_t el = globaltimer_ns() - t0;
                if (a.deadline_ns > 0 && el > (uint64_t)a.deadline_ns && count > 0) break;   // interval policy
                if (el > PDNN_PS_HARD_TIMEOUT_NS) { err = 1; break; }
                __nanosleep(200);
            }
            // straggler kill: tell every worker outside the set that this step is closed
            uint32_t kill = 0;
            for (int w = 0; w < a.n_workers; ++w) {
                if (!(mask & (1u << w))) {
                    kill |= 1u << w;
                    if (a.worker_closed[w] != nullptr) st_release_sys(a.worker_closed[w], cur_step);
                }
            }
            ctl[CTL_NSEL] = count;
            ctl[CTL_ERR] = err;
            margs.result[0] = mask;
            margs.result[1] = kill;
            margs.result[2] = count;
            margs.result[3] = err ? 0xDEADu : cur_step;
            // device-side timeline (globaltimer, microseconds): how long the master waited for the k-th arrival
            const uint64_t t1 = globaltimer_ns();
            margs.result[4] = (uint32_t)((t1 - t0) / 1000ull);
            ctl[5] = (uint32_t)(t1 / 1000ull);
            ctl[CTL_MASK] = mask;
            __threadfence();
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(ctl + CTL_MASK_STEP), "r"(cur_step) : "memory");
            s_mask = mask;
            s_nsel = count;
        } else {
            while (ld_acquire_gpu(ctl + CTL_MASK_STEP) != cur_step) __nanosleep(100);
            s_mask = ctl[CTL_MASK];
            s_nsel = ctl[CTL_NSEL];
        }
    }
    __syncthreads();
    const uint32_t mask = s_mask;
    const uint32_t nsel = s_nsel;

    // -------- phase B: reduce selected slots in rank order, 1/k, momentum SGD, cast, broadcast
    if (nsel > 0) {
        float scale;
        if (a.avg_mode == 1) scale = 1.f / (float)a.n_workers;
        else scale = 1.f / (float)nsel;
        const long long n4 = a.n / 4;
        const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const long long nth = (long long)gridDim.x * blockDim.x;
        for (long long i = tid; i < n4; i += nth) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a.multicast_slots != nullptr && nsel == (uint32_t)a.n_workers) {
                // all n sets contribute: one load-reduce through the NVLS multicast address sums the workers' local
                // slots inside the switch (1x ingress instead of n x)
                if (a.slot_is_bf16) {
                    uint32_t r0, r1;
                    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.bf16x2 {%0, %1}, [%2];"
                                 : "=r"(r0), "=r"(r1)
                                 : "l"(reinterpret_cast<const __nv_bfloat16*>(a.multicast_slots) + i * 4)
                                 : "memory");
                    const float2 lo = unpack_bf16x2(r0), hi = unpack_bf16x2(r1);
                    acc = make_float4(lo.x, lo.y, hi.x, hi.y);
                } else {
                    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(acc.x), "=f"(acc.y), "=f"(acc.z), "=f"(acc.w)
                                 : "l"(reinterpret_cast<const float*>(a.multicast_slots) + i * 4)
                                 : "memory");
                }
                if (a.poison_slots) {      // one multicast store poisons this element in every worker's local slot
                    if (a.slot_is_bf16)
                        asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(
                                         const_cast<__nv_bfloat16*>(reinterpret_cast<const __nv_bfloat16*>(a.multicast_slots)) + i * 4),
                                     "r"(0x7FC07FC0u), "r"(0x7FC07FC0u) : "memory");
                    else
                        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(
                                         const_cast<float*>(reinterpret_cast<const float*>(a.multicast_slots)) + i * 4),
                                     "r"(0x7FC00000u), "r"(0x7FC00000u), "r"(0x7FC00000u), "r"(0x7FC00000u) : "memory");
                }
            } else
            for (int w = 0; w < a.n_workers; ++w) {
                if (!(mask & (1u << w))) continue;
                if (a.slot_is_bf16) {
                    const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(a.slots[w]) + i * 4);
                    const float2 lo = unpack_bf16x2(u.x), hi = unpack_bf16x2(u.y);
                    acc.x += lo.x; acc.y += lo.y; acc.z += hi.x; acc.w += hi.y;
                    if (a.poison_slots)
                        *reinterpret_cast<uint2*>(const_cast<__nv_bfloat16*>(reinterpret_cast<const __nv_bfloat16*>(a.slots[w])) + i * 4) =
                            make_uint2(0x7FC07FC0u, 0x7FC07FC0u);
                } else {
                    const float4 g = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.slots[w]) + i * 4);
                    acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
                    if (a.poison_slots)
                        *reinterpret_cast<uint4*>(const_cast<float*>(reinterpret_cast<const float*>(a.slots[w])) + i * 4) =
                            make_uint4(0x7FC00000u, 0x7FC00000u, 0x7FC00000u, 0x7FC00000u);
                }
            }
            float4 m = reinterpret_cast<float4*>(a.momentum)[i];
            float4 wv = reinterpret_cast<float4*>(a.master_w)[i];
            m.x = a.mu * m.x + scale * acc.x; m.y = a.mu * m.y + scale * acc.y;
            m.z = a.mu * m.z + scale * acc.z; m.w = a.mu * m.w + scale * acc.w;
            wv.x -= a.lr * m.x; wv.y -= a.lr * m.y; wv.z -= a.lr * m.z; wv.w -= a.lr * m.w;
            reinterpret_cast<float4*>(a.momentum)[i] = m;
            reinterpret_cast<float4*>(a.master_w)[i] = wv;
            if (a.weights_are_bf16) {
                uint2 u;
                u.x = pack_bf16x2(wv.x, wv.y); u.y = pack_bf16x2(wv.z, wv.w);
                if (a.multicast_weights != nullptr) {
                    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(
                                     reinterpret_cast<__nv_bfloat16*>(a.multicast_weights) + i * 4),
                                 "r"(u.x), "r"(u.y)
                                 : "memory");
                } else {
                    for (int w = 0; w < a.n_workers; ++w)
                        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(a.worker_weights[w]) + i * 4) = u;
                }
            } else {
                uint4 u = *reinterpret_cast<uint4*>(&wv);
                if (a.multicast_weights != nullptr) {
                    multimem_st_v4(reinterpret_cast<float*>(a.multicast_weights) + i * 4, u);
                } else {
                    for (int w = 0; w < a.n_workers; ++w) st_na_v4(reinterpret_cast<float*>(a.worker_weights[w]) + i * 4, u);
                }
            }
        }
        if (tid == 0) {
            for (long long j = n4 * 4; j < a.n; ++j) {
                float acc = 0.f;
                for (int w = 0; w < a.n_workers; ++w) {
                    if (!(mask & (1u << w))) continue;
                    acc += a.slot_is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(a.slots[w])[j])
                                          : reinterpret_cast<const float*>(a.slots[w])[j];
                }
                float m = a.mu * a.momentum[j] + scale * acc;
                a.momentum[j] = m;
                const float wv = a.master_w[j] - a.lr * m;
                a.master_w[j] = wv;
                for (int w = 0; w < a.n_workers; ++w) {
                    if (a.weights_are_bf16) reinterpret_cast<__nv_bfloat16*>(a.worker_weights[w])[j] = __float2bfloat16(wv);
                    else reinterpret_cast<float*>(a.worker_weights[w])[j] = wv;
                }
            }
        }
    }

    // -------- phase C: grid completion, then release every worker with step+1
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const uint32_t done = atom_add_acqrel_gpu(ctl + CTL_DONE, 1u) + 1u;
        if (done == gridDim.x) {
            ctl[CTL_DONE] = 0;
            margs.result[5] = (uint32_t)(globaltimer_ns() / 1000ull) - ctl[5];   // reduce + update + broadcast, microseconds
            if (a.step_ptr != nullptr) st_relaxed_sys(a.step_ptr, cur_step + 1u);
            __threadfence_system();
            for (int w = 0; w < a.n_workers; ++w)
                if (a.worker_ready[w] != nullptr) st_release_sys(a.worker_ready[w], cur_step + 1u);
        }
    }
}

__global__ void ps_wait_flag_kernel(const uint32_t* flag, uint32_t value, uint32_t* cur_step_word, long long timeout_ns,
                                    uint32_t* status) {
    const uint64_t t0 = globaltimer_ns();
    uint32_t ok = 1;
    while (ld_acquire_sys(flag) < value) {
        if (timeout_ns > 0 && globaltimer_ns() - t0 > (uint64_t)timeout_ns) { ok = 0; break; }
        __nanosleep(100);
    }
    if (cur_step_word != nullptr) st_release_sys(cur_step_word, ok ? value : 0xFFFFFFFFu);
    if (status != nullptr) *status = ok;
}

// CUDA-graph mode: advance the device-side step word and wait for that step's weights.
__global__ void ps_wait_next_kernel(const uint32_t* flag, uint32_t* step_word, long long timeout_ns) {
    const uint32_t want = ld_relaxed_sys(step_word) + 1u;
    const uint64_t t0 = globaltimer_ns();
    while (ld_acquire_sys(flag) < want) {
        if (timeout_ns > 0 && globaltimer_ns() - t0 > (uint64_t)timeout_ns) break;
        __nanosleep(100);
    }
    st_release_sys(step_word, want);
}

__global__ void ps_delay_kernel(long long ns, const uint32_t* guard_closed, const uint32_t* guard_step) {
    const uint64_t t0 = globaltimer_ns();
    while (globaltimer_ns() - t0 < (uint64_t)ns) {
        if (guard_closed != nullptr && ld_relaxed_sys(guard_closed) >= ld_relaxed_sys(guard_step)) break;
        __nanosleep(500);
    }
}

__global__ void ps_set_word_kernel(uint32_t* p, uint32_t v) {
    __threadfence_system();
    st_release_sys(p, v);
}

__global__ void __launch_bounds__(512) p2p_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, long long n16) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    long long i = tid;
    for (; i + 3 * nth < n16; i += 4 * nth) {
        const uint4 a = ld_nc_v4(src + i), b = ld_nc_v4(src + i + nth), c = ld_nc_v4(src + i + 2 * nth), d = ld_nc_v4(src + i + 3 * nth);
        st_na_v4(dst + i, a); st_na_v4(dst + i + nth, b); st_na_v4(dst + i + 2 * nth, c); st_na_v4(dst + i + 3 * nth, d);
    }
    for (; i < n16; i += nth) st_na_v4(dst + i, ld_nc_v4(src + i));
}

}  // namespace pdnn

using namespace pdnn;
#define ST(s) reinterpret_cast<cudaStream_t>(s)

static int ps_grid(long long n_items_per_thread_unit, int threads, int max_ctas) {
    long long b = (n_items_per_thread_unit + threads - 1) / threads;
    int cap = max_ctas > 0 ? max_ctas : 148;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

// The push kernel needs one worker-local counter word for its grid-completion handshake; it lives in
// a small per-device pool so that callers do not have to thread scratch through every push.
static uint32_t* g_push_counter[64] = {nullptr};
static int push_counter(uint32_t** out) {
    int dev = 0;
    PDNN_CUDA_CHECK(cudaGetDevice(&dev));
    if (g_push_counter[dev] == nullptr) {
        PDNN_CUDA_CHECK(cudaMalloc(&g_push_counter[dev], 256));
        PDNN_CUDA_CHECK(cudaMemset(g_push_counter[dev], 0, 256));
    }
    *out = g_push_counter[dev];
    return 0;
}

extern "C" int pdnn_ps_push(const float* grad, void* slot_peer, long long n, int slot_is_bf16, uint32_t* flag_peer,
                            uint32_t step, const uint32_t* guard_closed, const uint32_t* guard_step, int max_ctas,
                            void* stream) {
    PDNN_REQUIRE(n > 0, "ps_push: empty bucket");
    PDNN_REQUIRE((reinterpret_cast<uintptr_t>(grad) & 15) == 0 && (reinterpret_cast<uintptr_t>(slot_peer) & 15) == 0,
                 "ps_push: buffers must be 16-byte aligned");
    uint32_t* counter = nullptr;
    if (int rc = push_counter(&counter)) return rc;
    const int grid = ps_grid(n / 8 + 1, 512, max_ctas);
    ps_push_kernel<<<grid, 512, 0, ST(stream)>>>(grad, slot_peer, n, slot_is_bf16, flag_peer, step, counter, guard_closed, guard_step, nullptr);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// Same push, but the step tag is read from device memory when the kernel runs: nothing in the launch depends on the
// step number, so the launch can be replayed from a CUDA graph.
extern "C" int pdnn_ps_push_dev(const float* grad, void* slot_peer, long long n, int slot_is_bf16, uint32_t* flag_peer,
                                const uint32_t* step_ptr, const uint32_t* guard_closed, const uint32_t* guard_step, int max_ctas,
                                void* stream) {
    PDNN_REQUIRE(n > 0 && step_ptr != nullptr, "ps_push_dev: empty bucket or missing step word");
    uint32_t* counter = nullptr;
    if (int rc = push_counter(&counter)) return rc;
    const int grid = ps_grid(n / 8 + 1, 512, max_ctas);
    ps_push_kernel<<<grid, 512, 0, ST(stream)>>>(grad, slot_peer, n, slot_is_bf16, flag_peer, 0u, counter, guard_closed, guard_step, step_ptr);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_ps_master_step(const pdnn_ps_master_t* args, uint32_t* result, void* stream) {
    PDNN_REQUIRE(args != nullptr && result != nullptr, "ps_master_step: null arguments");
    PDNN_REQUIRE(args->n_workers >= 1 && args->n_workers <= PDNN_MAX_WORKERS, "ps_master_step: n_workers=%d out of range", args->n_workers);
    PDNN_REQUIRE(args->k >= 1 && args->k <= args->n_workers, "ps_master_step: k=%d must be in [1, %d]", args->k, args->n_workers);
    PDNN_REQUIRE(args->control != nullptr && args->master_w != nullptr && args->momentum != nullptr, "ps_master_step: missing buffers");
    PDNN_REQUIRE(args->n % 4 == 0 || (args->multicast_weights == nullptr && args->multicast_slots == nullptr),
                 "ps_master_step: multicast paths need n %% 4 == 0");
    MasterArgs m;
    m.a = *args;
    m.result = result;
    const int grid = ps_grid(args->n / 4 + 1, 512, args->max_ctas);
    ps_master_step_kernel<<<grid, 512, 0, ST(stream)>>>(m);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_ps_wait_flag(const uint32_t* flag, uint32_t value, uint32_t* cur_step_word, long long timeout_ns, void* stream) {
    ps_wait_flag_kernel<<<1, 1, 0, ST(stream)>>>(flag, value, cur_step_word, timeout_ns, nullptr);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_ps_wait_next(const uint32_t* flag, uint32_t* step_word, long long timeout_ns, void* stream) {
    ps_wait_next_kernel<<<1, 1, 0, ST(stream)>>>(flag, step_word, timeout_ns);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_delay(long long ns, const uint32_t* guard_closed, const uint32_t* guard_step, void* stream) {
    if (ns <= 0) return 0;
    ps_delay_kernel<<<1, 1, 0, ST(stream)>>>(ns, guard_closed, guard_step);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_set_word(uint32_t* p, uint32_t v, void* stream) {
    ps_set_word_kernel<<<1, 1, 0, ST(stream)>>>(p, v);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_p2p_copy(const void* src, void* dst, long long nbytes, int max_ctas, void* stream) {
    PDNN_REQUIRE(nbytes % 16 == 0, "p2p_copy: size must be a multiple of 16 bytes");
    const int grid = ps_grid(nbytes / 16 / 4 + 1, 512, max_ctas > 0 ? max_ctas : 148 * 2);
    p2p_copy_kernel<<<grid, 512, 0, ST(stream)>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), nbytes / 16);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// ---- raw allocations + CUDA IPC (peer-mapped communication buffers) ---------------------------
extern "C" int pdnn_malloc(void** out, long long nbytes) {
    PDNN_CUDA_CHECK(cudaMalloc(out, (size_t)nbytes));
    PDNN_CUDA_CHECK(cudaMemset(*out, 0, (size_t)nbytes));
    return 0;
}
extern "C" int pdnn_free(void* p) {
    PDNN_CUDA_CHECK(cudaFree(p));
    return 0;
}
extern "C" int pdnn_ipc_get_handle(void* dev_ptr, void* handle_out64) {
    cudaIpcMemHandle_t h;
    PDNN_CUDA_CHECK(cudaIpcGetMemHandle(&h, dev_ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out64, &h, 64);
    return 0;
}
extern "C" int pdnn_ipc_open_handle(const void* handle64, void** dev_ptr_out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    PDNN_CUDA_CHECK(cudaIpcOpenMemHandle(dev_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int pdnn_ipc_close(void* dev_ptr) {
    PDNN_CUDA_CHECK(cudaIpcCloseMemHandle(dev_ptr));
    return 0;
}
extern "C" int pdnn_enable_peer_access(int peer_device) {
    int dev = 0, can = 0;
    PDNN_CUDA_CHECK(cudaGetDevice(&dev));
    if (peer_device == dev) return 0;
    PDNN_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, dev, peer_device));
    PDNN_REQUIRE(can, "device %d cannot access peer %d", dev, peer_device);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
    PDNN_CUDA_CHECK(e);
    return 0;
}
