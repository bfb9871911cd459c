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

// Zero-copy form of the push: the gradient bucket already sits where the master reads it (the worker's fp32 gradient
// buffer IS its slot, summed in the switch), so only the step-tagged arrival flag has to be published. The kernels that
// wrote the bucket precede this one in the stream; the fence orders their global writes before the flag for a reader
// on another GPU.
__global__ void ps_publish_kernel(uint32_t* flag_peer, uint32_t step, const uint32_t* step_ptr, const uint32_t* guard_closed,
                                  const uint32_t* guard_step) {
    if (step_ptr != nullptr) step = ld_relaxed_sys(step_ptr);
    if (guard_closed != nullptr && ld_relaxed_sys(guard_closed) >= ld_relaxed_sys(guard_step)) return;   // killed: nothing to announce
    __threadfence_system();
    st_release_sys(flag_peer, step);
}

// ------------------------------------------------------------------------------------------------
// master: per gradient bucket, in arrival order: take the first k arrivals, reduce + 1/k + momentum SGD +
// cast + broadcast; the last bucket also decides the kill set and releases the workers.
//
// The reference sums every message the moment Waitany returns it (sync_replicas_master_nn.py:148-166; C++
// sync_replicas_master_nn.h:71-101) and its kill decision is taken on the k-th arrival of parameter 0, the last
// gradient a worker produces (:172-187). Here a step is n_buckets x {wait kernel, reduce kernel} on the master's
// own stream: bucket b (deepest layers first) is reduced, applied and broadcast while the workers are still
// running the backward pass of the shallower layers, so only the last (shallowest) bucket sits on the critical
// path after the k-th worker finishes. New weights land in a parity buffer (step & 1) that no worker reads
// before the step is released, so updating bucket b mid-backward cannot disturb a worker that still needs the
// old weights for its data gradients.
// ------------------------------------------------------------------------------------------------
enum { CTL_DONE = 2, CTL_ERR = 4, CTL_TOTAL = 6, CTL_T0 = 8, CTL_TK = 10, CTL_BMASK = 64, CTL_BCNT = 128, CTL_SELECTED = 192 };

struct MasterArgs {
    pdnn_ps_master_t a;
    uint32_t* result;
};

__device__ __forceinline__ void multimem_st_v4(void* mc_addr, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void multimem_st_v2(void* mc_addr, uint2 v) {
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc_addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ uint64_t ld_u64(const uint32_t* p) { return (uint64_t)p[0] | ((uint64_t)p[1] << 32); }
__device__ __forceinline__ void st_u64(uint32_t* p, uint64_t v) { p[0] = (uint32_t)v; p[1] = (uint32_t)(v >> 32); }

// One warp: lane w polls worker w's arrival word of bucket b; arrivals are admitted in the order they are seen
// (simultaneous ones by rank) until k are in. <<<1, 32>>>
__global__ void __launch_bounds__(32) ps_bucket_wait_kernel(const __grid_constant__ MasterArgs margs, int b, int is_first, int is_last) {
    const pdnn_ps_master_t& a = margs.a;
    const uint32_t cur_step = a.step_ptr != nullptr ? ld_relaxed_sys(a.step_ptr) : a.step;
    uint32_t* ctl = a.control;
    const int lane = threadIdx.x;
    uint64_t t0;
    if (is_first) {
        t0 = globaltimer_ns();
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (lane == 0) st_u64(ctl + CTL_T0, t0);
    } else {
        t0 = ld_u64(ctl + CTL_T0);
    }
    const uint32_t* flag = a.arrival_flags + (long long)lane * a.arrival_stride + b;
    const uint32_t want = (uint32_t)a.k;
    uint32_t mask = 0, err = 0, polls = 0;
    const uint32_t sleep_ns = a.poll_sleep_ns > 0 ? (uint32_t)a.poll_sleep_ns : 400u;
    uint64_t now = t0;
    for (;;) {
        // relaxed polling (served by L2, no L1 invalidation on this SM, which other streams' CTAs share); one system-scope
        // fence after the loop orders the slot reads of the reduce kernel behind the flags seen here
        const bool arrived = lane < a.n_workers && !((mask >> lane) & 1u) && ld_relaxed_sys(flag) == cur_step;
        uint32_t m = __ballot_sync(0xffffffffu, arrived);
        while (m != 0 && (uint32_t)__popc(mask) < want) {
            mask |= m & (0u - m);     // lowest set bit
            m &= m - 1;
        }
        if ((uint32_t)__popc(mask) >= want) { now = __shfl_sync(0xffffffffu, globaltimer_ns(), 0); break; }
        if ((++polls & 15u) != 0u && a.deadline_ns == 0) { __nanosleep(sleep_ns); continue; }   // the clock is read every 16th poll
        now = __shfl_sync(0xffffffffu, globaltimer_ns(), 0);
        const uint64_t el = now - t0;
        if (a.deadline_ns > 0 && el > (uint64_t)a.deadline_ns && mask != 0) break;      // interval policy
        if (el > (a.hard_timeout_ns > 0 ? (uint64_t)a.hard_timeout_ns : PDNN_PS_HARD_TIMEOUT_NS)) { err = 1; break; }
        __nanosleep(sleep_ns);
    }
    __threadfence_system();
    const uint32_t count = (uint32_t)__popc(mask);
    if (is_last && lane < a.n_workers && !((mask >> lane) & 1u) && a.worker_closed[lane] != nullptr)
        st_release_sys(a.worker_closed[lane], cur_step);          // straggler kill: this step is closed for you
    if (is_last && lane < a.n_workers && lane < 16) {      // how often each worker was among the first k (since the run began)
        ctl[CTL_SELECTED + lane] += (mask >> lane) & 1u;
        margs.result[56 + lane] = ctl[CTL_SELECTED + lane];
    }
    if (lane == 0) {
        ctl[CTL_BMASK + b] = err ? 0u : mask;
        ctl[CTL_BCNT + b] = err ? 0u : count;
        if (err) ctl[CTL_ERR] = 1;
        margs.result[8 + b] = mask;
        margs.result[40 + b] = (uint32_t)((now - t0) / 1000ull);
        if (is_last) {
            const uint32_t all = a.n_workers >= 32 ? 0xffffffffu : ((1u << a.n_workers) - 1u);
            const uint32_t bad = ctl[CTL_ERR];
            margs.result[0] = mask;
            margs.result[1] = all & ~mask;
            margs.result[2] = count;
            margs.result[3] = (err || bad) ? 0xDEADu : cur_step;
            margs.result[4] = (uint32_t)((now - t0) / 1000ull);   // step start -> k-th arrival of the last bucket
            margs.result[7] = (uint32_t)a.n_buckets;
            ctl[CTL_TOTAL] += count;
            margs.result[6] = ctl[CTL_TOTAL];                     // gradient sets applied since the start of the run
            st_u64(ctl + CTL_TK, now);
        }
    }
}

template <bool BF16_SLOT>
__device__ __forceinline__ float4 ps_load_sum(const pdnn_ps_master_t& a, uint32_t mask, bool in_switch, long long i) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in_switch) {
        // every worker contributes: one load-reduce through the NVLS multicast address sums the workers' local slots
        // inside the switch (1x ingress instead of n x)
        if (BF16_SLOT) {
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
        return acc;
    }
    for (int w = 0; w < a.n_workers; ++w) {       // rank order: deterministic sum
        if (!((mask >> w) & 1u)) continue;
        if (BF16_SLOT) {
            const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(a.slots[w]) + i * 4);
            const float2 lo = unpack_bf16x2(u.x), hi = unpack_bf16x2(u.y);
            acc.x += lo.x; acc.y += lo.y; acc.z += hi.x; acc.w += hi.y;
        } else {
            const float4 g = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.slots[w]) + i * 4);
            acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
        }
    }
    return acc;
}

template <bool BF16_SLOT>
__device__ __forceinline__ void ps_poison(const pdnn_ps_master_t& a, uint32_t mask, bool in_switch, long long i) {
    if (in_switch) {
        if (BF16_SLOT) multimem_st_v2(const_cast<__nv_bfloat16*>(reinterpret_cast<const __nv_bfloat16*>(a.multicast_slots)) + i * 4,
                                      make_uint2(0x7FC07FC0u, 0x7FC07FC0u));
        else multimem_st_v4(const_cast<float*>(reinterpret_cast<const float*>(a.multicast_slots)) + i * 4,
                            make_uint4(0x7FC00000u, 0x7FC00000u, 0x7FC00000u, 0x7FC00000u));
        return;
    }
    for (int w = 0; w < a.n_workers; ++w) {
        if (!((mask >> w) & 1u)) continue;
        if (BF16_SLOT)
            *reinterpret_cast<uint2*>(const_cast<__nv_bfloat16*>(reinterpret_cast<const __nv_bfloat16*>(a.slots[w])) + i * 4) =
                make_uint2(0x7FC07FC0u, 0x7FC07FC0u);
        else
            *reinterpret_cast<uint4*>(const_cast<float*>(reinterpret_cast<const float*>(a.slots[w])) + i * 4) =
                make_uint4(0x7FC00000u, 0x7FC00000u, 0x7FC00000u, 0x7FC00000u);
    }
}

// Reduce + update + broadcast of elements [lo, hi) (multiples of 4). Each thread keeps PS_UNROLL independent
// gradient loads in flight (one multimem.ld_reduce round trip through the switch is ~2 us).
#define PS_UNROLL 4
template <bool BF16_SLOT>
__global__ void __launch_bounds__(512, 2) ps_bucket_reduce_kernel(const __grid_constant__ MasterArgs margs, int b, int is_last) {
    const pdnn_ps_master_t& a = margs.a;
    const uint32_t cur_step = a.step_ptr != nullptr ? ld_relaxed_sys(a.step_ptr) : a.step;
    uint32_t* ctl = a.control;
    const uint32_t mask = ctl[CTL_BMASK + b];      // written by the wait kernel that precedes this launch in the stream
    const uint32_t nsel = ctl[CTL_BCNT + b];
    if (nsel > 0) {
        const float scale = a.avg_mode == 1 ? 1.f / (float)a.n_workers : 1.f / (float)nsel;
        const bool in_switch = a.multicast_slots != nullptr && nsel == (uint32_t)a.n_workers;
        const int par = (int)((cur_step + 1u) & 1u);            // weights of step s+1 land in parity (s+1) & 1
        void* mcw = a.multicast_weights[par];
        const long long i0 = a.bucket_lo[b] / 4, i1 = a.bucket_hi[b] / 4;
        const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const long long nth = (long long)gridDim.x * blockDim.x;
        const float lr = a.lr, mu = a.mu;
        for (long long base = i0 + tid; base < i1; base += nth * PS_UNROLL) {
            // all loads of the PS_UNROLL elements are issued before any is used: the gradients come through the switch
            // (microseconds), momentum and master weights from HBM; issued back to back their latencies overlap
            float4 acc[PS_UNROLL], mom[PS_UNROLL], wgt[PS_UNROLL];
#pragma unroll
            for (int u = 0; u < PS_UNROLL; ++u) {
                const long long i = base + u * nth;
                if (i < i1) {
                    acc[u] = ps_load_sum<BF16_SLOT>(a, mask, in_switch, i);
                    mom[u] = __ldcs(reinterpret_cast<const float4*>(a.momentum) + i);
                    wgt[u] = __ldcs(reinterpret_cast<const float4*>(a.master_w) + i);
                }
            }
#pragma unroll
            for (int u = 0; u < PS_UNROLL; ++u) {
                const long long i = base + u * nth;
                if (i >= i1) continue;
                if (a.poison_slots) ps_poison<BF16_SLOT>(a, mask, in_switch, i);
                float4 m = mom[u];
                float4 wv = wgt[u];
                m.x = mu * m.x + scale * acc[u].x; m.y = mu * m.y + scale * acc[u].y;
                m.z = mu * m.z + scale * acc[u].z; m.w = mu * m.w + scale * acc[u].w;
                wv.x -= lr * m.x; wv.y -= lr * m.y; wv.z -= lr * m.z; wv.w -= lr * m.w;
                __stcs(reinterpret_cast<float4*>(a.momentum) + i, m);
                __stcs(reinterpret_cast<float4*>(a.master_w) + i, wv);
                if (a.weights_are_bf16) {
                    uint2 u2;
                    u2.x = pack_bf16x2(wv.x, wv.y); u2.y = pack_bf16x2(wv.z, wv.w);
                    if (mcw != nullptr) {
                        multimem_st_v2(reinterpret_cast<__nv_bfloat16*>(mcw) + i * 4, u2);
                    } else {
                        for (int r = 0; r < a.n_receivers; ++r)
                            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(a.recv_weights[par][r]) + i * 4) = u2;
                    }
                } else {
                    const uint4 u4 = *reinterpret_cast<uint4*>(&wv);
                    if (mcw != nullptr) {
                        multimem_st_v4(reinterpret_cast<float*>(mcw) + i * 4, u4);
                    } else {
                        for (int r = 0; r < a.n_receivers; ++r) st_na_v4(reinterpret_cast<float*>(a.recv_weights[par][r]) + i * 4, u4);
                    }
                }
            }
        }
    }
    if (!is_last) return;
    // -------- grid completion of the last bucket, then release every receiver with step+1
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const uint32_t done = atom_add_acqrel_gpu(ctl + CTL_DONE, 1u) + 1u;
        if (done == gridDim.x) {
            ctl[CTL_DONE] = 0;
            ctl[CTL_ERR] = 0;
            margs.result[5] = (uint32_t)((globaltimer_ns() - ld_u64(ctl + CTL_TK)) / 1000ull);   // k-th arrival -> release
            if (a.step_ptr != nullptr) st_relaxed_sys(a.step_ptr, cur_step + 1u);
            __threadfence_system();
            for (int r = 0; r < a.n_receivers; ++r)
                if (a.recv_ready[r] != nullptr) st_release_sys(a.recv_ready[r], cur_step + 1u);
        }
    }
}

// Worker: copy the landing buffer of this step (parity step & 1) into the private compute copy.
__global__ void __launch_bounds__(512) ps_fetch_kernel(const uint4* __restrict__ land0, const uint4* __restrict__ land1,
                                                       uint4* __restrict__ dst, long long n16, const uint32_t* step_word, uint32_t step) {
    if (step_word != nullptr) step = ld_relaxed_sys(step_word);
    const uint4* src = (step & 1u) ? land1 : land0;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    long long i = tid;
    for (; i + 3 * nth < n16; i += 4 * nth) {
        const uint4 a = src[i], b = src[i + nth], c = src[i + 2 * nth], d = src[i + 3 * nth];
        dst[i] = a; dst[i + nth] = b; dst[i + 2 * nth] = c; dst[i + 3 * nth] = d;
    }
    for (; i < n16; i += nth) dst[i] = src[i];
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

// Straggler injection sampled ON the device: the draw is a pure function of (seed, worker, step, layer) through
// Philox4x32-10, the step comes from the worker's device-side step word, so the launch is identical every step and can
// be replayed from a CUDA graph (the reference sleeps on the host: pure_py_code/distributed_worker.py:130-132 before
// every backward layer; distributed_TF/src/distributed_train.py:288-290 uniform 300..1200 ms per step).
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

__global__ void ps_delay_sampled_kernel(const __grid_constant__ pdnn_delay_spec_t sp, int worker, int layer, const uint32_t* step_word,
                                        const uint32_t* guard_closed, uint32_t* log_ns) {
    const uint32_t step = ld_relaxed_sys(step_word);
    uint32_t c[4] = {step, (uint32_t)layer, 0u, 0u};
    philox4x32_10(c, (uint32_t)sp.seed, (uint32_t)(sp.seed >> 32) ^ ((uint32_t)worker * 0x9E3779B9u));
    const float u0 = (float)(c[0] >> 8) * (1.0f / 16777216.0f), u1 = (float)(c[1] >> 8) * (1.0f / 16777216.0f);
    const float u2 = (float)(c[2] >> 8) * (1.0f / 16777216.0f);
    float ms = 0.f;
    const bool hit = (sp.rank_mask == 0ull || ((sp.rank_mask >> (worker + 1)) & 1ull)) && (sp.prob >= 1.0f || u0 < sp.prob);
    if (hit) {
        switch (sp.kind) {
            case 1: ms = sp.a; break;
            case 2: ms = sp.a + (sp.b - sp.a) * u1; break;
            case 3: ms = sp.a - sp.b * logf(1.0f - u1); break;
            case 4: ms = expf(sp.a + sp.b * sqrtf(-2.0f * logf(1.0f - u1)) * cospif(2.0f * u2)); break;
            case 5: ms = sp.a * powf(1.0f - u1, -1.0f / sp.b); break;
            default: break;
        }
    }
    const float nsf = fmaxf(ms, 0.f) * sp.scale * 1e6f;
    const uint64_t ns = (uint64_t)nsf;
    if (log_ns != nullptr) log_ns[layer] = (uint32_t)(ns > 0xffffffffull ? 0xffffffffull : ns);
    const uint64_t t0 = globaltimer_ns();
    while (globaltimer_ns() - t0 < ns) {
        if (guard_closed != nullptr && ld_relaxed_sys(guard_closed) >= step) break;      // the step was closed: stop waiting
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

// ---- link microbenchmarks: multicast store, in-switch load-reduce, flag round trip -------------------
__global__ void __launch_bounds__(512) mc_store_kernel(const uint4* __restrict__ src, uint4* mc_dst, long long n16) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    for (long long i = tid; i < n16; i += nth) multimem_st_v4(mc_dst + i, src[i]);
}
__global__ void __launch_bounds__(512) mc_ld_reduce_kernel(const float* mc_src, float4* __restrict__ dst, long long n4) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    for (long long base = tid; base < n4; base += nth * 4) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long i = base + u * nth;
            if (i < n4)
                asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(mc_src + i * 4) : "memory");
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long i = base + u * nth;
            if (i < n4) dst[i] = v[u];
        }
    }
}
// Two GPUs bounce a counter: the initiator stores i into the peer's word and waits for the echo in its own.
__global__ void flag_pingpong_kernel(uint32_t* mine, uint32_t* peers, int iters, int initiator, uint32_t base, unsigned long long* out_ns) {
    const uint64_t t0 = globaltimer_ns();
    for (int i = 1; i <= iters; ++i) {
        const uint32_t v = base + (uint32_t)i;
        if (initiator) {
            st_release_sys(peers, v);
            while (ld_acquire_sys(mine) != v) {}
        } else {
            while (ld_acquire_sys(mine) != v) {}
            st_release_sys(peers, v);
        }
    }
    if (out_ns != nullptr) *out_ns = globaltimer_ns() - t0;
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

extern "C" int pdnn_ps_publish(uint32_t* flag_peer, uint32_t step, const uint32_t* step_ptr, const uint32_t* guard_closed,
                               const uint32_t* guard_step, void* stream) {
    PDNN_REQUIRE(flag_peer != nullptr, "ps_publish: missing flag");
    ps_publish_kernel<<<1, 1, 0, ST(stream)>>>(flag_peer, step, step_ptr, guard_closed, guard_step);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_ps_master_step(const pdnn_ps_master_t* args, uint32_t* result, void* stream) {
    PDNN_REQUIRE(args != nullptr && result != nullptr, "ps_master_step: null arguments");
    PDNN_REQUIRE(args->n_workers >= 1 && args->n_workers <= PDNN_MAX_WORKERS, "ps_master_step: n_workers=%d out of range", args->n_workers);
    PDNN_REQUIRE(args->n_receivers >= 0 && args->n_receivers <= PDNN_MAX_RECEIVERS, "ps_master_step: n_receivers=%d out of range", args->n_receivers);
    PDNN_REQUIRE(args->k >= 1 && args->k <= args->n_workers, "ps_master_step: k=%d must be in [1, %d]", args->k, args->n_workers);
    PDNN_REQUIRE(args->control != nullptr && args->master_w != nullptr && args->momentum != nullptr, "ps_master_step: missing buffers");
    PDNN_REQUIRE(args->n_buckets >= 1 && args->n_buckets <= PDNN_MAX_BUCKETS, "ps_master_step: n_buckets=%d out of range", args->n_buckets);
    PDNN_REQUIRE(args->arrival_stride >= args->n_buckets, "ps_master_step: arrival_stride %d < n_buckets %d", args->arrival_stride, args->n_buckets);
    for (int b = 0; b < args->n_buckets; ++b)
        PDNN_REQUIRE(args->bucket_lo[b] >= 0 && args->bucket_hi[b] > args->bucket_lo[b] && args->bucket_hi[b] <= args->n &&
                         args->bucket_lo[b] % 4 == 0 && args->bucket_hi[b] % 4 == 0,
                     "ps_master_step: bucket %d = [%lld, %lld) must be a non-empty 4-element-aligned range inside [0, %lld)", b,
                     args->bucket_lo[b], args->bucket_hi[b], args->n);
    for (int b = 0; b < args->n_buckets; ++b)
        if (int rc = pdnn_ps_master_bucket(args, result, b, stream)) return rc;
    return 0;
}

// One bucket of a master step (see pdnn_ps_master_step, which is the loop over all of them; arguments validated there).
// A caller that knows when bucket b can first be complete (the colocated worker has just pushed its own share) can
// enqueue the buckets one by one behind that point, so that the polling warp is resident for the skew between the
// workers only, not for the whole backward pass.
extern "C" int pdnn_ps_master_bucket(const pdnn_ps_master_t* args, uint32_t* result, int b, void* stream) {
    PDNN_REQUIRE(args != nullptr && result != nullptr && b >= 0 && b < args->n_buckets, "ps_master_bucket: bad arguments");
    MasterArgs m;
    m.a = *args;
    m.result = result;
    const int last = b == args->n_buckets - 1;
    ps_bucket_wait_kernel<<<1, 32, 0, ST(stream)>>>(m, b, b == 0, last);
    PDNN_COUNT_LAUNCH();
    const int grid = ps_grid((args->bucket_hi[b] - args->bucket_lo[b]) / 4 / PS_UNROLL + 1, 512, args->max_ctas > 0 ? args->max_ctas : 296);
    if (args->slot_is_bf16) ps_bucket_reduce_kernel<true><<<grid, 512, 0, ST(stream)>>>(m, b, last);
    else ps_bucket_reduce_kernel<false><<<grid, 512, 0, ST(stream)>>>(m, b, last);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pdnn_ps_fetch_weights(const void* land0, const void* land1, void* dst, long long nbytes, const uint32_t* step_word,
                                     uint32_t step, void* stream) {
    PDNN_REQUIRE(nbytes > 0 && nbytes % 16 == 0, "ps_fetch_weights: size must be a positive multiple of 16 bytes");
    const int grid = ps_grid(nbytes / 16 / 4 + 1, 512, 148 * 2);
    ps_fetch_kernel<<<grid, 512, 0, ST(stream)>>>(reinterpret_cast<const uint4*>(land0), reinterpret_cast<const uint4*>(land1),
                                                   reinterpret_cast<uint4*>(dst), nbytes / 16, step_word, step);
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

extern "C" int pdnn_delay_sampled(const pdnn_delay_spec_t* spec, int worker, int layer, const uint32_t* step_word,
                                  const uint32_t* guard_closed, uint32_t* log_ns, void* stream) {
    PDNN_REQUIRE(spec != nullptr && step_word != nullptr, "delay_sampled: spec and step word are required");
    PDNN_REQUIRE(spec->kind >= 0 && spec->kind <= 5, "delay_sampled: unknown kind %d", spec->kind);
    PDNN_REQUIRE(worker >= 0 && worker < 63 && layer >= 0, "delay_sampled: worker / layer out of range");
    if (spec->kind == 0) return 0;
    ps_delay_sampled_kernel<<<1, 1, 0, ST(stream)>>>(*spec, worker, layer, step_word, guard_closed, log_ns);
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

extern "C" int pdnn_mc_store(const void* src, void* mc_dst, long long nbytes, int max_ctas, void* stream) {
    PDNN_REQUIRE(nbytes > 0 && nbytes % 16 == 0 && mc_dst != nullptr, "mc_store: needs a multicast address and a multiple of 16 bytes");
    mc_store_kernel<<<ps_grid(nbytes / 16 + 1, 512, max_ctas > 0 ? max_ctas : 148 * 2), 512, 0, ST(stream)>>>(
        reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(mc_dst), nbytes / 16);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}
extern "C" int pdnn_mc_ld_reduce_f32(const void* mc_src, float* dst, long long n, int max_ctas, void* stream) {
    PDNN_REQUIRE(n > 0 && n % 4 == 0 && mc_src != nullptr, "mc_ld_reduce: needs a multicast address and a multiple of 4 elements");
    mc_ld_reduce_kernel<<<ps_grid(n / 16 + 1, 512, max_ctas > 0 ? max_ctas : 148 * 2), 512, 0, ST(stream)>>>(
        reinterpret_cast<const float*>(mc_src), reinterpret_cast<float4*>(dst), n / 4);
    PDNN_COUNT_LAUNCH();
    PDNN_CUDA_CHECK(cudaGetLastError());
    return 0;
}
extern "C" int pdnn_flag_pingpong(uint32_t* mine, uint32_t* peers, int iters, int initiator, uint32_t base, unsigned long long* out_ns,
                                  void* stream) {
    PDNN_REQUIRE(mine != nullptr && peers != nullptr && iters > 0, "flag_pingpong: bad arguments");
    flag_pingpong_kernel<<<1, 1, 0, ST(stream)>>>(mine, peers, iters, initiator, base, out_ns);
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
