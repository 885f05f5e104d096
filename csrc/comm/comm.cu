// NVLink peer-memory communication fused with compute (BASELINE.json north-star).
//
// Every rank owns one slab (gradient arena | fp32 masters | bf16 shadows | exchange slots | flags) that all peers of the node
// map: cudaMalloc + CUDA IPC, or torch's symmetric-memory allocator when available (which adds an NVSwitch multicast view).  Kernels then dereference peer pointers
// directly; NVSwitch gives every pair full bandwidth, so there is no ring and no NCCL call on the hot path.
//
//   fused_allreduce_update_kernel  (one launch per gradient bucket, side stream, overlapped with backward)
//       barrier-in  : flag words, st.release.sys / ld.acquire.sys, epoch numbered (never reset)
//       two-shot    : rank r sums slice r of the bucket from all W gradient arenas over NVLink (fixed rank order),
//                     scales by 1/W, runs SGD-momentum / Adam / AdamW on the fp32 master (optimizer state is
//                     touched only by the slice owner), and PUSHES the new fp32 master + bf16 shadow to all peers
//       one-shot    : every rank sums the whole (small, latency-bound) bucket and updates its own replica
//       NVLS        : two-shot through the NVSwitch multicast mapping of the slab (when the host provides one): the slice
//                     owner pulls the in-switch reduced gradients with multimem.ld_reduce and broadcasts master + shadow with
//                     multimem.st (mode 3), or keeps the unicast pushes (mode 4, faster above ~512 MB)
//       bf16 wire   : (default) before the barrier-in every rank packs its fp32 gradients of the bucket to bf16 into a staging
//                     region of its slab; peers read THAT (2 B per element instead of 4), and with sharded masters the owner
//                     keeps its fp32 master slice to itself and pushes only the bf16 compute copy (2 B instead of 6):
//                     2.5x fewer NVLink bytes per bucket
//       barrier-out : once every peer has read my gradients and pushed my weights, re-zero my gradient bucket
//   bn_finalize_sync_kernel        SyncBatchNorm statistics: publish (sum, sumsq, count), wait, sum all peers'
//                                   slots in rank order, finalize scale/shift/running stats -- one launch, no NCCL
//   p2p_allreduce_small_kernel     the same exchange for the two backward BN reductions
#include "../common/picd_common.cuh"

#include <cstring>

namespace picd {

constexpr int MAX_PEERS = 8;

struct PeerTable {
    float* grad[MAX_PEERS];
    float* master[MAX_PEERS];
    __nv_bfloat16* shadow[MAX_PEERS];
    float* xchg[MAX_PEERS];          // exchange slots for small collectives
    unsigned* flags[MAX_PEERS];
    float* bufs[MAX_PEERS];          // module buffers (BatchNorm running statistics) kept identical to rank 0's, or nullptr
    __nv_bfloat16* gbf[MAX_PEERS];   // bf16 staging copies of the gradient arenas (the wire format of the reduction), or nullptr
    // NVSwitch multicast (NVLS) views of the same regions, or nullptr: a multimem.ld_reduce on mc_grad returns the sum over all
    // ranks computed inside the switch, a multimem.st on mc_master / mc_shadow lands in every rank's copy
    float* mc_grad;
    float* mc_master;
    __nv_bfloat16* mc_shadow;
    __nv_bfloat16* mc_gbf;
};

struct OptArgs {
    int kind;            // 0 sgd-momentum, 1 adam, 2 adamw
    float lr, momentum, beta1, beta2, eps, weight_decay;
    float bc1, bc2;      // 1 - beta^step (adam bias corrections)
    int nesterov;
    float grad_scale;    // 1 / world
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
// in-switch sum of four bf16 values from every rank (accumulated in fp32 inside the switch, delivered as bf16)
__device__ __forceinline__ uint2 multimem_ld_reduce_bf16x4(const __nv_bfloat16* mc) {
    uint2 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.bf16x2 {%0, %1}, [%2];"
                 : "=r"(v.x), "=r"(v.y) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ uint2 ld_peer_u2(const void* p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ void multimem_st_b64(void* mc, const uint2& v) {     // 8 raw bytes (four bf16 shadow weights)
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y))
                 : "memory");
}

// Spin until every peer's flag word (in MY flag array) has reached `epoch`; bounded so a lost peer traps instead of hanging
// the GPU forever.  The bound (nanoseconds of %globaltimer) comes from the host: PICD_FLAG_TIMEOUT_S, default 600 s -- the
// same order as the process-group timeout, so a rank that is merely slow (first batch of an epoch, a filesystem stall, a
// debugger) does not take its peers' contexts down.
__device__ long long g_flag_timeout_ns = 600LL * 1000000000LL;
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void wait_flags(const unsigned* my_flags, int slot, int W, unsigned epoch) {
    if (threadIdx.x < W) {
        const unsigned* f = my_flags + slot * MAX_PEERS + threadIdx.x;
        if (static_cast<int>(ld_acquire_sys(f) - epoch) < 0) {
            const unsigned long long t0 = globaltimer_ns();
            const unsigned long long budget = static_cast<unsigned long long>(g_flag_timeout_ns);
            while (static_cast<int>(ld_acquire_sys(f) - epoch) < 0) {
                if (globaltimer_ns() - t0 > budget) __trap();
            }
        }
    }
    __syncthreads();
}
__device__ __forceinline__ void signal_peers(const PeerTable& pt, int slot, int rank, int W, unsigned epoch) {
    __threadfence_system();
    if (threadIdx.x < W) st_release_sys(pt.flags[threadIdx.x] + slot * MAX_PEERS + rank, epoch);
}

__device__ __forceinline__ float opt_update(float p, float g, float& s0, float& s1, const OptArgs& o) {
    g *= o.grad_scale;
    if (o.kind == 0) {
        g = fmaf(o.weight_decay, p, g);
        s0 = fmaf(o.momentum, s0, g);
        const float upd = o.nesterov ? fmaf(o.momentum, s0, g) : s0;
        return fmaf(-o.lr, upd, p);
    }
    if (o.kind == 2) p *= (1.f - o.lr * o.weight_decay);
    else g = fmaf(o.weight_decay, p, g);
    s0 = fmaf(o.beta1, s0, (1.f - o.beta1) * g);
    s1 = fmaf(o.beta2, s1, (1.f - o.beta2) * g * g);
    const float denom = sqrtf(s1) / sqrtf(o.bc2) + o.eps;
    return p - (o.lr / o.bc1) * (s0 / denom);
}

// mode: 0 = local (W == 1, no flags), 1 = one-shot, 2 = two-shot, 3 = two-shot through NVSwitch multicast (NVLS; 4 = NVLS
// reduction with unicast pushes): the slice
// owner pulls the reduced gradients with ONE multimem.ld_reduce per vector instead of W peer loads and broadcasts the updated
// master / shadow with ONE multimem.st each instead of W peer stores
__global__ void __launch_bounds__(512) fused_allreduce_update_kernel(PeerTable pt, int rank, int W, int mode,
                                                                     long long start, long long count,
                                                                     float* __restrict__ state0,
                                                                     float* __restrict__ state1, OptArgs opt,
                                                                     int slot_in, int slot_out, unsigned epoch,
                                                                     unsigned* block_counter,
                                                                     const float* __restrict__ dyn, long long bcast_n,
                                                                     int wire_bf16, int shard_master,
                                                                     unsigned* pack_counter) {
    if (dyn != nullptr) {
        // step-varying scalars come from device memory so that a captured CUDA graph can be replayed step after step:
        // (lr, 1 - beta1^t, 1 - beta2^t, flag epoch as raw bits), rewritten by the host before each replay
        opt.lr = dyn[0]; opt.bc1 = dyn[1]; opt.bc2 = dyn[2];
        epoch = __float_as_uint(dyn[3]);
    }
    const unsigned* my_flags = pt.flags[rank];
    if (mode != 0) {
        if (wire_bf16) {
            // wire format: pack my fp32 gradients of this bucket to bf16 (what the peers will read); the LAST block of this
            // grid to finish tells the peers (self-resetting ticket, like the barrier-out below)
            const long long n8 = count >> 3;
            const float* src = pt.grad[rank] + start;
            __nv_bfloat16* dst = pt.gbf[rank] + start;
            for (long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v < n8;
                 v += static_cast<long long>(gridDim.x) * blockDim.x) {
                const float4 a = *reinterpret_cast<const float4*>(src + (v << 3));
                const float4 b = *reinterpret_cast<const float4*>(src + (v << 3) + 4);
                *reinterpret_cast<uint4*>(dst + (v << 3)) =
                    make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
            }
            __shared__ unsigned s_packed;
            __threadfence_system();
            __syncthreads();
            if (threadIdx.x == 0) {
                const unsigned old = atomicAdd(pack_counter, 1u);
                s_packed = (old + 1u == gridDim.x) ? 1u : 0u;
                if (s_packed) atomicExch(pack_counter, 0u);
            }
            __syncthreads();
            if (s_packed) signal_peers(pt, slot_in, rank, W, epoch);
        } else if (blockIdx.x == 0) {
            signal_peers(pt, slot_in, rank, W, epoch);                    // my gradients for this bucket are final
        }
        wait_flags(my_flags, slot_in, W, epoch);                          // ... and so are everybody else's
    }
    long long lo = start, hi = start + count;
    if (mode >= 2) {
        const long long per = count / W;
        lo = start + per * rank;
        hi = lo + per;
    }
    const long long nvec = (hi - lo) >> 2;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    constexpr int UNR = 2;     // independent 16-byte vectors per thread per iteration (2 x W peer loads in flight)
    for (long long v0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < nvec; v0 += stride * UNR) {
        float4 g[UNR], w[UNR], s0[UNR], s1[UNR];
        bool live[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const long long v = v0 + u * stride;
            live[u] = v < nvec;
            g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            s1[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (!live[u]) continue;
            const long long i = lo + (v << 2);
            if (mode == 0) {
                g[u] = *reinterpret_cast<const float4*>(pt.grad[rank] + i);
            } else if (wire_bf16) {
                if (mode >= 3) {
                    const uint2 r = multimem_ld_reduce_bf16x4(pt.mc_gbf + i);  // the switch adds the W bf16 copies in fp32
                    const float2 lo2 = unpack_bf16x2(r.x), hi2 = unpack_bf16x2(r.y);
                    g[u] = make_float4(lo2.x, lo2.y, hi2.x, hi2.y);
                } else {
                    uint2 part[MAX_PEERS];
#pragma unroll
                    for (int p = 0; p < MAX_PEERS; ++p)
                        if (p < W) part[p] = ld_peer_u2(pt.gbf[p] + i);       // 8 bytes per peer instead of 16
#pragma unroll
                    for (int p = 0; p < MAX_PEERS; ++p)                       // fixed rank order: bitwise reproducible
                        if (p < W) {
                            const float2 lo2 = unpack_bf16x2(part[p].x), hi2 = unpack_bf16x2(part[p].y);
                            g[u].x += lo2.x; g[u].y += lo2.y; g[u].z += hi2.x; g[u].w += hi2.y;
                        }
                }
            } else if (mode >= 3) {
                g[u] = multimem_ld_reduce_f4(pt.mc_grad + i);                 // the switch adds the W copies
            } else {
                float4 part[MAX_PEERS];
#pragma unroll
                for (int p = 0; p < MAX_PEERS; ++p)
                    if (p < W) part[p] = ld_peer_f4(pt.grad[p] + i);          // all W loads in flight together
#pragma unroll
                for (int p = 0; p < MAX_PEERS; ++p)                           // fixed rank order: bitwise reproducible
                    if (p < W) { g[u].x += part[p].x; g[u].y += part[p].y; g[u].z += part[p].z; g[u].w += part[p].w; }
            }
            w[u] = *reinterpret_cast<const float4*>(pt.master[rank] + i);
            s0[u] = *reinterpret_cast<const float4*>(state0 + i);
            if (opt.kind != 0) s1[u] = *reinterpret_cast<const float4*>(state1 + i);
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            if (!live[u]) continue;
            const long long i = lo + ((v0 + u * stride) << 2);
            w[u].x = opt_update(w[u].x, g[u].x, s0[u].x, s1[u].x, opt);
            w[u].y = opt_update(w[u].y, g[u].y, s0[u].y, s1[u].y, opt);
            w[u].z = opt_update(w[u].z, g[u].z, s0[u].z, s1[u].z, opt);
            w[u].w = opt_update(w[u].w, g[u].w, s0[u].w, s1[u].w, opt);
            *reinterpret_cast<float4*>(state0 + i) = s0[u];
            if (opt.kind != 0) *reinterpret_cast<float4*>(state1 + i) = s1[u];
            const uint2 sh = make_uint2(pack_bf16x2(w[u].x, w[u].y), pack_bf16x2(w[u].z, w[u].w));
            if (mode == 3) {
                // one store, delivered to every rank; with sharded masters only the bf16 compute copy travels
                if (shard_master) *reinterpret_cast<float4*>(pt.master[rank] + i) = w[u];
                else multimem_st_f4(pt.mc_master + i, w[u]);
                multimem_st_b64(pt.mc_shadow + i, sh);
            } else if (mode == 2 || mode == 4) {       // mode 4: in-switch reduction, unicast pushes
                if (shard_master) *reinterpret_cast<float4*>(pt.master[rank] + i) = w[u];
#pragma unroll
                for (int p = 0; p < MAX_PEERS; ++p)
                    if (p < W) {
                        if (!shard_master) *reinterpret_cast<float4*>(pt.master[p] + i) = w[u];     // push over NVLink (all-gather phase)
                        *reinterpret_cast<uint2*>(pt.shadow[p] + i) = sh;
                    }
            } else {
                *reinterpret_cast<float4*>(pt.master[rank] + i) = w[u];
                *reinterpret_cast<uint2*>(pt.shadow[rank] + i) = sh;
            }
            if (mode == 0) *reinterpret_cast<float4*>(pt.grad[rank] + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    if (mode == 0) return;

    // DDP's per-forward buffer broadcast (torch/nn/parallel/distributed.py:2214-2257) without a collective
    // call: the step's LAST bucket carries it.  Past the barrier-in every rank has finished this step's forward, so rank 0's
    // running statistics are final and nobody touches theirs before the next forward; rank 0 pushes its copy into every
    // peer's slab, and the barrier-out below publishes it.
    if (bcast_n > 0 && rank == 0) {
        const float* src = pt.bufs[0];
        for (long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v < (bcast_n >> 2); v += stride) {
            const float4 b = *reinterpret_cast<const float4*>(src + (v << 2));
#pragma unroll
            for (int p = 1; p < MAX_PEERS; ++p)
                if (p < W) *reinterpret_cast<float4*>(pt.bufs[p] + (v << 2)) = b;
        }
    }

    // barrier-out: the last block of this grid tells the peers that this rank is done reading and pushing
    __shared__ unsigned s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        // self-resetting ticket: launches of one bucket are stream-ordered, so the counter is 0 whenever a grid starts, for
        // any grid size (the calibration sweep reuses a bucket with different sizes and modes)
        const unsigned old = atomicAdd(block_counter, 1u);
        s_last = (old + 1u == gridDim.x) ? 1u : 0u;
        if (s_last) atomicExch(block_counter, 0u);
    }
    __syncthreads();
    if (s_last) signal_peers(pt, slot_out, rank, W, epoch);
    wait_flags(my_flags, slot_out, W, epoch);
    // every peer has consumed my gradients: leave the bucket zeroed for the next accumulation
    const long long zvec = count >> 2;
    for (long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v < zvec; v += stride)
        *reinterpret_cast<float4*>(pt.grad[rank] + start + (v << 2)) = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ---- small collectives over the exchange slots ----------------------------------------------------------
// PUSH exchange: every rank stores its `n` floats into sub-slot `rank` of the slot in EVERY peer's slab (fire-and-forget NVLink
// stores), signals, waits for the W flags -- and then sums W sub-slots of its OWN slab.  The pull form (publish locally, peers
// load remotely after the flags) paid one more NVLink round trip per exchange, 2 x 53 times per SyncBatchNorm step.
// A slot half holds W sub-slots of `sub` floats; single block.
__device__ __forceinline__ void xchg_publish_and_wait(const PeerTable& pt, int rank, int W, const float* in, int n,
                                                      int in_replicas, long long slot_off, int sub, int flag_slot,
                                                      unsigned epoch, float extra, int extra_at) {
    const long long mine = slot_off + static_cast<long long>(rank) * sub;
    for (int i = threadIdx.x; i < n + (extra_at >= 0 ? 1 : 0); i += blockDim.x) {
        float acc = 0.f;                                   // fold the per-CTA-group replicas while publishing
        if (i < n) {
            for (int r = 0; r < in_replicas; ++r) acc += in[static_cast<size_t>(r) * n + i];
        } else {
            acc = extra;                                   // one more scalar rides along (the local element count)
        }
        const int at = i < n ? i : extra_at;
#pragma unroll
        for (int p = 0; p < MAX_PEERS; ++p)
            if (p < W) pt.xchg[p][mine + at] = acc;
    }
    __syncthreads();
    signal_peers(pt, flag_slot, rank, W, epoch);
    wait_flags(pt.flags[rank], flag_slot, W, epoch);
}
__device__ __forceinline__ float xchg_sum(const PeerTable& pt, int rank, int W, long long slot_off, int sub, int i) {
    float acc = 0.f;
    const float* base = pt.xchg[rank] + slot_off + i;
    for (int p = 0; p < W; ++p) {                          // fixed rank order: every rank computes bit-identical sums
        float v;
        asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(base + static_cast<long long>(p) * sub));
        acc += v;
    }
    return acc;
}

// CUDA-graph replay: the launch was recorded with the epoch of the capture step and the EVEN half of its double-buffered
// slot; *dyn_delta (rewritten by the host before every replay) says how many steps later this replay runs.
__device__ __forceinline__ void apply_replay_delta(const unsigned* dyn_delta, int slot_floats, unsigned& epoch,
                                                   long long& slot_off) {
    if (dyn_delta != nullptr) {
        epoch += *dyn_delta;
        slot_off += (epoch & 1u) ? slot_floats : 0;
    }
}

__global__ void __launch_bounds__(1024) p2p_allreduce_small_kernel(PeerTable pt, int rank, int W, const float* in,
                                                                   int in_replicas, float* out, int n,
                                                                   long long slot_off, int flag_slot, unsigned epoch,
                                                                   const unsigned* dyn_delta, int slot_floats, int sub) {
    pdl_sync();
    apply_replay_delta(dyn_delta, slot_floats, epoch, slot_off);
    xchg_publish_and_wait(pt, rank, W, in, n, in_replicas, slot_off, sub, flag_slot, epoch, 0.f, -1);
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = xchg_sum(pt, rank, W, slot_off, sub, i);
}

// sums = local [sum(C), sumsq(C)], count = local element count.  Cross-replica merge + finalize in one launch.
__global__ void __launch_bounds__(1024) bn_finalize_sync_kernel(PeerTable pt, int rank, int W, const float* sums,
                                                                int replicas, float count, const float* gamma, const float* beta,
                                                                float* running_mean, float* running_var,
                                                                float momentum, float eps, float* scale, float* shift,
                                                                float* mean_out, float* invstd_out, float* total_count,
                                                                int C, long long slot_off, int flag_slot,
                                                                unsigned epoch, const unsigned* dyn_delta, int slot_floats,
                                                                int sub) {
    pdl_sync();
    apply_replay_delta(dyn_delta, slot_floats, epoch, slot_off);
    xchg_publish_and_wait(pt, rank, W, sums, 2 * C, replicas, slot_off, sub, flag_slot, epoch, count, 2 * C);
    const float n = xchg_sum(pt, rank, W, slot_off, sub, 2 * C);
    if (threadIdx.x == 0) total_count[0] = n;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float mean = xchg_sum(pt, rank, W, slot_off, sub, c) / n;
        float var = xchg_sum(pt, rank, W, slot_off, sub, C + c) / n - mean * mean;
        var = fmaxf(var, 0.f);
        const float invstd = rsqrtf(var + eps);
        const float sc = gamma[c] * invstd;
        scale[c] = sc;
        shift[c] = beta[c] - mean * sc;
        mean_out[c] = mean;
        invstd_out[c] = invstd;
        if (running_mean != nullptr) {
            const float unbiased = n > 1.f ? var * (n / (n - 1.f)) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
    }
}

struct CommCtx {
    PeerTable pt;
    int rank, world;
    unsigned* counters;     // per-bucket block counters (device)
    int num_buckets;
    long long bufs_n;       // floats in each rank's module-buffer region (0 = none registered)
    int wire_bf16;          // peers read bf16 staging copies of the gradients (picd_comm_set_wire)
    int shard_master;       // slice owners keep their fp32 master slices to themselves; only the bf16 copy is pushed
};

}  // namespace picd

using namespace picd;

// ---- slab + IPC -----------------------------------------------------------------------------------------
extern "C" int picd_slab_alloc(size_t bytes, void** out) {
    PICD_CUDA(cudaMalloc(out, bytes));
    PICD_CUDA(cudaMemset(*out, 0, bytes));
    return 0;
}
extern "C" int picd_slab_free(void* p) {
    PICD_CUDA(cudaFree(p));
    return 0;
}
extern "C" int picd_ipc_handle_bytes() { return static_cast<int>(sizeof(cudaIpcMemHandle_t)); }
extern "C" int picd_ipc_get_handle(void* p, void* handle_out) {
    cudaIpcMemHandle_t h;
    PICD_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}
extern "C" int picd_ipc_open(const void* handle, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    PICD_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int picd_ipc_close(void* p) {
    PICD_CUDA(cudaIpcCloseMemHandle(p));
    return 0;
}

// ---- communicator context -------------------------------------------------------------------------------
// ptrs: for each of `world` ranks the base pointers (as seen from THIS process) of grad, master, shadow, xchg, flags.
extern "C" int picd_comm_create(int rank, int world, void* const* grad, void* const* master, void* const* shadow,
                                void* const* xchg, void* const* flags, int num_buckets, void** ctx_out) {
    PICD_REQUIRE(world >= 1 && world <= MAX_PEERS, "comm: world=%d unsupported (1..%d ranks of one node)", world, MAX_PEERS);
    CommCtx* c = new CommCtx();
    memset(&c->pt, 0, sizeof(c->pt));
    for (int i = 0; i < world; ++i) {
        c->pt.grad[i] = static_cast<float*>(grad[i]);
        c->pt.master[i] = static_cast<float*>(master[i]);
        c->pt.shadow[i] = static_cast<__nv_bfloat16*>(shadow[i]);
        c->pt.xchg[i] = static_cast<float*>(xchg ? xchg[i] : nullptr);
        c->pt.flags[i] = static_cast<unsigned*>(flags ? flags[i] : nullptr);
    }
    c->rank = rank;
    c->world = world;
    c->bufs_n = 0;
    c->wire_bf16 = 0;
    c->shard_master = 0;
    PICD_CUDA(cudaMalloc(&c->counters, sizeof(unsigned) * 2 * (num_buckets + 1)));      // [barrier-out | pack] tickets per bucket
    PICD_CUDA(cudaMemset(c->counters, 0, sizeof(unsigned) * 2 * (num_buckets + 1)));
    c->num_buckets = num_buckets;
    *ctx_out = c;
    return 0;
}
// Registers the NVSwitch multicast views of the gradient / master / shadow regions (enables mode 3 of the bucket kernel).
extern "C" int picd_comm_set_multicast(void* ctx, void* mc_grad, void* mc_master, void* mc_shadow) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    c->pt.mc_grad = static_cast<float*>(mc_grad);
    c->pt.mc_master = static_cast<float*>(mc_master);
    c->pt.mc_shadow = static_cast<__nv_bfloat16*>(mc_shadow);
    return 0;
}
// Registers the per-rank module-buffer regions (n floats each, a multiple of 4, inside the peer-mapped slabs): a bucket launch
// with bcast_buffers != 0 then copies rank 0's region into every peer's.
extern "C" int picd_comm_set_buffers(void* ctx, void* const* bufs, long long n) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(n >= 0 && (n % 4) == 0, "comm_set_buffers: n=%lld must be a non-negative multiple of 4", n);
    for (int i = 0; i < c->world; ++i) c->pt.bufs[i] = static_cast<float*>(bufs[i]);
    c->bufs_n = n;
    return 0;
}
// Wire format of the gradient exchange.  gbf: per-rank bf16 staging regions (same element offsets as the gradient arenas), mc_gbf
// their NVSwitch multicast view or null; with them registered and wire_bf16 != 0 peers read 2 B per gradient element instead
// of 4.  shard_master != 0: in the two-shot / NVLS modes the slice owner no longer pushes its fp32 master slice (peers' copies
// of it go stale; the host gathers them when it needs a full state dict), only the bf16 compute copy: 2 B instead of 6.
extern "C" int picd_comm_set_wire(void* ctx, void* const* gbf, void* mc_gbf, int wire_bf16, int shard_master) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(!wire_bf16 || gbf != nullptr, "comm_set_wire: bf16 wire format needs the staging regions");
    for (int i = 0; i < c->world; ++i) c->pt.gbf[i] = gbf ? static_cast<__nv_bfloat16*>(gbf[i]) : nullptr;
    c->pt.mc_gbf = static_cast<__nv_bfloat16*>(mc_gbf);
    c->wire_bf16 = wire_bf16;
    c->shard_master = shard_master;
    return 0;
}
// Upper bound, in seconds, on how long a kernel spins for a peer's flag before it traps (default 600; env PICD_FLAG_TIMEOUT_S
// is read by the Python engine and passed here).  Applies to every communicator of the current device.
extern "C" int picd_comm_set_flag_timeout(double seconds) {
    PICD_REQUIRE(seconds > 0.0, "comm_set_flag_timeout: %f s", seconds);
    const long long ns = static_cast<long long>(seconds * 1e9);
    PICD_CUDA(cudaMemcpyToSymbol(g_flag_timeout_ns, &ns, sizeof(ns)));
    return 0;
}
extern "C" int picd_comm_destroy(void* ctx) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    if (c) {
        cudaFree(c->counters);
        delete c;
    }
    return 0;
}

// One gradient bucket: all-reduce over peer memory + 1/W + optimizer + bf16 shadow + gradient re-zero.
// mode: 0 local, 1 one-shot, 2 two-shot.  `epoch` must increase by one per call per bucket on every rank.
static int fused_update_impl(void* ctx, int bucket, int mode, long long start, long long count,
                             float* state0, float* state1, int kind, float lr, float momentum,
                             float beta1, float beta2, float eps, float weight_decay, float bc1,
                             float bc2, int nesterov, unsigned epoch, int blocks, const float* dyn, int bcast_buffers,
                             cudaStream_t stream) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(!bcast_buffers || c->bufs_n == 0 || c->pt.bufs[0] != nullptr, "fused_allreduce_update: buffers not registered");
    PICD_REQUIRE(count % (4 * c->world) == 0, "fused_allreduce_update: bucket of %lld elements does not split into "
                 "16-byte slices over %d ranks", count, c->world);
    if (c->world == 1) mode = 0;
    OptArgs o;
    o.kind = kind; o.lr = lr; o.momentum = momentum; o.beta1 = beta1; o.beta2 = beta2; o.eps = eps;
    o.weight_decay = weight_decay; o.bc1 = bc1; o.bc2 = bc2; o.nesterov = nesterov;
    o.grad_scale = 1.f / c->world;
    PICD_REQUIRE(mode < 3 || c->pt.mc_grad != nullptr, "fused_allreduce_update: NVLS mode without a multicast mapping");
    PICD_REQUIRE(!c->wire_bf16 || count % 8 == 0, "fused_allreduce_update: bf16 wire format needs count %% 8 == 0 (%lld)", count);
    PICD_REQUIRE(!(c->wire_bf16 && mode >= 3) || c->pt.mc_gbf != nullptr, "fused_allreduce_update: NVLS + bf16 wire without a multicast staging view");
    const long long work_vecs = (mode >= 2 ? count / c->world : count) / 4;
    long long want = (work_vecs + 511) / 512;
    // Peer kernels spin on flags, so the grid must be able to become co-resident: at most 2 CTAs of 512 threads per SM.
    // Small buckets (the usual <= 25 MiB, overlapped with backward) use few CTAs to leave the SMs to the wgrads.
    // (a CTA of this kernel holds 59K registers, so it evicts a persistent conv CTA from its SM: measured +2 ms/step
    //  on one GPU when the local update used 4 CTAs per SM.)
    if (blocks <= 0) blocks = count * 4 >= (32LL << 20) ? 2 * num_sms() : (count * 4 >= (4LL << 20) ? 64 : 32);
    if (want < blocks) blocks = static_cast<int>(want < 1 ? 1 : want);
    if (mode != 0) {
        // CTAs of this grid wait (through the peers) on each other, so the whole grid must fit on the GPU at once
        static int max_resident = 0;
        if (max_resident == 0) {
            int per_sm = 0;
            PICD_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_allreduce_update_kernel, 512, 0));
            max_resident = (per_sm > 0 ? per_sm : 1) * num_sms();
        }
        if (blocks > max_resident) blocks = max_resident;
    }
    fused_allreduce_update_kernel<<<blocks, 512, 0, stream>>>(c->pt, c->rank, c->world, mode, start, count, state0, state1, o,
                                                              2 * bucket, 2 * bucket + 1, epoch, c->counters + bucket, dyn,
                                                              (bcast_buffers && c->world > 1) ? c->bufs_n : 0LL,
                                                              c->wire_bf16, c->shard_master,
                                                              c->counters + c->num_buckets + 1 + bucket);
    PICD_LAUNCH_CHECK();
    return 0;
}

extern "C" int picd_fused_allreduce_update(void* ctx, int bucket, int mode, long long start, long long count,
                                           float* state0, float* state1, int kind, float lr, float momentum,
                                           float beta1, float beta2, float eps, float weight_decay, float bc1,
                                           float bc2, int nesterov, unsigned epoch, int blocks, int bcast_buffers,
                                           cudaStream_t stream) {
    return fused_update_impl(ctx, bucket, mode, start, count, state0, state1, kind, lr, momentum, beta1, beta2, eps,
                             weight_decay, bc1, bc2, nesterov, epoch, blocks, nullptr, bcast_buffers, stream);
}

// Graph-replay variant: lr, the two Adam bias corrections and the flag epoch are read from the device array
// dyn[4] = {lr, bc1, bc2, epoch bits} at run time instead of being frozen into the launch.
extern "C" int picd_fused_allreduce_update_dyn(void* ctx, int bucket, int mode, long long start, long long count,
                                               float* state0, float* state1, int kind, float momentum, float beta1,
                                               float beta2, float eps, float weight_decay, int nesterov, int blocks,
                                               const float* dyn, int bcast_buffers, cudaStream_t stream) {
    PICD_REQUIRE(dyn != nullptr, "fused_allreduce_update_dyn: dyn must point at 4 device floats");
    return fused_update_impl(ctx, bucket, mode, start, count, state0, state1, kind, 0.f, momentum, beta1, beta2, eps,
                             weight_decay, 1.f, 1.f, nesterov, 0u, blocks, dyn, bcast_buffers, stream);
}

// First kernel of a captured training step: advances the step-varying words of the dyn array ON THE DEVICE, so a replay needs
// no host-to-device copy at all.  dyn = {lr, 1 - beta1^t, 1 - beta2^t, t (flag epoch / dropout step, raw bits), replay index
// (raw bits)}: t and the replay index are incremented, the two Adam bias corrections recomputed from t; lr is only written by
// the host when the schedule changes it.
__global__ void dyn_advance_kernel(float* dyn, float beta1, float beta2) {
    if (threadIdx.x == 0) {
        const unsigned t = __float_as_uint(dyn[3]) + 1u;
        dyn[3] = __uint_as_float(t);
        dyn[4] = __uint_as_float(__float_as_uint(dyn[4]) + 1u);
        dyn[1] = 1.f - powf(beta1, static_cast<float>(t));
        dyn[2] = 1.f - powf(beta2, static_cast<float>(t));
    }
}
extern "C" int picd_dyn_advance(float* dyn, float beta1, float beta2, cudaStream_t stream) {
    PICD_REQUIRE(dyn != nullptr, "dyn_advance: dyn must point at 8 device floats");
    dyn_advance_kernel<<<1, 32, 0, stream>>>(dyn, beta1, beta2);
    PICD_LAUNCH_CHECK();
    return 0;
}

// slot_off: float offset of the slot HALF to use inside the exchange region; the half holds `world` sub-slots of `sub` floats
// (sub >= n, a multiple of 4), so it spans world * sub floats.  Launched with programmatic stream serialization.
extern "C" int picd_p2p_allreduce_small(void* ctx, const float* in, int in_replicas, float* out, int n,
                                        long long slot_off, int flag_slot, unsigned epoch, int sub, cudaStream_t stream) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(sub >= n, "p2p_allreduce_small: sub-slot of %d floats for a payload of %d", sub, n);
    PICD_CUDA(launch_pdl<PDL_BN>(p2p_allreduce_small_kernel, dim3(1), dim3(1024), 0, stream, c->pt, c->rank, c->world, in,
                                 in_replicas, out, n, slot_off, flag_slot, epoch, static_cast<const unsigned*>(nullptr), 0, sub));
    PICD_LAUNCH_CHECK();
    return 0;
}
// Graph-replay variants of the two small collectives: `slot_off` is the EVEN half of the slot, `epoch` the epoch at capture
// time; the kernel adds *dyn_delta (device memory) and picks the half from the resulting parity.
extern "C" int picd_p2p_allreduce_small_dyn(void* ctx, const float* in, int in_replicas, float* out, int n,
                                            long long slot_off, int flag_slot, unsigned epoch, const void* dyn_delta,
                                            int slot_floats, int sub, cudaStream_t stream) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(dyn_delta != nullptr, "p2p_allreduce_small_dyn: dyn_delta must point at a device word");
    PICD_REQUIRE(sub >= n, "p2p_allreduce_small_dyn: sub-slot of %d floats for a payload of %d", sub, n);
    PICD_CUDA(launch_pdl<PDL_BN>(p2p_allreduce_small_kernel, dim3(1), dim3(1024), 0, stream, c->pt, c->rank, c->world, in,
                                 in_replicas, out, n, slot_off, flag_slot, epoch, static_cast<const unsigned*>(dyn_delta),
                                 slot_floats, sub));
    PICD_LAUNCH_CHECK();
    return 0;
}

extern "C" int picd_bn_finalize_sync(void* ctx, const float* sums, int replicas, float count, const float* gamma, const float* beta,
                                     float* running_mean, float* running_var, float momentum, float eps, float* scale,
                                     float* shift, float* mean, float* invstd, float* total_count, int C,
                                     long long slot_off, int flag_slot, unsigned epoch, int sub, cudaStream_t stream) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(sub >= 2 * C + 1, "bn_finalize_sync: sub-slot of %d floats for %d channels", sub, C);
    PICD_CUDA(launch_pdl<PDL_BN>(bn_finalize_sync_kernel, dim3(1), dim3(1024), 0, stream, c->pt, c->rank, c->world, sums, replicas,
                                 count, gamma, beta, running_mean, running_var, momentum, eps, scale, shift, mean, invstd,
                                 total_count, C, slot_off, flag_slot, epoch, static_cast<const unsigned*>(nullptr), 0, sub));
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_bn_finalize_sync_dyn(void* ctx, const float* sums, int replicas, float count, const float* gamma,
                                         const float* beta, float* running_mean, float* running_var, float momentum,
                                         float eps, float* scale, float* shift, float* mean, float* invstd,
                                         float* total_count, int C, long long slot_off, int flag_slot, unsigned epoch,
                                         const void* dyn_delta, int slot_floats, int sub, cudaStream_t stream) {
    CommCtx* c = static_cast<CommCtx*>(ctx);
    PICD_REQUIRE(dyn_delta != nullptr, "bn_finalize_sync_dyn: dyn_delta must point at a device word");
    PICD_REQUIRE(sub >= 2 * C + 1, "bn_finalize_sync_dyn: sub-slot of %d floats for %d channels", sub, C);
    PICD_CUDA(launch_pdl<PDL_BN>(bn_finalize_sync_kernel, dim3(1), dim3(1024), 0, stream, c->pt, c->rank, c->world, sums, replicas,
                                 count, gamma, beta, running_mean, running_var, momentum, eps, scale, shift, mean, invstd,
                                 total_count, C, slot_off, flag_slot, epoch, static_cast<const unsigned*>(dyn_delta), slot_floats,
                                 sub));
    PICD_LAUNCH_CHECK();
    return 0;
}

// NVLS probe: does the device expose NVSwitch multicast objects (cuMulticastCreate & friends)?  The multicast MAPPING used by
// mode 3 / 4 of the bucket kernel is created by the host (picd_comm_set_multicast); this only reports the device capability.
// Returns 1 / 0, or -1 if unknown.
extern "C" int picd_nvls_supported() {
    typedef CUresult (*GetAttrFn)(int*, CUdevice_attribute, CUdevice);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &fn, cudaEnableDefault, &q) != cudaSuccess || fn == nullptr) return -1;
    int dev = 0, val = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (reinterpret_cast<GetAttrFn>(fn)(&val, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return -1;
    return val ? 1 : 0;
}
