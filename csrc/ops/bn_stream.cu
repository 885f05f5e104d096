// TMA-streamed BatchNorm kernels (apply, backward-reduce, backward-apply) for NHWC bf16 activations.
//
// The register-streaming versions in bn.cu top out at ~2.2-4.8 TB/s: a thread can only keep a handful of 16-byte loads
// in flight, and the loads of the next iteration are not issued while the current one computes.  Here the copy engine
// does the streaming: one producer thread issues 1-D bulk copies (cp.async.bulk, completion on an mbarrier) of 16 KB
// chunks of every input tensor into a multi-stage shared-memory ring, so 100-200 KB per SM are always in flight no
// matter what the consumers do; 16 consumer warps read the ring with conflict-free 16-byte shared loads.  A thread owns
// one vector of 8 channels for the whole kernel (C/8 divides the consumer count), so its per-channel coefficients stay
// in registers.  Persistent grid: one CTA per SM walks chunks blockIdx.x, blockIdx.x + grid, ...
#include "../common/picd_common.cuh"

namespace picd {

constexpr int SB_CONSUMERS = 512;
constexpr int SB_THREADS = SB_CONSUMERS + 64;     // + producer warp + store warp
constexpr int SB_MAX_CHUNK_BYTES = 32768;   // per tensor per stage: 32 KB when one tensor is streamed, 16 KB otherwise
constexpr int SB_MAX_STAGES = 8;

enum { SB_APPLY = 0, SB_BWD_REDUCE = 1, SB_BWD_APPLY = 2 };

struct StreamArgs {
    const __nv_bfloat16* in[3];      // APPLY: y, residual, -   BWD_*: dz, y, z
    int n_in;
    __nv_bfloat16* out0;             // APPLY: z              BWD_APPLY: dy
    __nv_bfloat16* out1;             // BWD_APPLY: dres (nullable)
    const float* scale;
    const float* shift;
    const float* mean;
    const float* invstd;
    const float* sums;               // BWD_APPLY: folded [2C]
    float* sums_out;                 // BWD_REDUCE: [STAT_REPLICAS][2C]
    const float* param_sums;         // BWD_APPLY: folded local sums for dgamma/dbeta
    float* dgamma;
    float* dbeta;
    float inv_count;
    const float* count_dev;          // BWD_APPLY, nullable: device-side global element count (SyncBN), replaces inv_count
    int act, use_z, eval_mode;
    long long P;
    int C, rows_per_chunk, stages, chunk_bytes;
    const uint8_t* mask;             // BWD_*: ReLU bit mask of a residual layer (one byte per 16-byte vector), replaces in[2]
    int use_mask, mask_chunk_bytes;  // bytes of mask per full chunk (rows_per_chunk * C/8, a multiple of 16)
    int stage_bytes;                 // n_in * chunk_bytes (+ 1 KB-rounded mask chunk)
    int n_active;                    // consumers that own a channel vector: the largest multiple of C/8 that is <= 512
    long long num_chunks;
};

__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, uint32_t smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(reinterpret_cast<uint64_t>(gdst)), "r"(smem_src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void unpack8(const uint4& raw, float (&f)[8]) {
    float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
__device__ __forceinline__ void ld8f(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}

template <int MODE>
__global__ void __launch_bounds__(SB_THREADS, 1) bn_stream_kernel(const StreamArgs a) {
    extern __shared__ uint8_t sb_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(sb_raw) + 127) & ~static_cast<uintptr_t>(127));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + SB_MAX_STAGES;
    uint64_t* done = empty + SB_MAX_STAGES;      // APPLY / BWD_APPLY: "results of this stage are in shared memory"
    uint8_t* ring = smem + 256;
    const int stages = a.stages;
    const int n_in = a.n_in;
    const int C = a.C;
    const int vecs = C >> 3;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            // outputs leave through bulk stores issued by the store warp, which then frees the stage on its own
            mbar_init(&empty[s], MODE == SB_BWD_REDUCE ? SB_CONSUMERS / 32 : 1);
            mbar_init(&done[s], SB_CONSUMERS / 32);
        }
        fence_mbar_init();
    }
    pdl_sync();        // the previous kernel (finalize / fold / the producer of the streamed tensors) has completed
    if (MODE == SB_BWD_APPLY && blockIdx.x == 0 && a.dgamma != nullptr) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            a.dbeta[c] += a.param_sums[c];
            a.dgamma[c] += a.param_sums[C + c];
        }
    }
    __syncthreads();

    if (warp == SB_CONSUMERS / 32) {
        // ---------------------------------------------------------------- producer
        if (lane == 0) {
            int s = 0;
            uint32_t phase = 0;
            for (long long c = blockIdx.x; c < a.num_chunks; c += gridDim.x) {
                const long long row0 = c * a.rows_per_chunk;
                long long rows = a.P - row0;
                if (rows > a.rows_per_chunk) rows = a.rows_per_chunk;
                const uint32_t bytes = static_cast<uint32_t>(rows * C * 2);
                mbar_wait(&empty[s], phase ^ 1);
                // mask bytes of this chunk, rounded up to the 16-byte bulk-copy granule (the mask buffer has slack at its end)
                const uint32_t mbytes = a.use_mask ? ((static_cast<uint32_t>(rows) * (C >> 3) + 15u) & ~15u) : 0u;
                mbar_arrive_expect_tx(&full[s], bytes * n_in + mbytes);
                uint8_t* stage = ring + static_cast<size_t>(s) * a.stage_bytes;
                for (int t = 0; t < n_in; ++t)
                    bulk_load(stage + static_cast<size_t>(t) * a.chunk_bytes, a.in[t] + row0 * C, bytes, &full[s]);
                if (a.use_mask)
                    bulk_load(stage + static_cast<size_t>(n_in) * a.chunk_bytes, a.mask + c * a.mask_chunk_bytes, mbytes, &full[s]);
                if (++s == stages) { s = 0; phase ^= 1; }
            }
        }
        return;
    }
    if (warp == SB_CONSUMERS / 32 + 1) {
        // ---------------------------------------------------------------- store warp
        // The consumers overwrite the input chunk in place with the results; one thread sends each finished stage to
        // global memory with 1-D bulk stores (full lines, no LSU traffic) and hands the stage back to the producer once
        // the copy engine has read it.
        if (MODE != SB_BWD_REDUCE && lane == 0) {
            int s = 0;
            uint32_t phase = 0;
            for (long long c = blockIdx.x; c < a.num_chunks; c += gridDim.x) {
                const long long row0 = c * a.rows_per_chunk;
                long long rows = a.P - row0;
                if (rows > a.rows_per_chunk) rows = a.rows_per_chunk;
                const uint32_t bytes = static_cast<uint32_t>(rows * C * 2);
                const uint32_t base = smem_u32(ring) + static_cast<uint32_t>(s) * a.stage_bytes;
                mbar_wait(&done[s], phase);
                bulk_store(a.out0 + row0 * C, base, bytes);
                if (MODE == SB_BWD_APPLY && a.out1 != nullptr) bulk_store(a.out1 + row0 * C, base + a.chunk_bytes, bytes);
                tma_store_commit();
                tma_store_wait_read();
                mbar_arrive(&empty[s]);
                if (++s == stages) { s = 0; phase ^= 1; }
            }
            tma_store_wait_all();
        }
        return;
    }

    // -------------------------------------------------------------------- consumers
    const int t = threadIdx.x;                 // 0..511
    const int cv = t % vecs;                   // the 8 channels this thread owns
    float sc[8], sh[8];
    ld8f(a.scale + cv * 8, sc);
    ld8f(a.shift + cv * 8, sh);
    float c1[8], c2[8];                        // BWD_REDUCE: mean, invstd;  BWD_APPLY: B, Cc of dy = sc*g + B*y + Cc
    float acc0[8], acc1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { c1[k] = 0.f; c2[k] = 0.f; acc0[k] = 0.f; acc1[k] = 0.f; }
    if (MODE == SB_BWD_REDUCE) {
        ld8f(a.mean + cv * 8, c1);
        ld8f(a.invstd + cv * 8, c2);
    } else if (MODE == SB_BWD_APPLY && !a.eval_mode) {
        float mu[8], is[8], s1[8], s2[8];
        ld8f(a.mean + cv * 8, mu); ld8f(a.invstd + cv * 8, is);
        ld8f(a.sums + cv * 8, s1); ld8f(a.sums + C + cv * 8, s2);
        const float inv_count = a.count_dev != nullptr ? 1.f / __ldg(a.count_dev) : a.inv_count;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            c1[k] = -sc[k] * is[k] * s2[k] * inv_count;
            c2[k] = -sc[k] * s1[k] * inv_count - c1[k] * mu[k];
        }
    }
    const uint32_t ring_u32 = smem_u32(ring);
    int s = 0;
    uint32_t phase = 0;
    for (long long c = blockIdx.x; c < a.num_chunks; c += gridDim.x) {
        const long long row0 = c * a.rows_per_chunk;
        long long rows = a.P - row0;
        if (rows > a.rows_per_chunk) rows = a.rows_per_chunk;
        const int nvec = static_cast<int>(rows) * vecs;
        mbar_wait(&full[s], phase);
        const uint32_t base = ring_u32 + static_cast<uint32_t>(s) * a.stage_bytes;
        const uint32_t cb = a.chunk_bytes;
        // vector idx of the chunk belongs to channel vector idx % vecs; stepping by n_active (a multiple of vecs) keeps every
        // thread on its own 8 channels for any C, not only powers of two
        const int n_active = a.n_active;
#pragma unroll 2
        for (int idx = t; idx < nvec; idx += n_active) {
            if (t < n_active) {
                if (MODE == SB_APPLY) {
                    float f[8], r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                    unpack8(lds128(base + idx * 16), f);
                    if (n_in > 1) unpack8(lds128(base + cb + idx * 16), r);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = apply_act(fmaf(f[j], sc[j], sh[j]) + r[j], a.act);
                    sts128(base + idx * 16, pack8(f));
                } else {
                    float g[8], yv[8];
                    unpack8(lds128(base + idx * 16), g);
                    unpack8(lds128(base + cb + idx * 16), yv);
                    if (a.act == ACT_RELU) {
                        if (a.use_mask) {
                            uint32_t mb;
                            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(mb) : "r"(base + n_in * cb + idx));
#pragma unroll
                            for (int j = 0; j < 8; ++j) g[j] = ((mb >> j) & 1u) ? g[j] : 0.f;
                        } else if (a.use_z) {
                            float zv[8];
                            unpack8(lds128(base + 2 * cb + idx * 16), zv);
#pragma unroll
                            for (int j = 0; j < 8; ++j) g[j] = zv[j] > 0.f ? g[j] : 0.f;
                        } else {
#pragma unroll
                            for (int j = 0; j < 8; ++j) g[j] = fmaf(yv[j], sc[j], sh[j]) > 0.f ? g[j] : 0.f;
                        }
                    } else if (a.act != ACT_NONE) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) g[j] *= act_grad(fmaf(yv[j], sc[j], sh[j]), a.act);
                    }
                    if (MODE == SB_BWD_REDUCE) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            acc0[j] += g[j];
                            acc1[j] = fmaf(g[j] * (yv[j] - c1[j]), c2[j], acc1[j]);
                        }
                    } else {
                        float d[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) d[j] = fmaf(sc[j], g[j], fmaf(c1[j], yv[j], c2[j]));
                        if (a.out1 != nullptr) sts128(base + cb + idx * 16, pack8(g));     // dres takes y's slot
                        sts128(base + idx * 16, pack8(d));                                  // dy takes dz's slot
                    }
                }
            }
        }
        if (MODE == SB_BWD_REDUCE) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        } else {
            fence_proxy_async_smem();          // the results were written through the generic proxy; the bulk store reads them
            __syncwarp();
            if (lane == 0) mbar_arrive(&done[s]);
        }
        if (++s == stages) { s = 0; phase ^= 1; }
    }

    if (MODE == SB_BWD_REDUCE) {
        // fold the threads that own the same channel vector (t, t + vecs, ...) through shared memory
        named_barrier_sync(2, SB_CONSUMERS);          // every consumer is done with the ring
        float* red = reinterpret_cast<float*>(ring);  // [512][16] floats = 32 KB
#pragma unroll
        for (int j = 0; j < 8; ++j) { red[t * 16 + j] = acc0[j]; red[t * 16 + 8 + j] = acc1[j]; }
        named_barrier_sync(2, SB_CONSUMERS);
        if (t < vecs) {
            float r0[8], r1[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { r0[j] = 0.f; r1[j] = 0.f; }
            for (int o = t; o < a.n_active; o += vecs)
#pragma unroll
                for (int j = 0; j < 8; ++j) { r0[j] += red[o * 16 + j]; r1[j] += red[o * 16 + 8 + j]; }
            float* rep = a.sums_out + static_cast<size_t>(blockIdx.x % STAT_REPLICAS) * 2 * C;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(rep + cv * 8 + j, r0[j]);
                atomicAdd(rep + C + cv * 8 + j, r1[j]);
            }
        }
    }
}

static bool stream_eligible(long long P, int C) {
    if (C < 64 || C > 2048 || (C % 8) != 0) return false;          // a thread owns one 16-byte vector; >= 2 rows per chunk
    return P * C * 2 >= (4LL << 20);                                // small tensors: launch-bound either way
}

template <int MODE>
static int launch_stream(StreamArgs& a, cudaStream_t stream) {
    static const int one_in_chunk = [] { const char* e = getenv("PICD_BN_CHUNK"); return e == nullptr ? SB_MAX_CHUNK_BYTES : atoi(e); }();
    // per-chunk handshakes bound the kernels, not bytes (two 16 KB tensors and three 16 KB tensors per chunk took the same
    // time): two streamed tensors get 32 KB chunks (3 stages), three keep 16 KB (4 stages)
    static const int two_in_chunk = [] { const char* e = getenv("PICD_BN_CHUNK_MULTI"); return e == nullptr ? 32768 : atoi(e); }();
    a.chunk_bytes = a.n_in == 1 ? one_in_chunk : (a.n_in == 2 ? two_in_chunk : 16384);
    a.rows_per_chunk = a.chunk_bytes / (a.C * 2);
    a.n_active = (SB_CONSUMERS / (a.C / 8)) * (a.C / 8);
    a.num_chunks = (a.P + a.rows_per_chunk - 1) / a.rows_per_chunk;
    a.mask_chunk_bytes = a.rows_per_chunk * (a.C / 8);
    if (a.use_mask && (a.mask_chunk_bytes % 16) != 0) return -1;       // bulk copies move 16-byte granules
    const int stage_bytes = a.n_in * a.chunk_bytes + (a.use_mask ? ((a.mask_chunk_bytes + 1023) & ~1023) : 0);
    a.stage_bytes = stage_bytes;
    int stages = (200 * 1024) / stage_bytes;
    if (stages > SB_MAX_STAGES) stages = SB_MAX_STAGES;
    a.stages = stages;
    int smem = stages * stage_bytes + 512;
    if (MODE == SB_BWD_REDUCE && smem < 33 * 1024) smem = 33 * 1024;
    static bool attr_set = false;
    if (!attr_set) {
        PICD_CUDA(cudaFuncSetAttribute(bn_stream_kernel<SB_APPLY>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        PICD_CUDA(cudaFuncSetAttribute(bn_stream_kernel<SB_BWD_REDUCE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        PICD_CUDA(cudaFuncSetAttribute(bn_stream_kernel<SB_BWD_APPLY>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    long long grid = num_sms();
    if (grid > a.num_chunks) grid = a.num_chunks;
    PICD_CUDA(launch_pdl<PDL_BN>(bn_stream_kernel<MODE>, dim3(static_cast<unsigned>(grid)), dim3(SB_THREADS), smem, stream, a));
    PICD_LAUNCH_CHECK();
    return 0;
}

// Each returns -1 when the shape is not eligible (caller falls back to the register-streaming kernel in bn.cu).
int bn_stream_apply(const void* y, const float* scale, const float* shift, const void* residual, int act, void* out,
                    long long P, int C, cudaStream_t stream) {
    if (!stream_eligible(P, C)) return -1;
    StreamArgs a{};
    a.in[0] = static_cast<const __nv_bfloat16*>(y);
    a.in[1] = static_cast<const __nv_bfloat16*>(residual);
    a.n_in = residual != nullptr ? 2 : 1;
    a.out0 = static_cast<__nv_bfloat16*>(out);
    a.scale = scale; a.shift = shift; a.act = act; a.P = P; a.C = C;
    return launch_stream<SB_APPLY>(a, stream);
}

// z / relu_mask: the ReLU mask source of a layer that added a residual before the activation -- the output tensor itself
// (a third streamed input) or, preferably, the bit mask written by the forward apply (1/16 of the bytes).
static void set_mask_source(StreamArgs& a, const void* z, const void* relu_mask, int act) {
    a.in[2] = static_cast<const __nv_bfloat16*>(z);
    a.mask = static_cast<const uint8_t*>(relu_mask);
    a.use_mask = (relu_mask != nullptr && act == ACT_RELU) ? 1 : 0;
    a.use_z = (!a.use_mask && z != nullptr && act == ACT_RELU) ? 1 : 0;
    a.n_in = a.use_z ? 3 : 2;
}

int bn_stream_bwd_reduce(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale,
                         const float* shift, const float* mean, const float* invstd, int act, float* sums, long long P,
                         int C, cudaStream_t stream) {
    if (!stream_eligible(P, C)) return -1;
    StreamArgs a{};
    a.in[0] = static_cast<const __nv_bfloat16*>(dz);
    a.in[1] = static_cast<const __nv_bfloat16*>(y);
    set_mask_source(a, z, relu_mask, act);
    a.scale = scale; a.shift = shift; a.mean = mean; a.invstd = invstd; a.act = act;
    a.sums_out = sums; a.P = P; a.C = C;
    return launch_stream<SB_BWD_REDUCE>(a, stream);
}

int bn_stream_bwd_apply(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale,
                        const float* shift, const float* mean, const float* invstd, int act, const float* sums,
                        float inv_count, const float* param_sums, float* dgamma, float* dbeta, void* dy, void* dres,
                        long long P, int C, int eval_mode, const float* count_dev, cudaStream_t stream) {
    if (!stream_eligible(P, C)) return -1;
    StreamArgs a{};
    a.in[0] = static_cast<const __nv_bfloat16*>(dz);
    a.in[1] = static_cast<const __nv_bfloat16*>(y);
    set_mask_source(a, z, relu_mask, act);
    a.out0 = static_cast<__nv_bfloat16*>(dy);
    a.out1 = static_cast<__nv_bfloat16*>(dres);
    a.scale = scale; a.shift = shift; a.mean = mean; a.invstd = invstd; a.act = act;
    a.sums = sums; a.inv_count = inv_count; a.param_sums = param_sums; a.dgamma = dgamma; a.dbeta = dbeta;
    a.eval_mode = eval_mode; a.count_dev = count_dev; a.P = P; a.C = C;
    return launch_stream<SB_BWD_APPLY>(a, stream);
}

}  // namespace picd
