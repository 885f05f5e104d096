// BatchNorm family for NHWC bf16 activations.
//
// The reference runs, per BN layer, batch_norm_stats -> (all_gather) -> gather_stats -> batch_norm_elemt -> ReLU
// forward and backward_reduce -> (all_reduce) -> backward_elemt (+ ReLU backward, + residual add) as separate
// ATen kernels.  Here:
//   forward : statistics come for free from the conv epilogue (conv_tc.cu) or from bn_stats;  bn_finalize turns
//             (sum, sumsq, count) into per-channel scale/shift and updates the running statistics;
//             bn_apply does y*scale+shift (+residual) (+activation) in ONE pass.
//   backward: bn_bwd_reduce folds the activation mask into the two reductions; bn_bwd_apply produces the conv
//             output gradient, the residual-branch gradient and the gamma/beta gradients in ONE pass.
// All kernels are bandwidth-bound: 16-byte vector accesses, grids sized to a few waves of 148 SMs.
#include "../common/picd_common.cuh"

#include <cstdlib>

namespace picd {

struct Vec8 {
    uint4 raw;
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
    __device__ __forceinline__ void unpack(float (&f)[8]) const {
        float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
        f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
    }
    __device__ __forceinline__ void pack(const float (&f)[8]) {
        raw.x = pack_bf16x2(f[0], f[1]); raw.y = pack_bf16x2(f[2], f[3]);
        raw.z = pack_bf16x2(f[4], f[5]); raw.w = pack_bf16x2(f[6], f[7]);
    }
};

// 2-D decomposition used by the column reductions: a block of 256 threads covers tx_count channel-vectors
// (8 channels each) by ty_count rows; consecutive threads read consecutive 16-byte vectors.
struct ReduceShape {
    int tx_count, ty_count, grid_x, grid_y, rows_per_block;
};
static ReduceShape reduce_shape(long long P, int C) {
    ReduceShape s;
    const int vecs = C / 8;
    s.tx_count = vecs < 256 ? vecs : 256;
    s.ty_count = 256 / s.tx_count;
    s.grid_x = (vecs + s.tx_count - 1) / s.tx_count;
    const int target_blocks = 4 * num_sms();
    long long gy = target_blocks / s.grid_x;
    if (gy < 1) gy = 1;
    long long rpb = (P + gy - 1) / gy;
    const long long min_rows = 4LL * s.ty_count;
    if (rpb < min_rows) rpb = min_rows;
    s.rows_per_block = static_cast<int>(rpb);
    s.grid_y = static_cast<int>((P + rpb - 1) / rpb);
    return s;
}

// Sum the 16 per-thread partials over the ty dimension and add them to two global [C] vectors.
__device__ __forceinline__ void block_column_reduce(float (&a)[8], float (&b)[8], int tx, int ty, int tx_count,
                                                    int ty_count, int cv, int vecs, float* out_a, float* out_b) {
    extern __shared__ float red_smem[];   // [ty_count][tx_count][16]
    const bool active = ty < ty_count;
    if (active) {
        float* slot = red_smem + (static_cast<size_t>(ty) * tx_count + tx) * 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) { slot[i] = a[i]; slot[8 + i] = b[i]; }
    }
    __syncthreads();
    if (active && ty == 0 && cv < vecs) {
        float sa[8], sb[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { sa[i] = 0.f; sb[i] = 0.f; }
        for (int y = 0; y < ty_count; ++y) {
            const float* slot = red_smem + (static_cast<size_t>(y) * tx_count + tx) * 16;
#pragma unroll
            for (int i = 0; i < 8; ++i) { sa[i] += slot[i]; sb[i] += slot[8 + i]; }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            atomicAdd(out_a + cv * 8 + i, sa[i]);
            atomicAdd(out_b + cv * 8 + i, sb[i]);
        }
    }
}

// ---------------------------------------------------------------------------------------------- stats
constexpr int ROW_UNROLL = 4;   // rows in flight per thread (memory-level parallelism of the streaming kernels)

__global__ void __launch_bounds__(256) bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ stats,
                                                       long long P, int C, int tx_count, int ty_count,
                                                       int rows_per_block) {
    pdl_sync();
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    float s[8], q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
    if (ty < ty_count && cv < vecs) {
        const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
        long long r1 = r0 + rows_per_block;
        if (r1 > P) r1 = P;
        for (long long r = r0 + ty; r < r1; r += static_cast<long long>(ty_count) * ROW_UNROLL) {
            Vec8 v[ROW_UNROLL];
#pragma unroll
            for (int u = 0; u < ROW_UNROLL; ++u) {
                const long long rr = r + static_cast<long long>(u) * ty_count;
                if (rr < r1) v[u].load(x + rr * C + cv * 8); else v[u].raw = make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < ROW_UNROLL; ++u) {
                float f[8];
                v[u].unpack(f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] += f[i] * f[i]; }
            }
        }
    }
    float* rep = stats + static_cast<size_t>(blockIdx.y % STAT_REPLICAS) * 2 * C;
    block_column_reduce(s, q, tx, ty, tx_count, ty_count, cv, vecs, rep, rep + C);
}

// ---------------------------------------------------------------------------------------------- finalize
// sums = [sum(C), sumsq(C)] over `count` elements per channel.  Produces scale = gamma*invstd and
// shift = beta - mean*scale for the fused apply, saves mean/invstd for backward and updates the running
// statistics with the unbiased variance (torch semantics, torch/nn/modules/_functions.py:106-115).
__device__ __forceinline__ float fold_replicas(const float* base, int idx, int stride, int replicas) {
    if (replicas == STAT_REPLICAS) {           // the usual case: all loads in flight at once instead of a dependent chain
        float v[STAT_REPLICAS];
#pragma unroll
        for (int r = 0; r < STAT_REPLICAS; ++r) v[r] = base[static_cast<size_t>(r) * stride + idx];
        float acc = 0.f;
#pragma unroll
        for (int r = 0; r < STAT_REPLICAS; ++r) acc += v[r];
        return acc;
    }
    float acc = 0.f;
    for (int r = 0; r < replicas; ++r) acc += base[static_cast<size_t>(r) * stride + idx];
    return acc;
}

__global__ void bn_finalize_kernel(const float* __restrict__ sums, int replicas, float count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* running_mean, float* running_var,
                                   float momentum, float eps, float* scale, float* shift, float* mean_out,
                                   float* invstd_out, int C) {
    pdl_sync();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float mean = fold_replicas(sums, c, 2 * C, replicas) / count;
    float var = fold_replicas(sums, C + c, 2 * C, replicas) / count - mean * mean;
    var = fmaxf(var, 0.f);
    const float invstd = rsqrtf(var + eps);
    const float sc = gamma[c] * invstd;
    scale[c] = sc;
    shift[c] = beta[c] - mean * sc;
    mean_out[c] = mean;
    invstd_out[c] = invstd;
    if (running_mean != nullptr) {
        const float unbiased = count > 1.f ? var * (count / (count - 1.f)) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
}

__global__ void bn_eval_coeffs_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ running_mean, const float* __restrict__ running_var,
                                      float eps, float* scale, float* shift, int C) {
    pdl_sync();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float sc = gamma[c] * rsqrtf(running_var[c] + eps);
    scale[c] = sc;
    shift[c] = beta[c] - running_mean[c] * sc;
}

// ---------------------------------------------------------------------------------------------- apply
// All streaming kernels below share one thread mapping: a thread owns ONE vector of 8 channels (its per-channel
// coefficients live in registers for the whole kernel) and walks over rows, ROW_UNROLL rows in flight.  Consecutive
// threads read consecutive 16-byte vectors, so a warp touches whole 128-byte lines.
__device__ __forceinline__ void load8f(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}

// TACT >= 0 fixes the activation at compile time (the generic apply_act switch drags exp/erf paths and ~50 registers along)
// Optional fused finalize: when fin.sums != nullptr every block first turns the (replicated) batch sums of ITS channel slab
// into scale / shift in shared memory -- 32 loads per thread -- and the blocks of the first row strip also publish scale,
// shift, mean, invstd (the backward needs them) and update the running statistics.  That removes the separate
// bn_finalize launch and one kernel-to-kernel dependency per BatchNorm layer.
struct BnFinalizeArgs {
    const float* sums;          // [replicas][2][C] or nullptr (then scale / shift are read from global memory)
    int replicas;
    float count, momentum, eps;
    const float* gamma;
    const float* beta;
    float* running_mean;
    float* running_var;
    float* mean_out;
    float* invstd_out;
};

template <int U, int MINB, int TACT>
__global__ void __launch_bounds__(256, MINB) bn_apply_kernel(const __nv_bfloat16* __restrict__ y,
                                                       float* __restrict__ scale,
                                                       float* __restrict__ shift,
                                                       const __nv_bfloat16* __restrict__ residual, int act,
                                                       __nv_bfloat16* __restrict__ out, long long P, int C,
                                                       int tx_count, int ty_count, int rows_per_block,
                                                       uint8_t* __restrict__ relu_mask, const BnFinalizeArgs fin) {
    pdl_sync();
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    __shared__ float s_coef[2][2048];         // scale / shift of the block's channel slab (tx_count * 8 <= 2048 channels)
    if (fin.sums != nullptr) {
        for (int cl = threadIdx.x; cl < tx_count * 8; cl += blockDim.x) {
            const int c = blockIdx.x * tx_count * 8 + cl;
            if (c >= C) break;
            const float mean = fold_replicas(fin.sums, c, 2 * C, fin.replicas) / fin.count;
            float var = fold_replicas(fin.sums, C + c, 2 * C, fin.replicas) / fin.count - mean * mean;
            var = fmaxf(var, 0.f);
            const float invstd = rsqrtf(var + fin.eps);
            const float scv = fin.gamma[c] * invstd;
            const float shv = fin.beta[c] - mean * scv;
            s_coef[0][cl] = scv;
            s_coef[1][cl] = shv;
            if (blockIdx.y == 0) {
                scale[c] = scv;
                shift[c] = shv;
                fin.mean_out[c] = mean;
                fin.invstd_out[c] = invstd;
                if (fin.running_mean != nullptr) {
                    const float unbiased = fin.count > 1.f ? var * (fin.count / (fin.count - 1.f)) : var;
                    fin.running_mean[c] = (1.f - fin.momentum) * fin.running_mean[c] + fin.momentum * mean;
                    fin.running_var[c] = (1.f - fin.momentum) * fin.running_var[c] + fin.momentum * unbiased;
                }
            }
        }
        __syncthreads();
    }
    if (ty >= ty_count || cv >= vecs) return;
    float sc[8], sh[8];
    if (fin.sums != nullptr) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { sc[k] = s_coef[0][tx * 8 + k]; sh[k] = s_coef[1][tx * 8 + k]; }
    } else {
        load8f(scale + cv * 8, sc);
        load8f(shift + cv * 8, sh);
    }
    const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
    long long r1 = r0 + rows_per_block;
    if (r1 > P) r1 = P;
    for (long long r = r0 + ty; r < r1; r += static_cast<long long>(ty_count) * U) {
        Vec8 v[U], rv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long rr = r + static_cast<long long>(u) * ty_count;
            if (rr < r1) {
                v[u].load(y + rr * C + cv * 8);
                if (residual != nullptr) rv[u].load(residual + rr * C + cv * 8);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long rr = r + static_cast<long long>(u) * ty_count;
            if (rr < r1) {
                float f[8], rs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                v[u].unpack(f);
                if (residual != nullptr) rv[u].unpack(rs);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float pre = fmaf(f[k], sc[k], sh[k]) + rs[k];
                    f[k] = TACT == ACT_NONE ? pre : (TACT == ACT_RELU ? fmaxf(pre, 0.f) : apply_act(pre, act));
                }
                Vec8 o;
                o.pack(f);
                o.store(out + rr * C + cv * 8);
                if (relu_mask != nullptr) {
                    // one bit per output: "the ReLU let it through".  The backward kernels of residual layers read this
                    // byte per 16-byte vector instead of streaming the whole output tensor again for its sign.
                    uint32_t bits = 0;
#pragma unroll
                    for (int k = 0; k < 8; ++k) bits |= (f[k] > 0.f ? 1u : 0u) << k;
                    relu_mask[rr * vecs + cv] = static_cast<uint8_t>(bits);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------- backward
// g = dz * act'(pre-activation).  For ReLU the mask is read from the saved output z (z > 0) when a residual was
// added before the activation, otherwise it is recomputed from y*scale+shift (one tensor less to read).
__device__ __forceinline__ void masked_grad(float (&g)[8], const float (&yv)[8], const float (&zv)[8], bool has_z,
                                            const float (&sc)[8], const float (&sh)[8], int act) {
    if (act == ACT_NONE) return;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (act == ACT_RELU) {
            const float pre = has_z ? zv[k] : fmaf(yv[k], sc[k], sh[k]);
            g[k] = pre > 0.f ? g[k] : 0.f;
        } else {
            g[k] *= act_grad(fmaf(yv[k], sc[k], sh[k]), act);
        }
    }
}

__global__ void __launch_bounds__(256, 2) bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dz,
                                                            const __nv_bfloat16* __restrict__ y,
                                                            const __nv_bfloat16* __restrict__ z,
                                                            const float* __restrict__ scale,
                                                            const float* __restrict__ shift,
                                                            const float* __restrict__ mean,
                                                            const float* __restrict__ invstd, int act,
                                                            float* __restrict__ sums, long long P, int C, int tx_count,
                                                            int ty_count, int rows_per_block) {
    pdl_sync();
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    float sg[8], sgx[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { sg[i] = 0.f; sgx[i] = 0.f; }
    if (ty < ty_count && cv < vecs) {
        float sc[8], sh[8], mu[8], is[8];
        load8f(scale + cv * 8, sc); load8f(shift + cv * 8, sh); load8f(mean + cv * 8, mu); load8f(invstd + cv * 8, is);
        const bool use_z = (z != nullptr) && (act == ACT_RELU);
        const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
        long long r1 = r0 + rows_per_block;
        if (r1 > P) r1 = P;
        for (long long r = r0 + ty; r < r1; r += static_cast<long long>(ty_count) * ROW_UNROLL) {
            Vec8 vd[ROW_UNROLL], vy[ROW_UNROLL], vz[ROW_UNROLL];
#pragma unroll
            for (int u = 0; u < ROW_UNROLL; ++u) {
                const long long rr = r + static_cast<long long>(u) * ty_count;
                if (rr < r1) {
                    const long long off = rr * C + cv * 8;
                    vd[u].load(dz + off);
                    vy[u].load(y + off);
                    if (use_z) vz[u].load(z + off);
                }
            }
#pragma unroll
            for (int u = 0; u < ROW_UNROLL; ++u) {
                const long long rr = r + static_cast<long long>(u) * ty_count;
                if (rr < r1) {
                    float g[8], yv[8], zv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                    vd[u].unpack(g);
                    vy[u].unpack(yv);
                    if (use_z) vz[u].unpack(zv);
                    masked_grad(g, yv, zv, use_z, sc, sh, act);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        sg[i] += g[i];
                        sgx[i] += g[i] * (yv[i] - mu[i]) * is[i];
                    }
                }
            }
        }
    }
    float* rep = sums + static_cast<size_t>(blockIdx.y % STAT_REPLICAS) * 2 * C;
    block_column_reduce(sg, sgx, tx, ty, tx_count, ty_count, cv, vecs, rep, rep + C);
}

// dy = gamma*invstd * (g - mean(g) - xhat * mean(g*xhat)) = A*g + B*y + Cc per channel;  dres = g (gradient of the
// residual branch).  Block (0,0) also accumulates the parameter gradients dgamma += sum(g*xhat), dbeta += sum(g) from
// `param_sums` (the LOCAL sums -- with SyncBN `sums` holds the cross-replica totals, `param_sums` this replica's).
// `eval_mode`: the statistics were frozen (running stats), so dy = scale * g.
__global__ void __launch_bounds__(256, 2) bn_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dz, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ z,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
    const float* __restrict__ invstd, int act, const float* __restrict__ sums, float inv_count,
    const float* __restrict__ param_sums, float* dgamma, float* dbeta, __nv_bfloat16* __restrict__ dy,
    __nv_bfloat16* __restrict__ dres, long long P, int C, int tx_count, int ty_count, int rows_per_block,
    int eval_mode, int sums_replicas, int param_replicas, const float* __restrict__ count_dev) {
    pdl_sync();
    const int vecs = C >> 3;
    if (count_dev != nullptr) inv_count = 1.f / __ldg(count_dev);     // SyncBN: the exchanged global element count
    if (blockIdx.x == 0 && blockIdx.y == 0 && dgamma != nullptr) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            dbeta[c] += fold_replicas(param_sums, c, 2 * C, param_replicas);
            dgamma[c] += fold_replicas(param_sums, C + c, 2 * C, param_replicas);
        }
    }
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    if (ty >= ty_count || cv >= vecs) return;
    float sc[8], sh[8], cb[8], cc[8];
    load8f(scale + cv * 8, sc);
    load8f(shift + cv * 8, sh);
    if (!eval_mode) {
        float mu[8], is[8], s1[8], s2[8];
        load8f(mean + cv * 8, mu); load8f(invstd + cv * 8, is);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s1[k] = fold_replicas(sums, cv * 8 + k, 2 * C, sums_replicas);
            s2[k] = fold_replicas(sums, C + cv * 8 + k, 2 * C, sums_replicas);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            cb[k] = -sc[k] * is[k] * s2[k] * inv_count;
            cc[k] = -sc[k] * s1[k] * inv_count - cb[k] * mu[k];
        }
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) { cb[k] = 0.f; cc[k] = 0.f; }
    }
    const bool use_z = (z != nullptr) && (act == ACT_RELU);
    const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
    long long r1 = r0 + rows_per_block;
    if (r1 > P) r1 = P;
    for (long long r = r0 + ty; r < r1; r += static_cast<long long>(ty_count) * ROW_UNROLL) {
        Vec8 vd[ROW_UNROLL], vy[ROW_UNROLL], vz[ROW_UNROLL];
#pragma unroll
        for (int u = 0; u < ROW_UNROLL; ++u) {
            const long long rr = r + static_cast<long long>(u) * ty_count;
            if (rr < r1) {
                const long long off = rr * C + cv * 8;
                vd[u].load(dz + off);
                vy[u].load(y + off);
                if (use_z) vz[u].load(z + off);
            }
        }
#pragma unroll
        for (int u = 0; u < ROW_UNROLL; ++u) {
            const long long rr = r + static_cast<long long>(u) * ty_count;
            if (rr < r1) {
                const long long off = rr * C + cv * 8;
                float g[8], yv[8], zv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                vd[u].unpack(g);
                vy[u].unpack(yv);
                if (use_z) vz[u].unpack(zv);
                masked_grad(g, yv, zv, use_z, sc, sh, act);
                if (dres != nullptr) {
                    Vec8 o;
                    o.pack(g);
                    o.store(dres + off);
                }
                float d[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] = fmaf(sc[k], g[k], fmaf(cb[k], yv[k], cc[k]));
                Vec8 o;
                o.pack(d);
                o.store(dy + off);
            }
        }
    }
}

// grid/block shape for the streaming (non-reducing) kernels: more blocks than the reductions, no smem needed
static ReduceShape stream_shape(long long P, int C) {
    ReduceShape s = reduce_shape(P, C);
    const int target_blocks = 8 * num_sms();
    long long gy = target_blocks / s.grid_x;
    if (gy < 1) gy = 1;
    long long rpb = (P + gy - 1) / gy;
    const long long min_rows = static_cast<long long>(ROW_UNROLL) * s.ty_count;
    if (rpb < min_rows) rpb = min_rows;
    s.rows_per_block = static_cast<int>(rpb);
    s.grid_y = static_cast<int>((P + rpb - 1) / rpb);
    return s;
}

// TMA-streamed variants (bn_stream.cu); each returns -1 when the shape is not eligible.
int bn_stream_apply(const void* y, const float* scale, const float* shift, const void* residual, int act, void* out,
                    long long P, int C, cudaStream_t stream);
int bn_stream_bwd_reduce(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale, const float* shift,
                         const float* mean, const float* invstd, int act, float* sums, long long P, int C,
                         cudaStream_t stream);
int bn_stream_bwd_apply(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale, const float* shift,
                        const float* mean, const float* invstd, int act, const float* sums, float inv_count,
                        const float* param_sums, float* dgamma, float* dbeta, void* dy, void* dres, long long P, int C,
                        int eval_mode, const float* count_dev, cudaStream_t stream);
static bool use_stream_kernels() {
    static const bool on = [] { const char* e = getenv("PICD_BN_STREAM"); return e == nullptr || e[0] != '0'; }();
    return on;
}

}  // namespace picd

using namespace picd;

extern "C" int picd_bn_stats(const void* x, float* stats, long long P, int C, cudaStream_t stream) {
    PICD_REQUIRE(C % 8 == 0, "bn_stats: C=%d must be a multiple of 8", C);
    const ReduceShape s = reduce_shape(P, C);
    const size_t smem = static_cast<size_t>(s.tx_count) * s.ty_count * 16 * sizeof(float);
    PICD_CUDA(launch_pdl<PDL_BN>(bn_stats_kernel, dim3(s.grid_x, s.grid_y), dim3(256), smem, stream,
                         static_cast<const __nv_bfloat16*>(x), stats, P, C, s.tx_count, s.ty_count, s.rows_per_block));
    PICD_LAUNCH_CHECK();
    return 0;
}

extern "C" int picd_stat_replicas() { return STAT_REPLICAS; }

// sums: `replicas` copies of [sum(C), sumsq(C)] (1 for an already folded buffer, STAT_REPLICAS straight from a kernel)
extern "C" int picd_bn_finalize(const float* sums, int replicas, float count, const float* gamma, const float* beta,
                                float* running_mean, float* running_var, float momentum, float eps, float* scale,
                                float* shift, float* mean, float* invstd, int C, cudaStream_t stream) {
    PICD_CUDA(launch_pdl<PDL_BN>(bn_finalize_kernel, dim3((C + 127) / 128), dim3(128), 0, stream, sums, replicas, count, gamma, beta,
                         running_mean, running_var, momentum, eps, scale, shift, mean, invstd, C));
    PICD_LAUNCH_CHECK();
    return 0;
}

extern "C" int picd_bn_eval_coeffs(const float* gamma, const float* beta, const float* running_mean,
                                   const float* running_var, float eps, float* scale, float* shift, int C,
                                   cudaStream_t stream) {
    PICD_CUDA(launch_pdl<PDL_BN>(bn_eval_coeffs_kernel, dim3((C + 127) / 128), dim3(128), 0, stream, gamma, beta, running_mean,
                         running_var, eps, scale, shift, C));
    PICD_LAUNCH_CHECK();
    return 0;
}

// relu_mask (optional, [P * C/8] bytes + one chunk of slack): bit k of byte v = output 8*v+k is positive.
static int bn_apply_impl(const void* y, float* scale, float* shift, const void* residual, int act, void* out,
                         void* relu_mask, long long P, int C, const BnFinalizeArgs& fin, cudaStream_t stream) {
    PICD_REQUIRE(C % 8 == 0, "bn_apply: C=%d must be a multiple of 8", C);
    // The TMA-streamed forward apply measured SLOWER (2.5 TB/s) than this register-streaming kernel (4.8 TB/s) --
    // one input tensor gives the ring too little work per handshake -- so it is opt-in (PICD_BN_STREAM_APPLY=1).
    static const bool stream_apply = [] { const char* e = getenv("PICD_BN_STREAM_APPLY"); return e != nullptr && e[0] == '1'; }();
    if (stream_apply && use_stream_kernels() && relu_mask == nullptr && fin.sums == nullptr) {
        const int rc = bn_stream_apply(y, scale, shift, residual, act, out, P, C, stream);
        if (rc >= 0) return rc;
    }
    static const int variant = [] { const char* e = getenv("PICD_BN_APPLY_VARIANT"); return e == nullptr ? 2 : atoi(e); }();   // 2 = 8 rows in flight (measured best)
    ReduceShape s = stream_shape(P, C);
#define PICD_BN_APPLY_LAUNCH_A(U, MINB, A)                                                                         \
    PICD_CUDA(launch_pdl<PDL_BN>(bn_apply_kernel<U, MINB, A>, dim3(s.grid_x, s.grid_y), dim3(256), 0, stream,               \
        static_cast<const __nv_bfloat16*>(y), scale, shift, static_cast<const __nv_bfloat16*>(residual), act,       \
        static_cast<__nv_bfloat16*>(out), P, C, s.tx_count, s.ty_count, s.rows_per_block,                           \
        static_cast<uint8_t*>(relu_mask), fin))
#define PICD_BN_APPLY_LAUNCH(U, MINB)                                                                              \
    do {                                                                                                            \
        if (act == ACT_RELU) PICD_BN_APPLY_LAUNCH_A(U, MINB, ACT_RELU);                                             \
        else if (act == ACT_NONE) PICD_BN_APPLY_LAUNCH_A(U, MINB, ACT_NONE);                                        \
        else PICD_BN_APPLY_LAUNCH_A(U, MINB, -1);                                                                   \
    } while (0)
    if (variant == 1) PICD_BN_APPLY_LAUNCH(4, 4);
    else if (variant == 2) {
        const long long min_rows = 8LL * s.ty_count;
        if (s.rows_per_block < min_rows) { s.rows_per_block = static_cast<int>(min_rows); s.grid_y = static_cast<int>((P + min_rows - 1) / min_rows); }
        PICD_BN_APPLY_LAUNCH(8, 2);
    } else if (variant == 3) PICD_BN_APPLY_LAUNCH(2, 6);
    else PICD_BN_APPLY_LAUNCH(4, 2);
#undef PICD_BN_APPLY_LAUNCH_A
#undef PICD_BN_APPLY_LAUNCH
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_bn_apply_mask(const void* y, const float* scale, const float* shift, const void* residual, int act,
                                  void* out, void* relu_mask, long long P, int C, cudaStream_t stream) {
    BnFinalizeArgs fin{};
    return bn_apply_impl(y, const_cast<float*>(scale), const_cast<float*>(shift), residual, act, out, relu_mask, P, C, fin, stream);
}
extern "C" int picd_bn_apply(const void* y, const float* scale, const float* shift, const void* residual, int act,
                             void* out, long long P, int C, cudaStream_t stream) {
    return picd_bn_apply_mask(y, scale, shift, residual, act, out, nullptr, P, C, stream);
}
// picd_bn_finalize + picd_bn_apply_mask in ONE launch: sums = `replicas` copies of [sum(C), sumsq(C)] over `count` elements
// per channel; scale / shift / mean / invstd are outputs (kept for the backward), running statistics are updated.
extern "C" int picd_bn_finalize_apply(const void* y, const float* sums, int replicas, float count, const float* gamma,
                                      const float* beta, float* running_mean, float* running_var, float momentum, float eps,
                                      float* scale, float* shift, float* mean, float* invstd, const void* residual, int act,
                                      void* out, void* relu_mask, long long P, int C, cudaStream_t stream) {
    PICD_REQUIRE(sums != nullptr, "bn_finalize_apply: sums must not be null");
    BnFinalizeArgs fin{};
    fin.sums = sums; fin.replicas = replicas; fin.count = count; fin.momentum = momentum; fin.eps = eps;
    fin.gamma = gamma; fin.beta = beta; fin.running_mean = running_mean; fin.running_var = running_var;
    fin.mean_out = mean; fin.invstd_out = invstd;
    return bn_apply_impl(y, scale, shift, residual, act, out, relu_mask, P, C, fin, stream);
}

// relu_mask (optional): the bit mask written by picd_bn_apply_mask; the streamed kernels read it instead of z.
extern "C" int picd_bn_bwd_reduce_mask(const void* dz, const void* y, const void* z, const void* relu_mask,
                                       const float* scale, const float* shift, const float* mean, const float* invstd,
                                       int act, float* sums, long long P, int C, cudaStream_t stream) {
    PICD_REQUIRE(C % 8 == 0, "bn_bwd_reduce: C=%d must be a multiple of 8", C);
    if (use_stream_kernels()) {
        const int rc = bn_stream_bwd_reduce(dz, y, z, relu_mask, scale, shift, mean, invstd, act, sums, P, C, stream);
        if (rc >= 0) return rc;
    }
    const ReduceShape s = reduce_shape(P, C);
    const size_t smem = static_cast<size_t>(s.tx_count) * s.ty_count * 16 * sizeof(float);
    PICD_CUDA(launch_pdl<PDL_BN>(bn_bwd_reduce_kernel, dim3(s.grid_x, s.grid_y), dim3(256), smem, stream,
        static_cast<const __nv_bfloat16*>(dz), static_cast<const __nv_bfloat16*>(y),
        static_cast<const __nv_bfloat16*>(z), scale, shift, mean, invstd, act, sums, P, C, s.tx_count, s.ty_count,
        s.rows_per_block));
    PICD_LAUNCH_CHECK();
    return 0;
}

extern "C" int picd_bn_bwd_reduce(const void* dz, const void* y, const void* z, const float* scale,
                                  const float* shift, const float* mean, const float* invstd, int act, float* sums,
                                  long long P, int C, cudaStream_t stream) {
    return picd_bn_bwd_reduce_mask(dz, y, z, nullptr, scale, shift, mean, invstd, act, sums, P, C, stream);
}

// In-place fold of `replicas` copies of n floats into copy 0.
__global__ void fold_replicas_kernel(float* buf, int n, int replicas) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    buf[i] = fold_replicas(buf, i, n, replicas);
}

// NOTE: folds `sums` / `param_sums` in place (copy 0 receives the totals) before the streaming kernel runs.
// count_dev (nullable): device pointer to the GLOBAL per-channel element count; when given it replaces inv_count.  SyncBatchNorm's
// forward exchange publishes the true total, so ranks with different batch sizes (a ragged last batch) scale dX correctly
// (torch/nn/modules/_functions.py:59-62,170-179).
extern "C" int picd_bn_bwd_apply_mask_cnt(const void* dz, const void* y, const void* z, const void* relu_mask,
                                          const float* scale, const float* shift, const float* mean, const float* invstd,
                                          int act, const float* sums, int sums_replicas, float inv_count,
                                          const float* count_dev, const float* param_sums, int param_replicas,
                                          float* dgamma, float* dbeta, void* dy, void* dres, long long P, int C,
                                          cudaStream_t stream) {
    PICD_REQUIRE(C % 8 == 0, "bn_bwd_apply: C=%d must be a multiple of 8", C);
    if (param_sums != nullptr && param_sums != sums && param_replicas > 1) {
        PICD_CUDA(launch_pdl<PDL_BN>(fold_replicas_kernel, dim3((2 * C + 255) / 256), dim3(256), 0, stream, const_cast<float*>(param_sums),
                             2 * C, param_replicas));
        param_replicas = 1;
    }
    if (sums_replicas > 1) {
        PICD_CUDA(launch_pdl<PDL_BN>(fold_replicas_kernel, dim3((2 * C + 255) / 256), dim3(256), 0, stream, const_cast<float*>(sums), 2 * C,
                             sums_replicas));
        if (param_sums == sums) param_replicas = 1;
        sums_replicas = 1;
    }
    if (use_stream_kernels()) {
        const int rc = bn_stream_bwd_apply(dz, y, z, relu_mask, scale, shift, mean, invstd, act, sums, inv_count, param_sums,
                                           dgamma, dbeta, dy, dres, P, C, 0, count_dev, stream);
        if (rc >= 0) return rc;
    }
    const ReduceShape s = stream_shape(P, C);
    PICD_CUDA(launch_pdl<PDL_BN>(bn_bwd_apply_kernel, dim3(s.grid_x, s.grid_y), dim3(256), 0, stream,
        static_cast<const __nv_bfloat16*>(dz), static_cast<const __nv_bfloat16*>(y),
        static_cast<const __nv_bfloat16*>(z), scale, shift, mean, invstd, act, sums, inv_count, param_sums, dgamma,
        dbeta, static_cast<__nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dres), P, C, s.tx_count, s.ty_count,
        s.rows_per_block, 0, sums_replicas, param_replicas, count_dev));
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_bn_bwd_apply_mask(const void* dz, const void* y, const void* z, const void* relu_mask,
                                      const float* scale, const float* shift, const float* mean, const float* invstd,
                                      int act, const float* sums, int sums_replicas, float inv_count,
                                      const float* param_sums, int param_replicas, float* dgamma, float* dbeta, void* dy,
                                      void* dres, long long P, int C, cudaStream_t stream) {
    return picd_bn_bwd_apply_mask_cnt(dz, y, z, relu_mask, scale, shift, mean, invstd, act, sums, sums_replicas, inv_count,
                                      nullptr, param_sums, param_replicas, dgamma, dbeta, dy, dres, P, C, stream);
}
extern "C" int picd_bn_bwd_apply(const void* dz, const void* y, const void* z, const float* scale,
                                 const float* shift, const float* mean, const float* invstd, int act,
                                 const float* sums, int sums_replicas, float inv_count, const float* param_sums,
                                 int param_replicas, float* dgamma, float* dbeta, void* dy, void* dres, long long P,
                                 int C, cudaStream_t stream) {
    return picd_bn_bwd_apply_mask(dz, y, z, nullptr, scale, shift, mean, invstd, act, sums, sums_replicas, inv_count, param_sums,
                                  param_replicas, dgamma, dbeta, dy, dres, P, C, stream);
}

extern "C" int picd_bn_bwd_eval(const void* dz, const void* y, const void* z, const float* scale, const float* shift,
                                int act, void* dy, void* dres, long long P, int C, cudaStream_t stream) {
    PICD_REQUIRE(C % 8 == 0, "bn_bwd_eval: C=%d must be a multiple of 8", C);
    if (use_stream_kernels()) {
        const int rc = bn_stream_bwd_apply(dz, y, z, nullptr, scale, shift, nullptr, nullptr, act, nullptr, 0.f, nullptr, nullptr,
                                           nullptr, dy, dres, P, C, 1, nullptr, stream);
        if (rc >= 0) return rc;
    }
    const ReduceShape s = stream_shape(P, C);
    bn_bwd_apply_kernel<<<dim3(s.grid_x, s.grid_y), 256, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(dz), static_cast<const __nv_bfloat16*>(y),
        static_cast<const __nv_bfloat16*>(z), scale, shift, nullptr, nullptr, act, nullptr, 0.f, nullptr, nullptr,
        nullptr, static_cast<__nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dres), P, C, s.tx_count, s.ty_count,
        s.rows_per_block, 1, 1, 1, nullptr);
    PICD_LAUNCH_CHECK();
    return 0;
}
