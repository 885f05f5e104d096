// Memory-bound helper kernels: pooling, global average pool, weighted softmax cross-entropy, activations,
// dropout, LayerNorm, small (SIMT) linear layers, first-layer packing, synthetic data.
// All activations are NHWC bf16; reductions accumulate in fp32.
#include "../common/picd_common.cuh"

#include <algorithm>

namespace picd {

static int grid_for(long long work, int block = 256, int waves = 8) {
    long long b = (work + block - 1) / block;
    const long long cap = static_cast<long long>(waves) * num_sms();
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

__device__ __forceinline__ void unpack8(const uint4& raw, float (&f)[8]) {
    float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}

// ------------------------------------------------------------------------------------------------ pooling
// mode 0 = max (stores the arg-max window offset as one byte per element), 1 = avg (count_include_pad).
__global__ void __launch_bounds__(256) pool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                       uint8_t* __restrict__ argmax, int N, int H, int W, int C, int OH,
                                                       int OW, int kh, int kw, int stride, int pad, int mode) {
    const int vecs = C >> 3;
    const long long total = static_cast<long long>(N) * OH * OW * vecs;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
        const int cv = static_cast<int>(i % vecs);
        long long t = i / vecs;
        const int ow = static_cast<int>(t % OW); t /= OW;
        const int oh = static_cast<int>(t % OH);
        const int n = static_cast<int>(t / OH);
        float acc[8];
        int arg[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { acc[k] = mode == 0 ? -INFINITY : 0.f; arg[k] = 0; }
        for (int a = 0; a < kh; ++a) {
            const int ih = oh * stride - pad + a;
            if (ih < 0 || ih >= H) continue;
            for (int b = 0; b < kw; ++b) {
                const int iw = ow * stride - pad + b;
                if (iw < 0 || iw >= W) continue;
                const uint4 raw = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * H + ih) * W + iw) * C + cv * 8);
                float f[8];
                unpack8(raw, f);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    if (mode == 0) {
                        if (f[k] > acc[k]) { acc[k] = f[k]; arg[k] = a * kw + b; }
                    } else {
                        acc[k] += f[k];
                    }
                }
            }
        }
        if (mode == 1) {
            const float inv = 1.f / (kh * kw);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] *= inv;
        }
        *reinterpret_cast<uint4*>(out + i * 8) = pack8(acc);
        if (mode == 0 && argmax != nullptr) {
            uint2 pk;
            pk.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
            pk.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
            *reinterpret_cast<uint2*>(argmax + i * 8) = pk;
        }
    }
}

// Gather formulation (one thread per INPUT vector, no atomics, deterministic): sum dy over the windows that
// contain this pixel; for max-pool only where the stored arg-max points back at it.
__global__ void __launch_bounds__(256) pool_bwd_kernel(const __nv_bfloat16* __restrict__ dy,
                                                       const uint8_t* __restrict__ argmax, __nv_bfloat16* __restrict__ dx,
                                                       int N, int H, int W, int C, int OH, int OW, int kh, int kw,
                                                       int stride, int pad, int mode) {
    const int vecs = C >> 3;
    const long long total = static_cast<long long>(N) * H * W * vecs;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    const float inv = 1.f / (kh * kw);
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
        const int cv = static_cast<int>(i % vecs);
        long long t = i / vecs;
        const int iw = static_cast<int>(t % W); t /= W;
        const int ih = static_cast<int>(t % H);
        const int n = static_cast<int>(t / H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // only window offsets congruent to (i + pad) mod stride can contain this pixel: step by `stride`
        for (int a = (ih + pad) % stride; a < kh; a += stride) {
            const int th = ih + pad - a;
            if (th < 0) break;
            const int oh = th / stride;
            if (oh >= OH) continue;
            for (int b = (iw + pad) % stride; b < kw; b += stride) {
                const int tw = iw + pad - b;
                if (tw < 0) break;
                const int ow = tw / stride;
                if (ow >= OW) continue;
                const long long o = ((static_cast<long long>(n) * OH + oh) * OW + ow) * vecs + cv;
                float g[8];
                unpack8(*reinterpret_cast<const uint4*>(dy + o * 8), g);
                if (mode == 0) {
                    const uint2 pk = *reinterpret_cast<const uint2*>(argmax + o * 8);
                    const int want = a * kw + b;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int am = ((k < 4 ? pk.x : pk.y) >> ((k & 3) * 8)) & 0xff;
                        if (am == want) acc[k] += g[k];
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[k] += g[k] * inv;
                }
            }
        }
        *reinterpret_cast<uint4*>(dx + i * 8) = pack8(acc);
    }
}

// Row-structured specialisations for the window shapes the backbones use (3x3/s2, 3x3/s1, 2x2/s2): one block per
// (image, output row), compile-time window loops, 32-bit index math.  The generic kernels above spend ~600 integer
// instructions per 16-byte vector on 64-bit div/mod and ran at a third of the memory roofline.
template <int KH, int KW, int S, int MODE>
__global__ void __launch_bounds__(256) pool_fwd_rows_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                            uint8_t* __restrict__ argmax, int H, int W, int C, int OH, int OW,
                                                            int pad) {
    const int vecs = C >> 3;
    const int n = blockIdx.x / OH, oh = blockIdx.x - n * OH;
    const uint4* xin = reinterpret_cast<const uint4*>(x) + static_cast<size_t>(n) * H * W * vecs;
    const size_t orow = (static_cast<size_t>(n) * OH + oh) * OW * vecs;
    const int items = OW * vecs;
    for (int t = threadIdx.x; t < items; t += blockDim.x) {
        const int ow = t / vecs, cv = t - ow * vecs;
        float acc[8];
        int arg[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { acc[k] = MODE == 0 ? -INFINITY : 0.f; arg[k] = 0; }
#pragma unroll
        for (int a = 0; a < KH; ++a) {
            const int ih = oh * S - pad + a;
#pragma unroll
            for (int b = 0; b < KW; ++b) {
                const int iw = ow * S - pad + b;
                if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
                float f[8];
                unpack8(xin[(static_cast<size_t>(ih) * W + iw) * vecs + cv], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    if (MODE == 0) {
                        if (f[k] > acc[k]) { acc[k] = f[k]; arg[k] = a * KW + b; }
                    } else {
                        acc[k] += f[k];
                    }
                }
            }
        }
        if (MODE == 1) {
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] *= 1.f / (KH * KW);
        }
        reinterpret_cast<uint4*>(out)[orow + t] = pack8(acc);
        if (MODE == 0 && argmax != nullptr) {
            uint2 pk;
            pk.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
            pk.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
            reinterpret_cast<uint2*>(argmax)[orow + t] = pk;
        }
    }
}

template <int KH, int KW, int S, int MODE>
__global__ void __launch_bounds__(256) pool_bwd_rows_kernel(const __nv_bfloat16* __restrict__ dy,
                                                            const uint8_t* __restrict__ argmax, __nv_bfloat16* __restrict__ dx,
                                                            int H, int W, int C, int OH, int OW, int pad) {
    const int vecs = C >> 3;
    const int n = blockIdx.x / H, ih = blockIdx.x - n * H;
    const size_t obase = static_cast<size_t>(n) * OH * OW * vecs;
    const size_t irow = (static_cast<size_t>(n) * H + ih) * W * vecs;
    const int items = W * vecs;
    for (int t = threadIdx.x; t < items; t += blockDim.x) {
        const int iw = t / vecs, cv = t - iw * vecs;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int a = 0; a < KH; ++a) {
            const int th = ih + pad - a;                    // = oh * S for the window whose tap `a` lands on this row
            if (th < 0 || (S > 1 && (th % S) != 0)) continue;
            const int oh = th / S;
            if (oh >= OH) continue;
#pragma unroll
            for (int b = 0; b < KW; ++b) {
                const int tw = iw + pad - b;
                if (tw < 0 || (S > 1 && (tw % S) != 0)) continue;
                const int ow = tw / S;
                if (ow >= OW) continue;
                const size_t o = obase + (static_cast<size_t>(oh) * OW + ow) * vecs + cv;
                float g[8];
                unpack8(reinterpret_cast<const uint4*>(dy)[o], g);
                if (MODE == 0) {
                    const uint2 pk = reinterpret_cast<const uint2*>(argmax)[o];
                    const uint32_t want = static_cast<uint32_t>(a * KW + b) * 0x01010101u;
                    const uint32_t ex = pk.x ^ want, ey = pk.y ^ want;      // a zero byte = this tap won that channel
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        if (((ex >> (k * 8)) & 0xff) == 0) acc[k] += g[k];
                        if (((ey >> (k * 8)) & 0xff) == 0) acc[4 + k] += g[4 + k];
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[k] += g[k] * (1.f / (KH * KW));
                }
            }
        }
        reinterpret_cast<uint4*>(dx)[irow + t] = pack8(acc);
    }
}

// [N, HW, C] -> [N, C] mean.  One block per (n, 256-channel slab); threads split HW, smem reduce.
__global__ void __launch_bounds__(256) gap_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                      int HW, int C) {
    const int vecs = C >> 3;
    const int n = blockIdx.y;
    const int tx_count = vecs < 32 ? vecs : 32;
    const int ty_count = 256 / tx_count;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    __shared__ float red[256 * 8];
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ty < ty_count && cv < vecs) {
        for (int p = ty; p < HW; p += ty_count) {
            float f[8];
            unpack8(*reinterpret_cast<const uint4*>(x + (static_cast<long long>(n) * HW + p) * C + cv * 8), f);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += f[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) red[threadIdx.x * 8 + k] = acc[k];
    __syncthreads();
    if (ty == 0 && cv < vecs) {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] += red[(yy * tx_count + tx) * 8 + k];
        const float inv = 1.f / HW;
#pragma unroll
        for (int k = 0; k < 8; ++k) s[k] *= inv;
        *reinterpret_cast<uint4*>(out + static_cast<long long>(n) * C + cv * 8) = pack8(s);
    }
}

// Few samples x few channels over many pixels (EfficientNet's early squeeze-excite layers: 64 x 144 channels over 150 x 150): the
// kernel above would run on N * ceil(C / 256) blocks only.  Here the pixels of a sample are split over blockIdx.z as well; partial
// sums meet in an fp32 scratch [N, C] (zeroed by the caller), gap_finish_kernel scales and rounds.
__global__ void __launch_bounds__(256) gap_fwd_part_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ acc_out, int HW, int C,
                                                           int rows_per_block) {
    pdl_sync();
    const int vecs = C >> 3;
    const int n = blockIdx.y;
    const int tx_count = vecs < 32 ? vecs : 32;
    const int ty_count = 256 / tx_count;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    __shared__ float red[256 * 8];
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ty < ty_count && cv < vecs) {
        const int p0 = blockIdx.z * rows_per_block;
        const int p1 = min(p0 + rows_per_block, HW);
        for (int p = p0 + ty; p < p1; p += 2 * ty_count) {          // two rows in flight
            const __nv_bfloat16* src = x + (static_cast<long long>(n) * HW + p) * C + cv * 8;
            const uint4 a = *reinterpret_cast<const uint4*>(src);
            const bool two = p + ty_count < p1;
            const uint4 b = two ? *reinterpret_cast<const uint4*>(src + static_cast<long long>(ty_count) * C) : make_uint4(0, 0, 0, 0);
            float f[8], g[8];
            unpack8(a, f); unpack8(b, g);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += f[k] + g[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) red[threadIdx.x * 8 + k] = acc[k];
    __syncthreads();
    if (ty == 0 && cv < vecs) {
        float t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] += red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(acc_out + static_cast<long long>(n) * C + cv * 8 + k, t[k]);
    }
}
__global__ void __launch_bounds__(256) gap_finish_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out, int total,
                                                         float inv) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < total) out[i] = __float2bfloat16(acc[i] * inv);
}

__global__ void __launch_bounds__(256) gap_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
                                                      long long total_vecs, int HW, int C) {
    const int vecs = C >> 3;
    const float inv = 1.f / HW;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total_vecs; i += step) {
        const int cv = static_cast<int>(i % vecs);
        const long long n = i / (static_cast<long long>(vecs) * HW);
        float g[8];
        unpack8(*reinterpret_cast<const uint4*>(dy + n * C + cv * 8), g);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] *= inv;
        *reinterpret_cast<uint4*>(dx + i * 8) = pack8(g);
    }
}

// ------------------------------------------------------------------------------------ per-(sample, channel) scaling
// out[n, p, c] = x[n, p, c] * s[n, c] * m[n]: the squeeze-excite gate of the MBConv blocks (reference: nn/classifier.py ->
// efficientnet_pytorch `x = torch.sigmoid(x_squeezed) * x`) with the stochastic-depth sample mask m (nullable) folded in.
// Backward in ONE pass over (dy, x): dx = dy * s * m and ds[n, c] += m * sum_p dy * x (fp32 atomics; a block owns one sample,
// <= 32 channel vectors and one chunk of the pixels).
__global__ void __launch_bounds__(256) channel_scale_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ s,
                                                                const float* __restrict__ m, __nv_bfloat16* __restrict__ out,
                                                                long long total_vecs, int HW, int C) {
    pdl_sync();
    const int vecs = C >> 3;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total_vecs; i += step) {
        const int cv = static_cast<int>(i % vecs);
        const long long n = i / (static_cast<long long>(vecs) * HW);
        float f[8], g[8];
        unpack8(*reinterpret_cast<const uint4*>(x + i * 8), f);
        unpack8(*reinterpret_cast<const uint4*>(s + n * C + cv * 8), g);
        const float mk = m != nullptr ? m[n] : 1.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] *= g[k] * mk;
        *reinterpret_cast<uint4*>(out + i * 8) = pack8(f);
    }
}

__global__ void __launch_bounds__(256) channel_scale_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                                const __nv_bfloat16* __restrict__ s, const float* __restrict__ m,
                                                                __nv_bfloat16* __restrict__ dx, float* __restrict__ ds, int HW, int C,
                                                                int rows_per_block) {
    pdl_sync();
    const int vecs = C >> 3;
    const int n = blockIdx.y;
    const int tx_count = vecs < 32 ? vecs : 32;
    const int ty_count = 256 / tx_count;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    __shared__ float red[256 * 8];
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const float mk = m != nullptr ? m[n] : 1.f;
    if (ty < ty_count && cv < vecs) {
        float g[8];
        unpack8(*reinterpret_cast<const uint4*>(s + static_cast<long long>(n) * C + cv * 8), g);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] *= mk;
        const int p0 = blockIdx.z * rows_per_block;
        const int p1 = min(p0 + rows_per_block, HW);
        for (int p = p0 + ty; p < p1; p += 2 * ty_count) {          // two rows in flight
            const long long o0 = (static_cast<long long>(n) * HW + p) * C + cv * 8;
            const bool two = p + ty_count < p1;
            const long long o1 = o0 + static_cast<long long>(ty_count) * C;
            const uint4 d0 = *reinterpret_cast<const uint4*>(dy + o0), x0 = *reinterpret_cast<const uint4*>(x + o0);
            uint4 d1 = make_uint4(0, 0, 0, 0), x1 = make_uint4(0, 0, 0, 0);
            if (two) { d1 = *reinterpret_cast<const uint4*>(dy + o1); x1 = *reinterpret_cast<const uint4*>(x + o1); }
            float a[8], b[8], o[8];
            unpack8(d0, a); unpack8(x0, b);
#pragma unroll
            for (int k = 0; k < 8; ++k) { acc[k] = fmaf(a[k], b[k], acc[k]); o[k] = a[k] * g[k]; }
            *reinterpret_cast<uint4*>(dx + o0) = pack8(o);
            if (two) {
                unpack8(d1, a); unpack8(x1, b);
#pragma unroll
                for (int k = 0; k < 8; ++k) { acc[k] = fmaf(a[k], b[k], acc[k]); o[k] = a[k] * g[k]; }
                *reinterpret_cast<uint4*>(dx + o1) = pack8(o);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) red[threadIdx.x * 8 + k] = acc[k];
    __syncthreads();
    if (ty == 0 && cv < vecs) {
        float t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] += red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(ds + static_cast<long long>(n) * C + cv * 8 + k, t[k] * mk);
    }
}

// ------------------------------------------------------------------------------------ weighted cross-entropy
// One block.  loss = sum_i w[y_i] * (lse_i - x_i[y_i]) / sum_i w[y_i]; dlogits written in the same launch
// (d loss / d logits, to be scaled by the incoming scalar gradient).  logits fp32 [B, nc] (tiny: the head output).
__global__ void __launch_bounds__(256) xent_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                                   const float* __restrict__ weight, float* __restrict__ loss,
                                                   float* __restrict__ dlogits, int B, int nc) {
    __shared__ float s_num[8], s_den[8];
    __shared__ float s_total[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float num = 0.f, den = 0.f;
    for (int r = warp; r < B; r += nwarps) {
        const float* row = logits + static_cast<long long>(r) * nc;
        float mx = -INFINITY;
        for (int j = lane; j < nc; j += 32) mx = fmaxf(mx, row[j]);
        mx = warp_max(mx);
        float se = 0.f;
        for (int j = lane; j < nc; j += 32) se += __expf(row[j] - mx);
        se = warp_sum(se);
        const int y = static_cast<int>(labels[r]);
        const float w = weight != nullptr ? weight[y] : 1.f;
        if (lane == 0) {
            num += w * (logf(se) + mx - row[y]);
            den += w;
        }
    }
    if (lane == 0) { s_num[warp] = num; s_den[warp] = den; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
        for (int i = 0; i < nwarps; ++i) { a += s_num[i]; b += s_den[i]; }
        s_total[0] = a; s_total[1] = b;
        loss[0] = a / b;
    }
    __syncthreads();
    const float inv_den = 1.f / s_total[1];
    for (int r = warp; r < B; r += nwarps) {
        const float* row = logits + static_cast<long long>(r) * nc;
        float mx = -INFINITY;
        for (int j = lane; j < nc; j += 32) mx = fmaxf(mx, row[j]);
        mx = warp_max(mx);
        float se = 0.f;
        for (int j = lane; j < nc; j += 32) se += __expf(row[j] - mx);
        se = warp_sum(se);
        const int y = static_cast<int>(labels[r]);
        const float w = (weight != nullptr ? weight[y] : 1.f) * inv_den;
        for (int j = lane; j < nc; j += 32) {
            const float pr = __expf(row[j] - mx) / se;
            dlogits[static_cast<long long>(r) * nc + j] = w * (pr - (j == y ? 1.f : 0.f));
        }
    }
}

// ------------------------------------------------------------------------------------ element-wise activation
// TACT >= 0 fixes the activation at compile time (no per-element switch); two vectors per thread are in flight.
constexpr int ACT_UNROLL = 2;      // independent 16-byte vectors (rows) in flight per thread in the activation kernels (4 measured slower: GELU backward is ALU-bound)
// TACT >= 0 fixes the activation at compile time (no per-element switch).
template <int TACT>
__global__ void __launch_bounds__(256) act_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ bias,
                                                      __nv_bfloat16* __restrict__ out, long long total_vecs, int C,
                                                      int act) {
    const int vecs = C >> 3;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < total_vecs; i0 += ACT_UNROLL * step) {
        uint4 raw[ACT_UNROLL];
#pragma unroll
        for (int u = 0; u < ACT_UNROLL; ++u) {
            const long long i = i0 + u * step;
            raw[u] = i < total_vecs ? *reinterpret_cast<const uint4*>(x + i * 8) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < ACT_UNROLL; ++u) {
            const long long i = i0 + u * step;
            if (i >= total_vecs) break;
            float f[8];
            unpack8(raw[u], f);
            if (bias != nullptr) {
                const int c0 = static_cast<int>(i % vecs) * 8;
#pragma unroll
                for (int k = 0; k < 8; ++k) f[k] += bias[c0 + k];
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = apply_act(f[k], TACT >= 0 ? TACT : act);
            *reinterpret_cast<uint4*>(out + i * 8) = pack8(f);
        }
    }
}
// dx = dy * act'(x + bias); optionally dbias[c] += sum over rows of dx[r, c] in the same pass (the bias gradient of the Linear
// layer that produced x: saves re-reading dx).  Thread = one vector of 8 channels, rows strided over the block, ACT_UNROLL rows in
// flight; shared-memory fold over the row dimension, one fp32 atomic per channel per block.
template <int TACT>
__global__ void __launch_bounds__(256) act_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                      const float* __restrict__ bias, __nv_bfloat16* __restrict__ dx,
                                                      float* __restrict__ dbias, long long P, int C, int tx_count,
                                                      int ty_count, int rows_per_block, int act) {
    extern __shared__ float ab_red[];     // [ty_count][tx_count][8] (only with dbias)
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, bs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const bool active = ty < ty_count && cv < vecs;
    if (active) {
        if (bias != nullptr) {
#pragma unroll
            for (int k = 0; k < 8; ++k) bs[k] = bias[cv * 8 + k];
        }
        const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
        long long r1 = r0 + rows_per_block;
        if (r1 > P) r1 = P;
        for (long long r = r0 + ty; r < r1; r += static_cast<long long>(ACT_UNROLL) * ty_count) {
            uint4 vg[ACT_UNROLL], vx[ACT_UNROLL];
#pragma unroll
            for (int u = 0; u < ACT_UNROLL; ++u) {
                const long long rr = r + static_cast<long long>(u) * ty_count;
                const bool ok = rr < r1;
                vg[u] = ok ? *reinterpret_cast<const uint4*>(dy + rr * C + cv * 8) : make_uint4(0, 0, 0, 0);
                vx[u] = ok ? *reinterpret_cast<const uint4*>(x + rr * C + cv * 8) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < ACT_UNROLL; ++u) {
                const long long rr = r + static_cast<long long>(u) * ty_count;
                if (rr >= r1) break;
                float g[8], f[8];
                unpack8(vg[u], g);
                unpack8(vx[u], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) g[k] *= act_grad(f[k] + bs[k], TACT >= 0 ? TACT : act);
                const uint4 o = pack8(g);
                *reinterpret_cast<uint4*>(dx + rr * C + cv * 8) = o;
                if (dbias != nullptr) {        // sum what the next kernels will actually read (the bf16-rounded values)
                    float q[8];
                    unpack8(o, q);
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[k] += q[k];
                }
            }
        }
    }
    if (dbias == nullptr) return;
    if (ty < ty_count) {
#pragma unroll
        for (int k = 0; k < 8; ++k) ab_red[(ty * tx_count + tx) * 8 + k] = acc[k];
    }
    __syncthreads();
    if (active && ty == 0) {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] += ab_red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(dbias + cv * 8 + k, s[k]);
    }
}

// out[c] += sum over rows of x[r, c] (bias gradients).  Thread = one vector of 8 channels, rows strided over the
// block, 4 rows in flight; shared-memory fold over the row dimension, one fp32 atomic per channel per block.
__global__ void __launch_bounds__(256) colsum_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                     long long P, int C, int tx_count, int ty_count, int rows_per_block) {
    extern __shared__ float cs_red[];     // [ty_count][tx_count][8]
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const bool active = ty < ty_count && cv < vecs;
    if (active) {
        const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
        long long r1 = r0 + rows_per_block;
        if (r1 > P) r1 = P;
        for (long long r = r0 + ty; r < r1; r += 4LL * ty_count) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long rr = r + static_cast<long long>(u) * ty_count;
                v[u] = rr < r1 ? *reinterpret_cast<const uint4*>(x + rr * C + cv * 8) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                float f[8];
                unpack8(v[u], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] += f[k];
            }
        }
    }
    if (ty < ty_count) {
#pragma unroll
        for (int k = 0; k < 8; ++k) cs_red[(ty * tx_count + tx) * 8 + k] = acc[k];
    }
    __syncthreads();
    if (active && ty == 0) {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] += cs_red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(out + cv * 8 + k, s[k]);
    }
}
// scalar fallback for C not a multiple of 8
__global__ void __launch_bounds__(256) colsum_scalar_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                            long long P, int C, int rows_per_block) {
    const int c = blockIdx.x * 64 + (threadIdx.x & 63);
    const int ty = threadIdx.x >> 6;
    __shared__ float red[256];
    float acc = 0.f;
    if (c < C) {
        const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_block;
        long long r1 = r0 + rows_per_block;
        if (r1 > P) r1 = P;
        for (long long r = r0 + ty; r < r1; r += 4) acc += __bfloat162float(x[r * C + c]);
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    if (ty == 0 && c < C) atomicAdd(out + c, red[threadIdx.x] + red[threadIdx.x + 64] + red[threadIdx.x + 128] + red[threadIdx.x + 192]);
}

// ------------------------------------------------------------------------------------ dropout (counter hash)
__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
    x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
    return x;
}
__global__ void __launch_bounds__(256) dropout_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                      long long n, float p, uint32_t seed,
                                                      const uint32_t* __restrict__ step_word) {
    if (step_word != nullptr) seed += step_word[0] * 0x632BE5ABu;   // graph replay: a fresh mask per step from a device counter
    const float scale = 1.f / (1.f - p);
    const uint32_t thr = static_cast<uint32_t>(p * 4294967296.0);
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += step) {
        const uint32_t h = hash_u32(static_cast<uint32_t>(i) * 0x9E3779B1u + seed);
        const float v = __bfloat162float(x[i]);
        out[i] = __float2bfloat16(h >= thr ? v * scale : 0.f);
    }
}

// ------------------------------------------------------------------------------------ LayerNorm (row = one warp)
// A lane owns columns lane*8 + 256*i (i < V) of every row its warp handles: gamma / beta sit in registers for the whole
// kernel and a row is read exactly once (it stays in registers between the statistics and the normalisation).  Warps walk
// the rows grid-stride, two rows per iteration so that the loads of the second row overlap the reduction of the first.
__device__ __forceinline__ void ld8f(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
// the same eight values in the permuted layout: halves 128 floats apart
__device__ __forceinline__ void ld8f_split(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 128);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
__device__ __forceinline__ void st8f_split(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 128) = make_float4(v[4], v[5], v[6], v[7]);
}
constexpr int LN_MAX_VECS = 4;     // D <= 4 * 256
// Shared-memory position of column j: a lane owns columns lane*8 .. lane*8+7 of every 256-column chunk and reads them as two
// float4; stored in place those sit 32 bytes from the neighbour lane's (2-way bank conflict on every access).  The permutation
// puts the first halves of all lanes side by side, then the second halves: [chunk][half][lane][4].
__device__ __forceinline__ int ln_perm(int j) { return (j & ~255) | (((j >> 2) & 1) << 7) | (((j & 255) >> 3) << 2) | (j & 3); }
// gamma / beta are staged in shared memory once per block and read back with 128-bit loads per row: as 2 x 8 x V registers
// per thread they cost an SM its third and fourth resident block, and latency hiding is all these kernels need.
template <int V>
__global__ void __launch_bounds__(256, 3) layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                                                               float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                               long long M, int D, float eps,
                                                               const __nv_bfloat16* __restrict__ res,
                                                               __nv_bfloat16* __restrict__ sum_out) {
    // res / sum_out (both or neither): the transformer's residual add fused in -- the row normalised is s = bf16(x + res),
    // which is also written to sum_out (the residual stream); one pass instead of an add kernel followed by this one
    __shared__ __align__(16) float sgb[2][V * 256];
    for (int j = threadIdx.x; j < V * 256; j += blockDim.x) {       // parameters do not depend on the previous kernel
        sgb[0][ln_perm(j)] = j < D ? gamma[j] : 0.f;
        sgb[1][ln_perm(j)] = j < D ? beta[j] : 0.f;
    }
    __syncthreads();
    pdl_sync();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float inv_d = 1.f / D;
    const long long nwarps = static_cast<long long>(gridDim.x) * (blockDim.x >> 5);
    for (long long row = (static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + warp) * 2; row < M; row += nwarps * 2) {
        uint4 raw[2][V];
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const int j = lane * 8 + i * 256;
                raw[r][i] = (row + r < M && j < D) ? *reinterpret_cast<const uint4*>(x + (row + r) * D + j) : make_uint4(0, 0, 0, 0);
            }
        if (res != nullptr) {
            uint4 rr[2][V];
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int i = 0; i < V; ++i) {
                    const int j = lane * 8 + i * 256;
                    rr[r][i] = (row + r < M && j < D) ? *reinterpret_cast<const uint4*>(res + (row + r) * D + j) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int i = 0; i < V; ++i) {
                    const int j = lane * 8 + i * 256;
                    if (row + r < M && j < D) {
                        float a[8], b[8];
                        unpack8(raw[r][i], a);
                        unpack8(rr[r][i], b);
#pragma unroll
                        for (int k = 0; k < 8; ++k) a[k] += b[k];
                        raw[r][i] = pack8(a);               // the bf16-rounded sum is what gets normalised (and re-read by the backward)
                        *reinterpret_cast<uint4*>(sum_out + (row + r) * D + j) = raw[r][i];
                    }
                }
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            if (row + r >= M) break;
            float s = 0.f, q = 0.f;
#pragma unroll
            for (int i = 0; i < V; ++i) {
                float f[8];
                unpack8(raw[r][i], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) { s += f[k]; q += f[k] * f[k]; }
            }
            s = warp_sum(s); q = warp_sum(q);
            const float mean = s * inv_d;
            const float rstd = rsqrtf(fmaxf(q * inv_d - mean * mean, 0.f) + eps);
            if (lane == 0) { mean_out[row + r] = mean; rstd_out[row + r] = rstd; }
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const int j = lane * 8 + i * 256;
                if (j < D) {
                    float f[8], gm[8], bt[8];
                    unpack8(raw[r][i], f);
                    ld8f_split(&sgb[0][ln_perm(j)], gm);
                    ld8f_split(&sgb[1][ln_perm(j)], bt);
#pragma unroll
                    for (int k = 0; k < 8; ++k) f[k] = (f[k] - mean) * rstd * gm[k] + bt[k];
                    *reinterpret_cast<uint4*>(y + (row + r) * D + j) = pack8(f);
                }
            }
        }
    }
}
// dx and (atomically accumulated) dgamma/dbeta.  A lane owns the same columns (lane*8 + 256*i) for every row its warp
// handles; the parameter-gradient partials of a warp live in ITS slice of shared memory (128-bit read-modify-write per row, no
// other warp touches it) instead of 2 x 8 x V registers per thread, and are folded across the block's warps once at the end
// (one fp32 atomic per column per block).  D <= 1024 (ViT-B: 768).
template <int V>
__global__ void __launch_bounds__(256, 3) layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                               const float* __restrict__ gamma, const float* __restrict__ mean,
                                                               const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx,
                                                               float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                               long long M, int D, int rows_per_block,
                                                               const __nv_bfloat16* __restrict__ add_in) {
    // add_in (nullable): gradient arriving over the residual stream; the kernel writes dx = LayerNorm-backward + add_in, i.e. the
    // gradient sum autograd would otherwise compute with a separate add kernel.
    // A row is read TWICE: pass A accumulates the two row sums and the parameter-gradient partials on the fly, pass B re-reads
    // dy / x (L1 hits: a row is 3 KB) and writes dx, so that no per-row array survives between the passes; the residual-stream
    // gradient is requested in pass A already (its DRAM latency would otherwise be exposed a second time per row).
    extern __shared__ __align__(16) float sm[];   // [warps][2][V*256] partial dgamma / dbeta, then [V*256] gamma
    constexpr int W2 = V * 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    float* mine = sm + static_cast<size_t>(warp) * 2 * W2;
    float* sgamma = sm + static_cast<size_t>(nw) * 2 * W2;
    for (int j = threadIdx.x; j < W2; j += blockDim.x) sgamma[ln_perm(j)] = j < D ? gamma[j] : 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
        const int j = lane * 8 + i * 256;
        const float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        st8f_split(mine + ln_perm(j), z);
        st8f_split(mine + W2 + ln_perm(j), z);
    }
    __syncthreads();
    pdl_sync();
    const float inv_d = 1.f / D;
    const long long r0 = static_cast<long long>(blockIdx.x) * rows_per_block;
    long long r1 = r0 + rows_per_block;
    if (r1 > M) r1 = M;
    for (long long row = r0 + warp; row < r1; row += nw) {
        const float mu = mean[row], rs = rstd[row];
        float a = 0.f, b = 0.f;
        uint4 ad[V];
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const int j = lane * 8 + i * 256;
            ad[i] = (add_in != nullptr && j < D) ? *reinterpret_cast<const uint4*>(add_in + row * D + j) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const int j = lane * 8 + i * 256;
            if (j < D) {
                float g[8], f[8], gm[8], pg[8], pb[8];
                unpack8(*reinterpret_cast<const uint4*>(dy + row * D + j), g);
                unpack8(*reinterpret_cast<const uint4*>(x + row * D + j), f);
                ld8f_split(sgamma + ln_perm(j), gm);
                ld8f_split(mine + ln_perm(j), pg);
                ld8f_split(mine + W2 + ln_perm(j), pb);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float xh = (f[k] - mu) * rs;
                    const float gg = g[k] * gm[k];
                    a += gg; b = fmaf(gg, xh, b);
                    pg[k] = fmaf(g[k], xh, pg[k]);
                    pb[k] += g[k];
                }
                st8f_split(mine + ln_perm(j), pg);
                st8f_split(mine + W2 + ln_perm(j), pb);
            }
        }
        a = warp_sum(a) * inv_d; b = warp_sum(b) * inv_d;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const int j = lane * 8 + i * 256;
            if (j < D) {
                float g[8], f[8], gm[8], o[8], adf[8];
                unpack8(*reinterpret_cast<const uint4*>(dy + row * D + j), g);
                unpack8(*reinterpret_cast<const uint4*>(x + row * D + j), f);
                ld8f_split(sgamma + ln_perm(j), gm);
                unpack8(ad[i], adf);
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = rs * (g[k] * gm[k] - a - (f[k] - mu) * rs * b) + adf[k];
                *reinterpret_cast<uint4*>(dx + row * D + j) = pack8(o);
            }
        }
    }
    // fold the block's warps: every thread sums a few columns over the warps and issues one global atomic per column
    __syncthreads();
    for (int j = threadIdx.x; j < D; j += blockDim.x) {
        float a = 0.f, b = 0.f;
        const int pj = ln_perm(j);
        for (int w = 0; w < nw; ++w) {
            a += sm[static_cast<size_t>(w) * 2 * W2 + pj];
            b += sm[static_cast<size_t>(w) * 2 * W2 + W2 + pj];
        }
        atomicAdd(dgamma + j, a);
        atomicAdd(dbeta + j, b);
    }
}

// ------------------------------------------------------------------------------------ small linear layers (SIMT)
// The classifier head (2048->128->64->32->nc, batch rows) is far below one tensor-core tile; a warp per output
// element keeps it simple and exact.  x bf16 [M,K], w fp32 [N,K] (the master weights), y bf16 or fp32 [M,N].
__global__ void __launch_bounds__(256) small_linear_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                               const float* __restrict__ bias, void* __restrict__ y,
                                                               int M, int N, int K, int act, int out_fp32) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= M * N) return;
    const int m = warp / N, n = warp % N;
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) acc += __bfloat162float(x[static_cast<long long>(m) * K + k]) * w[static_cast<long long>(n) * K + k];
    acc = warp_sum(acc);
    if (lane == 0) {
        acc = apply_act(acc + (bias != nullptr ? bias[n] : 0.f), act);
        if (out_fp32) static_cast<float*>(y)[static_cast<long long>(m) * N + n] = acc;
        else static_cast<__nv_bfloat16*>(y)[static_cast<long long>(m) * N + n] = __float2bfloat16(acc);
    }
}
// dpre[m,n] = dy[m,n] * act'(y[m,n]) for relu (y>0) / none; then dx = dpre W, dW += dpre^T x, db += sum dpre.
__global__ void __launch_bounds__(256) small_linear_dpre_kernel(const void* __restrict__ dy, const void* __restrict__ y,
                                                                float* __restrict__ dpre, int total, int act, int dy_fp32,
                                                                int y_fp32) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    float g = dy_fp32 ? static_cast<const float*>(dy)[i] : __bfloat162float(static_cast<const __nv_bfloat16*>(dy)[i]);
    if (act == ACT_RELU) {
        const float v = y_fp32 ? static_cast<const float*>(y)[i] : __bfloat162float(static_cast<const __nv_bfloat16*>(y)[i]);
        g = v > 0.f ? g : 0.f;
    }
    dpre[i] = g;
}
__global__ void __launch_bounds__(256) small_linear_dx_kernel(const float* __restrict__ dpre, const float* __restrict__ w,
                                                              __nv_bfloat16* __restrict__ dx, int M, int N, int K) {
    // eight lanes share one output element and split the reduction over n (squeeze-excite layers: M * K is a few thousand
    // outputs over N up to ~1400 -- one thread per output was a 1400-step dependent chain on 15 blocks)
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = t >> 3, sub = t & 7;
    const bool live = i < M * K;
    const int m = live ? i / K : 0, k = live ? i % K : 0;
    float acc = 0.f;
    if (live)
        for (int n = sub; n < N; n += 8) acc = fmaf(dpre[m * N + n], w[static_cast<long long>(n) * K + k], acc);
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (live && sub == 0) dx[i] = __float2bfloat16(acc);
}
__global__ void __launch_bounds__(256) small_linear_dw_kernel(const float* __restrict__ dpre, const __nv_bfloat16* __restrict__ x,
                                                              float* __restrict__ dw, float* __restrict__ db, int M, int N,
                                                              int K) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N * K) {
        const int n = i / K, k = i % K;
        float acc = 0.f;
        for (int m = 0; m < M; ++m) acc += dpre[m * N + n] * __bfloat162float(x[static_cast<long long>(m) * K + k]);
        dw[i] += acc;
    }
    if (db != nullptr && i < N) {
        float acc = 0.f;
        for (int m = 0; m < M; ++m) acc += dpre[m * N + i];
        db[i] += acc;
    }
}

// ------------------------------------------------------------------------------------ first-layer packing
// out[n, i, j, (a*s + b)*C + c] = x[n, i*s + a - pad_h, j*s + b - pad_w, c]  (zero outside the image / padding
// channels).  Turns a stride-s RxS convolution over 3 channels into a stride-1 ceil(R/s) x ceil(S/s) convolution
// over s*s*3 (padded to Cp) channels, i.e. into tensor-core friendly K (ResNet 7x7/s2 stem, ViT 16x16/s16 patches).
// Input may be bf16 NHWC or uint8 NHWC (then ImageNet normalisation is fused: dp/loader.py:86-91).
template <typename T>
__global__ void __launch_bounds__(256) pack_s2d_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int H,
                                                       int W, int C, int Hp, int Wp, int Cp, int s, int pad_h, int pad_w,
                                                       int normalize) {
    // one thread = 8 consecutive packed channels of one cell (a 16-byte store); Cp is a multiple of 16
    const int vecs = Cp >> 3;
    const long long total = static_cast<long long>(N) * Hp * Wp * vecs;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    const float mean[3] = {0.485f, 0.456f, 0.406f}, stdv[3] = {0.229f, 0.224f, 0.225f};
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
        const int cv = static_cast<int>(i % vecs);
        long long t = i / vecs;
        const int j = static_cast<int>(t % Wp); t /= Wp;
        const int ii = static_cast<int>(t % Hp);
        const int n = static_cast<int>(t / Hp);
        float f[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int cp = cv * 8 + k;
            float v = 0.f;
            if (cp < s * s * C) {
                const int c = cp % C, ab = cp / C, a = ab / s, b = ab % s;
                const int ih = ii * s + a - pad_h, iw = j * s + b - pad_w;
                if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
                    v = static_cast<float>(x[((static_cast<long long>(n) * H + ih) * W + iw) * C + c]);
                    if (normalize) v = (v / 255.f - mean[c % 3]) / stdv[c % 3];
                }
            }
            f[k] = v;
        }
        *reinterpret_cast<uint4*>(out + i * 8) = pack8(f);
    }
}

// ------------------------------------------------------------------------------------ synthetic / normalise
__global__ void __launch_bounds__(256) synth_kernel(void* __restrict__ img, long long* __restrict__ labels, long long n_el,
                                                    int batch, int C, int num_classes, uint32_t base, int out_fp32) {
    const float mean[3] = {0.485f, 0.456f, 0.406f}, stdv[3] = {0.229f, 0.224f, 0.225f};
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_el; i += step) {
        const uint32_t h = hash_u32(static_cast<uint32_t>(i) + base);
        const int c = static_cast<int>(i % C);
        const float px = static_cast<float>(h & 255u) / 255.0f;
        const float v = (px - mean[c % 3]) / stdv[c % 3];
        if (out_fp32) static_cast<float*>(img)[i] = v;
        else static_cast<__nv_bfloat16*>(img)[i] = __float2bfloat16(v);
        if (i < batch) labels[i] = static_cast<long long>(hash_u32(static_cast<uint32_t>(i) + (base ^ 0xA5A5A5A5u)) % static_cast<uint32_t>(num_classes));
    }
}
__global__ void __launch_bounds__(256) normalize_u8_kernel(const uint8_t* __restrict__ in, void* __restrict__ out, long long n_el,
                                                           int C, int out_fp32) {
    const float mean[3] = {0.485f, 0.456f, 0.406f}, stdv[3] = {0.229f, 0.224f, 0.225f};
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_el; i += step) {
        const int c = static_cast<int>(i % C);
        const float v = (static_cast<float>(in[i]) / 255.0f - mean[c % 3]) / stdv[c % 3];
        if (out_fp32) static_cast<float*>(out)[i] = v;
        else static_cast<__nv_bfloat16*>(out)[i] = __float2bfloat16(v);
    }
}

// fp32 -> bf16 cast of a flat buffer (shadow-weight refresh)
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += step) out[i] = __float2bfloat16(in[i]);
}


// ------------------------------------------------------------------------------------ channel concat / split (NHWC)
// Inception blocks concatenate 3-4 branch outputs along channels.  forward: dst[p, off_i + c] = src_i[p, c];
// backward (split = 1): src_i[p, c] = dst[p, off_i + c].  One launch for all branches, 16-byte vectors.
constexpr int CAT_MAX = 8;
struct CatArgs {
    __nv_bfloat16* part[CAT_MAX];
    int c_off[CAT_MAX + 1];      // channel offset of every part inside the wide tensor (in elements); c_off[n] = Ctot
    int n;
};
__global__ void __launch_bounds__(256) concat_nhwc_kernel(const CatArgs a, __nv_bfloat16* __restrict__ wide, long long P,
                                                          int split) {
    const int ctot = a.c_off[a.n];
    const int vtot = ctot >> 3;
    const long long total = P * vtot;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
        const long long row = i / vtot;
        const int v = static_cast<int>(i - row * vtot);
        const int c = v << 3;
        int k = 0;
#pragma unroll
        for (int j = 1; j < CAT_MAX; ++j)
            if (j < a.n && c >= a.c_off[j]) k = j;
        const int cw = a.c_off[k + 1] - a.c_off[k];
        uint4* pw = reinterpret_cast<uint4*>(wide + row * ctot + c);
        uint4* pp = reinterpret_cast<uint4*>(a.part[k] + row * cw + (c - a.c_off[k]));
        if (split) *pp = *pw; else *pw = *pp;
    }
}

// ------------------------------------------------------------------------------------ top-1 hits
// out[b] = 1 if argmax(logits[b, :]) == label[b] else 0 (first maximum wins, like torch.argmax); one thread per row.
template <typename T>
__global__ void __launch_bounds__(256) top1_hits_kernel(const T* __restrict__ logits, const long long* __restrict__ labels,
                                                        long long* __restrict__ out, int B, int NC) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const T* row = logits + static_cast<size_t>(b) * NC;
    float best = static_cast<float>(row[0]);
    int arg = 0;
    for (int c = 1; c < NC; ++c) {
        const float v = static_cast<float>(row[c]);
        if (v > best) { best = v; arg = c; }
    }
    out[b] = (arg == static_cast<int>(labels[b])) ? 1 : 0;
}

// Specialisation for the stems of the CNN backbones (3 input channels, stride 2 -> 12 packed channels padded to 16): one
// thread = one packed cell, all index arithmetic compile-time or 32-bit (the generic kernel spends ~80 integer instructions
// per element on run-time div/mod by C and s).
template <typename T>
__global__ void __launch_bounds__(256) pack_s2d_c3s2_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ out, int H, int W,
                                                            int Hp, int Wp, int pad_h, int pad_w, int normalize) {
    const int n = blockIdx.x / Hp, ii = blockIdx.x - n * Hp;
    const float mean[3] = {0.485f, 0.456f, 0.406f}, inv_std[3] = {1.f / 0.229f, 1.f / 0.224f, 1.f / 0.225f};
    const T* xn = x + static_cast<size_t>(n) * H * W * 3;
    __nv_bfloat16* orow = out + (static_cast<size_t>(n) * Hp + ii) * Wp * 16;
    for (int j = threadIdx.x; j < Wp; j += blockDim.x) {
        float f[16];
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const int ih = ii * 2 + a - pad_h;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int iw = j * 2 + b - pad_w;
                const bool ok = ih >= 0 && ih < H && iw >= 0 && iw < W;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float v = 0.f;
                    if (ok) {
                        v = static_cast<float>(xn[(static_cast<size_t>(ih) * W + iw) * 3 + c]);
                        if (normalize) v = (v * (1.f / 255.f) - mean[c]) * inv_std[c];
                    }
                    f[(a * 2 + b) * 3 + c] = v;
                }
            }
        }
        f[12] = f[13] = f[14] = f[15] = 0.f;
        float lo[8], hi[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { lo[k] = f[k]; hi[k] = f[8 + k]; }
        uint4* dst = reinterpret_cast<uint4*>(orow + static_cast<size_t>(j) * 16);
        dst[0] = pack8(lo);
        dst[1] = pack8(hi);
    }
}

}  // namespace picd

using namespace picd;
typedef __nv_bfloat16 bf16;

extern "C" int picd_pool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C, int OH, int OW, int kh,
                             int kw, int stride, int pad, int mode, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "pool: C=%d must be a multiple of 8", C);
    PICD_REQUIRE(kh * kw <= 255, "pool: window too large");
    const long long total = 1LL * N * OH * OW * (C / 8);
#define PICD_POOL_FWD(KH, KW, S, M)                                                                                   \
    if (kh == KH && kw == KW && stride == S && mode == M) {                                                            \
        pool_fwd_rows_kernel<KH, KW, S, M><<<N * OH, 256, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out), \
                                                                   static_cast<uint8_t*>(argmax), H, W, C, OH, OW, pad); \
        PICD_LAUNCH_CHECK();                                                                                            \
        return 0;                                                                                                       \
    }
    PICD_POOL_FWD(3, 3, 2, 0) PICD_POOL_FWD(3, 3, 1, 0) PICD_POOL_FWD(2, 2, 2, 0)
    PICD_POOL_FWD(3, 3, 2, 1) PICD_POOL_FWD(3, 3, 1, 1) PICD_POOL_FWD(2, 2, 2, 1)
#undef PICD_POOL_FWD
    pool_fwd_kernel<<<grid_for(total), 256, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out),
                                                     static_cast<uint8_t*>(argmax), N, H, W, C, OH, OW, kh, kw, stride, pad, mode);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_pool_bwd(const void* dy, const void* argmax, void* dx, int N, int H, int W, int C, int OH, int OW,
                             int kh, int kw, int stride, int pad, int mode, cudaStream_t st) {
    const long long total = 1LL * N * H * W * (C / 8);
#define PICD_POOL_BWD(KH, KW, S, M)                                                                                     \
    if (kh == KH && kw == KW && stride == S && mode == M) {                                                              \
        pool_bwd_rows_kernel<KH, KW, S, M><<<N * H, 256, 0, st>>>(static_cast<const bf16*>(dy),                           \
                                                                  static_cast<const uint8_t*>(argmax),                    \
                                                                  static_cast<bf16*>(dx), H, W, C, OH, OW, pad);          \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_POOL_BWD(3, 3, 2, 0) PICD_POOL_BWD(3, 3, 1, 0) PICD_POOL_BWD(2, 2, 2, 0)
    PICD_POOL_BWD(3, 3, 2, 1) PICD_POOL_BWD(3, 3, 1, 1) PICD_POOL_BWD(2, 2, 2, 1)
#undef PICD_POOL_BWD
    pool_bwd_kernel<<<grid_for(total), 256, 0, st>>>(static_cast<const bf16*>(dy), static_cast<const uint8_t*>(argmax),
                                                     static_cast<bf16*>(dx), N, H, W, C, OH, OW, kh, kw, stride, pad, mode);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_gap_fwd(const void* x, void* out, int N, int HW, int C, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "gap: C=%d must be a multiple of 8", C);
    const int vecs = C / 8, tx = vecs < 32 ? vecs : 32;
    gap_fwd_kernel<<<dim3((vecs + tx - 1) / tx, N), 256, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out), HW, C);
    PICD_LAUNCH_CHECK();
    return 0;
}
// Same result as picd_gap_fwd with the pixels of a sample split over blocks; scratch = fp32 [N, C], ZEROED by the caller.
// Returns through `use_split` semantics on the Python side: worth it when N * ceil(C / 256) blocks would leave SMs idle.
extern "C" int picd_gap_fwd_split(const void* x, float* scratch, void* out, int N, int HW, int C, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "gap_fwd: C=%d must be a multiple of 8", C);
    const int vecs = C / 8;
    const int tx_count = vecs < 32 ? vecs : 32;
    const int ty_count = 256 / tx_count;
    const int xblocks = (vecs + tx_count - 1) / tx_count;
    int z = std::max(1, (4 * num_sms() + N * xblocks - 1) / (N * xblocks));
    z = std::min(z, std::max(1, HW / (2 * ty_count)));
    const int rpb = (HW + z - 1) / z;
    z = (HW + rpb - 1) / rpb;
    PICD_CUDA(launch_pdl<PDL_ROWWISE>(gap_fwd_part_kernel, dim3(xblocks, N, z), dim3(256), 0, st, static_cast<const bf16*>(x), scratch, HW,
                                      C, rpb));
    PICD_CUDA(launch_pdl<PDL_ROWWISE>(gap_finish_kernel, dim3((N * C + 255) / 256), dim3(256), 0, st, static_cast<const float*>(scratch),
                                      static_cast<bf16*>(out), N * C, 1.f / HW));
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_gap_bwd(const void* dy, void* dx, int N, int HW, int C, cudaStream_t st) {
    const long long total = 1LL * N * HW * (C / 8);
    gap_bwd_kernel<<<grid_for(total), 256, 0, st>>>(static_cast<const bf16*>(dy), static_cast<bf16*>(dx), total, HW, C);
    PICD_LAUNCH_CHECK();
    return 0;
}

// out = x * s[n, c] * m[n] (x, out [N, HW, C] bf16; s [N, C] bf16; m [N] fp32 or null)
extern "C" int picd_channel_scale_fwd(const void* x, const void* s, const float* m, void* out, int N, int HW, int C, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "channel_scale: C=%d must be a multiple of 8", C);
    const long long total = static_cast<long long>(N) * HW * (C / 8);
    const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 16LL * num_sms()));
    PICD_CUDA(launch_pdl<PDL_ROWWISE>(channel_scale_fwd_kernel, dim3(blocks), dim3(256), 0, st, static_cast<const bf16*>(x),
                                      static_cast<const bf16*>(s), m, static_cast<bf16*>(out), total, HW, C));
    PICD_LAUNCH_CHECK();
    return 0;
}
// dx = dy * s * m;  ds[n, c] += m[n] * sum_p dy * x   (ds fp32 [N, C], zeroed by the caller)
extern "C" int picd_channel_scale_bwd(const void* dy, const void* x, const void* s, const float* m, void* dx, float* ds, int N,
                                      int HW, int C, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "channel_scale: C=%d must be a multiple of 8", C);
    const int vecs = C / 8;
    const int tx_count = vecs < 32 ? vecs : 32;
    const int ty_count = 256 / tx_count;
    const int xblocks = (vecs + tx_count - 1) / tx_count;
    // enough blocks to fill the GPU: split the pixels of a sample when N * channel slabs alone are too few
    int z = std::max(1, (4 * num_sms() + N * xblocks - 1) / (N * xblocks));
    z = std::min(z, std::max(1, HW / (2 * ty_count)));
    const int rpb = (HW + z - 1) / z;
    z = (HW + rpb - 1) / rpb;
    PICD_CUDA(launch_pdl<PDL_ROWWISE>(channel_scale_bwd_kernel, dim3(xblocks, N, z), dim3(256), 0, st, static_cast<const bf16*>(dy),
                                      static_cast<const bf16*>(x), static_cast<const bf16*>(s), m, static_cast<bf16*>(dx), ds, HW, C, rpb));
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_cross_entropy(const float* logits, const long long* labels, const float* weight, float* loss,
                                  float* dlogits, int B, int nc, cudaStream_t st) {
    xent_kernel<<<1, 256, 0, st>>>(logits, labels, weight, loss, dlogits, B, nc);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_act_fwd(const void* x, const float* bias, void* out, long long rows, int C, int act, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "act: C=%d must be a multiple of 8", C);
    const long long total = rows * (C / 8);
#define PICD_ACT_FWD(A) act_fwd_kernel<A><<<grid_for(total / ACT_UNROLL + 1), 256, 0, st>>>(static_cast<const bf16*>(x), bias, static_cast<bf16*>(out), total, C, act)
    switch (act) {
        case ACT_RELU: PICD_ACT_FWD(ACT_RELU); break;
        case ACT_GELU: PICD_ACT_FWD(ACT_GELU); break;
        case ACT_SWISH: PICD_ACT_FWD(ACT_SWISH); break;
        case ACT_SIGMOID: PICD_ACT_FWD(ACT_SIGMOID); break;
        default: PICD_ACT_FWD(-1); break;
    }
#undef PICD_ACT_FWD
    PICD_LAUNCH_CHECK();
    return 0;
}
// dx = dy * act'(x + bias) and, when dbias != nullptr, dbias[c] += sum_r dx[r, c] in the same pass.
extern "C" int picd_act_bwd_bias(const void* dy, const void* x, const float* bias, void* dx, float* dbias, long long rows,
                                 int C, int act, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "act_bwd: C=%d must be a multiple of 8", C);
    const int vecs = C / 8;
    const int tx = vecs < 32 ? vecs : 32, ty = 256 / tx;     // a warp covers 512 contiguous bytes of a row
    const int gx = (vecs + tx - 1) / tx;
    long long gy = (8LL * num_sms()) / gx;
    if (gy < 1) gy = 1;
    long long rpb = (rows + gy - 1) / gy;
    if (rpb < static_cast<long long>(ACT_UNROLL) * ty) rpb = static_cast<long long>(ACT_UNROLL) * ty;
    gy = (rows + rpb - 1) / rpb;
    const size_t smem = dbias != nullptr ? static_cast<size_t>(tx) * ty * 8 * sizeof(float) : 0;
#define PICD_ACT_BWD(A)                                                                                                     \
    act_bwd_kernel<A><<<dim3(gx, static_cast<unsigned>(gy)), 256, smem, st>>>(static_cast<const bf16*>(dy),                  \
        static_cast<const bf16*>(x), bias, static_cast<bf16*>(dx), dbias, rows, C, tx, ty, static_cast<int>(rpb), act)
    switch (act) {
        case ACT_RELU: PICD_ACT_BWD(ACT_RELU); break;
        case ACT_GELU: PICD_ACT_BWD(ACT_GELU); break;
        case ACT_SWISH: PICD_ACT_BWD(ACT_SWISH); break;
        case ACT_SIGMOID: PICD_ACT_BWD(ACT_SIGMOID); break;
        default: PICD_ACT_BWD(-1); break;
    }
#undef PICD_ACT_BWD
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_act_bwd(const void* dy, const void* x, const float* bias, void* dx, long long rows, int C, int act,
                            cudaStream_t st) {
    return picd_act_bwd_bias(dy, x, bias, dx, nullptr, rows, C, act, st);
}
extern "C" int picd_colsum(const void* x, float* out, long long P, int C, cudaStream_t st) {
    if (C % 8 == 0) {
        const int vecs = C / 8;
        const int tx = vecs < 32 ? vecs : 32, ty = 256 / tx;     // a warp covers 512 contiguous bytes of a row
        const int gx = (vecs + tx - 1) / tx;
        long long gy = (4LL * num_sms()) / gx;
        if (gy < 1) gy = 1;
        long long rpb = (P + gy - 1) / gy;
        if (rpb < 4LL * ty) rpb = 4LL * ty;
        gy = (P + rpb - 1) / rpb;
        colsum_kernel<<<dim3(gx, static_cast<unsigned>(gy)), 256, tx * ty * 8 * sizeof(float), st>>>(
            static_cast<const bf16*>(x), out, P, C, tx, ty, static_cast<int>(rpb));
    } else {
        const int gx = (C + 63) / 64;
        long long gy = (4LL * num_sms()) / gx;
        if (gy < 1) gy = 1;
        long long rpb = (P + gy - 1) / gy;
        if (rpb < 16) rpb = 16;
        gy = (P + rpb - 1) / rpb;
        colsum_scalar_kernel<<<dim3(gx, static_cast<unsigned>(gy)), 256, 0, st>>>(static_cast<const bf16*>(x), out, P, C,
                                                                                  static_cast<int>(rpb));
    }
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_dropout(const void* x, void* out, long long n, float p, unsigned seed, cudaStream_t st) {
    dropout_kernel<<<grid_for(n), 256, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out), n, p, seed, nullptr);
    PICD_LAUNCH_CHECK();
    return 0;
}
// Same mask generator, but the seed is mixed with a 32-bit step word read from device memory when the kernel runs, so a
// launch recorded in a CUDA graph draws a different mask at every replay (forward and backward read the same word).
extern "C" int picd_dropout_dyn(const void* x, void* out, long long n, float p, unsigned seed, const void* step_word,
                                cudaStream_t st) {
    dropout_kernel<<<grid_for(n), 256, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out), n, p, seed,
                                                static_cast<const uint32_t*>(step_word));
    PICD_LAUNCH_CHECK();
    return 0;
}
// s = x + res (bf16) -> sum_out, y = LayerNorm(s): the residual add of a transformer block fused into the LayerNorm that
// follows it.  res == nullptr: plain LayerNorm of x (sum_out unused).
extern "C" int picd_add_layernorm_fwd(const void* x, const void* res, const float* gamma, const float* beta, void* sum_out,
                                      void* y, float* mean, float* rstd, long long M, int D, float eps, cudaStream_t st) {
    PICD_REQUIRE(D % 8 == 0 && D <= 256 * LN_MAX_VECS, "layernorm: D=%d must be a multiple of 8 and <= 1024", D);
    PICD_REQUIRE((res == nullptr) == (sum_out == nullptr), "add_layernorm: res and sum_out go together");
    const int V = (D + 255) / 256;
    long long blocks = std::min<long long>((M + 15) / 16, 3LL * num_sms());     // resident blocks only: the parameter staging
    if (blocks < 1) blocks = 1;                                                 // is paid once per block
#define PICD_LN_FWD(VV)                                                                                              \
    PICD_CUDA(launch_pdl<PDL_ROWWISE>(layernorm_fwd_kernel<VV>, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st,            \
                         static_cast<const bf16*>(x), gamma, beta, static_cast<bf16*>(y), mean, rstd, M, D, eps,     \
                         static_cast<const bf16*>(res), static_cast<bf16*>(sum_out)))
    if (V == 1) PICD_LN_FWD(1); else if (V == 2) PICD_LN_FWD(2); else if (V == 3) PICD_LN_FWD(3); else PICD_LN_FWD(4);
#undef PICD_LN_FWD
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                                  long long M, int D, float eps, cudaStream_t st) {
    return picd_add_layernorm_fwd(x, nullptr, gamma, beta, nullptr, y, mean, rstd, M, D, eps, st);
}
// dx = LayerNorm-backward(dy, x) + add_in (nullable: the gradient arriving over the residual stream).
extern "C" int picd_layernorm_bwd_add(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                                      const void* add_in, void* dx, float* dgamma, float* dbeta, long long M, int D,
                                      cudaStream_t st) {
    PICD_REQUIRE(D % 8 == 0 && D <= 256 * LN_MAX_VECS, "layernorm: D=%d must be a multiple of 8 and <= 1024", D);
    static bool attr_set = false;
    if (!attr_set) {       // V = 4 parks 8 warps x 2 x 1024 floats + gamma = 68 KB
        PICD_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
        PICD_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
        attr_set = true;
    }
    long long blocks = 3LL * num_sms();
    long long rpb = (M + blocks - 1) / blocks;
    if (rpb < 16) rpb = 16;
    blocks = (M + rpb - 1) / rpb;
    const int V = (D + 255) / 256;
#define PICD_LN_BWD(VV)                                                                                                   \
    PICD_CUDA(launch_pdl<PDL_ROWWISE>(layernorm_bwd_kernel<VV>, dim3(static_cast<unsigned>(blocks)), dim3(256),                         \
        (8 * 2 + 1) * VV * 256 * sizeof(float), st,                                                                             \
        static_cast<const bf16*>(dy), static_cast<const bf16*>(x), gamma, mean, rstd, static_cast<bf16*>(dx), dgamma, dbeta, \
        M, D, static_cast<int>(rpb), static_cast<const bf16*>(add_in)))
    if (V == 1) PICD_LN_BWD(1); else if (V == 2) PICD_LN_BWD(2); else if (V == 3) PICD_LN_BWD(3); else PICD_LN_BWD(4);
#undef PICD_LN_BWD
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                                  void* dx, float* dgamma, float* dbeta, long long M, int D, cudaStream_t st) {
    return picd_layernorm_bwd_add(dy, x, gamma, mean, rstd, nullptr, dx, dgamma, dbeta, M, D, st);
}
extern "C" int picd_small_linear_fwd(const void* x, const float* w, const float* bias, void* y, int M, int N, int K, int act,
                                     int out_fp32, cudaStream_t st) {
    const long long threads = 32LL * M * N;
    small_linear_fwd_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, st>>>(static_cast<const bf16*>(x), w, bias, y,
                                                                                         M, N, K, act, out_fp32);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_small_linear_bwd(const void* dy, const void* y, const void* x, const float* w, float* dpre_scratch,
                                     void* dx, float* dw, float* db, int M, int N, int K, int act, int dy_fp32, int y_fp32,
                                     cudaStream_t st) {
    small_linear_dpre_kernel<<<(M * N + 255) / 256, 256, 0, st>>>(dy, y, dpre_scratch, M * N, act, dy_fp32, y_fp32);
    if (dx != nullptr)
        small_linear_dx_kernel<<<(M * K * 8 + 255) / 256, 256, 0, st>>>(dpre_scratch, w, static_cast<bf16*>(dx), M, N, K);
    const int work = N * K > N ? N * K : N;
    small_linear_dw_kernel<<<(work + 255) / 256, 256, 0, st>>>(dpre_scratch, static_cast<const bf16*>(x), dw, db, M, N, K);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_pack_s2d(const void* x, void* out, int N, int H, int W, int C, int Hp, int Wp, int Cp, int s, int pad_h,
                             int pad_w, int src_u8, cudaStream_t st) {
    PICD_REQUIRE(Cp % 8 == 0, "pack_s2d: packed channel count %d must be a multiple of 8", Cp);
    const long long total = 1LL * N * Hp * Wp * (Cp / 8);
    if (C == 3 && s == 2 && Cp == 16) {
        if (src_u8)
            pack_s2d_c3s2_kernel<uint8_t><<<N * Hp, 128, 0, st>>>(static_cast<const uint8_t*>(x), static_cast<bf16*>(out), H, W, Hp, Wp,
                                                                  pad_h, pad_w, 1);
        else
            pack_s2d_c3s2_kernel<bf16><<<N * Hp, 128, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out), H, W, Hp, Wp, pad_h,
                                                               pad_w, 0);
        PICD_LAUNCH_CHECK();
        return 0;
    }
    if (src_u8)
        pack_s2d_kernel<uint8_t><<<grid_for(total), 256, 0, st>>>(static_cast<const uint8_t*>(x), static_cast<bf16*>(out), N, H, W, C,
                                                                  Hp, Wp, Cp, s, pad_h, pad_w, 1);
    else
        pack_s2d_kernel<bf16><<<grid_for(total), 256, 0, st>>>(static_cast<const bf16*>(x), static_cast<bf16*>(out), N, H, W, C, Hp,
                                                               Wp, Cp, s, pad_h, pad_w, 0);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_synthetic_batch(void* img, long long* labels, int batch, int H, int W, int C, int num_classes,
                                    unsigned base, int out_fp32, cudaStream_t st) {
    const long long n_el = 1LL * batch * H * W * C;
    synth_kernel<<<grid_for(n_el), 256, 0, st>>>(img, labels, n_el, batch, C, num_classes, base, out_fp32);
    PICD_LAUNCH_CHECK();
    return 0;
}
// parts: array of n device pointers (host memory), channels: array of n channel counts (multiples of 8)
extern "C" int picd_concat_nhwc(void* const* parts, const int* channels, int n, void* wide, long long P, int split,
                                cudaStream_t st) {
    PICD_REQUIRE(n >= 1 && n <= CAT_MAX, "concat: %d parts (max %d)", n, CAT_MAX);
    CatArgs a;
    int off = 0;
    for (int i = 0; i < n; ++i) {
        PICD_REQUIRE(channels[i] % 8 == 0, "concat: part %d has %d channels (must be a multiple of 8)", i, channels[i]);
        a.part[i] = static_cast<bf16*>(parts[i]);
        a.c_off[i] = off;
        off += channels[i];
    }
    for (int i = n; i <= CAT_MAX; ++i) a.c_off[i] = off;
    for (int i = n; i < CAT_MAX; ++i) a.part[i] = nullptr;
    a.n = n;
    concat_nhwc_kernel<<<grid_for(P * (off / 8)), 256, 0, st>>>(a, static_cast<bf16*>(wide), P, split);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_top1_hits(const void* logits, int is_bf16, const long long* labels, long long* out, int B, int NC,
                              cudaStream_t st) {
    if (B <= 0) return 0;
    if (is_bf16) top1_hits_kernel<bf16><<<(B + 255) / 256, 256, 0, st>>>(static_cast<const bf16*>(logits), labels, out, B, NC);
    else top1_hits_kernel<float><<<(B + 255) / 256, 256, 0, st>>>(static_cast<const float*>(logits), labels, out, B, NC);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_normalize_u8(const void* in, void* out, long long n_el, int C, int out_fp32, cudaStream_t st) {
    normalize_u8_kernel<<<grid_for(n_el), 256, 0, st>>>(static_cast<const uint8_t*>(in), out, n_el, C, out_fp32);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t st) {
    cast_f32_bf16_kernel<<<grid_for(n), 256, 0, st>>>(in, static_cast<bf16*>(out), n);
    PICD_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of picd_small_linear_bwd.
//
// Who needs it: a host that is not allowed to create device memory at the point where it runs the classifier-head
// backward -- for example a serving or training process that handed its whole memory budget to its own pool at
// start-up and forbids cudaMalloc afterwards, a plug-in living under another framework's allocator, or a
// latency-critical loop that cannot afford cudaMalloc's device-wide synchronisation.  Such a caller still owns the
// inputs and outputs, but it has nowhere to get the `dpre` scratch (M*N floats) the ordinary entry point asks for.
//
// Here that scratch is a fixed array compiled into the library image.  Consequences a user must know:
//   * capacity: M*N <= PICD_SMALL_LINEAR_NOALLOC_MAX_FLOATS (262144 floats = 1 MiB); larger problems are rejected
//     with an error, nothing is launched.  Every GPU that loads the library carries this 1 MiB until exit.
//   * the array is shared by ALL calls on a device and keeps whatever the previous call left in it.  Nothing in it
//     has to start at zero: the first kernel of each call overwrites the M*N floats it uses before anything reads
//     them.  Two calls whose kernels overlap on the GPU (different streams without an ordering between them) race
//     on the array; issue them on one stream or order them with events.
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device
//     current).  That first call should therefore happen before any stream capture begins.
//   * no cudaMalloc / cudaFree / allocating library routine is called here, on any call.
// Results are bit-identical to picd_small_linear_bwd given the same inputs (same kernels, same launch shapes).
#include <mutex>

#define PICD_SMALL_LINEAR_NOALLOC_MAX_FLOATS (1 << 18)

static __device__ __align__(256) float g_small_linear_noalloc_scratch[PICD_SMALL_LINEAR_NOALLOC_MAX_FLOATS];

// In host code the array's name is only a handle for the runtime, not a device address; the address a kernel can use
// has to be asked for, once per device, while that device is current.
static float* g_noalloc_scratch_addr[64] = {nullptr};
static std::mutex g_noalloc_mutex;

static int small_linear_noalloc_scratch(float** out) {
    int dev = -1;
    PICD_CUDA(cudaGetDevice(&dev));
    PICD_REQUIRE(dev >= 0 && dev < 64, "small_linear_bwd_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        PICD_CUDA(cudaGetSymbolAddress(&p, g_small_linear_noalloc_scratch));
        PICD_REQUIRE(p != nullptr, "small_linear_bwd_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_scratch_addr[dev] = static_cast<float*>(p);
    }
    *out = g_noalloc_scratch_addr[dev];
    return 0;
}

extern "C" int picd_small_linear_noalloc_max_floats() { return PICD_SMALL_LINEAR_NOALLOC_MAX_FLOATS; }

extern "C" int picd_small_linear_bwd_noalloc(const void* dy, const void* y, const void* x, const float* w, void* dx,
                                             float* dw, float* db, int M, int N, int K, int act, int dy_fp32,
                                             int y_fp32, cudaStream_t st) {
    PICD_REQUIRE(M > 0 && N > 0 && K > 0, "small_linear_bwd_noalloc: M, N, K must be positive (got %d, %d, %d)", M, N, K);
    const long long need = static_cast<long long>(M) * N;
    PICD_REQUIRE(need <= PICD_SMALL_LINEAR_NOALLOC_MAX_FLOATS,
                 "small_linear_bwd_noalloc: M*N = %lld floats of scratch exceed the built-in %d; use "
                 "picd_small_linear_bwd with caller-provided scratch", need, PICD_SMALL_LINEAR_NOALLOC_MAX_FLOATS);
    float* scratch = nullptr;
    if (int rc = small_linear_noalloc_scratch(&scratch)) return rc;
    small_linear_dpre_kernel<<<(M * N + 255) / 256, 256, 0, st>>>(dy, y, scratch, M * N, act, dy_fp32, y_fp32);
    if (dx != nullptr)
        small_linear_dx_kernel<<<(M * K * 8 + 255) / 256, 256, 0, st>>>(scratch, w, static_cast<bf16*>(dx), M, N, K);
    const int work = N * K > N ? N * K : N;
    small_linear_dw_kernel<<<(work + 255) / 256, 256, 0, st>>>(scratch, static_cast<const bf16*>(x), dw, db, M, N, K);
    PICD_LAUNCH_CHECK();
    return 0;
}
