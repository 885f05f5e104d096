// ResNet stem: BatchNorm apply + ReLU + 3x3 / stride-2 / pad-1 max-pool as ONE pass, and its backward as two.
//
// The reference reaches torchvision's `bn1 -> relu -> maxpool` (nn/classifier.py:36-37 -> torchvision resnet.py) as three
// library kernels forward and three backward, each a full pass over the 112x112x64 activation (411 MB at batch 256) -- the
// largest tensor of the network.  Here the normalised activation z is never materialised:
//   forward : pooled[oh, ow] = max over the window of relu(y * scale + shift), arg-max offset stored as one byte per element
//             (reads y once, writes the 4x smaller pooled tensor)
//   backward: g[ih, iw] = relu'(.) * sum of dpool over the (<= 4) windows whose arg-max points at (ih, iw) is recomputed on
//             the fly from dpool + arg-max (both 4x smaller than z) by the reduction kernel (BatchNorm sums) and again by
//             the apply kernel (dy = scale * g + B * y + C): the 411 MB dz tensor never exists.
// NHWC bf16; one block per (image, row); a thread owns one 8-channel vector position of the row.
#include "../common/picd_common.cuh"

namespace picd {

__device__ __forceinline__ void stem_unpack8(const uint4& raw, float (&f)[8]) {
    float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 stem_pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
__device__ __forceinline__ void stem_ld8(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// ------------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(256) bn_relu_maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ y,
                                                                  const float* __restrict__ scale,
                                                                  const float* __restrict__ shift,
                                                                  __nv_bfloat16* __restrict__ out, uint8_t* __restrict__ argmax,
                                                                  int H, int W, int C, int OH, int OW) {
    pdl_sync();
    const int vecs = C >> 3;
    const int n = blockIdx.x / OH, oh = blockIdx.x - n * OH;
    const uint4* yin = reinterpret_cast<const uint4*>(y) + static_cast<size_t>(n) * H * W * vecs;
    const size_t orow = (static_cast<size_t>(n) * OH + oh) * OW * vecs;
    const int items = OW * vecs;
    for (int t = threadIdx.x; t < items; t += blockDim.x) {
        const int ow = t / vecs, cv = t - ow * vecs;
        float sc[8], sh[8];
        stem_ld8(scale + cv * 8, sc);
        stem_ld8(shift + cv * 8, sh);
        float best[8];
        int arg[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; arg[k] = 0; }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int ih = oh * 2 - 1 + a;
#pragma unroll
            for (int b = 0; b < 3; ++b) {
                const int iw = ow * 2 - 1 + b;
                if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
                float f[8];
                stem_unpack8(yin[(static_cast<size_t>(ih) * W + iw) * vecs + cv], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    // the value the unfused path would have stored: relu(bn(y)) rounded to bf16
                    const float z = bf16_round(fmaxf(fmaf(f[k], sc[k], sh[k]), 0.f));
                    if (z > best[k]) { best[k] = z; arg[k] = a * 3 + b; }
                }
            }
        }
        reinterpret_cast<uint4*>(out)[orow + t] = stem_pack8(best);
        uint2 pk;
        pk.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
        pk.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
        reinterpret_cast<uint2*>(argmax)[orow + t] = pk;
    }
}

// ------------------------------------------------------------------------------------------------ backward
// A thread owns one 2x2 input QUAD (rows 2m, 2m+1; columns 2j, 2j+1) of one 8-channel vector.  With a 3x3 / stride-2 / pad-1 pool
// those four pixels are touched by exactly the windows (m, j), (m, j+1), (m+1, j), (m+1, j+1), and each (pixel, window) pair
// has one fixed tap index -- so the pooled gradient of the quad is four (dpool, arg-max) loads and nine byte compares, without a
// data-dependent branch, and all eight loads of the quad (4 x y, 4 x window) are independent.
struct StemQuad {
    uint4 y[2][2];      // raw bf16 y vectors (zero outside the image)
    uint4 d[2][2];      // dpool of window (m + i, j + k) (zero outside)
    uint2 a[2][2];      // its arg-max bytes (0xff outside: matches no tap)
};

__device__ __forceinline__ void stem_load_quad(StemQuad& q, const uint4* __restrict__ yimg, const uint4* __restrict__ dimg,
                                               const uint2* __restrict__ aimg, int m, int j, int cv, int vecs, int H, int W, int OH,
                                               int OW) {
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int ih = 2 * m + i, iw = 2 * j + k;
            q.y[i][k] = (ih < H && iw < W) ? yimg[(static_cast<size_t>(ih) * W + iw) * vecs + cv] : make_uint4(0, 0, 0, 0);
            const int oh = m + i, ow = j + k;
            const bool in = oh < OH && ow < OW;
            const size_t o = (static_cast<size_t>(oh) * OW + ow) * vecs + cv;
            q.d[i][k] = in ? dimg[o] : make_uint4(0, 0, 0, 0);
            q.a[i][k] = in ? aimg[o] : make_uint2(0xffffffffu, 0xffffffffu);
        }
}

// g[k] += d[k] where the window's stored tap equals `tap`
__device__ __forceinline__ void stem_take(float (&g)[8], const uint4& d, const uint2& a, uint32_t tap) {
    float f[8];
    stem_unpack8(d, f);
    const uint32_t want = tap * 0x01010101u;
    const uint32_t ex = a.x ^ want, ey = a.y ^ want;                     // a zero byte = this tap won that channel
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        g[k] += ((ex >> (k * 8)) & 0xffu) == 0 ? f[k] : 0.f;
        g[4 + k] += ((ey >> (k * 8)) & 0xffu) == 0 ? f[4 + k] : 0.f;
    }
}

// pooled gradient of quad pixel (i, k) before the ReLU derivative; taps are a * 3 + b with a = ih + 1 - 2 oh, b = iw + 1 - 2 ow
template <int I, int K>
__device__ __forceinline__ void stem_quad_g(const StemQuad& q, float (&g)[8]) {
#pragma unroll
    for (int c = 0; c < 8; ++c) g[c] = 0.f;
    stem_take(g, q.d[0][0], q.a[0][0], (I + 1) * 3 + (K + 1));                       // window (m, j)
    if (K == 1) stem_take(g, q.d[0][1], q.a[0][1], (I + 1) * 3 + 0);                 // window (m, j + 1)
    if (I == 1) stem_take(g, q.d[1][0], q.a[1][0], 0 * 3 + (K + 1));                 // window (m + 1, j)
    if (I == 1 && K == 1) stem_take(g, q.d[1][1], q.a[1][1], 0);                     // window (m + 1, j + 1)
}

// Both kernels walk `rows_per_block` quad rows per block as one flat index space; 256 % (C / 8) == 0 keeps a thread on one
// channel vector.  REDUCE: sums[STAT_REPLICAS][2C] += (sum g, sum g * xhat).  Otherwise: dy = scale * g + cb * y + cc.
template <bool REDUCE>
__global__ void __launch_bounds__(256, 2) bn_relu_maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dpool,
                                                                     const uint8_t* __restrict__ argmax,
                                                                     const __nv_bfloat16* __restrict__ y,
                                                                     const float* __restrict__ scale, const float* __restrict__ shift,
                                                                     const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                     float* __restrict__ sums, float inv_count,
                                                                     const float* __restrict__ count_dev,
                                                                     const float* __restrict__ param_sums, float* dgamma, float* dbeta,
                                                                     __nv_bfloat16* __restrict__ dy, int H, int W, int C, int OH, int OW,
                                                                     int rows_per_block, int total_qrows) {
    pdl_sync();
    __shared__ float red[REDUCE ? 256 : 1][17];
    const int vecs = C >> 3;
    const int cv = threadIdx.x % vecs;
    const int Hq = (H + 1) >> 1, Wq = (W + 1) >> 1;
    float sc[8], sh[8], p0[8], p1[8];        // REDUCE: p0 = mean, p1 = invstd; apply: p0 = cb, p1 = cc
    stem_ld8(scale + cv * 8, sc);
    stem_ld8(shift + cv * 8, sh);
    stem_ld8(mean + cv * 8, p0);
    stem_ld8(invstd + cv * 8, p1);
    float s1[8], s2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { s1[k] = 0.f; s2[k] = 0.f; }
    if (!REDUCE) {
        if (count_dev != nullptr) inv_count = 1.f / __ldg(count_dev);
        if (blockIdx.x == 0 && dgamma != nullptr) {
            for (int c = threadIdx.x; c < C; c += blockDim.x) {
                dbeta[c] += param_sums[c];
                dgamma[c] += param_sums[C + c];
            }
        }
        float t1[8], t2[8];
        stem_ld8(sums + cv * 8, t1);
        stem_ld8(sums + C + cv * 8, t2);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float cb = -sc[k] * p1[k] * t2[k] * inv_count;
            const float cc = -sc[k] * t1[k] * inv_count - cb * p0[k];
            p0[k] = cb;
            p1[k] = cc;
        }
    }
    const int q_begin = blockIdx.x * rows_per_block;
    const int q_end = min(q_begin + rows_per_block, total_qrows);
    const int per_row = Wq * vecs;
    const int items = (q_end - q_begin) * per_row;
    for (int t = threadIdx.x; t < items; t += blockDim.x) {
        const int qr = q_begin + t / per_row;
        const int j = (t % per_row) / vecs;
        const int n = qr / Hq, m = qr - n * Hq;
        const uint4* yimg = reinterpret_cast<const uint4*>(y) + static_cast<size_t>(n) * H * W * vecs;
        const size_t obase = static_cast<size_t>(n) * OH * OW * vecs;
        StemQuad q;
        stem_load_quad(q, yimg, reinterpret_cast<const uint4*>(dpool) + obase, reinterpret_cast<const uint2*>(argmax) + obase, m, j, cv,
                       vecs, H, W, OH, OW);
        auto pixel = [&](const uint4& yraw, float (&g)[8], int ih, int iw) {
            if (ih >= H || iw >= W) return;
            float yv[8];
            stem_unpack8(yraw, yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = fmaf(yv[k], sc[k], sh[k]) > 0.f ? g[k] : 0.f;
            if (REDUCE) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    s1[k] += g[k];
                    s2[k] = fmaf(g[k] * (yv[k] - p0[k]), p1[k], s2[k]);
                }
            } else {
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = fmaf(sc[k], g[k], fmaf(p0[k], yv[k], p1[k]));
                reinterpret_cast<uint4*>(dy)[(static_cast<size_t>(n) * H * W + static_cast<size_t>(ih) * W + iw) * vecs + cv] =
                    stem_pack8(o);
            }
        };
        float g[8];
        stem_quad_g<0, 0>(q, g); pixel(q.y[0][0], g, 2 * m, 2 * j);
        stem_quad_g<0, 1>(q, g); pixel(q.y[0][1], g, 2 * m, 2 * j + 1);
        stem_quad_g<1, 0>(q, g); pixel(q.y[1][0], g, 2 * m + 1, 2 * j);
        stem_quad_g<1, 1>(q, g); pixel(q.y[1][1], g, 2 * m + 1, 2 * j + 1);
    }
    if (REDUCE) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { red[threadIdx.x][k] = s1[k]; red[threadIdx.x][8 + k] = s2[k]; }
        __syncthreads();
        if (threadIdx.x < vecs * 16) {                    // thread -> (channel vector, one of its 16 partial sums)
            const int v = threadIdx.x / 16, jj = threadIdx.x % 16;
            float acc = 0.f;
            for (int o = v; o < 256; o += vecs) acc += red[o][jj];
            float* rep = sums + static_cast<size_t>(blockIdx.x % STAT_REPLICAS) * 2 * C;
            atomicAdd(rep + (jj < 8 ? 0 : C) + v * 8 + (jj & 7), acc);
        }
    }
}

__global__ void stem_fold_kernel(float* buf, int n, int replicas) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float acc = 0.f;
    for (int r = 0; r < replicas; ++r) acc += buf[static_cast<size_t>(r) * n + i];
    buf[i] = acc;
}

}  // namespace picd

using namespace picd;

static int stem_check(int H, int W, int C, int OH, int OW) {
    PICD_REQUIRE(C % 8 == 0 && 256 % (C / 8) == 0 && C <= 2048, "bn_relu_maxpool: C=%d must be a multiple of 8 with C/8 dividing 256", C);
    PICD_REQUIRE(OH == (H + 2 - 3) / 2 + 1 && OW == (W + 2 - 3) / 2 + 1, "bn_relu_maxpool: output %dx%d does not match a 3x3/s2/p1 "
                 "pool of %dx%d", OH, OW, H, W);
    return 0;
}

// out[N,OH,OW,C] = maxpool3x3/s2/p1(relu(y * scale + shift)), argmax[N,OH,OW,C] bytes (window offset 0..8) for the backward.
extern "C" int picd_bn_relu_maxpool_fwd(const void* y, const float* scale, const float* shift, void* out, void* argmax, int N,
                                        int H, int W, int C, int OH, int OW, cudaStream_t st) {
    if (int rc = stem_check(H, W, C, OH, OW)) return rc;
    PICD_CUDA(launch_pdl<PDL_BN>(bn_relu_maxpool_fwd_kernel, dim3(N * OH), dim3(256), 0, st, static_cast<const __nv_bfloat16*>(y),
                                 scale, shift, static_cast<__nv_bfloat16*>(out), static_cast<uint8_t*>(argmax), H, W, C, OH, OW));
    PICD_LAUNCH_CHECK();
    return 0;
}

// sums[STAT_REPLICAS][2C] += (sum g, sum g * xhat) with g = relu'(y*scale+shift) * maxpool_backward(dpool, argmax): the BatchNorm
// backward reduction of the stem without the 4x larger dz tensor.
extern "C" int picd_bn_relu_maxpool_bwd_reduce(const void* dpool, const void* argmax, const void* y, const float* scale,
                                               const float* shift, const float* mean, const float* invstd, float* sums, int N,
                                               int H, int W, int C, int OH, int OW, cudaStream_t st) {
    if (int rc = stem_check(H, W, C, OH, OW)) return rc;
    const int total_q = N * ((H + 1) / 2);
    const int blocks = std::min(total_q, 8 * num_sms());
    const int rpb = (total_q + blocks - 1) / blocks;
    PICD_CUDA(launch_pdl<PDL_BN>(bn_relu_maxpool_bwd_kernel<true>, dim3((total_q + rpb - 1) / rpb), dim3(256), 0, st,
                                 static_cast<const __nv_bfloat16*>(dpool), static_cast<const uint8_t*>(argmax),
                                 static_cast<const __nv_bfloat16*>(y), scale, shift, mean, invstd, sums, 0.f,
                                 static_cast<const float*>(nullptr), static_cast<const float*>(nullptr), static_cast<float*>(nullptr),
                                 static_cast<float*>(nullptr), static_cast<__nv_bfloat16*>(nullptr), H, W, C, OH, OW, rpb, total_q));
    PICD_LAUNCH_CHECK();
    return 0;
}

// dy = BatchNorm-backward(g) with g recomputed as above.  `sums` ([sums_replicas][2C], folded in place) are the (global) sums,
// `param_sums` ([param_replicas][2C], folded in place; may alias sums) this replica's, added into dgamma / dbeta.  count_dev
// (nullable): device pointer to the global element count, replaces inv_count (SyncBatchNorm).
extern "C" int picd_bn_relu_maxpool_bwd_apply(const void* dpool, const void* argmax, const void* y, const float* scale,
                                              const float* shift, const float* mean, const float* invstd, float* sums,
                                              int sums_replicas, float inv_count, const float* count_dev, float* param_sums,
                                              int param_replicas, float* dgamma, float* dbeta, void* dy, int N, int H, int W, int C,
                                              int OH, int OW, cudaStream_t st) {
    if (int rc = stem_check(H, W, C, OH, OW)) return rc;
    if (param_sums != nullptr && param_sums != sums && param_replicas > 1)
        PICD_CUDA(launch_pdl<PDL_BN>(stem_fold_kernel, dim3((2 * C + 255) / 256), dim3(256), 0, st, param_sums, 2 * C, param_replicas));
    if (sums_replicas > 1)
        PICD_CUDA(launch_pdl<PDL_BN>(stem_fold_kernel, dim3((2 * C + 255) / 256), dim3(256), 0, st, sums, 2 * C, sums_replicas));
    const int total_q = N * ((H + 1) / 2);
    const int rpb = 4;                                    // W = 112, C = 64: 4 * 56 * 8 items = 7 full sweeps of the block
    PICD_CUDA(launch_pdl<PDL_BN>(bn_relu_maxpool_bwd_kernel<false>, dim3((total_q + rpb - 1) / rpb), dim3(256), 0, st,
                                 static_cast<const __nv_bfloat16*>(dpool), static_cast<const uint8_t*>(argmax),
                                 static_cast<const __nv_bfloat16*>(y), scale, shift, mean, invstd, sums, inv_count, count_dev,
                                 static_cast<const float*>(param_sums), dgamma, dbeta, static_cast<__nv_bfloat16*>(dy), H, W, C, OH, OW,
                                 rpb, total_q));
    PICD_LAUNCH_CHECK();
    return 0;
}
