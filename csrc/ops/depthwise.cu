// Depthwise convolution (EfficientNet MBConv) for NHWC bf16: one input channel per output channel, so there is no GEMM
// and no tensor-core shape -- it is a bandwidth-bound stencil.  Thread = 8 channels (one 16-byte vector) of one pixel;
// weights are passed tap-major, fp32: wt[tap][C].  fprop and dgrad are gathers (no atomics); wgrad keeps the per-tap
// partial sums of a thread's 8 channels in registers over a strip of pixels and folds them with fp32 atomics.
#include "../common/picd_common.cuh"

namespace picd {

__device__ __forceinline__ void dw_unpack8(const uint4& raw, float (&f)[8]) {
    float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 dw_pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}

__global__ void __launch_bounds__(256) dw_fprop_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ wt,
                                                       __nv_bfloat16* __restrict__ y, int N, int H, int W, int C, int OH,
                                                       int OW, int K, int stride, int pad_t, int pad_l) {
    const int vecs = C >> 3;
    const long long total = static_cast<long long>(N) * OH * OW * vecs;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
        const int cv = static_cast<int>(i % vecs);
        long long t = i / vecs;
        const int ow = static_cast<int>(t % OW); t /= OW;
        const int oh = static_cast<int>(t % OH);
        const int n = static_cast<int>(t / OH);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int r = 0; r < K; ++r) {
            const int ih = oh * stride - pad_t + r;
            if (ih < 0 || ih >= H) continue;
            for (int q = 0; q < K; ++q) {
                const int iw = ow * stride - pad_l + q;
                if (iw < 0 || iw >= W) continue;
                float f[8];
                dw_unpack8(*reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * H + ih) * W + iw) * C + cv * 8), f);
                const float* wp = wt + static_cast<long long>(r * K + q) * C + cv * 8;
                const float4 w0 = *reinterpret_cast<const float4*>(wp), w1 = *reinterpret_cast<const float4*>(wp + 4);
                acc[0] = fmaf(f[0], w0.x, acc[0]); acc[1] = fmaf(f[1], w0.y, acc[1]);
                acc[2] = fmaf(f[2], w0.z, acc[2]); acc[3] = fmaf(f[3], w0.w, acc[3]);
                acc[4] = fmaf(f[4], w1.x, acc[4]); acc[5] = fmaf(f[5], w1.y, acc[5]);
                acc[6] = fmaf(f[6], w1.z, acc[6]); acc[7] = fmaf(f[7], w1.w, acc[7]);
            }
        }
        *reinterpret_cast<uint4*>(y + i * 8) = dw_pack8(acc);
    }
}

__global__ void __launch_bounds__(256) dw_dgrad_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ wt,
                                                       __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int OH,
                                                       int OW, int K, int stride, int pad_t, int pad_l) {
    const int vecs = C >> 3;
    const long long total = static_cast<long long>(N) * H * W * vecs;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
        const int cv = static_cast<int>(i % vecs);
        long long t = i / vecs;
        const int iw = static_cast<int>(t % W); t /= W;
        const int ih = static_cast<int>(t % H);
        const int n = static_cast<int>(t / H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int r = (ih + pad_t) % stride; r < K; r += stride) {
            const int th = ih + pad_t - r;
            if (th < 0) break;
            const int oh = th / stride;
            if (oh >= OH) continue;
            for (int q = (iw + pad_l) % stride; q < K; q += stride) {
                const int tw = iw + pad_l - q;
                if (tw < 0) break;
                const int ow = tw / stride;
                if (ow >= OW) continue;
                float g[8];
                dw_unpack8(*reinterpret_cast<const uint4*>(dy + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + cv * 8), g);
                const float* wp = wt + static_cast<long long>(r * K + q) * C + cv * 8;
                const float4 w0 = *reinterpret_cast<const float4*>(wp), w1 = *reinterpret_cast<const float4*>(wp + 4);
                acc[0] = fmaf(g[0], w0.x, acc[0]); acc[1] = fmaf(g[1], w0.y, acc[1]);
                acc[2] = fmaf(g[2], w0.z, acc[2]); acc[3] = fmaf(g[3], w0.w, acc[3]);
                acc[4] = fmaf(g[4], w1.x, acc[4]); acc[5] = fmaf(g[5], w1.y, acc[5]);
                acc[6] = fmaf(g[6], w1.z, acc[6]); acc[7] = fmaf(g[7], w1.w, acc[7]);
            }
        }
        *reinterpret_cast<uint4*>(dx + i * 8) = dw_pack8(acc);
    }
}

constexpr int DW_TAPS_PER_PASS = 9;
// dwt[tap][C] += sum over output pixels of dy * x(shifted).  blockIdx = (channel-vector block, pixel strip, tap group).
__global__ void __launch_bounds__(256) dw_wgrad_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                       float* __restrict__ dwt, int N, int H, int W, int C, int OH, int OW,
                                                       int K, int stride, int pad_t, int pad_l, int tx_count, int ty_count,
                                                       long long pix_per_block) {
    extern __shared__ float dw_red[];          // [ty_count][tx_count][8]
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    const int tap0 = blockIdx.z * DW_TAPS_PER_PASS;
    const int ntap = min(DW_TAPS_PER_PASS, K * K - tap0);
    float acc[DW_TAPS_PER_PASS][8];
#pragma unroll
    for (int t = 0; t < DW_TAPS_PER_PASS; ++t)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[t][k] = 0.f;
    const bool active = ty < ty_count && cv < vecs;
    if (active) {
        const long long P = static_cast<long long>(N) * OH * OW;
        const long long p0 = blockIdx.y * pix_per_block;
        long long p1 = p0 + pix_per_block;
        if (p1 > P) p1 = P;
        for (long long pix = p0 + ty; pix < p1; pix += ty_count) {
            long long t = pix;
            const int ow = static_cast<int>(t % OW); t /= OW;
            const int oh = static_cast<int>(t % OH);
            const int n = static_cast<int>(t / OH);
            float g[8];
            dw_unpack8(*reinterpret_cast<const uint4*>(dy + pix * C + cv * 8), g);
#pragma unroll
            for (int tt = 0; tt < DW_TAPS_PER_PASS; ++tt) {
                if (tt < ntap) {
                    const int tap = tap0 + tt;
                    const int ih = oh * stride - pad_t + tap / K, iw = ow * stride - pad_l + tap % K;
                    if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
                        float f[8];
                        dw_unpack8(*reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * H + ih) * W + iw) * C + cv * 8), f);
#pragma unroll
                        for (int k = 0; k < 8; ++k) acc[tt][k] = fmaf(g[k], f[k], acc[tt][k]);
                    }
                }
            }
        }
    }
    // fold the row dimension of the block through shared memory, one tap at a time
#pragma unroll
    for (int tt = 0; tt < DW_TAPS_PER_PASS; ++tt) {
        if (tt >= ntap) break;                 // uniform over the block
        __syncthreads();
        if (ty < ty_count) {
#pragma unroll
            for (int k = 0; k < 8; ++k) dw_red[(ty * tx_count + tx) * 8 + k] = acc[tt][k];
        }
        __syncthreads();
        if (active && ty == 0) {
            float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] += dw_red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(dwt + static_cast<long long>(tap0 + tt) * C + cv * 8 + k, s[k]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Row-structured specialisations (K = 3 / 5, stride 1 / 2: every depthwise layer of EfficientNet).  The generic kernels
// above spend most of their instructions on 64-bit div/mod and run-time tap arithmetic (the wgrad ran at ~8% of the
// memory roofline).  Here blockIdx.y walks (image, row) pairs, a thread owns 8 channels and strides over the pixels of a
// row, the tap loops are compile-time, and the block's weights sit in shared memory.
template <int K, int S>
__global__ void __launch_bounds__(256) dw_fprop_rows_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ wt,
                                                            __nv_bfloat16* __restrict__ y, int NR, int H, int W, int C, int OH,
                                                            int OW, int pad_t, int pad_l, int tx_count, int ty_count,
                                                            int rows_per_block) {
    extern __shared__ float dw_sw[];           // [K*K][tx_count*8]
    const int vecs = C >> 3, cw = tx_count * 8;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    for (int i = threadIdx.x; i < K * K * cw; i += blockDim.x) {
        const int tap = i / cw, c = blockIdx.x * cw + (i - tap * cw);
        dw_sw[i] = c < C ? wt[static_cast<size_t>(tap) * C + c] : 0.f;
    }
    __syncthreads();
    if (ty >= ty_count || cv >= vecs) return;
    const float* swc = dw_sw + tx * 8;
    int row1 = (blockIdx.y + 1) * rows_per_block;
    if (row1 > NR) row1 = NR;
    for (int row = blockIdx.y * rows_per_block; row < row1; ++row) {
        const int n = row / OH, oh = row - n * OH;
        const __nv_bfloat16* xn = x + static_cast<size_t>(n) * H * W * C + cv * 8;
        __nv_bfloat16* yr = y + static_cast<size_t>(row) * OW * C + cv * 8;
        for (int ow = ty; ow < OW; ow += ty_count) {
            float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int r = 0; r < K; ++r) {
                const int ih = oh * S - pad_t + r;
                if (ih < 0 || ih >= H) continue;
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    const int iw = ow * S - pad_l + q;
                    if (iw < 0 || iw >= W) continue;
                    float f[8];
                    dw_unpack8(*reinterpret_cast<const uint4*>(xn + (static_cast<size_t>(ih) * W + iw) * C), f);
                    const float4 w0 = *reinterpret_cast<const float4*>(swc + (r * K + q) * cw);
                    const float4 w1 = *reinterpret_cast<const float4*>(swc + (r * K + q) * cw + 4);
                    acc[0] = fmaf(f[0], w0.x, acc[0]); acc[1] = fmaf(f[1], w0.y, acc[1]);
                    acc[2] = fmaf(f[2], w0.z, acc[2]); acc[3] = fmaf(f[3], w0.w, acc[3]);
                    acc[4] = fmaf(f[4], w1.x, acc[4]); acc[5] = fmaf(f[5], w1.y, acc[5]);
                    acc[6] = fmaf(f[6], w1.z, acc[6]); acc[7] = fmaf(f[7], w1.w, acc[7]);
                }
            }
            *reinterpret_cast<uint4*>(yr + static_cast<size_t>(ow) * C) = dw_pack8(acc);
        }
    }
}

// dx[n, ih, iw] = sum over taps (r, q) with (ih + pad_t - r) and (iw + pad_l - q) divisible by S of dy[n, ./S, ./S] * w[r, q]
template <int K, int S>
__global__ void __launch_bounds__(256) dw_dgrad_rows_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ wt,
                                                            __nv_bfloat16* __restrict__ dx, int NR, int H, int W, int C, int OH,
                                                            int OW, int pad_t, int pad_l, int tx_count, int ty_count,
                                                            int rows_per_block) {
    extern __shared__ float dw_sw[];
    const int vecs = C >> 3, cw = tx_count * 8;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    for (int i = threadIdx.x; i < K * K * cw; i += blockDim.x) {
        const int tap = i / cw, c = blockIdx.x * cw + (i - tap * cw);
        dw_sw[i] = c < C ? wt[static_cast<size_t>(tap) * C + c] : 0.f;
    }
    __syncthreads();
    if (ty >= ty_count || cv >= vecs) return;
    const float* swc = dw_sw + tx * 8;
    int row1 = (blockIdx.y + 1) * rows_per_block;
    if (row1 > NR) row1 = NR;
    for (int row = blockIdx.y * rows_per_block; row < row1; ++row) {     // row = n * H + ih
        const int n = row / H, ih = row - n * H;
        const __nv_bfloat16* gn = dy + static_cast<size_t>(n) * OH * OW * C + cv * 8;
        __nv_bfloat16* dr = dx + static_cast<size_t>(row) * W * C + cv * 8;
        for (int iw = ty; iw < W; iw += ty_count) {
            float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int r = 0; r < K; ++r) {
                const int th = ih + pad_t - r;
                if (th < 0 || (S > 1 && (th % S) != 0)) continue;
                const int oh = th / S;
                if (oh >= OH) continue;
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    const int tw = iw + pad_l - q;
                    if (tw < 0 || (S > 1 && (tw % S) != 0)) continue;
                    const int ow = tw / S;
                    if (ow >= OW) continue;
                    float g[8];
                    dw_unpack8(*reinterpret_cast<const uint4*>(gn + (static_cast<size_t>(oh) * OW + ow) * C), g);
                    const float4 w0 = *reinterpret_cast<const float4*>(swc + (r * K + q) * cw);
                    const float4 w1 = *reinterpret_cast<const float4*>(swc + (r * K + q) * cw + 4);
                    acc[0] = fmaf(g[0], w0.x, acc[0]); acc[1] = fmaf(g[1], w0.y, acc[1]);
                    acc[2] = fmaf(g[2], w0.z, acc[2]); acc[3] = fmaf(g[3], w0.w, acc[3]);
                    acc[4] = fmaf(g[4], w1.x, acc[4]); acc[5] = fmaf(g[5], w1.y, acc[5]);
                    acc[6] = fmaf(g[6], w1.z, acc[6]); acc[7] = fmaf(g[7], w1.w, acc[7]);
                }
            }
            *reinterpret_cast<uint4*>(dr + static_cast<size_t>(iw) * C) = dw_pack8(acc);
        }
    }
}

// dwt[(r, q)][c] += sum over output pixels of dy[n, oh, ow, c] * x[n, oh*S - pad_t + r, ow*S - pad_l + q, c].
// blockIdx.z = filter row r: a thread keeps the K taps of that row for its 8 channels in registers over a strip of output
// rows, folds them across the block through shared memory and adds them with one fp32 atomic per (tap, channel) per block.
template <int K, int S>
__global__ void __launch_bounds__(256) dw_wgrad_rows_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                            float* __restrict__ dwt, int NR, int H, int W, int C, int OH,
                                                            int OW, int pad_t, int pad_l, int tx_count, int ty_count,
                                                            int rows_per_block) {
    extern __shared__ float dw_red[];          // [ty_count][tx_count][8]
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    const int r = blockIdx.z;
    float acc[K][8];
#pragma unroll
    for (int q = 0; q < K; ++q)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[q][k] = 0.f;
    const bool active = ty < ty_count && cv < vecs;
    if (active) {
        int row1 = (blockIdx.y + 1) * rows_per_block;
        if (row1 > NR) row1 = NR;
        for (int row = blockIdx.y * rows_per_block; row < row1; ++row) {     // row = n * OH + oh
            const int n = row / OH, oh = row - n * OH;
            const int ih = oh * S - pad_t + r;
            if (ih < 0 || ih >= H) continue;
            const __nv_bfloat16* xr = x + (static_cast<size_t>(n) * H + ih) * W * C + cv * 8;
            const __nv_bfloat16* gr = dy + static_cast<size_t>(row) * OW * C + cv * 8;
            for (int ow = ty; ow < OW; ow += ty_count) {
                float g[8];
                dw_unpack8(*reinterpret_cast<const uint4*>(gr + static_cast<size_t>(ow) * C), g);
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    const int iw = ow * S - pad_l + q;
                    if (iw < 0 || iw >= W) continue;
                    float f[8];
                    dw_unpack8(*reinterpret_cast<const uint4*>(xr + static_cast<size_t>(iw) * C), f);
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[q][k] = fmaf(g[k], f[k], acc[q][k]);
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < K; ++q) {
        __syncthreads();
        if (ty < ty_count) {
#pragma unroll
            for (int k = 0; k < 8; ++k) dw_red[(ty * tx_count + tx) * 8 + k] = acc[q][k];
        }
        __syncthreads();
        if (active && ty == 0) {
            float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] += dw_red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(dwt + static_cast<size_t>(r * K + q) * C + cv * 8 + k, s[k]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Stride-1 STRIP kernels (every 3x3 / 5x5 stride-1 depthwise layer): a thread owns FOUR consecutive output pixels of a row for
// its 8 channels.  Per filter row it loads the K + 3 input vectors those four outputs need ONCE (the row kernels above loaded
// and unpacked K per output, 2.5x as many for K = 5) and all loads of a filter row are issued together.  dgrad is the same
// kernel on dy with the filter flipped (kFlip) and the complementary padding.
template <int K, bool kFlip>
__global__ void __launch_bounds__(256, 2) dw_conv_strip_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ wt,
                                                               __nv_bfloat16* __restrict__ y, int NR, int H, int W, int C, int OH,
                                                               int OW, int pad_t, int pad_l, int tx_count, int ty_count,
                                                               int rows_per_block) {
    extern __shared__ float dw_sw[];           // [K*K][tx_count*8]
    const int vecs = C >> 3, cw = tx_count * 8;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    // shared weights as [tap][half][tx][4]: a thread's two float4 loads per tap then sit 16 bytes from its neighbour's (the
    // [tap][tx][8] layout put them 32 bytes apart: a 2-way bank conflict on every load)
    const int hw = cw >> 1;
    for (int i = threadIdx.x; i < K * K * cw; i += blockDim.x) {
        const int tap = i / cw, cl = i - tap * cw, c = blockIdx.x * cw + cl;
        const int src_tap = kFlip ? (K * K - 1 - tap) : tap;
        dw_sw[tap * cw + ((cl >> 2) & 1) * hw + (cl >> 3) * 4 + (cl & 3)] = c < C ? wt[static_cast<size_t>(src_tap) * C + c] : 0.f;
    }
    __syncthreads();
    if (ty >= ty_count || cv >= vecs) return;
    const float* swc = dw_sw + tx * 4;
    const int row0 = blockIdx.y * rows_per_block;
    const int nrows = min(rows_per_block, NR - row0);
    const int nstrips = (OW + 3) >> 2;
    for (int it = ty; it < nrows * nstrips; it += ty_count) {       // (row, strip) pairs of the block, flattened
        const int lr = it / nstrips, st = it - lr * nstrips;
        const int row = row0 + lr;
        const int n = row / OH, oh = row - n * OH;
        const int ow0 = st << 2;
        const __nv_bfloat16* xn = x + static_cast<size_t>(n) * H * W * C + cv * 8;
        float acc[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[j][k] = 0.f;
#pragma unroll
        for (int r = 0; r < K; ++r) {
            const int ih = oh - pad_t + r;
            if (ih < 0 || ih >= H) continue;
            const __nv_bfloat16* xr = xn + static_cast<size_t>(ih) * W * C;
            uint4 raw[K + 3];
#pragma unroll
            for (int j = 0; j < K + 3; ++j) {
                const int iw = ow0 - pad_l + j;
                raw[j] = (iw >= 0 && iw < W) ? *reinterpret_cast<const uint4*>(xr + static_cast<size_t>(iw) * C) : make_uint4(0, 0, 0, 0);
            }
            float f[K + 3][8];
#pragma unroll
            for (int j = 0; j < K + 3; ++j) dw_unpack8(raw[j], f[j]);
#pragma unroll
            for (int q = 0; q < K; ++q) {
                const float4 w0 = *reinterpret_cast<const float4*>(swc + (r * K + q) * cw);
                const float4 w1 = *reinterpret_cast<const float4*>(swc + (r * K + q) * cw + hw);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    acc[j][0] = fmaf(f[j + q][0], w0.x, acc[j][0]); acc[j][1] = fmaf(f[j + q][1], w0.y, acc[j][1]);
                    acc[j][2] = fmaf(f[j + q][2], w0.z, acc[j][2]); acc[j][3] = fmaf(f[j + q][3], w0.w, acc[j][3]);
                    acc[j][4] = fmaf(f[j + q][4], w1.x, acc[j][4]); acc[j][5] = fmaf(f[j + q][5], w1.y, acc[j][5]);
                    acc[j][6] = fmaf(f[j + q][6], w1.z, acc[j][6]); acc[j][7] = fmaf(f[j + q][7], w1.w, acc[j][7]);
                }
            }
        }
        __nv_bfloat16* yr = y + static_cast<size_t>(row) * OW * C + cv * 8;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (ow0 + j < OW) *reinterpret_cast<uint4*>(yr + static_cast<size_t>(ow0 + j) * C) = dw_pack8(acc[j]);
    }
}

// wgrad, stride 1: blockIdx.z = filter row r; a thread takes strips of four output pixels: 4 dy + (K + 3) x vectors feed
// 4 * K * 8 FMAs (the row kernel: 1 + K vectors per K * 8 FMAs).
template <int K>
__global__ void __launch_bounds__(256, 2) dw_wgrad_strip_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                                float* __restrict__ dwt, int NR, int H, int W, int C, int OH, int OW,
                                                                int pad_t, int pad_l, int tx_count, int ty_count, int rows_per_block) {
    extern __shared__ float dw_red[];          // [ty_count][tx_count][8]
    const int vecs = C >> 3;
    const int tx = threadIdx.x % tx_count, ty = threadIdx.x / tx_count;
    const int cv = blockIdx.x * tx_count + tx;
    const int r = blockIdx.z;
    float acc[K][8];
#pragma unroll
    for (int q = 0; q < K; ++q)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[q][k] = 0.f;
    const bool active = ty < ty_count && cv < vecs;
    if (active) {
        const int row0 = blockIdx.y * rows_per_block;
        const int nrows = min(rows_per_block, NR - row0);
        const int nstrips = (OW + 3) >> 2;
        for (int it = ty; it < nrows * nstrips; it += ty_count) {
            const int lr = it / nstrips, st = it - lr * nstrips;
            const int row = row0 + lr;                                   // row = n * OH + oh
            const int n = row / OH, oh = row - n * OH;
            const int ih = oh - pad_t + r;
            if (ih < 0 || ih >= H) continue;
            const int ow0 = st << 2;
            const __nv_bfloat16* xr = x + (static_cast<size_t>(n) * H + ih) * W * C + cv * 8;
            const __nv_bfloat16* gr = dy + static_cast<size_t>(row) * OW * C + cv * 8;
            uint4 graw[4], xraw[K + 3];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                graw[j] = (ow0 + j < OW) ? *reinterpret_cast<const uint4*>(gr + static_cast<size_t>(ow0 + j) * C) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int j = 0; j < K + 3; ++j) {
                const int iw = ow0 - pad_l + j;
                xraw[j] = (iw >= 0 && iw < W) ? *reinterpret_cast<const uint4*>(xr + static_cast<size_t>(iw) * C) : make_uint4(0, 0, 0, 0);
            }
            float f[K + 3][8];
#pragma unroll
            for (int j = 0; j < K + 3; ++j) dw_unpack8(xraw[j], f[j]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float g[8];
                dw_unpack8(graw[j], g);
#pragma unroll
                for (int q = 0; q < K; ++q)
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[q][k] = fmaf(g[k], f[j + q][k], acc[q][k]);
            }
        }
    }
#pragma unroll
    for (int q = 0; q < K; ++q) {
        __syncthreads();
        if (ty < ty_count) {
#pragma unroll
            for (int k = 0; k < 8; ++k) dw_red[(ty * tx_count + tx) * 8 + k] = acc[q][k];
        }
        __syncthreads();
        if (active && ty == 0) {
            float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int yy = 0; yy < ty_count; ++yy)
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] += dw_red[(yy * tx_count + tx) * 8 + k];
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(dwt + static_cast<size_t>(r * K + q) * C + cv * 8 + k, s[k]);
        }
    }
}

struct DwShape { int tx, ty, gx, gy, rpb; };
// thread block = tx channel vectors x ty pixel lanes; blockIdx.y strips of rows sized for ~`waves` blocks per SM
static DwShape dw_shape(int C, int rows, int waves, int gz) {
    DwShape d;
    const int vecs = C / 8;
    d.tx = vecs < 32 ? vecs : 32;
    d.ty = 256 / d.tx;
    d.gx = (vecs + d.tx - 1) / d.tx;
    long long gy = (static_cast<long long>(waves) * num_sms()) / (static_cast<long long>(d.gx) * gz);
    if (gy < 1) gy = 1;
    if (gy > rows) gy = rows;
    d.rpb = static_cast<int>((rows + gy - 1) / gy);
    d.gy = (rows + d.rpb - 1) / d.rpb;
    return d;
}

static int dw_grid(long long work) {
    long long b = (work + 255) / 256;
    const long long cap = 16LL * num_sms();
    if (b > cap) b = cap;
    return static_cast<int>(b < 1 ? 1 : b);
}

}  // namespace picd

using namespace picd;

// y[N,OH,OW,C] = depthwise KxK conv of x[N,H,W,C]; wt = weights as fp32 [K*K][C]; padding may be asymmetric
// (pad_t/pad_l given, bottom/right implied by OH/OW), which is how TF-"same" stride-2 layers are expressed.
extern "C" int picd_depthwise_fprop(const void* x, const float* wt, void* y, int N, int H, int W, int C, int OH, int OW, int K,
                                    int stride, int pad_t, int pad_l, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "depthwise: C=%d must be a multiple of 8", C);
    static const bool strips = [] { const char* e = getenv("PICD_DW_STRIPS"); return e == nullptr || atoi(e) != 0; }();
#define PICD_DW_FPROP_STRIP(KK)                                                                                        \
    if (strips && K == KK && stride == 1) {                                                                              \
        const DwShape d = dw_shape(C, N * OH, 8, 1);                                                                      \
        dw_conv_strip_kernel<KK, false><<<dim3(d.gx, d.gy), 256, KK * KK * d.tx * 8 * sizeof(float), st>>>(                \
            static_cast<const __nv_bfloat16*>(x), wt, static_cast<__nv_bfloat16*>(y), N * OH, H, W, C, OH, OW, pad_t, pad_l, \
            d.tx, d.ty, d.rpb);                                                                                           \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_DW_FPROP_STRIP(3) PICD_DW_FPROP_STRIP(5)
#undef PICD_DW_FPROP_STRIP
#define PICD_DW_FPROP(KK, SS)                                                                                          \
    if (K == KK && stride == SS) {                                                                                       \
        const DwShape d = dw_shape(C, N * OH, 8, 1);                                                                      \
        dw_fprop_rows_kernel<KK, SS><<<dim3(d.gx, d.gy), 256, KK * KK * d.tx * 8 * sizeof(float), st>>>(                   \
            static_cast<const __nv_bfloat16*>(x), wt, static_cast<__nv_bfloat16*>(y), N * OH, H, W, C, OH, OW, pad_t, pad_l, \
            d.tx, d.ty, d.rpb);                                                                                           \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_DW_FPROP(3, 1) PICD_DW_FPROP(3, 2) PICD_DW_FPROP(5, 1) PICD_DW_FPROP(5, 2)
#undef PICD_DW_FPROP
    const long long total = 1LL * N * OH * OW * (C / 8);
    dw_fprop_kernel<<<dw_grid(total), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), wt, static_cast<__nv_bfloat16*>(y), N, H, W,
                                                   C, OH, OW, K, stride, pad_t, pad_l);
    PICD_LAUNCH_CHECK();
    return 0;
}
extern "C" int picd_depthwise_dgrad(const void* dy, const float* wt, void* dx, int N, int H, int W, int C, int OH, int OW, int K,
                                    int stride, int pad_t, int pad_l, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "depthwise: C=%d must be a multiple of 8", C);
    static const bool strips = [] { const char* e = getenv("PICD_DW_STRIPS"); return e == nullptr || atoi(e) != 0; }();
    // stride 1: dx = dy convolved with the flipped filter under the complementary padding; "input" of that convolution is dy
    // [OH, OW], its output dx [H, W]
#define PICD_DW_DGRAD_STRIP(KK)                                                                                        \
    if (strips && K == KK && stride == 1) {                                                                              \
        const DwShape d = dw_shape(C, N * H, 8, 1);                                                                       \
        dw_conv_strip_kernel<KK, true><<<dim3(d.gx, d.gy), 256, KK * KK * d.tx * 8 * sizeof(float), st>>>(                 \
            static_cast<const __nv_bfloat16*>(dy), wt, static_cast<__nv_bfloat16*>(dx), N * H, OH, OW, C, H, W,            \
            KK - 1 - pad_t, KK - 1 - pad_l, d.tx, d.ty, d.rpb);                                                           \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_DW_DGRAD_STRIP(3) PICD_DW_DGRAD_STRIP(5)
#undef PICD_DW_DGRAD_STRIP
#define PICD_DW_DGRAD(KK, SS)                                                                                          \
    if (K == KK && stride == SS) {                                                                                       \
        const DwShape d = dw_shape(C, N * H, 8, 1);                                                                       \
        dw_dgrad_rows_kernel<KK, SS><<<dim3(d.gx, d.gy), 256, KK * KK * d.tx * 8 * sizeof(float), st>>>(                   \
            static_cast<const __nv_bfloat16*>(dy), wt, static_cast<__nv_bfloat16*>(dx), N * H, H, W, C, OH, OW, pad_t, pad_l, \
            d.tx, d.ty, d.rpb);                                                                                           \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_DW_DGRAD(3, 1) PICD_DW_DGRAD(3, 2) PICD_DW_DGRAD(5, 1) PICD_DW_DGRAD(5, 2)
#undef PICD_DW_DGRAD
    const long long total = 1LL * N * H * W * (C / 8);
    dw_dgrad_kernel<<<dw_grid(total), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dy), wt, static_cast<__nv_bfloat16*>(dx), N, H, W,
                                                   C, OH, OW, K, stride, pad_t, pad_l);
    PICD_LAUNCH_CHECK();
    return 0;
}
// dwt (fp32 [K*K][C]) += ...
extern "C" int picd_depthwise_wgrad(const void* x, const void* dy, float* dwt, int N, int H, int W, int C, int OH, int OW, int K,
                                    int stride, int pad_t, int pad_l, cudaStream_t st) {
    PICD_REQUIRE(C % 8 == 0, "depthwise: C=%d must be a multiple of 8", C);
    static const bool strips = [] { const char* e = getenv("PICD_DW_STRIPS"); return e == nullptr || atoi(e) != 0; }();
#define PICD_DW_WGRAD_STRIP(KK)                                                                                        \
    if (strips && K == KK && stride == 1) {                                                                              \
        const DwShape d = dw_shape(C, N * OH, 4, KK);                                                                     \
        dw_wgrad_strip_kernel<KK><<<dim3(d.gx, d.gy, KK), 256, d.tx * d.ty * 8 * sizeof(float), st>>>(                     \
            static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), dwt, N * OH, H, W, C, OH, OW, pad_t,  \
            pad_l, d.tx, d.ty, d.rpb);                                                                                    \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_DW_WGRAD_STRIP(3) PICD_DW_WGRAD_STRIP(5)
#undef PICD_DW_WGRAD_STRIP
#define PICD_DW_WGRAD(KK, SS)                                                                                          \
    if (K == KK && stride == SS) {                                                                                       \
        const DwShape d = dw_shape(C, N * OH, 4, KK);                                                                     \
        dw_wgrad_rows_kernel<KK, SS><<<dim3(d.gx, d.gy, KK), 256, d.tx * d.ty * 8 * sizeof(float), st>>>(                  \
            static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), dwt, N * OH, H, W, C, OH, OW, pad_t,  \
            pad_l, d.tx, d.ty, d.rpb);                                                                                    \
        PICD_LAUNCH_CHECK();                                                                                              \
        return 0;                                                                                                         \
    }
    PICD_DW_WGRAD(3, 1) PICD_DW_WGRAD(3, 2) PICD_DW_WGRAD(5, 1) PICD_DW_WGRAD(5, 2)
#undef PICD_DW_WGRAD
    const int vecs = C / 8;
    const int tx = vecs < 32 ? vecs : 32, ty = 256 / tx;
    const int gx = (vecs + tx - 1) / tx;
    const long long P = 1LL * N * OH * OW;
    long long gy = (4LL * num_sms()) / gx;
    if (gy < 1) gy = 1;
    long long ppb = (P + gy - 1) / gy;
    if (ppb < ty) ppb = ty;
    gy = (P + ppb - 1) / ppb;
    const int gz = (K * K + DW_TAPS_PER_PASS - 1) / DW_TAPS_PER_PASS;
    dw_wgrad_kernel<<<dim3(gx, static_cast<unsigned>(gy), gz), 256, tx * ty * 8 * sizeof(float), st>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), dwt, N, H, W, C, OH, OW, K, stride, pad_t, pad_l,
        tx, ty, ppb);
    PICD_LAUNCH_CHECK();
    return 0;
}
