// Multi-head self-attention for short sequences (ViT: 197 tokens, head dim 64) on tcgen05 + TMEM + TMA.
//
// The reference has no attention; BASELINE.json config 4 (ViT-B/16) asks for it on the tensor cores.  The whole key/value
// set of one head fits in shared memory, so there is no online-softmax loop: per (batch, head, 128-query tile)
//   forward : S = Q K^T (one UMMA chain into TMEM) -> exact row softmax out of TMEM (thread = row) -> P (bf16) into a
//             128B-swizzled smem tile -> O = P V (V read through an MN-major descriptor exactly as TMA landed it) -> O / l
//   backward: recompute S, dP = dO V^T, dS = P o (dP - delta) * scale, then
//             attn_bwd_dq   : dQ = dS K          (one CTA per query tile)
//             attn_bwd_dkdv : dK = dS^T Q, dV = P^T dO accumulated over the query tiles (one CTA per key tile)
// Q, K, V are read straight out of the packed QKV projection output [B*S, 3*H*64] with 2-D TMA boxes (no head-split
// copies) and O / dQKV are written in the layouts the surrounding GEMMs consume ([B*S, H*64] and [B*S, 3*H*64]).
// Out-of-range key columns are masked in the softmax, out-of-range query rows are simply not stored.
#include "../common/picd_common.cuh"

#include <algorithm>
#include <cstring>

namespace picd {

constexpr int ATT_D = 64;          // head dimension (fixed: one 128-byte swizzle row)
constexpr int ATT_TILE = 128;      // queries (or keys) per CTA
constexpr int ATT_MAX_S = 256;

struct AttnParams {
    CUtensorMap qkv_map;     // [B*S rows, 3*H*64 cols] bf16, box (64, 128)
    CUtensorMap do_map;      // [B*S rows, H*64 cols]  bf16, box (64, 128)   (backward)
    __nv_bfloat16* o;        // fwd out / bwd in   [B*S, H*64]
    const __nv_bfloat16* dout;
    __nv_bfloat16* dqkv;     // [B*S, 3*H*64]
    float* lse;              // [B, H, S]  log2-domain: m*scale*log2e + log2(l)
    float* delta;            // [B, H, S]  rowsum(dO o O)
    int B, H, S, n_keys;     // n_keys = S rounded up to 16
    float scale_log2e, scale;
};

struct AttBars {
    uint64_t load, mma0, mma1, mma2;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint8_t* att_align(uint8_t* p) {
    return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~static_cast<uintptr_t>(1023));
}
// thread = row: write 32 consecutive columns (bf16) of a K-major SW128 tile (64-column chunks of 16 KB)
__device__ __forceinline__ void store_row_chunk32(uint32_t tile_u32, int row, int col0, const float (&f)[32]) {
    const uint32_t dst = tile_u32 + (col0 >> 6) * 16384 + row * 128;
    const uint32_t half = static_cast<uint32_t>((col0 >> 5) & 1) * 64;
    const uint32_t rx = static_cast<uint32_t>(row & 7) << 4;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const uint32_t a0 = pack_bf16x2(f[8 * c + 0], f[8 * c + 1]), a1 = pack_bf16x2(f[8 * c + 2], f[8 * c + 3]);
        const uint32_t a2 = pack_bf16x2(f[8 * c + 4], f[8 * c + 5]), a3 = pack_bf16x2(f[8 * c + 6], f[8 * c + 7]);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + ((half + c * 16) ^ rx)), "r"(a0), "r"(a1), "r"(a2),
                     "r"(a3)
                     : "memory");
    }
}
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_u32, int kstep) {   // 16 K-elements per step
    return make_smem_desc(tile_u32 + (kstep >> 2) * 16384 + (kstep & 3) * 32, 16, 1024);
}
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_u32, int kstep) {  // rows = K, 128 B of N per row
    return make_smem_desc(tile_u32 + kstep * 2048, 8192, 1024);
}

// ------------------------------------------------------------------------------------------------ forward
// 256 threads: two threads per query row (warps w and w+4 own the same TMEM lanes and take alternate 32-column chunks).
// P overwrites the Q/K tiles once S is complete, so a CTA needs 97 KB and 256 TMEM columns: two CTAs per SM overlap
// one another's load / MMA / softmax phases.
__global__ void __launch_bounds__(256, 2) attn_fwd_kernel(const __grid_constant__ AttnParams p) {
    extern __shared__ uint8_t att_raw[];
    uint8_t* smem = att_align(att_raw);
    AttBars* bars = reinterpret_cast<AttBars*>(smem);
    float* s_red = reinterpret_cast<float*>(smem + 128);     // [2][128] row partials (max, then sum)
    uint8_t* sP = smem + 2048;            // 64 KB; Q lives in its first 16 KB and K in the next 32 KB until S is done
    uint8_t* sQ = sP;
    uint8_t* sK = sP + 16384;
    uint8_t* sV = sP + 65536;             // 32 KB
    const int warp = uniform_warp_id(), lane = threadIdx.x & 31;
    const int half = warp >> 2;
    const int mt = blockIdx.x & 1;
    const int bh = blockIdx.x >> 1;
    const int h = bh % p.H, b = bh / p.H;
    const int q0 = mt * ATT_TILE;
    if (q0 >= p.S) return;

    if (threadIdx.x == 0) {
        mbar_init(&bars->load, 1); mbar_init(&bars->mma0, 1); mbar_init(&bars->mma1, 1);
        fence_mbar_init();
        tma_prefetch_desc(&p.qkv_map);
    }
    if (warp == 0) { tmem_alloc(&bars->tmem_base, 256); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    const int HD = p.H * ATT_D;
    const int row_base = b * p.S;

    if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
        mbar_arrive_expect_tx_warp(&bars->load, 16384 * 5);
        tma_load_2d_warp(sQ, &p.qkv_map, &bars->load, h * ATT_D, row_base + q0);
        tma_load_2d_warp(sK, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base);
        tma_load_2d_warp(sK + 16384, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base + 128);
        tma_load_2d_warp(sV, &p.qkv_map, &bars->load, 2 * HD + h * ATT_D, row_base);
        tma_load_2d_warp(sV + 16384, &p.qkv_map, &bars->load, 2 * HD + h * ATT_D, row_base + 128);
        mbar_wait(&bars->load, 0);
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(ATT_TILE, p.n_keys, 0, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            umma_bf16_warp(tmem, desc_kmajor(smem_u32(sQ), k), desc_kmajor(smem_u32(sK), k), idesc, k != 0);
        umma_commit_warp(&bars->mma0);
    }
    // ---- softmax: thread = (query row, column half)
    const int row = (warp & 3) * 32 + lane;
    const uint32_t t_row = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    mbar_wait(&bars->mma0, 0);
    tc_fence_after();
    const int nchunks = (p.n_keys + 31) >> 5;
    float mx = -INFINITY;
    for (int g = half; g < nchunks; g += 2) {
        uint32_t v[32];
        tmem_ld_32x32(t_row + g * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (g * 32 + i < p.S) mx = fmaxf(mx, __uint_as_float(v[i]));
    }
    s_red[half * 128 + row] = mx;
    __syncthreads();                       // also: every S read of pass 1 is done
    mx = fmaxf(s_red[row], s_red[128 + row]);
    __syncthreads();
    const float m2 = mx * p.scale_log2e;
    float l = 0.f;
    const uint32_t sP_u32 = smem_u32(sP);
    for (int g = half; g < nchunks; g += 2) {
        uint32_t v[32];
        tmem_ld_32x32(t_row + g * 32, v);
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const float e = (g * 32 + i < p.S) ? exp2f(fmaf(__uint_as_float(v[i]), p.scale_log2e, -m2)) : 0.f;
            f[i] = e;
            l += e;
        }
        store_row_chunk32(sP_u32, row, g * 32, f);       // overwrites Q/K: the S = Q K^T chain has completed (mma0)
    }
    s_red[half * 128 + row] = l;
    tc_fence_before();
    fence_proxy_async_smem();
    __syncthreads();
    l = s_red[row] + s_red[128 + row];
    if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(ATT_TILE, ATT_D, 0, 1);
        const int ksteps = p.n_keys >> 4;
        for (int k = 0; k < ksteps; ++k)
            umma_bf16_warp(tmem, desc_kmajor(sP_u32, k), desc_mnmajor(smem_u32(sV), k), idesc, k != 0);   // O reuses S's columns
        umma_commit_warp(&bars->mma1);
    }
    mbar_wait(&bars->mma1, 0);
    tc_fence_after();
    const int q = q0 + row;
    const float inv_l = 1.f / l;
    {
        const int g = half;                 // each half stores 32 of the 64 output columns
        uint32_t v[32];
        tmem_ld_32x32(t_row + g * 32, v);
        tmem_ld_wait();
        if (q < p.S) {
            uint4* dst = reinterpret_cast<uint4*>(p.o + static_cast<size_t>(row_base + q) * HD + h * ATT_D + g * 32);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint4 w;
                w.x = pack_bf16x2(__uint_as_float(v[8 * c + 0]) * inv_l, __uint_as_float(v[8 * c + 1]) * inv_l);
                w.y = pack_bf16x2(__uint_as_float(v[8 * c + 2]) * inv_l, __uint_as_float(v[8 * c + 3]) * inv_l);
                w.z = pack_bf16x2(__uint_as_float(v[8 * c + 4]) * inv_l, __uint_as_float(v[8 * c + 5]) * inv_l);
                w.w = pack_bf16x2(__uint_as_float(v[8 * c + 6]) * inv_l, __uint_as_float(v[8 * c + 7]) * inv_l);
                dst[c] = w;
            }
        }
    }
    if (q < p.S && half == 0) p.lse[(static_cast<size_t>(b) * p.H + h) * p.S + q] = m2 + log2f(l);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------------ backward: dQ (+ delta)
// Persistent: one CTA per SM walks (batch, head, query tile) items; the Q / dO / K / V tiles of item i+1 are prefetched by
// TMA into the other half of shared memory while item i computes, and barrier setup / TMEM allocation happen once per CTA
// (they and the load latency were ~40% of the former one-item-per-CTA kernel).  dS overwrites the Q / dO / V tiles of its own
// item (they are dead once S and dP sit in TMEM), which is what makes two 96 KB item buffers fit.
constexpr int ATT_DQ_ITEM_BYTES = 32768 + 16384 + 16384 + 32768;      // K | Q | dO | V   (dS aliases Q..V: 64 KB)
struct AttDqBars {
    uint64_t load[2], mma0, mma1;
    uint32_t tmem_base;
};
__global__ void __launch_bounds__(256, 1) attn_bwd_dq_kernel(const __grid_constant__ AttnParams p) {
    extern __shared__ uint8_t att_raw[];
    uint8_t* smem = att_align(att_raw);
    AttDqBars* bars = reinterpret_cast<AttDqBars*>(smem);
    uint8_t* buf0 = smem + 1024;
    const int warp = uniform_warp_id(), lane = threadIdx.x & 31;
    const int half = warp >> 2;          // two threads per query row: alternate 32-column chunks
    const int n_tiles = (p.S + ATT_TILE - 1) / ATT_TILE;
    const int items = p.B * p.H * n_tiles;

    if (threadIdx.x == 0) {
        mbar_init(&bars->load[0], 1); mbar_init(&bars->load[1], 1); mbar_init(&bars->mma0, 1); mbar_init(&bars->mma1, 1);
        fence_mbar_init();
    }
    if (warp == 0) { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    const int HD = p.H * ATT_D;

    auto issue_loads = [&](int item, int b) {          // whole warp 0; one elected lane issues
        const int mt = item % n_tiles, bh = item / n_tiles;
        const int h = bh % p.H, bb = bh / p.H;
        const int row_base = bb * p.S, q0 = mt * ATT_TILE;
        uint8_t* sK = buf0 + b * ATT_DQ_ITEM_BYTES;
        uint8_t* sQ = sK + 32768;
        uint8_t* sdO = sQ + 16384;
        uint8_t* sV = sdO + 16384;
        mbar_arrive_expect_tx_warp(&bars->load[b], 16384 * 6);
        tma_load_2d_warp(sQ, &p.qkv_map, &bars->load[b], h * ATT_D, row_base + q0);
        tma_load_2d_warp(sdO, &p.do_map, &bars->load[b], h * ATT_D, row_base + q0);
        tma_load_2d_warp(sK, &p.qkv_map, &bars->load[b], HD + h * ATT_D, row_base);
        tma_load_2d_warp(sK + 16384, &p.qkv_map, &bars->load[b], HD + h * ATT_D, row_base + 128);
        tma_load_2d_warp(sV, &p.qkv_map, &bars->load[b], 2 * HD + h * ATT_D, row_base);
        tma_load_2d_warp(sV + 16384, &p.qkv_map, &bars->load[b], 2 * HD + h * ATT_D, row_base + 128);
    };

    if (warp == 0 && static_cast<int>(blockIdx.x) < items) issue_loads(blockIdx.x, 0);
    int it = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
        const int b = it & 1;
        const uint32_t ld_phase = (it >> 1) & 1, mma_phase = it & 1;
        const int mt = item % n_tiles, bh = item / n_tiles;
        const int h = bh % p.H, bb = bh / p.H;
        const int row_base = bb * p.S, q0 = mt * ATT_TILE;
        uint8_t* sK = buf0 + b * ATT_DQ_ITEM_BYTES;
        uint8_t* sQ = sK + 32768;
        uint8_t* sdO = sQ + 16384;
        uint8_t* sV = sdO + 16384;
        uint8_t* sdS = sQ;                              // 64 KB over Q | dO | V
        if (warp == 0) {
            // the other buffer was last touched by item it-1, whose dQ MMA and stores completed before the loop came back here
            if (item + static_cast<int>(gridDim.x) < items) issue_loads(item + gridDim.x, b ^ 1);
            mbar_wait(&bars->load[b], ld_phase);
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(ATT_TILE, p.n_keys, 0, 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // S = Q K^T  -> columns [0, 256)
                umma_bf16_warp(tmem, desc_kmajor(smem_u32(sQ), k), desc_kmajor(smem_u32(sK), k), idesc, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // dP = dO V^T -> columns [256, 512)
                umma_bf16_warp(tmem + 256, desc_kmajor(smem_u32(sdO), k), desc_kmajor(smem_u32(sV), k), idesc, k != 0);
            umma_commit_warp(&bars->mma0);
        }
        const int row = (warp & 3) * 32 + lane;
        const int q = q0 + row;
        const bool qvalid = q < p.S;
        // delta = rowsum(dO o O), straight from global memory (128 B per tensor per thread)
        float delta = 0.f;
        if (qvalid) {
            const uint4* po = reinterpret_cast<const uint4*>(p.o + static_cast<size_t>(row_base + q) * HD + h * ATT_D);
            const uint4* pd = reinterpret_cast<const uint4*>(p.dout + static_cast<size_t>(row_base + q) * HD + h * ATT_D);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint4 a = po[c], d = pd[c];
                const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
                const float2 d0 = unpack_bf16x2(d.x), d1 = unpack_bf16x2(d.y), d2 = unpack_bf16x2(d.z), d3 = unpack_bf16x2(d.w);
                delta += a0.x * d0.x + a0.y * d0.y + a1.x * d1.x + a1.y * d1.y + a2.x * d2.x + a2.y * d2.y + a3.x * d3.x + a3.y * d3.y;
            }
            if (half == 0) p.delta[(static_cast<size_t>(bb) * p.H + h) * p.S + q] = delta;
        }
        const float lse = qvalid ? p.lse[(static_cast<size_t>(bb) * p.H + h) * p.S + q] : 0.f;
        const uint32_t t_row = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
        mbar_wait(&bars->mma0, mma_phase);
        tc_fence_after();
        const uint32_t sdS_u32 = smem_u32(sdS);
        for (int g = half; g < (p.n_keys + 31) >> 5; g += 2) {
            uint32_t s[32], dp[32];
            tmem_ld_32x32(t_row + g * 32, s);
            tmem_ld_32x32(t_row + 256 + g * 32, dp);
            tmem_ld_wait();
            float f[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float pr = (g * 32 + i < p.S) ? exp2f(fmaf(__uint_as_float(s[i]), p.scale_log2e, -lse)) : 0.f;
                f[i] = qvalid ? pr * (__uint_as_float(dp[i]) - delta) * p.scale : 0.f;
            }
            store_row_chunk32(sdS_u32, row, g * 32, f);     // over Q / dO / V: the S and dP chains have completed (mma0)
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncthreads();
        if (warp == 0) {
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(ATT_TILE, ATT_D, 0, 1);
            const int ksteps = p.n_keys >> 4;
            for (int k = 0; k < ksteps; ++k)         // dQ = dS K  (K read MN-major), into columns [0, 64)
                umma_bf16_warp(tmem, desc_kmajor(sdS_u32, k), desc_mnmajor(smem_u32(sK), k), idesc, k != 0);
            umma_commit_warp(&bars->mma1);
        }
        mbar_wait(&bars->mma1, mma_phase);
        tc_fence_after();
        {
            const int g = half;
            uint32_t v[32];
            tmem_ld_32x32(t_row + g * 32, v);
            tmem_ld_wait();
            if (qvalid) {
                uint4* dst = reinterpret_cast<uint4*>(p.dqkv + static_cast<size_t>(row_base + q) * 3 * HD + h * ATT_D + g * 32);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 w;
                    w.x = pack_bf16x2(__uint_as_float(v[8 * c + 0]), __uint_as_float(v[8 * c + 1]));
                    w.y = pack_bf16x2(__uint_as_float(v[8 * c + 2]), __uint_as_float(v[8 * c + 3]));
                    w.z = pack_bf16x2(__uint_as_float(v[8 * c + 4]), __uint_as_float(v[8 * c + 5]));
                    w.w = pack_bf16x2(__uint_as_float(v[8 * c + 6]), __uint_as_float(v[8 * c + 7]));
                    dst[c] = w;
                }
            }
        }
        tc_fence_before();
        __syncthreads();          // every thread has read dQ out of TMEM and is done with this item's buffer
        tc_fence_after();
    }
    if (warp == 0) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ backward: dK, dV
__global__ void __launch_bounds__(256, 1) attn_bwd_dkdv_kernel(const __grid_constant__ AttnParams p) {
    extern __shared__ uint8_t att_raw[];
    uint8_t* smem = att_align(att_raw);
    AttBars* bars = reinterpret_cast<AttBars*>(smem);
    float* s_lse = reinterpret_cast<float*>(smem + 128);     // 128 floats
    float* s_delta = s_lse + 128;                            // 128 floats  (ends at 128 + 1024 = 1152 < 2048)
    uint8_t* sK = smem + 2048;            // 16 KB  (this CTA's key tile)
    uint8_t* sV = sK + 16384;             // 16 KB
    uint8_t* sQ = sV + 16384;             // 16 KB  (current query tile)
    uint8_t* sdO = sQ + 16384;            // 16 KB
    uint8_t* sPT = sdO + 16384;           // 32 KB  P^T  [keys x queries]
    uint8_t* sdST = sPT + 32768;          // 32 KB  dS^T [keys x queries]
    const int warp = uniform_warp_id(), lane = threadIdx.x & 31;
    const int half = warp >> 2;          // two threads per key row: alternate 32-query chunks
    const int kt = blockIdx.x & 1;
    const int bh = blockIdx.x >> 1;
    const int h = bh % p.H, b = bh / p.H;
    const int k0 = kt * ATT_TILE;
    if (k0 >= p.S) return;

    if (threadIdx.x == 0) {
        mbar_init(&bars->load, 1); mbar_init(&bars->mma0, 1); mbar_init(&bars->mma1, 1);
        fence_mbar_init();
    }
    if (warp == 0) { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    const int HD = p.H * ATT_D;
    const int row_base = b * p.S;
    const int row = (warp & 3) * 32 + lane;      // key index inside the tile
    const int key = k0 + row;
    const uint32_t t_row = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    const int n_qtiles = (p.S + ATT_TILE - 1) / ATT_TILE;
    uint32_t ph = 0;

    for (int qt = 0; qt < n_qtiles; ++qt) {
        const int q0 = qt * ATT_TILE;
        if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
            mbar_arrive_expect_tx_warp(&bars->load, qt == 0 ? 16384 * 4 : 16384 * 2);
            if (qt == 0) {
                tma_load_2d_warp(sK, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base + k0);
                tma_load_2d_warp(sV, &p.qkv_map, &bars->load, 2 * HD + h * ATT_D, row_base + k0);
            }
            tma_load_2d_warp(sQ, &p.qkv_map, &bars->load, h * ATT_D, row_base + q0);
            tma_load_2d_warp(sdO, &p.do_map, &bars->load, h * ATT_D, row_base + q0);
        }
        // per-query log-sum-exp and delta of this query tile (written by the forward / dQ kernels)
        if (threadIdx.x < ATT_TILE) {
            const int qq = q0 + threadIdx.x;
            const size_t off = (static_cast<size_t>(b) * p.H + h) * p.S + qq;
            s_lse[threadIdx.x] = qq < p.S ? p.lse[off] : 0.f;
            s_delta[threadIdx.x] = qq < p.S ? p.delta[off] : 0.f;
        }
        __syncthreads();
        if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
            mbar_wait(&bars->load, ph);
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(ATT_TILE, ATT_TILE, 0, 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // S^T = K Q^T   -> columns [0, 128)
                umma_bf16_warp(tmem, desc_kmajor(smem_u32(sK), k), desc_kmajor(smem_u32(sQ), k), idesc, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // dP^T = V dO^T -> columns [128, 256)
                umma_bf16_warp(tmem + 128, desc_kmajor(smem_u32(sV), k), desc_kmajor(smem_u32(sdO), k), idesc, k != 0);
            umma_commit_warp(&bars->mma0);
        }
        mbar_wait(&bars->mma0, ph);
        tc_fence_after();
        const uint32_t sPT_u32 = smem_u32(sPT), sdST_u32 = smem_u32(sdST);
#pragma unroll 1
        for (int g = half; g < 4; g += 2) {
            uint32_t s[32], dp[32];
            tmem_ld_32x32(t_row + g * 32, s);
            tmem_ld_32x32(t_row + 128 + g * 32, dp);
            tmem_ld_wait();
            float fp[32], fd[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int qq = g * 32 + i;
                const bool ok = (q0 + qq < p.S) && (key < p.S);
                const float pr = ok ? exp2f(fmaf(__uint_as_float(s[i]), p.scale_log2e, -s_lse[qq])) : 0.f;
                fp[i] = pr;
                fd[i] = pr * (__uint_as_float(dp[i]) - s_delta[qq]) * p.scale;
            }
            store_row_chunk32(sPT_u32, row, g * 32, fp);
            store_row_chunk32(sdST_u32, row, g * 32, fd);
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncthreads();
        if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(ATT_TILE, ATT_D, 0, 1);
#pragma unroll
            for (int k = 0; k < 8; ++k)      // dV += P^T dO   (dO read MN-major) -> columns [256, 320)
                umma_bf16_warp(tmem + 256, desc_kmajor(sPT_u32, k), desc_mnmajor(smem_u32(sdO), k), idesc, (qt | k) != 0);
#pragma unroll
            for (int k = 0; k < 8; ++k)      // dK += dS^T Q   (Q read MN-major)  -> columns [320, 384)
                umma_bf16_warp(tmem + 320, desc_kmajor(sdST_u32, k), desc_mnmajor(smem_u32(sQ), k), idesc, (qt | k) != 0);
            umma_commit_warp(&bars->mma1);
        }
        mbar_wait(&bars->mma1, ph);      // sQ / sdO / sPT / sdST may be overwritten by the next query tile
        tc_fence_after();
        ph ^= 1;
        __syncthreads();
    }
    // ---- store dK (slot 1 of dqkv) and dV (slot 2)
#pragma unroll
    for (int which = 0; which < 2; ++which) {
        {
            const int g = half;
            uint32_t v[32];
            tmem_ld_32x32(t_row + (which == 0 ? 320 : 256) + g * 32, v);
            tmem_ld_wait();
            if (key < p.S) {
                uint4* dst = reinterpret_cast<uint4*>(p.dqkv + static_cast<size_t>(row_base + key) * 3 * HD + (which + 1) * HD +
                                                      h * ATT_D + g * 32);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 w;
                    w.x = pack_bf16x2(__uint_as_float(v[8 * c + 0]), __uint_as_float(v[8 * c + 1]));
                    w.y = pack_bf16x2(__uint_as_float(v[8 * c + 2]), __uint_as_float(v[8 * c + 3]));
                    w.z = pack_bf16x2(__uint_as_float(v[8 * c + 4]), __uint_as_float(v[8 * c + 5]));
                    w.w = pack_bf16x2(__uint_as_float(v[8 * c + 6]), __uint_as_float(v[8 * c + 7]));
                    dst[c] = w;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

static int make_2d_map(CUtensorMap* m, const void* base, long long rows, int cols) {
    uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
    uint64_t strides[1] = {(uint64_t)cols * 2};
    uint32_t box[2] = {64, 128};
    return make_tensor_map(m, base, 2, 2, dims, strides, box, true);
}

static int fill_params(AttnParams& p, const void* qkv, int B, int S, int H) {
    memset(&p, 0, sizeof(p));
    p.B = B; p.H = H; p.S = S;
    p.n_keys = (S + 15) / 16 * 16;
    p.scale = 0.125f;                              // 1/sqrt(64)
    p.scale_log2e = 0.125f * 1.4426950408889634f;
    return make_2d_map(&p.qkv_map, qkv, static_cast<long long>(B) * S, 3 * H * ATT_D);
}

}  // namespace picd

using namespace picd;

// o[B*S, H*64] = softmax(q k^T / 8) v per head, q/k/v packed as qkv[B*S, 3*H*64]; lse[B,H,S] saved for backward.
extern "C" int picd_attention_fwd(const void* qkv, void* o, float* lse, int B, int S, int H, cudaStream_t stream) {
    PICD_REQUIRE(S >= 1 && S <= ATT_MAX_S, "attention: sequence length %d not in [1, %d]", S, ATT_MAX_S);
    AttnParams p;
    if (fill_params(p, qkv, B, S, H)) return 3;
    p.o = static_cast<__nv_bfloat16*>(o);
    p.lse = lse;
    const int smem = 2048 + 65536 + 32768 + 1024;       // 99 KB: two CTAs per SM
    static bool attr = false;
    if (!attr) {
        PICD_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr = true;
    }
    attn_fwd_kernel<<<B * H * 2, 256, smem, stream>>>(p);
    PICD_LAUNCH_CHECK();
    return 0;
}

// dqkv[B*S, 3*H*64] from dout[B*S, H*64]; delta[B,H,S] is scratch written by the first kernel and read by the second.
extern "C" int picd_attention_bwd(const void* qkv, const void* o, const void* dout, const float* lse, float* delta,
                                  void* dqkv, int B, int S, int H, cudaStream_t stream) {
    PICD_REQUIRE(S >= 1 && S <= ATT_MAX_S, "attention: sequence length %d not in [1, %d]", S, ATT_MAX_S);
    AttnParams p;
    if (fill_params(p, qkv, B, S, H)) return 3;
    if (make_2d_map(&p.do_map, dout, static_cast<long long>(B) * S, H * ATT_D)) return 3;
    p.o = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(o));
    p.dout = static_cast<const __nv_bfloat16*>(dout);
    p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
    p.lse = const_cast<float*>(lse);
    p.delta = delta;
    const int smem_dq = 1024 + 2 * ATT_DQ_ITEM_BYTES + 1024;
    const int smem_kv = 2048 + 16384 * 4 + 32768 * 2 + 1024;
    static bool attr = false;
    if (!attr) {
        PICD_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq));
        PICD_CUDA(cudaFuncSetAttribute(attn_bwd_dkdv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kv));
        attr = true;
    }
    attn_bwd_dq_kernel<<<std::min(B * H * ((S + ATT_TILE - 1) / ATT_TILE), num_sms()), 256, smem_dq, stream>>>(p);
    PICD_LAUNCH_CHECK();
    attn_bwd_dkdv_kernel<<<B * H * 2, 256, smem_kv, stream>>>(p);
    PICD_LAUNCH_CHECK();
    return 0;
}
