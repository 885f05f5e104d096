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
#include <cstdlib>
#include <cstring>

namespace picd {

constexpr int ATT_D = 64;          // head dimension (fixed: one 128-byte swizzle row)
constexpr int ATT_TILE = 128;      // queries (or keys) per CTA
constexpr int ATT_MAX_S = 256;

struct AttnParams {
    CUtensorMap qkv_map;     // [B*S rows, 3*H*64 cols] bf16, box (64, 128)
    CUtensorMap do_map;      // [B*S rows, H*64 cols]  bf16, box (64, 128)   (backward)
    CUtensorMap qkv_map64;   // the same tensors with 64-row boxes (query sub-tiles of the dK / dV kernel)
    CUtensorMap do_map64;
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

// 2^x on the SFU in ONE instruction.  exp2f() without fast-math expands to a range check + scaling around ex2.approx (denormal
// results), which the softmax does not need: every argument is <= 0 after the row maximum / log-sum-exp is subtracted, and
// results below 2^-126 flush to zero, which is what a bf16 probability becomes anyway.  ncu showed these kernels issue-bound
// (34 % issue utilisation with 16 warps, the rest latency), so instruction count matters.
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
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
        mbar_init(&bars->load, 1); mbar_init(&bars->mma0, 1); mbar_init(&bars->mma1, 1); mbar_init(&bars->mma2, 1);
        fence_mbar_init();
        tma_prefetch_desc(&p.qkv_map);
    }
    if (warp == 0) { tmem_alloc(&bars->tmem_base, 256); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    pdl_sync();
    const int HD = p.H * ATT_D;
    const int row_base = b * p.S;

    if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
        // two barriers: S = Q K^T starts as soon as Q and K (48 of the 80 KB) have landed; V (bars->mma2 doubles as its load
        // barrier here) is only needed after the softmax
        mbar_arrive_expect_tx_warp(&bars->load, 16384 * 3);
        tma_load_2d_warp(sQ, &p.qkv_map, &bars->load, h * ATT_D, row_base + q0);
        tma_load_2d_warp(sK, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base);
        tma_load_2d_warp(sK + 16384, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base + 128);
        mbar_arrive_expect_tx_warp(&bars->mma2, 16384 * 2);
        tma_load_2d_warp(sV, &p.qkv_map, &bars->mma2, 2 * HD + h * ATT_D, row_base);
        tma_load_2d_warp(sV + 16384, &p.qkv_map, &bars->mma2, 2 * HD + h * ATT_D, row_base + 128);
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
    // a warp whose 32 query rows all lie beyond the sequence (second tile of S = 197: rows 96..127) skips both softmax passes:
    // its P rows stay whatever the Q / K bytes were, which only reaches output rows that are never stored
    const int nchunks = (q0 + (warp & 3) * 32 < p.S) ? (p.n_keys + 31) >> 5 : 0;
    float mx = -INFINITY;
    for (int g = half; g < nchunks; g += 2) {
        uint32_t v[32];
        tmem_ld_32x32(t_row + g * 32, v);
        tmem_ld_wait();
        if (g * 32 + 32 <= p.S) {
#pragma unroll
            for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (g * 32 + i < p.S) mx = fmaxf(mx, __uint_as_float(v[i]));
        }
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
        if (g * 32 + 32 <= p.S) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                f[i] = fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -m2));
                l += f[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                f[i] = (g * 32 + i < p.S) ? fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -m2)) : 0.f;
                l += f[i];
            }
        }
        store_row_chunk32(sP_u32, row, g * 32, f);       // overwrites Q/K: the S = Q K^T chain has completed (mma0)
    }
    s_red[half * 128 + row] = l;
    tc_fence_before();
    fence_proxy_async_smem();
    __syncthreads();
    l = s_red[row] + s_red[128 + row];
    if (warp == 0) {      // whole warp, one elected lane issues (picd_common.cuh)
        mbar_wait(&bars->mma2, 0);           // V has landed (long ago)
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

// ------------------------------------------------------------------------------------------------ forward, pipelined
// One PERSISTENT CTA per SM walks (batch, head, query tile) items: a TMA warp keeps the loads of the next two items in flight
// (two Q | K | V buffers), one warp issues the S = Q K^T chain and another the O = P V chain, and TWO softmax teams of eight warps take
// alternate items, each with its own shared-memory buffer, its own 256 TMEM columns and its own named barrier -- while one team
// waits (for its scores, for its P V product) the other one runs its exponentials.  This is the overlap the one-item-per-CTA
// kernel above gets from a second resident CTA, without that kernel's per-item set-up, TMEM allocation and exposed first load
// (a third of its stall samples in ncu).
struct AttPipeBars {
    uint64_t qk_full[2], v_full[2], s_full[2], p_full[2], o_full[2], t_free[2];
    uint32_t tmem_base;
};
constexpr int ATTP_TEAM_WARPS = 8;
constexpr int ATTP_PV_WARP = 2 + 2 * ATTP_TEAM_WARPS;            // the last warp issues the O = P V chain
constexpr int ATTP_THREADS = 32 * (ATTP_PV_WARP + 1);
constexpr int ATTP_HEADER = 8192;                // barriers (256 B) | row partials [team][max / sum][half][128] at +1024
constexpr int ATTP_BUF = 65536 + 32768;          // P (over Q | K) + V

__global__ void __launch_bounds__(ATTP_THREADS, 1) attn_fwd_pipe_kernel(const __grid_constant__ AttnParams p) {
    extern __shared__ uint8_t att_raw[];
    uint8_t* smem = att_align(att_raw);
    AttPipeBars* bars = reinterpret_cast<AttPipeBars*>(smem);
    const int warp = uniform_warp_id(), lane = threadIdx.x & 31;
    const int n_tiles = (p.S + ATT_TILE - 1) / ATT_TILE;
    const int total = p.B * p.H * n_tiles;
    const int n_items = static_cast<int>(blockIdx.x) < total ? (total - static_cast<int>(blockIdx.x) + gridDim.x - 1) / gridDim.x : 0;
    const int HD = p.H * ATT_D;

    if (threadIdx.x == 0) {
        for (int b = 0; b < 2; ++b) {
            mbar_init(&bars->qk_full[b], 1); mbar_init(&bars->v_full[b], 1); mbar_init(&bars->s_full[b], 1);
            mbar_init(&bars->p_full[b], 1); mbar_init(&bars->o_full[b], 1); mbar_init(&bars->t_free[b], ATTP_TEAM_WARPS);
        }
        fence_mbar_init();
        tma_prefetch_desc(&p.qkv_map);
    }
    if (warp == 1) { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    pdl_sync();

    auto decode = [&](int k, int& bb, int& h, int& q0) {
        const int it = blockIdx.x + k * gridDim.x;
        const int mt = it % n_tiles, bh = it / n_tiles;
        h = bh % p.H; bb = bh / p.H; q0 = mt * ATT_TILE;
    };

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        for (int k = 0; k < n_items; ++k) {
            const int b = k & 1;
            int bb, h, q0;
            decode(k, bb, h, q0);
            const int row_base = bb * p.S;
            uint8_t* buf = smem + ATTP_HEADER + b * ATTP_BUF;
            if (k >= 2) mbar_wait(&bars->o_full[b], ((k >> 1) - 1) & 1);     // P V of item k - 2 has consumed this buffer
            mbar_arrive_expect_tx_warp(&bars->qk_full[b], 16384 * 3);
            tma_load_2d_warp(buf, &p.qkv_map, &bars->qk_full[b], h * ATT_D, row_base + q0);
            tma_load_2d_warp(buf + 16384, &p.qkv_map, &bars->qk_full[b], HD + h * ATT_D, row_base);
            tma_load_2d_warp(buf + 32768, &p.qkv_map, &bars->qk_full[b], HD + h * ATT_D, row_base + 128);
            mbar_arrive_expect_tx_warp(&bars->v_full[b], 16384 * 2);
            tma_load_2d_warp(buf + 65536, &p.qkv_map, &bars->v_full[b], 2 * HD + h * ATT_D, row_base);
            tma_load_2d_warp(buf + 65536 + 16384, &p.qkv_map, &bars->v_full[b], 2 * HD + h * ATT_D, row_base + 128);
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer 1: S = Q K^T, item after item
        // (two issuing warps, one per chain: a single in-order issuer let an S waiting for its operands hold up a ready P V and
        //  vice versa; one warp polling both chains cost more issue slots than it saved)
        const uint32_t idesc_s = make_idesc_bf16(ATT_TILE, p.n_keys, 0, 0);
        for (int k = 0; k < n_items; ++k) {
            const int b = k & 1;
            const uint32_t buf = smem_u32(smem + ATTP_HEADER + b * ATTP_BUF);
            mbar_wait(&bars->qk_full[b], (k >> 1) & 1);
            if (k >= 2) mbar_wait(&bars->t_free[b], ((k >> 1) - 1) & 1);       // O of item k - 2 has left these TMEM columns
            tc_fence_after();
#pragma unroll
            for (int j = 0; j < 4; ++j)
                umma_bf16_warp(tmem + b * 256, desc_kmajor(buf, j), desc_kmajor(buf + 16384, j), idesc_s, j != 0);
            umma_commit_warp(&bars->s_full[b]);
        }
    } else if (warp == ATTP_PV_WARP) {
        // ------------------------------------------------------------------ MMA issuer 2: O = P V
        const uint32_t idesc_o = make_idesc_bf16(ATT_TILE, ATT_D, 0, 1);
        const int ksteps = p.n_keys >> 4;
        for (int k = 0; k < n_items; ++k) {
            const int b = k & 1;
            const uint32_t par = (k >> 1) & 1;
            const uint32_t buf = smem_u32(smem + ATTP_HEADER + b * ATTP_BUF);
            mbar_wait(&bars->v_full[b], par);
            mbar_wait(&bars->p_full[b], par);
            tc_fence_after();
            for (int kk = 0; kk < ksteps; ++kk)           // O overwrites the first 64 columns of the item's own scores (read by now)
                umma_bf16_warp(tmem + b * 256, desc_kmajor(buf, kk), desc_mnmajor(buf + 65536, kk), idesc_o, kk != 0);
            umma_commit_warp(&bars->o_full[b]);
        }
    } else {
        // ------------------------------------------------------------------ softmax + read-out: two teams of 8 warps
        const int team = (warp - 2) >> 3;            // items k = team, team + 2, ... ; buffer / TMEM half b = team
        const int half = ((warp - 2) >> 2) & 1;      // which half of the score columns (chunks half, half + 2, ...)
        const int q = warp & 3;                      // TMEM lane quarter this warp may touch
        const int row = q * 32 + lane;
        const int b = team;
        float* s_max = reinterpret_cast<float*>(smem + 1024) + team * 512;      // [half][128]
        float* s_sum = s_max + 256;
        const uint32_t t_row = tmem + b * 256 + (static_cast<uint32_t>(q * 32) << 16);
        const uint32_t sP_u32 = smem_u32(smem + ATTP_HEADER + b * ATTP_BUF);
        const uint32_t bar_id = 1 + team;
        for (int k = team; k < n_items; k += 2) {
            int bb, h, q0;
            decode(k, bb, h, q0);
            const uint32_t par = (k >> 1) & 1;
            // a warp whose 32 query rows all lie beyond the sequence skips both passes (its P rows only reach unstored rows)
            const int nchunks = (q0 + q * 32 < p.S) ? (p.n_keys + 31) >> 5 : 0;
            mbar_wait(&bars->s_full[b], par);
            tc_fence_after();
            float mx = -INFINITY;
            for (int g = half; g < nchunks; g += 2) {
                uint32_t v[32];
                tmem_ld_32x32(t_row + g * 32, v);
                tmem_ld_wait();
                if (g * 32 + 32 <= p.S) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (g * 32 + i < p.S) mx = fmaxf(mx, __uint_as_float(v[i]));
                }
            }
            s_max[half * 128 + row] = mx;
            named_barrier_sync(bar_id, 32 * ATTP_TEAM_WARPS);
            mx = fmaxf(s_max[row], s_max[128 + row]);
            const float m2 = mx * p.scale_log2e;
            float l = 0.f;
            for (int g = half; g < nchunks; g += 2) {
                uint32_t v[32];
                tmem_ld_32x32(t_row + g * 32, v);
                tmem_ld_wait();
                float f[32];
                if (g * 32 + 32 <= p.S) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        f[i] = fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -m2));
                        l += f[i];
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        f[i] = (g * 32 + i < p.S) ? fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -m2)) : 0.f;
                        l += f[i];
                    }
                }
                store_row_chunk32(sP_u32, row, g * 32, f);       // over Q / K: S = Q K^T has completed (s_full)
            }
            s_sum[half * 128 + row] = l;
            tc_fence_before();
            fence_proxy_async_smem();
            named_barrier_sync(bar_id, 32 * ATTP_TEAM_WARPS);     // every P chunk is written and every S column read
            if (((warp - 2) & 7) == 0 && lane == 0) mbar_arrive(&bars->p_full[b]);
            l = s_sum[row] + s_sum[128 + row];
            const int qq = q0 + row;
            if (half == 0 && qq < p.S) p.lse[(static_cast<size_t>(bb) * p.H + h) * p.S + qq] = m2 + log2f(l);
            // ---- read-out (the other team computes meanwhile)
            mbar_wait(&bars->o_full[b], par);
            tc_fence_after();
            {
                uint32_t v[32];
                tmem_ld_32x32(t_row + half * 32, v);          // each half stores 32 of the 64 output columns
                tmem_ld_wait();
                if (qq < p.S) {
                    const float inv_l = 1.f / l;
                    uint4* dst = reinterpret_cast<uint4*>(p.o + static_cast<size_t>(bb * p.S + qq) * HD + h * ATT_D + half * 32);
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
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->t_free[b]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ backward: dQ (+ delta)
// One CTA per (batch, head, 128-query tile), TWO CTAs per SM.  The former version kept S = Q K^T and dP = dO V^T side by side in
// TMEM (2 x 256 columns = the whole TMEM, one CTA per SM) and ran its phases -- load, MMA, exp, MMA, store: ~10 us of latency
// per item -- back to back with nothing to overlap them but a prefetch.  Here S and dP share ONE set of columns: the
// probabilities P = exp2(S * c - lse) are packed to bf16 and parked in registers (two threads per row, <= 64 registers each)
// while the dP MMA overwrites S; dS = P o (dP - delta) * scale then goes to shared memory over the dead Q / dO / V tiles and
// dQ = dS K lands in the same columns again.  256 TMEM columns and 98 KB of shared memory per CTA: two CTAs per SM hide
// each other's latency chain.
struct AttDqBars {
    uint64_t load, mma0, mma1, mma2, load2;
    uint32_t tmem_base;
};
__device__ __forceinline__ void store_row_chunk32_bits(uint32_t tile_u32, int row, int col0, const uint32_t (&f)[32]) {
    const uint32_t dst = tile_u32 + (col0 >> 6) * 16384 + row * 128;
    const uint32_t half = static_cast<uint32_t>((col0 >> 5) & 1) * 64;
    const uint32_t rx = static_cast<uint32_t>(row & 7) << 4;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const uint32_t a0 = pack_bf16x2(__uint_as_float(f[8 * c + 0]), __uint_as_float(f[8 * c + 1]));
        const uint32_t a1 = pack_bf16x2(__uint_as_float(f[8 * c + 2]), __uint_as_float(f[8 * c + 3]));
        const uint32_t a2 = pack_bf16x2(__uint_as_float(f[8 * c + 4]), __uint_as_float(f[8 * c + 5]));
        const uint32_t a3 = pack_bf16x2(__uint_as_float(f[8 * c + 6]), __uint_as_float(f[8 * c + 7]));
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + ((half + c * 16) ^ rx)), "r"(a0), "r"(a1), "r"(a2),
                     "r"(a3)
                     : "memory");
    }
}
__global__ void __launch_bounds__(256, 2) attn_bwd_dq_kernel(const __grid_constant__ AttnParams p) {
    extern __shared__ uint8_t att_raw[];
    uint8_t* smem = att_align(att_raw);
    AttDqBars* bars = reinterpret_cast<AttDqBars*>(smem);
    uint8_t* sK = smem + 1024;            // 32 KB
    uint8_t* sQ = sK + 32768;             // 16 KB
    uint8_t* sdO = sQ + 16384;            // 16 KB
    uint8_t* sV = sdO + 16384;            // 32 KB
    uint8_t* sdS = sQ;                    // 64 KB over Q | dO | V once S and dP have been computed
    const int warp = uniform_warp_id(), lane = threadIdx.x & 31;
    const int half = warp >> 2;          // two threads per query row: alternate 32-column chunks
    const int n_tiles = (p.S + ATT_TILE - 1) / ATT_TILE;
    const int item = blockIdx.x;
    const int mt = item % n_tiles, bh = item / n_tiles;
    const int h = bh % p.H, bb = bh / p.H;
    const int row_base = bb * p.S, q0 = mt * ATT_TILE;

    if (threadIdx.x == 0) {
        mbar_init(&bars->load, 1); mbar_init(&bars->mma0, 1); mbar_init(&bars->mma1, 1); mbar_init(&bars->mma2, 1);
        mbar_init(&bars->load2, 1);
        fence_mbar_init();
        tma_prefetch_desc(&p.qkv_map); tma_prefetch_desc(&p.do_map);
    }
    if (warp == 0) { tmem_alloc(&bars->tmem_base, 256); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    pdl_sync();
    const int HD = p.H * ATT_D;

    if (warp == 0) {          // whole warp, one elected lane issues
        // two barriers: S = Q K^T starts when Q and K (48 of the 96 KB) have landed; dO and V are needed for dP only
        mbar_arrive_expect_tx_warp(&bars->load, 16384 * 3);
        tma_load_2d_warp(sQ, &p.qkv_map, &bars->load, h * ATT_D, row_base + q0);
        tma_load_2d_warp(sK, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base);
        tma_load_2d_warp(sK + 16384, &p.qkv_map, &bars->load, HD + h * ATT_D, row_base + 128);
        mbar_arrive_expect_tx_warp(&bars->load2, 16384 * 3);
        tma_load_2d_warp(sdO, &p.do_map, &bars->load2, h * ATT_D, row_base + q0);
        tma_load_2d_warp(sV, &p.qkv_map, &bars->load2, 2 * HD + h * ATT_D, row_base);
        tma_load_2d_warp(sV + 16384, &p.qkv_map, &bars->load2, 2 * HD + h * ATT_D, row_base + 128);
        mbar_wait(&bars->load, 0);
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(ATT_TILE, p.n_keys, 0, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)      // S = Q K^T
            umma_bf16_warp(tmem, desc_kmajor(smem_u32(sQ), k), desc_kmajor(smem_u32(sK), k), idesc, k != 0);
        umma_commit_warp(&bars->mma0);
    }
    const int row = (warp & 3) * 32 + lane;
    const int q = q0 + row;
    const bool qvalid = q < p.S;
    // delta = rowsum(dO o O), straight from global memory (128 B per tensor per thread); overlaps the loads and the S MMA
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
    // (a warp whose 32 query rows all lie beyond the sequence skips both phases: dQ rows do not depend on one another and
    //  those rows are never stored)
    const int nchunks = (q0 + (warp & 3) * 32 < p.S) ? (p.n_keys + 31) >> 5 : 0;

    // ---- phase 1: P = exp2(S * c - lse), packed to bf16 and kept in registers
    uint32_t pk[4][16];
    mbar_wait(&bars->mma0, 0);
    tc_fence_after();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int g = half + 2 * j;
        if (g < nchunks) {
            uint32_t sv[32];
            tmem_ld_32x32(t_row + g * 32, sv);
            tmem_ld_wait();
            if (g * 32 + 32 <= p.S) {            // every key of this chunk exists: no per-element predicate
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    pk[j][i] = pack_bf16x2(fast_exp2(fmaf(__uint_as_float(sv[2 * i]), p.scale_log2e, -lse)),
                                           fast_exp2(fmaf(__uint_as_float(sv[2 * i + 1]), p.scale_log2e, -lse)));
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float p0 = (g * 32 + 2 * i < p.S) ? fast_exp2(fmaf(__uint_as_float(sv[2 * i]), p.scale_log2e, -lse)) : 0.f;
                    const float p1 = (g * 32 + 2 * i + 1 < p.S) ? fast_exp2(fmaf(__uint_as_float(sv[2 * i + 1]), p.scale_log2e, -lse)) : 0.f;
                    pk[j][i] = pack_bf16x2(p0, p1);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();                     // every S column has been read: the dP MMA may overwrite them
    if (warp == 0) {
        mbar_wait(&bars->load2, 0);      // dO and V have landed
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(ATT_TILE, p.n_keys, 0, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)      // dP = dO V^T, into the columns S occupied
            umma_bf16_warp(tmem, desc_kmajor(smem_u32(sdO), k), desc_kmajor(smem_u32(sV), k), idesc, k != 0);
        umma_commit_warp(&bars->mma1);
    }
    // ---- phase 2: dS = P o (dP - delta) * scale -> shared memory (over Q | dO | V: every MMA that read them has completed)
    mbar_wait(&bars->mma1, 0);
    tc_fence_after();
    const uint32_t sdS_u32 = smem_u32(sdS);
    const float sc_row = qvalid ? p.scale : 0.f;         // rows beyond the sequence produce dS = 0
    const float nds = -delta * sc_row;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int g = half + 2 * j;
        if (g < nchunks) {
            uint32_t dp[32];
            tmem_ld_32x32(t_row + g * 32, dp);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) {       // dS = P * (dP - delta) * scale = P * fma(dP, scale, -delta * scale)
                const float2 pr = unpack_bf16x2(pk[j][i]);
                dp[2 * i] = __float_as_uint(pr.x * fmaf(__uint_as_float(dp[2 * i]), sc_row, nds));
                dp[2 * i + 1] = __float_as_uint(pr.y * fmaf(__uint_as_float(dp[2 * i + 1]), sc_row, nds));
            }
            store_row_chunk32_bits(sdS_u32, row, g * 32, dp);
        }
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
        umma_commit_warp(&bars->mma2);
    }
    mbar_wait(&bars->mma2, 0);
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
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------------ backward: dK, dV
// One CTA per (batch, head, 128-key tile); the queries are walked in SUB-TILES OF 64.  With 128-query tiles the kernel needed
// 384 TMEM columns (S^T, dP^T, dV, dK) and 134 KB of shared memory, i.e. one CTA per SM running its load -> MMA -> exp -> MMA
// phases strictly one after the other (2.9 ms of a 23 ms ViT-B/16 step).  64-query sub-tiles need 4 x 64 = 256 columns and
// 100 KB, so TWO CTAs share an SM and fill each other's phase gaps; the Q / dO sub-tile of the next iteration is prefetched
// into a second 16 KB buffer while the current one computes.
constexpr int ATT_QSUB = 64;
struct AttKvBars {
    uint64_t load[2], mma0, mma1;
    uint32_t tmem_base;
};
__global__ void __launch_bounds__(256, 2) attn_bwd_dkdv_kernel(const __grid_constant__ AttnParams p) {
    extern __shared__ uint8_t att_raw[];
    uint8_t* smem = att_align(att_raw);
    AttKvBars* bars = reinterpret_cast<AttKvBars*>(smem);
    float* s_lse = reinterpret_cast<float*>(smem + 128);     // [2][64]
    float* s_delta = s_lse + 2 * ATT_QSUB;                   // [2][64]  (ends at 128 + 1024 < 2048)
    uint8_t* sK = smem + 2048;            // 16 KB  (this CTA's key tile)
    uint8_t* sV = sK + 16384;             // 16 KB
    uint8_t* sQ = sV + 16384;             // 2 x 8 KB  (query sub-tile, double buffered)
    uint8_t* sdO = sQ + 16384;            // 2 x 8 KB
    uint8_t* sPT = sdO + 16384;           // 16 KB  P^T  [128 keys x 64 queries]
    uint8_t* sdST = sPT + 16384;          // 16 KB  dS^T [128 keys x 64 queries]
    const int warp = uniform_warp_id(), lane = threadIdx.x & 31;
    const int half = warp >> 2;          // two threads per key row: one 32-query chunk each
    const int kt = blockIdx.x & 1;
    const int bh = blockIdx.x >> 1;
    const int h = bh % p.H, b = bh / p.H;
    const int k0 = kt * ATT_TILE;
    if (k0 >= p.S) return;

    if (threadIdx.x == 0) {
        mbar_init(&bars->load[0], 1); mbar_init(&bars->load[1], 1); mbar_init(&bars->mma0, 1); mbar_init(&bars->mma1, 1);
        fence_mbar_init();
        tma_prefetch_desc(&p.qkv_map); tma_prefetch_desc(&p.qkv_map64); tma_prefetch_desc(&p.do_map64);
    }
    if (warp == 0) { tmem_alloc(&bars->tmem_base, 256); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = uniform_u32(bars->tmem_base);
    pdl_sync();
    const int HD = p.H * ATT_D;
    const int row_base = b * p.S;
    const int row = (warp & 3) * 32 + lane;      // key index inside the tile
    const int key = k0 + row;
    const uint32_t t_row = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    const int n_sub = (p.S + ATT_QSUB - 1) / ATT_QSUB;
    const size_t stat_base = (static_cast<size_t>(b) * p.H + h) * p.S;

    auto issue_sub = [&](int qt) {       // whole warp 0; one elected lane issues
        const int bsel = qt & 1;
        mbar_arrive_expect_tx_warp(&bars->load[bsel], qt == 0 ? 16384 * 2 + 8192 * 2 : 8192 * 2);
        if (qt == 0) {
            tma_load_2d_warp(sK, &p.qkv_map, &bars->load[0], HD + h * ATT_D, row_base + k0);
            tma_load_2d_warp(sV, &p.qkv_map, &bars->load[0], 2 * HD + h * ATT_D, row_base + k0);
        }
        tma_load_2d_warp(sQ + bsel * 8192, &p.qkv_map64, &bars->load[bsel], h * ATT_D, row_base + qt * ATT_QSUB);
        tma_load_2d_warp(sdO + bsel * 8192, &p.do_map64, &bars->load[bsel], h * ATT_D, row_base + qt * ATT_QSUB);
    };
    if (warp == 0) issue_sub(0);

    for (int qt = 0; qt < n_sub; ++qt) {
        const int bsel = qt & 1;
        const int q0 = qt * ATT_QSUB;
        const uint32_t ld_phase = (qt >> 1) & 1, ph = qt & 1;
        // the other Q / dO buffer was last read by the dV / dK MMAs of sub-tile qt - 1, which completed (mma1) before this
        // iteration started: prefetch sub-tile qt + 1 into it
        if (warp == 0 && qt + 1 < n_sub) issue_sub(qt + 1);
        if (threadIdx.x < ATT_QSUB) {
            const int qq = q0 + threadIdx.x;
            s_lse[bsel * ATT_QSUB + threadIdx.x] = qq < p.S ? p.lse[stat_base + qq] : INFINITY;
            s_delta[bsel * ATT_QSUB + threadIdx.x] = qq < p.S ? -p.delta[stat_base + qq] * p.scale : 0.f;     // -delta * scale
        }
        if (warp == 0) {
            mbar_wait(&bars->load[bsel], ld_phase);
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(ATT_TILE, ATT_QSUB, 0, 0);
            const uint32_t q_u32 = smem_u32(sQ + bsel * 8192), do_u32 = smem_u32(sdO + bsel * 8192);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // S^T = K Q^T   -> columns [0, 64)
                umma_bf16_warp(tmem, desc_kmajor(smem_u32(sK), k), desc_kmajor(q_u32, k), idesc, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // dP^T = V dO^T -> columns [64, 128)
                umma_bf16_warp(tmem + 64, desc_kmajor(smem_u32(sV), k), desc_kmajor(do_u32, k), idesc, k != 0);
            umma_commit_warp(&bars->mma0);
        }
        __syncthreads();                     // s_lse / s_delta of this sub-tile are visible
        mbar_wait(&bars->mma0, ph);
        tc_fence_after();
        const uint32_t sPT_u32 = smem_u32(sPT), sdST_u32 = smem_u32(sdST);
        {
            const int g = half;              // 32 queries per thread
            uint32_t sv[32], dp[32];
            tmem_ld_32x32(t_row + g * 32, sv);
            tmem_ld_32x32(t_row + 64 + g * 32, dp);
            tmem_ld_wait();
            const uint32_t lse_u32 = smem_u32(s_lse + bsel * ATT_QSUB + g * 32);
            const uint32_t del_u32 = smem_u32(s_delta + bsel * ATT_QSUB + g * 32);
            // s_lse holds +inf for queries beyond the sequence (exp2(-inf) = 0) and this row's scale is 0 for keys beyond
            // it, so the loop needs no per-element predicate; per-query constants come as 16-byte broadcast loads
            const float kscale = key < p.S ? p.scale_log2e : 0.f;
            const float koff = key < p.S ? 0.f : -INFINITY;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                float l4[4], d4[4];
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(l4[0]), "=f"(l4[1]), "=f"(l4[2]), "=f"(l4[3]) : "r"(lse_u32 + i * 4));
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(d4[0]), "=f"(d4[1]), "=f"(d4[2]), "=f"(d4[3]) : "r"(del_u32 + i * 4));
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float pr = fast_exp2(fmaf(__uint_as_float(sv[i + e]), kscale, koff - l4[e]));
                    sv[i + e] = __float_as_uint(pr);
                    dp[i + e] = __float_as_uint(pr * fmaf(__uint_as_float(dp[i + e]), p.scale, d4[e]));
                }
            }
            store_row_chunk32_bits(sPT_u32, row, g * 32, sv);
            store_row_chunk32_bits(sdST_u32, row, g * 32, dp);
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncthreads();
        if (warp == 0) {
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(ATT_TILE, ATT_D, 0, 1);
            const uint32_t q_u32 = smem_u32(sQ + bsel * 8192), do_u32 = smem_u32(sdO + bsel * 8192);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // dV += P^T dO   (dO read MN-major) -> columns [128, 192)
                umma_bf16_warp(tmem + 128, desc_kmajor(sPT_u32, k), desc_mnmajor(do_u32, k), idesc, (qt | k) != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)      // dK += dS^T Q   (Q read MN-major)  -> columns [192, 256)
                umma_bf16_warp(tmem + 192, desc_kmajor(sdST_u32, k), desc_mnmajor(q_u32, k), idesc, (qt | k) != 0);
            umma_commit_warp(&bars->mma1);
        }
        mbar_wait(&bars->mma1, ph);      // sPT / sdST and this Q / dO buffer may be overwritten from here on
        tc_fence_after();
        __syncthreads();
    }
    // ---- store dK (slot 1 of dqkv) and dV (slot 2)
#pragma unroll
    for (int which = 0; which < 2; ++which) {
        const int g = half;
        uint32_t v[32];
        tmem_ld_32x32(t_row + (which == 0 ? 192 : 128) + g * 32, v);
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
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

static int make_2d_map(CUtensorMap* m, const void* base, long long rows, int cols, int box_rows = 128) {
    uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
    uint64_t strides[1] = {(uint64_t)cols * 2};
    uint32_t box[2] = {64, (uint32_t)box_rows};
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
    const int smem_pipe = ATTP_HEADER + 2 * ATTP_BUF + 1024;
    static bool attr = false;
    if (!attr) {
        PICD_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        PICD_CUDA(cudaFuncSetAttribute(attn_fwd_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_pipe));
        attr = true;
    }
    // default: the persistent two-team kernel (B256 S197 H12: 0.140 -> 0.124 ms; library SDPA 0.104); PICD_ATTN_PIPE=0 = one item per CTA
    static const bool pipe = [] { const char* e = getenv("PICD_ATTN_PIPE"); return e == nullptr || atoi(e) != 0; }();
    if (pipe) {
        const int total = B * H * ((S + ATT_TILE - 1) / ATT_TILE);
        PICD_CUDA(launch_pdl<PDL_ATTN>(attn_fwd_pipe_kernel, dim3(std::min(total, num_sms())), dim3(ATTP_THREADS), smem_pipe, stream, p));
        PICD_LAUNCH_CHECK();
        return 0;
    }
    PICD_CUDA(launch_pdl<PDL_ATTN>(attn_fwd_kernel, dim3(B * H * 2), dim3(256), smem, stream, p));
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
    if (make_2d_map(&p.do_map64, dout, static_cast<long long>(B) * S, H * ATT_D, ATT_QSUB)) return 3;
    if (make_2d_map(&p.qkv_map64, qkv, static_cast<long long>(B) * S, 3 * H * ATT_D, ATT_QSUB)) return 3;
    p.o = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(o));
    p.dout = static_cast<const __nv_bfloat16*>(dout);
    p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
    p.lse = const_cast<float*>(lse);
    p.delta = delta;
    const int smem_dq = 1024 + 32768 + 16384 + 16384 + 32768 + 1024;       // 98 KB: two CTAs per SM
    const int smem_kv = 2048 + 16384 * 6 + 1024;       // 99 KB: two CTAs per SM
    static bool attr = false;
    if (!attr) {
        PICD_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq));
        PICD_CUDA(cudaFuncSetAttribute(attn_bwd_dkdv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kv));
        attr = true;
    }
    PICD_CUDA(launch_pdl<PDL_ATTN>(attn_bwd_dq_kernel, dim3(B * H * ((S + ATT_TILE - 1) / ATT_TILE)), dim3(256), smem_dq, stream, p));
    PICD_LAUNCH_CHECK();
    PICD_CUDA(launch_pdl<PDL_ATTN>(attn_bwd_dkdv_kernel, dim3(B * H * 2), dim3(256), smem_kv, stream, p));
    PICD_LAUNCH_CHECK();
    return 0;
}
