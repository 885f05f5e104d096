// Implicit-GEMM convolution / GEMM engine on tcgen05 + TMEM + TMA (sm_100a).
//
// Replaces the cuDNN conv fprop/dgrad/wgrad and cuBLAS Linear calls the reference reaches through
// torchvision.  One design covers every filter shape:
//
//   * activations are NHWC bf16; an output tile is a box of 128 pixels [bn images x bh rows x bw cols];
//   * a convolution is a list of TAPS.  For tap (r,s) the A operand of the GEMM is the SAME box shifted by
//     (r-pad, s-pad): one 4-D TMA load whose out-of-bounds elements are zero-filled by the hardware, so
//     padding, 1x7/7x1/5x5 filters and image borders need no special code and no im2col buffer exists;
//   * stride-2 convolutions address one of four "parity views" of the tensor (x[:, ph::2, pw::2, :], encoded as
//     plain strided tensor maps), which turns them into stride-1 tap lists; stride-2 dgrad is four such
//     launches writing through the parity views of dx;
//   * weights stay in their channels_last layout [Cout, R*S, Cin]: K-major B operand for fprop, MN-major B
//     operand (same memory, different UMMA descriptor) for dgrad -- no transposed weight copy;
//   * wgrad is a GEMM over the pixel dimension with BOTH operands MN-major (dy and x tiles exactly as TMA
//     lands them), split-K over pixel ranges, fp32 accumulation in TMEM, vector red.add into the fp32
//     gradient arena.
//
// conv_gemm_kernel: persistent, warp-specialised (warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue), 4-8 stage
// smem ring, double-buffered 128xBN fp32 accumulator in TMEM, epilogue converts to bf16 into a 128B-swizzled staging tile
// that one thread stores with TMA (or TMA reduce-add) while all eight warps accumulate per-channel sum / sum-of-squares for
// BatchNorm from the same staging tile (fused BN statistics).  Variants selected per launch:
//   * halo mode (3x3, stride 1): one activation box with a halo row above/below per column shift, the three dh taps are
//     descriptor shifts inside it; tiny weight sets stay resident in shared memory;
//   * CTA pairs (tcgen05 cta_group::2): two SMs execute one M=256 UMMA, each loading its own pixel rows and half of B;
//   * producer and MMA loops are warp-convergent (elect.sync inside the asm) so that operands live in uniform registers.
#include "../common/picd_common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace picd {

constexpr int MAX_TAPS = 64;
constexpr int TILE_M = 128;
constexpr int KCHUNK = 64;                   // bf16 elements per 128B swizzle row
constexpr int A_STAGE_BYTES = TILE_M * 128;  // 16 KB
constexpr int SMEM_HEADER = 2048;           // mbarriers + TMEM base (first 512 B), bias tile of the CTA (1 KB at +512)
constexpr int SMEM_BIAS_OFF = 512;
constexpr int SMEM_HEADER_BNR = 5120;       // bn_reduce launches: + scale / shift / mean of the CTA's 256 channels (3 x 1 KB at +2048)
constexpr int SMEM_COEF_OFF = 2048;
constexpr int MAX_DYN_SMEM = 232448;         // 227 KB

struct Tap {
    int16_t amap;   // which A tensor map (parity view)
    int16_t dh, dw; // pixel offset added to the tile origin
    int16_t btap;   // tap index inside the weight tensor
};

struct ConvGemmParams {
    CUtensorMap amap[4];
    CUtensorMap bmap;
    CUtensorMap dmap;
    Tap taps[MAX_TAPS];
    int ntaps, kchunks, b_mn;
    int BN, numN, num_m_tiles, stages;
    int tiles_w, tiles_h;
    int bw_log2, bh_log2;
    int W, H, N;          // output extents (for masking over-hanging rows)
    int Cout;
    float* stats;         // [STAT_REPLICAS][2][Cout] (sum, sumsq) accumulators or nullptr
    const float* bias;    // [Cout] or nullptr
    int act;
    int half_w;           // columns drained per staging pass: 128 (deep operand ring) or 256 (short-K tiles, one pass)
    int cluster;          // 1: launched as CTA pairs (clusters of 2, cta_group::2): each CTA loads its own half of every B tile
    int reduce_add;       // 1: the output tile is ADDED to the destination (TMA reduce) instead of overwriting it
    int quad;             // pairs only, K-major B, free schedule: clusters of FOUR CTAs = two pairs on vertically adjacent 256-row
                          // tiles of the same column tile.  The pairs need the same B tile: each CTA fetches a QUARTER of it and
                          // TMA-multicasts that quarter to its twin in the other pair (tensor map over this GPU's memory), so a
                          // k-iteration pulls 16 + 8 KB instead of 16 + 16 KB per CTA through L2.
    int nbw;              // free schedule: column tiles are walked in blocks of nbw (all row tiles of a block before the next block),
                          // so a block's B rows stay in L2 while the A rows stream past once per block
    int free_sched;       // pairs only, no per-channel statistics: pair-tiles are dealt round-robin over ALL pairs (the column
                          // tile of a CTA changes from tile to tile), so every SM pair works whatever numN is
    // halo mode (3x3, stride 1): one activation box with a one-row halo above and below is loaded per (k-chunk, dw) and
    // the three dh taps read it through descriptors shifted by whole pixel rows -> 3 instead of 9 A loads per k-chunk.
    // taps[] is then ordered [dw][dh] with dh = -1, 0, +1.
    int halo;
    int a_stage_bytes;    // bytes of one A stage: 16 KB, or (bh + 2) * bw * 128 in halo mode
    int b_resident;       // halo mode, tiny weights: all 9 B taps are loaded once per CTA and stay in shared memory
    int header_bytes;     // shared-memory header: SMEM_HEADER, or SMEM_HEADER_BNR when BatchNorm coefficient vectors are staged
    // ---- fused BatchNorm-backward reduction (dgrad launches) --------------------------------------------------
    // The tile this kernel stores is dz of the UPSTREAM conv+BN(+ReLU) layer.  With that layer's raw conv output y (ymap,
    // same geometry as dmap: one extra TMA load per output pass) and its BN coefficients the epilogue accumulates
    // sum(g) and sum(g * xhat), g = dz * [y*scale+shift > 0], xhat = (y - mean) * invstd, from the bf16-rounded values it
    // stores -- exactly what bn_bwd_reduce would compute from a second pass over dz and y, which therefore disappears.
    int bn_reduce;        // 0 / 1; the sums go to `stats` ([STAT_REPLICAS][2][Cout], same layout as the forward statistics)
    int r_relu;           // upstream activation is ReLU (mask recomputed from y) or identity
    int ybufs;            // y tiles in flight + 1: the tile of pass k + ybufs - 1 is requested when pass k starts (1..4)
    CUtensorMap ymap;
    const float* r_scale; const float* r_shift; const float* r_mean; const float* r_invstd;
};

struct WgradParams {
    CUtensorMap dymap;
    CUtensorMap xmap[4];
    Tap taps[MAX_TAPS];
    int ntaps;
    float* dw;            // fp32 [Cout][taps_total][Cin], accumulated with red.add
    int Cout, Cin, taps_total;
    int BN, numN, numM, ksplit, num_p_tiles, stages;
    int tiles_w, tiles_h, bw_log2, bh_log2;
    int G, cin_blocks, ngroups;   // taps per CTA (their B tiles sit side by side along N), 64-channel blocks per tap
    // bias gradient db[cout] += sum over pixels of dy[pixel, cout] on the tensor cores: the CTAs that own the first tap / first
    // input-channel tile multiply their dy tiles by a constant tile of ONES as well (one extra N = 16 MMA per k-step into 16
    // spare TMEM columns) -- the separate column-sum pass over dy disappears.
    float* dbias;                 // [Cout] fp32, accumulated with atomics; nullptr = no bias
    int tmem_cols;                // 256, or 512 when BN + 16 spare columns do not fit
};

struct SmemBars {
    uint64_t full[8];
    uint64_t empty[8];
    uint64_t tfull[2];
    uint64_t tempty[2];
    uint64_t bres;        // halo mode with resident weights: "all taps have landed"
    uint64_t ybar[4];     // bn_reduce: "the upstream y tile of an output pass has landed" (one per y buffer)
    uint32_t tmem_base;
};

// (w, h, n) coordinates of the tile a CTA works on, without per-tile integer divisions: the tile index advances by a constant step
// whose digits in the mixed radix (tiles_w, tiles_h) are computed once; an advance is three adds and two conditional carries.
// (ncu on the 56x56 64-channel layers: four division sequences -- I2F / MUFU.RCP / F2I plus ~20 integer ops each -- per tile and
// warp sat on the critical path of an epilogue that is the bottleneck of those layers.)
struct TileCoord {
    int tw, th, tn, dw, dh, dn;
    __device__ __forceinline__ void set(int m, int tiles_w, int tiles_h) {
        tw = m % tiles_w;
        const int q = m / tiles_w;
        th = q % tiles_h;
        tn = q / tiles_h;
    }
    __device__ __forceinline__ void init(int m, int step, int tiles_w, int tiles_h) {
        set(m, tiles_w, tiles_h);
        dw = step % tiles_w;
        const int q = step / tiles_w;
        dh = q % tiles_h;
        dn = q / tiles_h;
    }
    __device__ __forceinline__ void advance(int tiles_w, int tiles_h) {
        tw += dw;
        int c = tw >= tiles_w ? 1 : 0;
        tw -= c ? tiles_w : 0;
        th += dh + c;
        c = th >= tiles_h ? 1 : 0;
        th -= c ? tiles_h : 0;
        tn += dn + c;
    }
};

__device__ __forceinline__ uint8_t* align_smem(uint8_t* p) {
    return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~static_cast<uintptr_t>(1023));
}

// =====================================================================================================
//  conv / GEMM forward-style kernel: D[pixels, Cout] = sum_taps A_tap[pixels, Cin] * B_tap[Cout, Cin]^T
// =====================================================================================================
// kCluster: CTA PAIRS (tcgen05 cta_group::2).  The two CTAs of a cluster work on adjacent pixel tiles of the SAME column
// tile as ONE 256-row UMMA: each CTA loads its own 128 pixel rows of A and HALF of every B (weight) tile, the leader
// (cluster rank 0) issues the MMAs, each CTA drains its own 128 accumulator rows from its own TMEM.  Per CTA a k-iteration
// then moves 16 + 16 KB instead of 16 + 32 KB through shared memory for the same tensor work (K-major B only).
template <bool kBiasAct, bool kCluster>
__global__ void __launch_bounds__(320, 1) conv_gemm_kernel(const __grid_constant__ ConvGemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem(smem_raw);
    SmemBars* bars = reinterpret_cast<SmemBars*>(smem);
    const int BN = p.BN;
    const int stages = p.stages;
    const int b_tap = (kCluster ? BN / 2 : BN) * 128;             // one weight tap tile: BN rows x 64 k (or 64 k x BN);
                                                                  // a CTA of a pair holds half the rows
    const int b_stage = p.halo ? (p.b_resident ? 0 : 3 * b_tap) : b_tap;
    const int a_stage = p.a_stage_bytes;
    uint8_t* sA = smem + p.header_bytes;
    uint8_t* sB = sA + stages * a_stage;
    uint8_t* sBres = sB + stages * b_stage;                       // 9 resident taps (halo + b_resident only)
    uint8_t* sOut = sBres + (p.b_resident ? 9 * b_tap : 0);
    uint8_t* sY = sOut + (BN < p.half_w ? BN : p.half_w) * 256;   // bn_reduce: upstream y tile of the current pass (same shape as sOut)

    const int warp = uniform_warp_id();      // provably warp-uniform: the role branches below stay convergent
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) {
            mbar_init(&bars->full[i], 1);
            mbar_init(&bars->empty[i], (kCluster && p.quad) ? 2 : 1);   // pair: the leader's commit arrives once in each CTA;
                                                                        // quad: a stage is free when BOTH pairs have consumed it
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bars->tfull[i], 1);
            mbar_init(&bars->tempty[i], kCluster ? 16 : 8);   // pair: the leader waits for both CTAs' epilogue warps
        }
        mbar_init(&bars->bres, 1);
        for (int i = 0; i < 4; ++i) mbar_init(&bars->ybar[i], 1);
        fence_mbar_init();
    }
    if (kCluster) cluster_sync();          // both CTAs are resident before the pair-wide TMEM allocation
    if (warp == 1) {
        if (kCluster) { tmem_alloc_pair(&bars->tmem_base, 512); tmem_relinquish_pair(); }
        else { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.amap[0]);
        tma_prefetch_desc(&p.bmap);
        tma_prefetch_desc(&p.dmap);
        if (p.bn_reduce) tma_prefetch_desc(&p.ymap);
    }
    tc_fence_before();
    __syncthreads();
    if (kCluster) cluster_sync();          // the peer's barriers exist before anything arrives on them
    tc_fence_after();
    const uint32_t tmem_base = uniform_u32(bars->tmem_base);
    pdl_sync();        // everything above overlapped the previous kernel's tail; from here on its results are visible

    const int numN = p.numN;
    const int crank = kCluster ? static_cast<int>(uniform_u32(cluster_ctarank())) : 0;
    // tile walk: (non-cluster) CTA b owns column tile b % numN and pixel tiles b/numN, +grid_m, ...
    //            (cluster)     cluster c owns column tile c % numN; its two CTAs take pixel tiles 2j and 2j+1 in lockstep
    const bool quad = kCluster && p.quad != 0;
    const int csh = quad ? 2 : 1;            // log2(cluster size)
    const int unit = kCluster ? (blockIdx.x >> csh) : blockIdx.x;
    const int units = kCluster ? (gridDim.x >> csh) : gridDim.x;
    const int prank = crank & 1;             // rank inside the CTA pair (0 = the pair's leader)
    const bool free_s = kCluster && p.free_sched != 0;
    int n_tile = unit % numN;                // (free schedule: the column tile of the CURRENT tile, set by tile_at)
    const int grid_m = kCluster ? 2 * (units / numN) : units / numN;
    const int m_first = kCluster ? 2 * (unit / numN) + crank : unit / numN;
    const int m_limit = kCluster ? p.num_m_tiles + crank : p.num_m_tiles;   // the pair iterates while its EVEN tile exists
    // i-th tile of this CTA: sets n_tile and m, false when the walk is over.  Free schedule: pair-tile t = unit + i * units
    // with the column tile fastest, so concurrently running pairs share their A rows through L2.
    TileCoord tc;                            // coordinates of the current tile (per thread: every role walks on its own)
    auto tile_at = [&](int i, int& m) -> bool {
        if (free_s) {
            const int t = unit + i * units;
            const int mgroups = (p.num_m_tiles + (1 << csh) - 1) >> csh;      // row groups (pairs / quads of 128-row tiles)
            if (t >= mgroups * numN) return false;
            const int per_block = p.nbw * mgroups;
            const int blk = t / per_block;
            const int r = t - blk * per_block;
            const int w = min(p.nbw, numN - blk * p.nbw);                     // the last block may be narrower
            const int mg = r / w;
            n_tile = blk * p.nbw + (r - mg * w);
            m = (mg << csh) + crank;
            tc.set(m, p.tiles_w, p.tiles_h);
            return true;
        }
        m = m_first + i * grid_m;
        if (i == 0) tc.init(m, grid_m, p.tiles_w, p.tiles_h);
        else tc.advance(p.tiles_w, p.tiles_h);
        return m < m_limit;
    };
    const int k_iters = p.halo ? 3 * p.kchunks : p.ntaps * p.kchunks;
    float* sBias = reinterpret_cast<float*>(smem + SMEM_BIAS_OFF);
    if (kBiasAct) {
        // a CTA keeps one column tile for its whole life, so its slice of the bias vector is staged once
        for (int i = threadIdx.x; i < BN; i += blockDim.x) {
            const int c = n_tile * BN + i;
            sBias[i] = (p.bias != nullptr && c < p.Cout) ? p.bias[c] : 0.f;
        }
        __syncthreads();
    }
    float* sCoef = reinterpret_cast<float*>(smem + SMEM_COEF_OFF);       // [3][256]: scale, shift, mean of the upstream BN
    if (p.bn_reduce) {
        for (int i = threadIdx.x; i < BN; i += blockDim.x) {
            const int c = n_tile * BN + i;
            const bool in = c < p.Cout;
            sCoef[i] = in ? p.r_scale[c] : 0.f;
            sCoef[256 + i] = in ? p.r_shift[c] : 0.f;
            sCoef[512 + i] = in ? p.r_mean[c] : 0.f;
        }
        __syncthreads();
    }
    const int bw = 1 << p.bw_log2, bh = 1 << p.bh_log2;
    const int bn = TILE_M >> (p.bw_log2 + p.bh_log2);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        {   // all 32 lanes walk the loop; the *_warp calls elect one lane to issue
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx_bytes = a_stage + b_stage;
            auto load_b_tap = [&](uint8_t* dst, uint64_t* bar, int kc, int btap) {
                if (!p.b_mn) {
                    tma_load_3d_warp(dst, &p.bmap, bar, kc * KCHUNK, btap, n_tile * BN);
                } else {
                    for (int j = 0; j < BN / 64; ++j) {
                        int c0 = n_tile * BN + j * 64;
                        if (c0 >= p.Cout) c0 = 0;   // never request a fully out-of-range box
                        tma_load_3d_warp(dst + j * 8192, &p.bmap, bar, c0, btap, kc * KCHUNK);
                    }
                }
            };
            if (p.halo && p.b_resident) {
                mbar_arrive_expect_tx_warp(&bars->bres, 9 * b_tap);
                for (int t = 0; t < 9; ++t) load_b_tap(sBres + t * b_tap, &bars->bres, 0, p.taps[t].btap);
            }
            for (int ti = 0, m = 0; tile_at(ti, m); ++ti) {
                const int w0 = tc.tw * bw, h0 = tc.th * bh, n0 = tc.tn * bn;
                if (p.halo) {
                    for (int kc = 0; kc < p.kchunks; ++kc) {
                        for (int dwi = 0; dwi < 3; ++dwi) {
                            const Tap tp = p.taps[dwi * 3];          // the dh = -1 tap of this column: origin of the halo box
                            mbar_wait(&bars->empty[stage], phase ^ 1);
                            mbar_arrive_expect_tx_warp(&bars->full[stage], tx_bytes);
                            tma_load_4d_warp(sA + stage * a_stage, &p.amap[0], &bars->full[stage], kc * KCHUNK, w0 + tp.dw,
                                        h0 + tp.dh, n0);
                            if (!p.b_resident) {
                                for (int dhi = 0; dhi < 3; ++dhi)
                                    load_b_tap(sB + stage * b_stage + dhi * b_tap, &bars->full[stage], kc,
                                               p.taps[dwi * 3 + dhi].btap);
                            }
                            if (++stage == stages) { stage = 0; phase ^= 1; }
                        }
                    }
                    continue;
                }
                for (int t = 0; t < p.ntaps; ++t) {
                    const Tap tp = p.taps[t];
                    for (int kc = 0; kc < p.kchunks; ++kc) {
                        mbar_wait(&bars->empty[stage], phase ^ 1);
                        uint8_t* dstB = sB + stage * b_stage;
                        if (kCluster) {
                            // pair: the leader's barrier collects the bytes of BOTH CTAs; this CTA fetches its own pixel
                            // rows and weight rows [crank * BN/2, +BN/2) into its own shared memory
                            if (prank == 0) mbar_arrive_expect_tx_warp(&bars->full[stage], 2 * tx_bytes);
                            tma_load_4d_pair_warp(sA + stage * a_stage, &p.amap[tp.amap], &bars->full[stage], kc * KCHUNK,
                                                  w0 + tp.dw, h0 + tp.dh, n0);
                            if (quad) {
                                // this CTA's quarter of the B tile, delivered to itself and to its twin (same pair rank) in
                                // the other pair; each delivery credits the leader of the pair it lands in
                                const int twin = crank >> 1;
                                tma_load_3d_pair_mc_warp(dstB + twin * (b_tap >> 1), &p.bmap, &bars->full[stage], kc * KCHUNK,
                                                         tp.btap, n_tile * BN + prank * (BN >> 1) + twin * (BN >> 2),
                                                         0x5u << prank);
                            } else if (!p.b_mn) {
                                tma_load_3d_pair_warp(dstB, &p.bmap, &bars->full[stage], kc * KCHUNK, tp.btap,
                                                      n_tile * BN + prank * (BN >> 1));
                            } else {                     // MN-major: this CTA's half of the columns, in 64-column boxes
                                for (int j = 0; j < BN / 128; ++j) {
                                    int c0 = n_tile * BN + prank * (BN >> 1) + j * 64;
                                    if (c0 >= p.Cout) c0 = 0;   // never request a fully out-of-range box
                                    tma_load_3d_pair_warp(dstB + j * 8192, &p.bmap, &bars->full[stage], c0, tp.btap, kc * KCHUNK);
                                }
                            }
                        } else {
                            mbar_arrive_expect_tx_warp(&bars->full[stage], tx_bytes);
                            tma_load_4d_warp(sA + stage * a_stage, &p.amap[tp.amap], &bars->full[stage], kc * KCHUNK,
                                             w0 + tp.dw, h0 + tp.dh, n0);
                            if (!p.b_mn) {
                                tma_load_3d_warp(dstB, &p.bmap, &bars->full[stage], kc * KCHUNK, tp.btap, n_tile * BN);
                            } else {
                                for (int j = 0; j < BN / 64; ++j) {
                                    int c0 = n_tile * BN + j * 64;
                                    if (c0 >= p.Cout) c0 = 0;   // never request a fully out-of-range box
                                    tma_load_3d_warp(dstB + j * 8192, &p.bmap, &bars->full[stage], c0, tp.btap, kc * KCHUNK);
                                }
                            }
                        }
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        // The whole warp walks the loop (uniform control flow keeps the descriptors in uniform registers); one elected
        // lane issues.  Descriptors are built once from templates: only the 14-bit start-address field changes.
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t accphase = 0;
        const uint32_t idesc = make_idesc_bf16(kCluster ? 2 * TILE_M : TILE_M, BN, 0, p.b_mn);
        const uint64_t a_tmpl = make_smem_desc(0, 16, 1024);
        const uint64_t b_tmpl = p.b_mn ? make_smem_desc(0, 8192, 1024) : make_smem_desc(0, 16, 1024);
        const uint32_t b_kstep = p.b_mn ? (2048u >> 4) : (32u >> 4);      // descriptor units (16 B) per K=16 step
        const uint32_t sA_lo = smem_u32(sA) >> 4, sB_lo = smem_u32(sB) >> 4, sBres_lo = smem_u32(sBres) >> 4;
        const uint32_t a_stage_lo = static_cast<uint32_t>(a_stage) >> 4, b_stage_lo = static_cast<uint32_t>(b_stage) >> 4;
        const uint32_t b_tap_lo = static_cast<uint32_t>(b_tap) >> 4;
        const uint32_t row_lo = (static_cast<uint32_t>(bw) * 128u) >> 4;   // one pixel row of the halo box
        if (p.halo && p.b_resident) { mbar_wait(&bars->bres, 0); tc_fence_after(); }
        for (int ti = 0, m = 0; tile_at(ti, m) && (!kCluster || prank == 0); ++ti) {     // pair: only the leader issues
            mbar_wait(&bars->tempty[acc], accphase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * 256;
            for (int it = 0; it < k_iters; ++it) {
                mbar_wait(&bars->full[stage], phase);
                tc_fence_after();
                const uint64_t adesc0 = a_tmpl | (sA_lo + stage * a_stage_lo);
                if (p.halo) {
                    // the stage holds rows h0-1 .. h0+bh of one column shift: tap dh reads it from row (dh + 1) on
                    const int dwi = it % 3;
                    const uint64_t bdesc0 = b_tmpl | (p.b_resident ? sBres_lo + dwi * 3 * b_tap_lo : sB_lo + stage * b_stage_lo);
#pragma unroll
                    for (int dhi = 0; dhi < 3; ++dhi) {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_warp(tmem_d, adesc0 + (dhi * row_lo + k * 2), bdesc0 + (dhi * b_tap_lo + k * b_kstep), idesc,
                                           (it | dhi | k) != 0);
                    }
                } else {
                    const uint64_t bdesc0 = b_tmpl | (sB_lo + stage * b_stage_lo);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        if (kCluster) umma_bf16_pair_warp(tmem_d, adesc0 + k * 2, bdesc0 + k * b_kstep, idesc, (it | k) != 0);
                        else umma_bf16_warp(tmem_d, adesc0 + k * 2, bdesc0 + k * b_kstep, idesc, (it | k) != 0);
                    }
                }
                if (kCluster) umma_commit_pair_warp(&bars->empty[stage], quad ? 0xFu : 3u);   // frees the stage in both CTAs (quad:
                                                                                                // one of two arrivals, in all four)
                else umma_commit_warp(&bars->empty[stage]);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
            if (kCluster) umma_commit_pair_warp(&bars->tfull[acc], 3u << (crank & 2));   // both CTAs' accumulator rows are complete
            else umma_commit_warp(&bars->tfull[acc]);
            acc ^= 1;
            if (acc == 0) accphase ^= 1;
        }
    } else {
        // ------------------------------------------------------------------ epilogue (8 warps)
        // Kept deliberately compact (explicit shared-space accesses, no per-element branches): an earlier version
        // unrolled into ~4800 SASS instructions and stalled on instruction fetch (profiles/ncu_fprop_64_256_v1.md).
        // Two warps share each TMEM lane quarter and split the columns of a pass between them: with one warp per SM
        // sub-partition nothing hid the tcgen05.ld / ld.shared latencies and short-K layers ran at the epilogue's pace.
        const int q = warp & 3;                  // TMEM lane quarter this warp may read
        const int colhalf = (warp - 2) >> 2;     // which 32-column groups of a pass this warp drains (even / odd)
        const int row = q * 32 + lane;           // accumulator row == pixel index inside the tile
        const int tid_e = threadIdx.x - 64;      // 0..255
        // The tile is drained in column passes of <= half_w channels through ONE staging buffer (the smem this saves buys a
        // 4th operand stage for 256-wide tiles: with 3 stages the TMA round trip was longer than the MMA time it had to hide).
        // statistics: per pass, thread -> (8-channel chunk, row group); partial sums live in registers across all tiles
        float stS[2][8], stQ[2][8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { stS[0][i] = stS[1][i] = 0.f; stQ[0][i] = stQ[1][i] = 0.f; }
        const int hw = p.half_w, hw_log2 = hw == 256 ? 8 : 7;
        const int n_halves = (BN + hw - 1) >> hw_log2;
        int acc = 0;
        uint32_t accphase = 0;
        // bn_reduce: the y tiles are requested ybufs - 1 passes ahead of their use (an output pass of a short-K layer lasts
        // less than one DRAM round trip).  Thread tid_e == 0 walks the same (tile, pass) sequence with a cursor of its own.
        const uint32_t ysize = static_cast<uint32_t>(BN < hw ? BN : hw) * 256u;
        int pf_m = m_first, pf_hf = 0;
        uint32_t pf_buf = 0, yc_buf = 0, yc_phase = 0;       // y-buffer cursors of the requests / of the consumer (no modulo per pass)
        TileCoord pfc;
        pfc.init(m_first, grid_m, p.tiles_w, p.tiles_h);
        auto issue_y = [&]() {
            if (pf_m >= m_limit) return;
            if (pf_m < p.num_m_tiles) {
                const int ptw = pfc.tw, pth = pfc.th, ptn = pfc.tn;
                const int pcols = (BN - pf_hf * hw) < hw ? (BN - pf_hf * hw) : hw;
                uint8_t* dst = sY + pf_buf * ysize;
                uint64_t* bar = &bars->ybar[pf_buf];
                uint32_t ybytes = 0;
                for (int j = 0; j < (pcols >> 6); ++j)
                    if (n_tile * BN + pf_hf * hw + j * 64 < p.Cout) ybytes += 16384;
                mbar_arrive_expect_tx(bar, ybytes);
                for (int j = 0; j < (pcols >> 6); ++j) {
                    const int c0 = n_tile * BN + pf_hf * hw + j * 64;
                    if (c0 < p.Cout) tma_load_4d(dst + j * 16384, &p.ymap, bar, c0, ptw * bw, pth * bh, ptn * bn);
                }
                if (++pf_buf == static_cast<uint32_t>(p.ybufs)) pf_buf = 0;
            }
            if (++pf_hf == n_halves) { pf_hf = 0; pf_m += grid_m; pfc.advance(p.tiles_w, p.tiles_h); }
        };
        if (p.bn_reduce && tid_e == 0)
            for (int i = 0; i + 1 < p.ybufs; ++i) issue_y();
        const int iw = row & (bw - 1);
        const int ih = (row >> p.bw_log2) & (bh - 1);
        const int in_ = row >> (p.bw_log2 + p.bh_log2);
        const uint32_t sOut_u32 = smem_u32(sOut);
        const uint32_t row_base = sOut_u32 + row * 128;
        const uint32_t row_xor = static_cast<uint32_t>(row & 7) << 4;
        const uint32_t bias_u32 = smem_u32(sBias);

        // converts one 32-column group to bf16 and writes it into the 128B-swizzled staging tile
        auto stage_group = [&](uint32_t (&v)[32], int g, int gg, uint32_t vmask) {
            if (kBiasAct) {
                const uint32_t bsrc = bias_u32 + g * 128;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    float b0, b1, b2, b3;     // same address in every lane: a shared-memory broadcast
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(b0), "=f"(b1), "=f"(b2), "=f"(b3) : "r"(bsrc + i * 4));
                    v[i + 0] = __float_as_uint(__uint_as_float(v[i + 0]) + b0);
                    v[i + 1] = __float_as_uint(__uint_as_float(v[i + 1]) + b1);
                    v[i + 2] = __float_as_uint(__uint_as_float(v[i + 2]) + b2);
                    v[i + 3] = __float_as_uint(__uint_as_float(v[i + 3]) + b3);
                }
                if (p.act == ACT_RELU) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(fmaxf(__uint_as_float(v[i]), 0.f));
                }   // other activations are separate kernels: a generic switch here forced v[] into local memory
            }
            const uint32_t dst = row_base + (gg >> 1) * 16384;
            const uint32_t chunk0 = static_cast<uint32_t>(gg & 1) * 64;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint32_t a0 = pack_bf16x2(__uint_as_float(v[8 * c + 0]), __uint_as_float(v[8 * c + 1])) & vmask;
                const uint32_t a1 = pack_bf16x2(__uint_as_float(v[8 * c + 2]), __uint_as_float(v[8 * c + 3])) & vmask;
                const uint32_t a2 = pack_bf16x2(__uint_as_float(v[8 * c + 4]), __uint_as_float(v[8 * c + 5])) & vmask;
                const uint32_t a3 = pack_bf16x2(__uint_as_float(v[8 * c + 6]), __uint_as_float(v[8 * c + 7])) & vmask;
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + ((chunk0 + c * 16) ^ row_xor)),
                             "r"(a0), "r"(a1), "r"(a2), "r"(a3)
                             : "memory");
            }
        };

        for (int ti = 0, m = 0; tile_at(ti, m); ++ti) {
            const int w0 = tc.tw * bw, h0 = tc.th * bh, n0 = tc.tn * bn;
            const bool tile_ok = m < p.num_m_tiles;          // the odd CTA of the last pair may have no tile of its own
            if (kBiasAct && free_s) {
                // the column tile moved: restage its bias slice.  Every reader of the old slice is past the barrier that closed the
                // previous tile's last pass; the first barrier of the pass below publishes the new one.
                for (int i = tid_e; i < BN; i += 256) {
                    const int c = n_tile * BN + i;
                    sBias[i] = (p.bias != nullptr && c < p.Cout) ? p.bias[c] : 0.f;
                }
            }
            const bool valid = tile_ok && (w0 + iw < p.W) && (h0 + ih < p.H) && (n0 + in_ < p.N);
            const uint32_t vmask = valid ? 0xffffffffu : 0u;     // rows hanging over the tensor edge are stored as 0
            mbar_wait(&bars->tfull[acc], accphase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + acc * 256 + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
            for (int hf = 0; hf < n_halves; ++hf) {
                const int hcols = (BN - hf * hw) < hw ? (BN - hf * hw) : hw;     // 64, 128 or 256 columns in this pass
                const int g0 = hf * (hw >> 5);
                const int ngt = hcols >> 6;                   // 32-column groups per thread: 1, 2 or 4
                uint32_t va[32], vb[32];
                tmem_ld_32x32(t_row + (g0 + colhalf) * 32, va);       // in flight while the staging buffer drains
                if (tid_e == 0) tma_store_wait_read();        // the previous TMA store has drained the staging buffer
                named_barrier_sync(1, 256);
                // everybody is past the statistics of the previous pass, so the y buffer it used is free: request the tile
                // of the pass ybufs - 1 ahead into it
                if (p.bn_reduce && tid_e == 0) issue_y();
#pragma unroll 1
                for (int j = 0; j < ngt; j += 2) {            // software pipeline: the next group loads while this one converts
                    const int gg = 2 * j + colhalf;
                    tmem_ld_wait();
                    if (j + 1 < ngt) tmem_ld_32x32(t_row + (g0 + gg + 2) * 32, vb);
                    stage_group(va, g0 + gg, gg, vmask);
                    if (j + 1 < ngt) {
                        tmem_ld_wait();
                        if (j + 2 < ngt) tmem_ld_32x32(t_row + (g0 + gg + 4) * 32, va);
                        stage_group(vb, g0 + gg + 2, gg + 2, vmask);
                    }
                }
                if (hf == n_halves - 1) {                     // every accumulator column has been read
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (kCluster) mbar_arrive_leader(&bars->tempty[acc]);      // the leader's MMA warp waits for both CTAs
                        else mbar_arrive(&bars->tempty[acc]);
                    }
                }
                fence_proxy_async_smem();
                named_barrier_sync(1, 256);
                if (tid_e == 0) {
                    for (int j = 0; j < (hcols >> 6) && tile_ok; ++j) {
                        const int c0 = n_tile * BN + hf * hw + j * 64;
                        if (c0 < p.Cout) {
                            if (p.reduce_add) tma_reduce_add_4d(&p.dmap, sOut + j * 16384, c0, w0, h0, n0);
                            else tma_store_4d(&p.dmap, sOut + j * 16384, c0, w0, h0, n0);
                        }
                    }
                    tma_store_commit();
                }
                if (p.bn_reduce && tile_ok) {
                    // BatchNorm-backward sums of the upstream layer: g = dz * relu'(y*scale+shift) from the staged bf16 tile,
                    // y from the tile TMA just delivered; thread -> (8-channel chunk, row group) as for the forward statistics
                    const int nch = hcols >> 3;                            // 8, 16 or 32
                    const int my_ch = tid_e & (nch - 1);
                    const int rows_per_rg = nch >> 1;
                    const int r0 = (tid_e >> (31 - __clz(nch))) * rows_per_rg;
                    const uint32_t chunk_off = (my_ch >> 3) * 16384;
                    const uint32_t stat_chunk = static_cast<uint32_t>(my_ch & 7);
                    const uint32_t y_u32 = smem_u32(sY) + yc_buf * ysize;
                    float sc[8], sh[8], mu[8];
                    {
                        const uint32_t cbase = smem_u32(sCoef) + (hf * hw + my_ch * 8) * 4;
#pragma unroll
                        for (int i = 0; i < 8; i += 4) {
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(sc[i]), "=f"(sc[i + 1]), "=f"(sc[i + 2]), "=f"(sc[i + 3]) : "r"(cbase + i * 4));
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(sh[i]), "=f"(sh[i + 1]), "=f"(sh[i + 2]), "=f"(sh[i + 3]) : "r"(cbase + 1024 + i * 4));
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(mu[i]), "=f"(mu[i + 1]), "=f"(mu[i + 2]), "=f"(mu[i + 3]) : "r"(cbase + 2048 + i * 4));
                        }
                    }
                    if (!p.r_relu) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { sc[i] = 0.f; sh[i] = 1.f; }      // mask always true
                    }
                    float s[8], qq[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] = 0.f; qq[i] = 0.f; }
                    mbar_wait(&bars->ybar[yc_buf], yc_phase);
                    if (++yc_buf == static_cast<uint32_t>(p.ybufs)) { yc_buf = 0; yc_phase ^= 1u; }
#pragma unroll 2
                    for (int r = r0; r < r0 + rows_per_rg; ++r) {
                        const uint32_t off = chunk_off + r * 128 + ((stat_chunk ^ (r & 7)) << 4);
                        uint32_t w4[4], y4[4];
                        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                     : "=r"(w4[0]), "=r"(w4[1]), "=r"(w4[2]), "=r"(w4[3]) : "r"(sOut_u32 + off));
                        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                     : "=r"(y4[0]), "=r"(y4[1]), "=r"(y4[2]), "=r"(y4[3]) : "r"(y_u32 + off));
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float2 f = unpack_bf16x2(w4[i]);
                            const float2 yv = unpack_bf16x2(y4[i]);
                            const float g0 = fmaf(yv.x, sc[2 * i], sh[2 * i]) > 0.f ? f.x : 0.f;
                            const float g1 = fmaf(yv.y, sc[2 * i + 1], sh[2 * i + 1]) > 0.f ? f.y : 0.f;
                            s[2 * i] += g0; qq[2 * i] = fmaf(g0, yv.x - mu[2 * i], qq[2 * i]);
                            s[2 * i + 1] += g1; qq[2 * i + 1] = fmaf(g1, yv.y - mu[2 * i + 1], qq[2 * i + 1]);
                        }
                    }
                    if (hf == 0) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { stS[0][i] += s[i]; stQ[0][i] += qq[i]; }
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { stS[1][i] += s[i]; stQ[1][i] += qq[i]; }
                    }
                } else if (p.stats != nullptr && !p.bn_reduce) {
                    // per-channel sum / sumsq of the bf16-rounded outputs (what BatchNorm will actually read): every thread
                    // owns one 16-byte chunk (8 channels) over a group of rows, read back from the staging tile with v4 loads
                    const int nch = hcols >> 3;                            // 8, 16 or 32 chunks per row
                    const int my_ch = tid_e & (nch - 1);
                    const int rows_per_rg = nch >> 1;                      // 4, 8 or 16 rows per thread
                    const int r0 = (tid_e >> (31 - __clz(nch))) * rows_per_rg;
                    const uint32_t stat_base = sOut_u32 + (my_ch >> 3) * 16384;
                    const uint32_t stat_chunk = static_cast<uint32_t>(my_ch & 7);
                    float s[8], qq[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] = 0.f; qq[i] = 0.f; }
#pragma unroll 2
                    for (int r = r0; r < r0 + rows_per_rg; ++r) {
                        uint32_t w4[4];
                        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                     : "=r"(w4[0]), "=r"(w4[1]), "=r"(w4[2]), "=r"(w4[3])
                                     : "r"(stat_base + r * 128 + ((stat_chunk ^ (r & 7)) << 4)));
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float2 f = unpack_bf16x2(w4[i]);
                            s[2 * i] += f.x; qq[2 * i] = fmaf(f.x, f.x, qq[2 * i]);
                            s[2 * i + 1] += f.y; qq[2 * i + 1] = fmaf(f.y, f.y, qq[2 * i + 1]);
                        }
                    }
                    if (hf == 0) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { stS[0][i] += s[i]; stQ[0][i] += qq[i]; }
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { stS[1][i] += s[i]; stQ[1][i] += qq[i]; }
                    }
                }
            }
            acc ^= 1;
            if (acc == 0) accphase ^= 1;
        }
        if (tid_e == 0) tma_store_wait_all();
        if (p.stats != nullptr) {
            // fold the row groups through the (now idle) staging buffer, then one atomic per channel and statistic
            float* st = p.stats + static_cast<size_t>((blockIdx.x / numN) % STAT_REPLICAS) * 2 * p.Cout;
            float* red = reinterpret_cast<float*>(sOut);       // [row group][2][hcols] floats: 16 KB for every pass width
            for (int hf = 0; hf < n_halves; ++hf) {
                const int hcols = (BN - hf * hw) < hw ? (BN - hf * hw) : hw;
                const int nch = hcols >> 3;
                const int my_ch = tid_e & (nch - 1);
                const int my_rg = tid_e / nch;
                const int nrg = 256 / nch;
                named_barrier_sync(1, 256);                    // staging buffer free (TMA drained / previous pass consumed)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    red[(my_rg * 2 + 0) * hcols + my_ch * 8 + i] = hf == 0 ? stS[0][i] : stS[1][i];
                    red[(my_rg * 2 + 1) * hcols + my_ch * 8 + i] = hf == 0 ? stQ[0][i] : stQ[1][i];
                }
                named_barrier_sync(1, 256);
                if (tid_e < hcols) {
                    float a = 0.f, b = 0.f;
                    for (int g = 0; g < nrg; ++g) {
                        a += red[(g * 2 + 0) * hcols + tid_e];
                        b += red[(g * 2 + 1) * hcols + tid_e];
                    }
                    const int c = n_tile * BN + hf * hw + tid_e;
                    if (c < p.Cout) {
                        if (p.bn_reduce) b *= p.r_invstd[c];          // sum(g * (y - mean)) -> sum(g * xhat)
                        atomicAdd(st + c, a);
                        atomicAdd(st + p.Cout + c, b);
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (kCluster) cluster_sync();          // nobody leaves while the pair's MMAs / barrier arrivals can still touch this CTA
    if (warp == 1) {
        if (kCluster) tmem_dealloc_pair(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

// =====================================================================================================
//  wgrad kernel: dW[cout, tap, cin] += sum_pixels dy[pixel, cout] * x[pixel + tap, cin]
//  One CTA = one (k-split, tap, cout-tile, cin-tile); both operands MN-major as landed by TMA.
// =====================================================================================================
__global__ void __launch_bounds__(192, 1) conv_wgrad_kernel(const __grid_constant__ WgradParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem(smem_raw);
    SmemBars* bars = reinterpret_cast<SmemBars*>(smem);
    const int BN = p.BN;
    const int stages = p.stages;
    const int b_stage = BN * 128;
    uint8_t* sA = smem + SMEM_HEADER;
    uint8_t* sB = sA + stages * A_STAGE_BYTES;
    uint8_t* sOnes = sB + stages * b_stage;          // 8 KB of bf16 1.0 (dbias only): a [64 x 64] B tile in any layout

    const int warp = uniform_warp_id();
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) {
            mbar_init(&bars->full[i], 1);
            mbar_init(&bars->empty[i], 1);
        }
        mbar_init(&bars->tfull[0], 1);
        fence_mbar_init();
    }
    if (p.dbias != nullptr) {
        for (int i = threadIdx.x; i < 2048; i += blockDim.x) reinterpret_cast<uint32_t*>(sOnes)[i] = 0x3F803F80u;
        fence_proxy_async_smem();
    }
    if (warp == 1) {
        tmem_alloc(&bars->tmem_base, p.tmem_cols);
        tmem_relinquish();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.dymap);
        tma_prefetch_desc(&p.xmap[0]);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = uniform_u32(bars->tmem_base);
    pdl_sync();

    const int inner = p.ngroups * p.numM * p.numN;
    const int split = blockIdx.x / inner;
    int rem = blockIdx.x % inner;
    const int nt = rem % p.numN; rem /= p.numN;
    const int mt = rem % p.numM; rem /= p.numM;
    const int group = rem;                      // taps [group*G, group*G + G) share this CTA's dy tiles
    const int pt_begin = static_cast<int>((static_cast<long long>(p.num_p_tiles) * split) / p.ksplit);
    const int pt_end = static_cast<int>((static_cast<long long>(p.num_p_tiles) * (split + 1)) / p.ksplit);
    // the CTAs of the first tap group share the bias work: input-channel tile nt sums dy over the pixel tiles pt with
    // pt % numN == nt (one CTA doing all of it finished ~10 % after its siblings and set the kernel's tail)
    const bool bias_cta = p.dbias != nullptr && group == 0;
    const int bias_first = pt_begin + ((nt - pt_begin % p.numN) + p.numN) % p.numN;      // first pt >= pt_begin with pt % numN == nt
    const bool do_bias = bias_cta && bias_first < pt_end;
    const int bw = 1 << p.bw_log2, bh = 1 << p.bh_log2;
    const int bn = 64 >> (p.bw_log2 + p.bh_log2);

    if (warp == 0) {
        {   // all 32 lanes walk the loop; the *_warp calls elect one lane to issue
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx_bytes = A_STAGE_BYTES + b_stage;
            int ca1 = mt * TILE_M + 64;
            if (ca1 >= p.Cout) ca1 = mt * TILE_M;       // duplicate rows instead of a fully out-of-range box
            // per 64-column block of B: which tap (shift + parity map) and which channel offset -- fixed for the CTA
            int b_dw[4], b_dh[4], b_c0[4];
            const CUtensorMap* b_map[4];
            const int nblk = BN / 64;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int tj = p.G == 1 ? group : group * p.G + j / p.cin_blocks;
                if (tj >= p.ntaps) tj = p.ntaps - 1;                // padding block of the last group (ignored later)
                const Tap tp = p.taps[tj];
                int c0 = p.G == 1 ? nt * BN + j * 64 : (j % p.cin_blocks) * 64;
                if (c0 >= p.Cin) c0 = 0;
                b_dw[j] = tp.dw; b_dh[j] = tp.dh; b_c0[j] = c0; b_map[j] = &p.xmap[tp.amap];
            }
            for (int pt = pt_begin; pt < pt_end; ++pt) {
                const int tw = pt % p.tiles_w;
                const int th = (pt / p.tiles_w) % p.tiles_h;
                const int tn = pt / (p.tiles_w * p.tiles_h);
                const int w0 = tw * bw, h0 = th * bh, n0 = tn * bn;
                mbar_wait(&bars->empty[stage], phase ^ 1);
                mbar_arrive_expect_tx_warp(&bars->full[stage], tx_bytes);
                uint8_t* dA = sA + stage * A_STAGE_BYTES;
                tma_load_4d_warp(dA, &p.dymap, &bars->full[stage], mt * TILE_M, w0, h0, n0);
                tma_load_4d_warp(dA + 8192, &p.dymap, &bars->full[stage], ca1, w0, h0, n0);
                uint8_t* dB = sB + stage * b_stage;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j < nblk)
                        tma_load_4d_warp(dB + j * 8192, b_map[j], &bars->full[stage], b_c0[j], w0 + b_dw[j], h0 + b_dh[j], n0);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        {   // warp-convergent issue (see picd_common.cuh): descriptors stay in uniform registers
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t idesc = make_idesc_bf16(TILE_M, BN, 1, 1);
            const uint32_t idesc_ones = make_idesc_bf16(TILE_M, 16, 1, 1);
            const uint64_t tmpl = make_smem_desc(0, 8192, 1024);
            const uint32_t sA_lo = smem_u32(sA) >> 4, sB_lo = smem_u32(sB) >> 4;
            const uint64_t onesdesc0 = tmpl | (smem_u32(sOnes) >> 4);
            const uint32_t a_stage_lo = A_STAGE_BYTES >> 4, b_stage_lo = static_cast<uint32_t>(b_stage) >> 4;
            for (int pt = pt_begin; pt < pt_end; ++pt) {
                mbar_wait(&bars->full[stage], phase);
                tc_fence_after();
                const uint64_t adesc0 = tmpl | (sA_lo + stage * a_stage_lo);
                const uint64_t bdesc0 = tmpl | (sB_lo + stage * b_stage_lo);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16_warp(tmem_base, adesc0 + k * (2048 >> 4), bdesc0 + k * (2048 >> 4), idesc,
                                   (pt > pt_begin) || (k != 0));
                if (do_bias && (pt - bias_first) % p.numN == 0 && pt >= bias_first) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)      // dy^T x ones: every one of the 16 result columns is the bias gradient
                        umma_bf16_warp(tmem_base + BN, adesc0 + k * (2048 >> 4), onesdesc0 + k * (2048 >> 4), idesc_ones,
                                       (pt > bias_first) || (k != 0));
                }
                umma_commit_warp(&bars->empty[stage]);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
            umma_commit_warp(&bars->tfull[0]);
        }
    } else {
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int cout = mt * TILE_M + row;
        if (pt_end > pt_begin) {
            mbar_wait(&bars->tfull[0], 0);
            tc_fence_after();
            const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
            if (do_bias) {
                uint32_t v[32];
                tmem_ld_32x32(t_row + BN, v);
                tmem_ld_wait();
                if (cout < p.Cout) atomicAdd(p.dbias + cout, __uint_as_float(v[0]));
            }
            for (int g = 0; g < BN / 32; ++g) {
                uint32_t v[32];
                tmem_ld_32x32(t_row + g * 32, v);
                tmem_ld_wait();
                const int j = g >> 1;
                const int tj = p.G == 1 ? group : group * p.G + j / p.cin_blocks;
                if (cout < p.Cout && tj < p.ntaps) {
                    float* drow = p.dw + (static_cast<size_t>(cout) * p.taps_total + p.taps[tj].btap) * p.Cin;
                    const int cin0 = (p.G == 1 ? nt * BN + j * 64 : (j % p.cin_blocks) * 64) + (g & 1) * 32;
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        const int cin = cin0 + i;
                        if (cin + 3 < p.Cin) {
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(drow + cin),
                                         "f"(__uint_as_float(v[i])), "f"(__uint_as_float(v[i + 1])),
                                         "f"(__uint_as_float(v[i + 2])), "f"(__uint_as_float(v[i + 3]))
                                         : "memory");
                        } else {
                            for (int e = 0; e < 4; ++e)
                                if (cin + e < p.Cin) atomicAdd(drow + cin + e, __uint_as_float(v[i + e]));
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

// =====================================================================================================
//  host side
// =====================================================================================================
struct Geometry {
    int bw_log2, bh_log2, bn;
    int tiles_w, tiles_h, tiles_n;
};

// Choose the [bn x bh x bw] pixel box (product = tile_pixels, all powers of two) with the least padding.
static Geometry choose_geometry(int N, int H, int W, int tile_pixels) {
    Geometry best{};
    double best_cost = 1e30;
    int lg = 0;
    while ((1 << lg) < tile_pixels) ++lg;
    for (int lw = 0; lw <= lg; ++lw) {
        for (int lh = 0; lh + lw <= lg; ++lh) {
            const int bw = 1 << lw, bh = 1 << lh, bn = tile_pixels >> (lw + lh);
            if (bw > 256 || bh > 256 || bn > 256) continue;
            const long long padded = 1LL * ceil_div(W, bw) * bw * ceil_div(H, bh) * bh * ceil_div(N, bn) * bn;
            // tie-break towards wide boxes (long contiguous runs per TMA row)
            const double cost = static_cast<double>(padded) * (1.0 + 0.001 * (lg - lw));
            if (cost < best_cost) {
                best_cost = cost;
                best = Geometry{lw, lh, bn, ceil_div(W, bw), ceil_div(H, bh), ceil_div(N, bn)};
            }
        }
    }
    return best;
}

static int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static int pos_mod(int a, int b) { return ((a % b) + b) % b; }

// 4-D NHWC activation map (optionally the (ph,pw) parity view for stride 2).  box = (64, bw, bh, bn).
static int make_act_map(CUtensorMap* m, const void* base, int N, int H, int W, int C, int stride, int ph, int pw,
                        int bw, int bh, int bn) {
    const __nv_bfloat16* b = static_cast<const __nv_bfloat16*>(base) + (static_cast<size_t>(ph) * W + pw) * C;
    const int Hv = (H - ph + stride - 1) / stride, Wv = (W - pw + stride - 1) / stride;
    uint64_t dims[4] = {(uint64_t)C, (uint64_t)std::max(Wv, 1), (uint64_t)std::max(Hv, 1), (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)stride * C * 2, (uint64_t)stride * W * C * 2, (uint64_t)H * W * C * 2};
    uint32_t box[4] = {64, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn};
    return make_tensor_map(m, b, 2, 4, dims, strides, box, true);
}


// ---- halo mode (3x3, stride 1, pad 1) -------------------------------------------------------------------------------
// Pixel tile = bw x bh pixels of ONE image with bw a multiple of 8, so that every 8-row UMMA group is 8 consecutive pixels
// of one image row and a dh shift is a whole number of 1024-byte groups inside the (bh + 2)-row box.
static Geometry choose_halo_geometry(int N, int H, int W) {
    Geometry best{};
    double best_cost = 1e30;
    for (int lw = 3; lw <= 7; ++lw) {
        const int bw = 1 << lw, bh = TILE_M >> lw, lh = 7 - lw;
        const long long tiles = 1LL * ceil_div(W, bw) * ceil_div(H, bh);
        const double cost = static_cast<double>(tiles) * (TILE_M + 0.5 * (bh + 2) * bw);   // MMA rows + operand bytes
        if (cost < best_cost) {
            best_cost = cost;
            best = Geometry{lw, lh, 1, ceil_div(W, bw), ceil_div(H, bh), N};
        }
    }
    return best;
}

// Reorders a 3x3 tap list to [dw][dh] (dh, dw = -1, 0, +1) if it is one; PICD_HALO=0 disables, =1 forces where legal.
static bool setup_halo(ConvGemmParams& p, int nt, int stride, int H, int W, int cout_gemm) {
    static const int mode = [] { const char* e = getenv("PICD_HALO"); return e == nullptr ? -1 : atoi(e); }();
    if (mode == 0 || nt != 9 || stride != 1) return false;
    if (mode != 1 && (std::min(H, W) < 28 || cout_gemm > 128)) return false;
    Tap sorted[9];
    bool seen[9] = {false, false, false, false, false, false, false, false, false};
    for (int t = 0; t < 9; ++t) {
        const Tap tp = p.taps[t];
        if (tp.amap != 0 || tp.dh < -1 || tp.dh > 1 || tp.dw < -1 || tp.dw > 1) return false;
        const int idx = (tp.dw + 1) * 3 + (tp.dh + 1);
        if (seen[idx]) return false;
        seen[idx] = true;
        sorted[idx] = tp;
    }
    for (int t = 0; t < 9; ++t) p.taps[t] = sorted[t];
    p.halo = 1;
    return true;
}

static int num_n_tiles(int C, int* BN) {
    const int numN = ceil_div(C, 256);
    int bn = ceil_div(ceil_div(C, numN), 64) * 64;
    *BN = bn;
    return ceil_div(C, bn);
}

// CTA pairs (cta_group::2) pay when operand traffic (not the output) limits the kernel: enough K work per tile.
// PICD_CLUSTER=0 disables it, =1 forces it wherever it is legal, unset = heuristic.
static bool want_cluster(int k_iters, int num_m_tiles, int BN, bool b_mn) {
    static const int mode = [] { const char* e = getenv("PICD_CLUSTER"); return e == nullptr ? -1 : atoi(e); }();
    if (mode == 0 || num_m_tiles < 2) return false;
    if (b_mn && (BN % 128) != 0) return false;          // MN-major B is split between the pair in 64-column boxes
    if (mode == 1) return true;
    // Pairs pay when the tile is operand-bound, not HBM-bound: per CTA a k-iteration costs 2 * BN tensor cycles but moves
    // 128 * (64 + BN/k_iters) * 2 bytes of activations -- below ~12 k-iterations of a 256-wide tile the memory system is the
    // limit and independent CTAs schedule better (measured: the 56x56 / 28x28 ResNet layers lose 10-20% in pair mode).
    return k_iters >= 12 && BN >= 192;
}

// Clusters of four (two pairs sharing their B tile by TMA multicast, see ConvGemmParams::quad): for plain GEMM-shaped launches
// that already want pairs.  The multicast load reads the weight operand, which must live in THIS GPU's memory.
static bool want_quad(const void* b_operand, int BN, int num_m_tiles, int k_iters) {
    static const int mode = [] { const char* e = getenv("PICD_QUAD"); return e == nullptr ? -1 : atoi(e); }();
    if (mode == 0 || (BN % 32) != 0 || num_m_tiles < 8) return false;
    // Opt-in (PICD_QUAD=1).  Measured on 8192^3 with the blocked tile walk: 1464 TFLOP/s against 1560 for plain pairs -- only 33
    // clusters of four are co-resident on the 148 SMs (132 CTAs), which costs more than the 25 % fewer L2 -> SM bytes buy; on
    // the ViT shapes (K = 768 / 3072, 197 row tiles) four CTAs in lockstep lose 4-6 % to two.
    (void)k_iters;
    if (mode != 1) return false;
    cudaPointerAttributes at{};
    int dev = -1;
    if (cudaPointerGetAttributes(&at, b_operand) != cudaSuccess || cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return false; }
    if (at.type != cudaMemoryTypeDevice || at.device != dev) return false;
    return true;
}

// halo mode keeps three weight taps per stage, so its column tiles are capped at 128 channels
static int num_n_tiles_halo(int C, int* BN) {
    const int numN = ceil_div(C, 128);
    int bn = ceil_div(ceil_div(C, numN), 64) * 64;
    *BN = bn;
    return ceil_div(C, bn);
}

static int launch_conv_gemm(ConvGemmParams& p, int Cout_gemm, cudaStream_t stream) {
    int BN;
    p.numN = p.halo ? num_n_tiles_halo(Cout_gemm, &BN) : num_n_tiles(Cout_gemm, &BN);
    p.BN = BN;
    p.Cout = Cout_gemm;
    if (!p.halo) p.a_stage_bytes = A_STAGE_BYTES;
    // tiny weight sets (64 -> 64 channels) stay resident: the ring then carries activations only
    p.b_resident = (p.halo && p.kchunks == 1 && 9 * BN * 128 <= 80 * 1024) ? 1 : 0;
    const int resident = p.b_resident ? 9 * BN * 128 : 0;
    const int stage_bytes = p.a_stage_bytes + (p.halo ? (p.b_resident ? 0 : 3 * BN * 128) : (p.cluster ? BN / 2 : BN) * 128);
    // short-K tiles are epilogue-bound: drain 256 columns in one pass; otherwise drain halves and spend the smem on stages
    p.half_w = (BN == 256 && p.ntaps * p.kchunks <= 2 && !p.bn_reduce) ? 256 : 128;
    const int staging = std::min(BN, p.half_w) * 256;   // 128 rows x half_w columns x 2 B
    p.header_bytes = p.bn_reduce ? SMEM_HEADER_BNR : SMEM_HEADER;
    // bn_reduce: 1-3 more staging-sized buffers receive the upstream y tiles, requested ybufs - 1 output passes ahead; deep
    // prefetch where shared memory allows it without starving the operand ring
    p.ybufs = 0;
    if (p.bn_reduce) {
        const char* fe = getenv("PICD_YBUFS");             // tuning override, read per call (benchmarks flip it in-process)
        const int forced = fe == nullptr ? 0 : atoi(fe);
        // measured (benchmarks/fused_reduce_layers.py): a second buffer pays only for short-K tiles, whose output pass is
        // shorter than a DRAM round trip (56x56 64 <- 256 1x1: 0.123 -> 0.103 ms); elsewhere the operand ring needs the
        // shared memory more (14x14 256 <- 256 3x3: 0.057 -> 0.059 -> 0.072 ms with 1 / 2 / 3 buffers)
        const int k_iters = p.halo ? 3 * p.kchunks : p.ntaps * p.kchunks;
        p.ybufs = 1;
        if (k_iters <= 4 && (MAX_DYN_SMEM - 1024 - p.header_bytes - staging * 3 - resident) / stage_bytes >= 3) p.ybufs = 2;
        if (forced >= 1 && forced <= 4) p.ybufs = forced;
    }
    int stages = (MAX_DYN_SMEM - 1024 - p.header_bytes - staging * (1 + p.ybufs) - resident) / stage_bytes;
    stages = std::min(stages, (p.halo || p.cluster) ? 8 : 6);
    PICD_REQUIRE(stages >= 2, "conv_gemm: tile does not fit shared memory (BN=%d)", BN);
    p.stages = stages;
    static bool attr_set = false;
    if (!attr_set) {
        PICD_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_DYN_SMEM));
        PICD_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_DYN_SMEM));
        PICD_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_DYN_SMEM));
        PICD_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_DYN_SMEM));
        attr_set = true;
    }
    const int sms = num_sms();
    const bool bias_act = p.bias != nullptr || p.act != ACT_NONE;
    if (p.cluster) {
        int nclusters = std::max(p.numN, (sms / 2) / p.numN * p.numN);
        const long long pairs = 1LL * ((p.num_m_tiles + 1) / 2) * p.numN;
        if (nclusters > pairs) nclusters = static_cast<int>(std::max<long long>(p.numN, pairs / p.numN * p.numN));
        // plain GEMMs (no per-channel statistics tied to a fixed column tile): when the column-tile count does not divide the
        // SM pairs (8192 columns: 32 tiles -> 64 of 74 pairs), deal the pair-tiles round-robin over every pair instead
        static const bool free_ok = [] { const char* e = getenv("PICD_FREE_SCHED"); return e == nullptr || atoi(e) != 0; }();
        p.free_sched = 0;
        int csize = 2;
        if (p.quad) {
            static const int max_quads = [] {
                cudaLaunchConfig_t q{};
                q.gridDim = dim3(4 * (num_sms() / 4)); q.blockDim = dim3(320); q.dynamicSmemBytes = MAX_DYN_SMEM;
                cudaLaunchAttribute a[1];
                a[0].id = cudaLaunchAttributeClusterDimension;
                a[0].val.clusterDim.x = 4; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
                q.attrs = a; q.numAttrs = 1;
                int n = 0;
                if (cudaOccupancyMaxActiveClusters(&n, conv_gemm_kernel<false, true>, &q) != cudaSuccess) { cudaGetLastError(); n = 0; }
                return n;
            }();
            const long long quads = 1LL * ((p.num_m_tiles + 3) / 4) * p.numN;
            const int cap = std::min(sms / 4, max_quads);
            static bool told = false;
            if (!told && getenv("PICD_DEBUG") != nullptr) {
                fprintf(stderr, "[picd] conv_gemm quad mode: %d clusters of 4 co-resident (device), using %d\n", max_quads, cap);
                told = true;
            }
            PICD_REQUIRE(cap >= 1, "conv_gemm: no cluster of four CTAs fits this device (%d)", cap);
            p.free_sched = 1;
            csize = 4;
            nclusters = static_cast<int>(std::min<long long>(cap, quads));
        } else if (free_ok && p.stats == nullptr && !p.bn_reduce && pairs > sms / 2) {
            const long long rounds_fixed = (pairs + nclusters - 1) / nclusters;
            const long long rounds_free = (pairs + sms / 2 - 1) / (sms / 2);
            // (a pair that cannot be co-resident would run as a second wave: never ask for more clusters than the device places)
            static const int max_pairs = [] {
                cudaLaunchConfig_t q{};
                q.gridDim = dim3(2 * (num_sms() / 2)); q.blockDim = dim3(320); q.dynamicSmemBytes = MAX_DYN_SMEM;
                cudaLaunchAttribute a[1];
                a[0].id = cudaLaunchAttributeClusterDimension;
                a[0].val.clusterDim.x = 2; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
                q.attrs = a; q.numAttrs = 1;
                int n = 0;
                if (cudaOccupancyMaxActiveClusters(&n, conv_gemm_kernel<false, true>, &q) != cudaSuccess) { cudaGetLastError(); n = 0; }
                return n;
            }();
            const int cap = std::min(sms / 2, max_pairs);
            if (cap > nclusters && (pairs + cap - 1) / cap < rounds_fixed) { p.free_sched = 1; nclusters = cap; }
            (void)rounds_free;
        }
        if (p.free_sched) {
            // column block = what keeps its B rows (BN x K bf16 each) within ~32 MB of the 126 MB L2 (8192^3: blocks of 8 column
            // tiles 1560 TFLOP/s, 16: 1522, 4: 1535, all 32 i.e. no blocking: 1419; DRAM reads drop from 2.1 GB to ~0.7 GB)
            const long long tile_bytes = 2LL * p.BN * p.ntaps * p.kchunks * KCHUNK;
            p.nbw = static_cast<int>(std::max<long long>(1, std::min<long long>(p.numN, (32LL << 20) / tile_bytes)));
            static const int forced_nbw = [] { const char* e = getenv("PICD_NBW"); return e == nullptr ? 0 : atoi(e); }();
            if (forced_nbw > 0) p.nbw = std::min(forced_nbw, p.numN);
        }
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(csize * nclusters);
        cfg.blockDim = dim3(320);
        cfg.dynamicSmemBytes = MAX_DYN_SMEM;
        cfg.stream = stream;
        cudaLaunchAttribute attr[2];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = csize; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl_enabled() ? 2 : 1;
        if (bias_act) PICD_CUDA(cudaLaunchKernelEx(&cfg, conv_gemm_kernel<true, true>, p));
        else PICD_CUDA(cudaLaunchKernelEx(&cfg, conv_gemm_kernel<false, true>, p));
        count_launch();
        return 0;
    }
    const long long total = 1LL * p.num_m_tiles * p.numN;
    int grid = static_cast<int>(std::min<long long>(total, sms));
    grid = std::max(p.numN, grid / p.numN * p.numN);
    // request the whole smem so that exactly one CTA (and one 512-column TMEM allocation) lives on an SM
    if (bias_act)
        PICD_CUDA(launch_pdl(conv_gemm_kernel<true, false>, dim3(grid), dim3(320), MAX_DYN_SMEM, stream, p));
    else
        PICD_CUDA(launch_pdl(conv_gemm_kernel<false, false>, dim3(grid), dim3(320), MAX_DYN_SMEM, stream, p));
    PICD_LAUNCH_CHECK();
    return 0;
}

}  // namespace picd

using namespace picd;

// y[N,OH,OW,Cout] = conv(x[N,H,W,Cin], w[Cout,R,S,Cin]); optional fused per-channel (sum, sumsq) for BatchNorm,
// optional bias + activation epilogue.  NHWC bf16, stride 1 or 2.
extern "C" int picd_conv_fprop(const void* x, const void* w, void* y, float* stats, const float* bias, int act,
                               int N, int H, int W, int Cin, int Cout, int R, int S, int stride, int pad_h,
                               int pad_w, int OH, int OW, cudaStream_t stream) {
    PICD_REQUIRE(act == ACT_NONE || act == ACT_RELU, "conv_fprop: only none/relu are fused into the epilogue (act=%d)", act);
    PICD_REQUIRE(stride == 1 || stride == 2, "conv_fprop: stride %d unsupported", stride);
    PICD_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0, "conv_fprop: channels must be multiples of 8 (Cin=%d Cout=%d)", Cin, Cout);
    PICD_REQUIRE(R * S <= MAX_TAPS, "conv_fprop: %dx%d filter has too many taps", R, S);
    ConvGemmParams p;
    std::memset(&p, 0, sizeof(p));
    int nt = 0;
    for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s) {
            const int dr = r - pad_h, ds = s - pad_w;
            Tap t;
            t.amap = static_cast<int16_t>(pos_mod(dr, stride) * stride + pos_mod(ds, stride));
            t.dh = static_cast<int16_t>(floor_div(dr, stride));
            t.dw = static_cast<int16_t>(floor_div(ds, stride));
            t.btap = static_cast<int16_t>(r * S + s);
            p.taps[nt++] = t;
        }
    const bool halo = setup_halo(p, nt, stride, OH, OW, Cout);
    const Geometry g = halo ? choose_halo_geometry(N, OH, OW) : choose_geometry(N, OH, OW, TILE_M);
    const int bw = 1 << g.bw_log2, bh = 1 << g.bh_log2;
    if (halo) {
        p.a_stage_bytes = (bh + 2) * bw * 128;
        if (make_act_map(&p.amap[0], x, N, H, W, Cin, 1, 0, 0, bw, bh + 2, 1)) return 3;
    } else {
        for (int ph = 0; ph < stride; ++ph)
            for (int pw = 0; pw < stride; ++pw)
                if (make_act_map(&p.amap[ph * stride + pw], x, N, H, W, Cin, stride, ph, pw, bw, bh, g.bn)) return 3;
    }
    {
        uint64_t dims[3] = {(uint64_t)Cin, (uint64_t)(R * S), (uint64_t)Cout};
        uint64_t strides[2] = {(uint64_t)Cin * 2, (uint64_t)R * S * Cin * 2};
        int BN;
        if (halo) num_n_tiles_halo(Cout, &BN); else num_n_tiles(Cout, &BN);
        p.cluster = (!halo && want_cluster(R * S * ceil_div(Cin, KCHUNK), g.tiles_w * g.tiles_h * g.tiles_n, BN, false)) ? 1 : 0;
        p.quad = (p.cluster && stats == nullptr && want_quad(w, BN, g.tiles_w * g.tiles_h * g.tiles_n, R * S * ceil_div(Cin, KCHUNK))) ? 1 : 0;
        uint32_t box[3] = {64, 1, (uint32_t)(p.quad ? BN / 4 : p.cluster ? BN / 2 : BN)};   // pair: each CTA fetches half the rows,
                                                                                             // quad: a quarter (and multicasts it)
        if (make_tensor_map(&p.bmap, w, 2, 3, dims, strides, box, true)) return 3;
    }
    if (make_act_map(&p.dmap, y, N, OH, OW, Cout, 1, 0, 0, bw, bh, g.bn)) return 3;
    p.ntaps = nt;
    p.kchunks = ceil_div(Cin, KCHUNK);
    p.b_mn = 0;
    p.num_m_tiles = g.tiles_w * g.tiles_h * g.tiles_n;
    p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h;
    p.bw_log2 = g.bw_log2; p.bh_log2 = g.bh_log2;
    p.W = OW; p.H = OH; p.N = N;
    p.stats = stats; p.bias = bias; p.act = act;
    return launch_conv_gemm(p, Cout, stream);
}

// Fused BatchNorm-backward reduction of a dgrad launch (see ConvGemmParams::bn_reduce): dx is dz of the upstream BN layer.
struct BnReduceArgs {
    const void* y;            // the upstream layer's raw conv output, [N,H,W,Cin] bf16 (same shape as dx)
    const float* scale; const float* shift; const float* mean; const float* invstd;     // [Cin]
    float* sums;              // [STAT_REPLICAS][2][Cin], zeroed by the caller
    int relu;                 // the upstream activation is ReLU (else identity)
};

// dx[N,H,W,Cin] = conv_transpose(dy[N,OH,OW,Cout], w[Cout,R,S,Cin]).  Weights are read through an MN-major
// descriptor, i.e. the same tensor as fprop.  Stride 2 = one launch per output parity class.
static int conv_dgrad_impl(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                           int R, int S, int stride, int pad_h, int pad_w, int OH, int OW, int accumulate,
                           cudaStream_t stream, const BnReduceArgs* bnr = nullptr) {
    PICD_REQUIRE(bnr == nullptr || !accumulate, "conv_dgrad: the fused BN reduction needs the FINAL gradient, not a partial sum");
    PICD_REQUIRE(stride == 1 || stride == 2, "conv_dgrad: stride %d unsupported", stride);
    PICD_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0, "conv_dgrad: channels must be multiples of 8");
    // If some output parity class receives no tap (e.g. 1x1 stride-2), zero dx once BEFORE any class is written.
    if (stride > 1) {
        bool any_empty = false;
        for (int ph = 0; ph < stride && !any_empty; ++ph)
            for (int pw = 0; pw < stride && !any_empty; ++pw) {
                int cnt = 0;
                for (int r = 0; r < R; ++r)
                    for (int s = 0; s < S; ++s)
                        if (pos_mod(ph + pad_h - r, stride) == 0 && pos_mod(pw + pad_w - s, stride) == 0) ++cnt;
                if (cnt == 0) any_empty = true;
            }
        if (any_empty && !accumulate) PICD_CUDA(cudaMemsetAsync(dx, 0, static_cast<size_t>(N) * H * W * Cin * 2, stream));
    }
    for (int ph = 0; ph < stride; ++ph) {
        for (int pw = 0; pw < stride; ++pw) {
            const int Hv = (H - ph + stride - 1) / stride, Wv = (W - pw + stride - 1) / stride;
            if (Hv <= 0 || Wv <= 0) continue;
            ConvGemmParams p;
            std::memset(&p, 0, sizeof(p));
            int nt = 0;
            for (int r = 0; r < R; ++r)
                for (int s = 0; s < S; ++s) {
                    const int eh = ph + pad_h - r, ew = pw + pad_w - s;
                    if (pos_mod(eh, stride) != 0 || pos_mod(ew, stride) != 0) continue;
                    Tap t;
                    t.amap = 0;
                    t.dh = static_cast<int16_t>(eh / stride);
                    t.dw = static_cast<int16_t>(ew / stride);
                    t.btap = static_cast<int16_t>(r * S + s);
                    p.taps[nt++] = t;
                }
            if (nt == 0) continue;   // zeroed above
            const bool halo = setup_halo(p, nt, stride, Hv, Wv, Cin);
            const Geometry g = halo ? choose_halo_geometry(N, Hv, Wv) : choose_geometry(N, Hv, Wv, TILE_M);
            const int bw = 1 << g.bw_log2, bh = 1 << g.bh_log2;
            if (halo) p.a_stage_bytes = (bh + 2) * bw * 128;
            if (make_act_map(&p.amap[0], dy, N, OH, OW, Cout, 1, 0, 0, bw, halo ? bh + 2 : bh, g.bn)) return 3;
            {
                uint64_t dims[3] = {(uint64_t)Cin, (uint64_t)(R * S), (uint64_t)Cout};
                uint64_t strides[2] = {(uint64_t)Cin * 2, (uint64_t)R * S * Cin * 2};
                uint32_t box[3] = {64, 1, 64};
                if (make_tensor_map(&p.bmap, w, 2, 3, dims, strides, box, true)) return 3;
            }
            if (make_act_map(&p.dmap, dx, N, H, W, Cin, stride, ph, pw, bw, bh, g.bn)) return 3;
            if (bnr != nullptr) {
                if (make_act_map(&p.ymap, bnr->y, N, H, W, Cin, stride, ph, pw, bw, bh, g.bn)) return 3;
                p.bn_reduce = 1;
                p.r_relu = bnr->relu;
                p.r_scale = bnr->scale; p.r_shift = bnr->shift; p.r_mean = bnr->mean; p.r_invstd = bnr->invstd;
            }
            p.ntaps = nt;
            p.kchunks = ceil_div(Cout, KCHUNK);
            p.b_mn = 1;
            p.num_m_tiles = g.tiles_w * g.tiles_h * g.tiles_n;
            p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h;
            p.bw_log2 = g.bw_log2; p.bh_log2 = g.bh_log2;
            p.W = Wv; p.H = Hv; p.N = N;
            p.stats = bnr != nullptr ? bnr->sums : nullptr; p.bias = nullptr; p.act = ACT_NONE;
            p.reduce_add = accumulate;
            {
                int BNd;
                num_n_tiles(Cin, &BNd);
                p.cluster = (!halo && want_cluster(nt * p.kchunks, p.num_m_tiles, BNd, true)) ? 1 : 0;
            }
            if (int rc = launch_conv_gemm(p, Cin, stream)) return rc;
        }
    }
    return 0;
}

extern "C" int picd_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                               int R, int S, int stride, int pad_h, int pad_w, int OH, int OW,
                               cudaStream_t stream) {
    return conv_dgrad_impl(dy, w, dx, N, H, W, Cin, Cout, R, S, stride, pad_h, pad_w, OH, OW, 0, stream);
}
// dx += conv_transpose(dy, w): the input gradient is accumulated into an existing bf16 tensor with TMA reduce-add
// (fuses the residual-branch gradient sum of a ResNet block into the dgrad store).
extern "C" int picd_conv_dgrad_accumulate(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                                          int R, int S, int stride, int pad_h, int pad_w, int OH, int OW,
                                          cudaStream_t stream) {
    return conv_dgrad_impl(dy, w, dx, N, H, W, Cin, Cout, R, S, stride, pad_h, pad_w, OH, OW, 1, stream);
}

// dx = conv_transpose(dy, w) where dx is the gradient dz flowing into an upstream conv + BatchNorm(+ReLU) layer: the epilogue
// also accumulates that layer's BatchNorm-backward sums sum(g), sum(g * xhat) (g = dz masked by the ReLU, recomputed from
// y_up * scale + shift) into sums[STAT_REPLICAS][2][Cin] -- the separate bn_bwd_reduce pass over dz and y_up disappears
// (torch/nn/modules/_functions.py:144-153 batch_norm_backward_reduce).  Stride-2 dgrads with an empty parity class leave zeros
// there, which contribute nothing.
extern "C" int picd_conv_dgrad_bnreduce(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                                        int R, int S, int stride, int pad_h, int pad_w, int OH, int OW,
                                        const void* y_up, const float* scale, const float* shift, const float* mean,
                                        const float* invstd, int relu, float* sums, cudaStream_t stream) {
    PICD_REQUIRE(y_up != nullptr && sums != nullptr, "conv_dgrad_bnreduce: y_up / sums must not be null");
    BnReduceArgs b{y_up, scale, shift, mean, invstd, sums, relu};
    return conv_dgrad_impl(dy, w, dx, N, H, W, Cin, Cout, R, S, stride, pad_h, pad_w, OH, OW, 0, stream, &b);
}
// dw[Cout,R,S,Cin] (fp32) += sum over pixels of dy (x) x.  The caller provides a zeroed (or accumulating) buffer.
static int conv_wgrad_impl(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W, int Cin, int Cout,
                           int R, int S, int stride, int pad_h, int pad_w, int OH, int OW, cudaStream_t stream) {
    PICD_REQUIRE(stride == 1 || stride == 2, "conv_wgrad: stride %d unsupported", stride);
    PICD_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0, "conv_wgrad: channels must be multiples of 8");
    PICD_REQUIRE(R * S <= MAX_TAPS, "conv_wgrad: too many taps");
    WgradParams p;
    std::memset(&p, 0, sizeof(p));
    const Geometry g = choose_geometry(N, OH, OW, 64);
    const int bw = 1 << g.bw_log2, bh = 1 << g.bh_log2;
    if (make_act_map(&p.dymap, dy, N, OH, OW, Cout, 1, 0, 0, bw, bh, g.bn)) return 3;
    for (int ph = 0; ph < stride; ++ph)
        for (int pw = 0; pw < stride; ++pw)
            if (make_act_map(&p.xmap[ph * stride + pw], x, N, H, W, Cin, stride, ph, pw, bw, bh, g.bn)) return 3;
    int nt = 0;
    for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s) {
            const int dr = r - pad_h, ds = s - pad_w;
            Tap t;
            t.amap = static_cast<int16_t>(pos_mod(dr, stride) * stride + pos_mod(ds, stride));
            t.dh = static_cast<int16_t>(floor_div(dr, stride));
            t.dw = static_cast<int16_t>(floor_div(ds, stride));
            t.btap = static_cast<int16_t>(r * S + s);
            p.taps[nt++] = t;
        }
    p.ntaps = nt;
    p.dw = dw; p.Cout = Cout; p.Cin = Cin; p.taps_total = R * S;
    p.dbias = dbias;
    int BN;
    p.numN = num_n_tiles(Cin, &BN);
    p.cin_blocks = ceil_div(Cin, 64);
    p.G = 1;
    p.ngroups = p.ntaps;
    static const bool group_taps = []{ const char* e = getenv("PICD_WGRAD_GROUP"); return e == nullptr || e[0] != '0'; }();
    if (group_taps && p.ntaps > 1 && p.cin_blocks <= 2) {
        // several taps per CTA: one dy tile (A) feeds G shifted x tiles (B) laid side by side along N
        const int maxG = std::min(p.ntaps, 4 / p.cin_blocks);
        int bestG = 1, best_cost = 1 << 30;
        for (int G = 1; G <= maxG; ++G) {
            const int cost = ceil_div(p.ntaps, G) * (2 + G * p.cin_blocks);     // TMA boxes per pixel tile
            if (cost <= best_cost) { best_cost = cost; bestG = G; }
        }
        p.G = bestG;
        if (p.G > 1) {
            p.ngroups = ceil_div(p.ntaps, p.G);
            p.numN = 1;
            BN = p.G * p.cin_blocks * 64;
        }
    }
    p.BN = BN;
    p.numM = ceil_div(Cout, TILE_M);
    p.num_p_tiles = g.tiles_w * g.tiles_h * g.tiles_n;
    p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h;
    p.bw_log2 = g.bw_log2; p.bh_log2 = g.bh_log2;
    const int stage_bytes = A_STAGE_BYTES + BN * 128;
    const int ones_bytes = dbias != nullptr ? 8192 : 0;
    p.tmem_cols = (dbias != nullptr && BN + 32 > 256) ? 512 : 256;      // the epilogue reads 32 columns from column BN on
    // small tiles: shallow ring, two CTAs per SM (2 x 256 TMEM columns) so one CTA's red.add epilogue overlaps
    // the other's main loop; wide tiles: one CTA per SM with a deep ring.
    int smem_launch;
    if (BN <= 128) {
        p.stages = (BN <= 64) ? 4 : 3;
        smem_launch = p.stages * stage_bytes + SMEM_HEADER + 1024 + ones_bytes;
    } else {
        p.stages = std::min(6, (MAX_DYN_SMEM - 1024 - SMEM_HEADER - ones_bytes) / stage_bytes);
        smem_launch = MAX_DYN_SMEM;
    }
    const int inner = p.ngroups * p.numM * p.numN;
    const int sms = num_sms();
    // split-K factor: minimise (waves) x (k-iterations per CTA + epilogue), i.e. avoid a nearly empty last wave
    const int capacity = sms * (BN <= 128 ? 2 : 1);
    const int max_split = std::max(1, std::min(p.num_p_tiles / 4, (4 * capacity) / inner));
    int ksplit = 1;
    long long best_cost = -1;
    for (int ks = 1; ks <= max_split; ++ks) {
        const long long waves = (1LL * inner * ks + capacity - 1) / capacity;
        const long long cost = waves * ((p.num_p_tiles + ks - 1) / ks + 8);      // 8 ~ epilogue + prologue in k-iterations
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; ksplit = ks; }
    }
    p.ksplit = ksplit;
    static bool attr_set = false;
    if (!attr_set) {
        PICD_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_DYN_SMEM));
        attr_set = true;
    }
    PICD_CUDA(launch_pdl(conv_wgrad_kernel, dim3(inner * ksplit), dim3(192), smem_launch, stream, p));
    PICD_LAUNCH_CHECK();
    return 0;
}

extern "C" int picd_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout,
                               int R, int S, int stride, int pad_h, int pad_w, int OH, int OW,
                               cudaStream_t stream) {
    return conv_wgrad_impl(x, dy, dw, nullptr, N, H, W, Cin, Cout, R, S, stride, pad_h, pad_w, OH, OW, stream);
}
// Same, plus the bias gradient db[Cout] (fp32, accumulated) = column sums of dy, computed by the same kernel with a ones-tile MMA.
extern "C" int picd_conv_wgrad_bias(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W, int Cin,
                                    int Cout, int R, int S, int stride, int pad_h, int pad_w, int OH, int OW,
                                    cudaStream_t stream) {
    return conv_wgrad_impl(x, dy, dw, dbias, N, H, W, Cin, Cout, R, S, stride, pad_h, pad_w, OH, OW, stream);
}

// D[M,N] = act(A[M,K] * B[N,K]^T + bias): the Linear forward (and, with swapped roles, every GEMM whose
// operands are K-major).  All bf16 row-major, fp32 accumulate.
extern "C" int picd_gemm_nt(const void* a, const void* b, void* d, const float* bias, int act, int M, int N_, int K,
                            cudaStream_t stream) {
    return picd_conv_fprop(a, b, d, nullptr, bias, act, 1, 1, M, K, N_, 1, 1, 1, 0, 0, 1, M, stream);
}
// D[M,K] = A[M,N] * B[N,K]  (Linear dgrad: dX = dY * W with W stored [N,K]).
extern "C" int picd_gemm_nn(const void* a, const void* b, void* d, int M, int N_, int K, cudaStream_t stream) {
    return conv_dgrad_impl(a, b, d, 1, 1, M, K, N_, 1, 1, 1, 0, 0, 1, M, 0, stream);
}
// D[N,K] (fp32) += A[M,N]^T * B[M,K]  (Linear wgrad: dW = dY^T X).
extern "C" int picd_gemm_tn_f32(const void* a_dy, const void* b_x, float* d, int M, int N_, int K,
                                cudaStream_t stream) {
    return picd_conv_wgrad(b_x, a_dy, d, 1, 1, M, K, N_, 1, 1, 1, 0, 0, 1, M, stream);
}
// Linear wgrad + bias gradient in one kernel: D[N,K] += A^T B and dbias[N] += column sums of A (= dY).
extern "C" int picd_gemm_tn_f32_bias(const void* a_dy, const void* b_x, float* d, float* dbias, int M, int N_, int K,
                                     cudaStream_t stream) {
    return conv_wgrad_impl(b_x, a_dy, d, dbias, 1, 1, M, K, N_, 1, 1, 1, 0, 0, 1, M, stream);
}
