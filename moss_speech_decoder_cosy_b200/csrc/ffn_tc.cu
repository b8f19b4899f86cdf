// Fused feed-forward of BasicTransformerBlock (matcha/models/components/transformer.py:299-314 with diffusers'
// FeedForward = Linear(256 -> 1024) + exact-erf GELU, Linear(1024 -> 256)) as ONE persistent tcgen05 kernel:
//
//   hres[r, :] += W2 . gelu(W1 . xn[r, :] + b1) + b2          (fp32 residual stream, in place)
//   out[r, :]   = LayerNorm_next(hres[r, :])  or  bf16(hres[r, :])          (operand of the next GEMM)
//
// The 1024-wide hidden activation never leaves the SM.  Per 128-row tile:
//   X tile (128 x 256 bf16, 64 KB) is loaded once by TMA and stays in shared memory;
//   for each of 8 hidden chunks of 128:  G1: acc1[b] = X . W1[chunk]^T        (tcgen05.mma M128 N128, K = 256)
//                                        E1: acc1[b] -> +b1 -> GELU -> bf16 -> P[b], written back with tcgen05.st over
//                                            the accumulator columns it was drained from
//                                        G2: acc2 += P[b] . W2[:, chunk]^T    (tcgen05.mma M128 N256, K = 128, A from TMEM)
//   acc2 is preloaded with hres + b2 through tcgen05.st at the start of the tile, so the MMAs accumulate the residual
//   stream themselves;  final: acc2 -> registers -> hres (coalesced stores in the tiled fp32 layout), LayerNorm ->
//   bf16 -> swizzled smem -> TMA store.
// acc1 (and with it P) is double buffered in TMEM (2 x 128 columns) next to acc2 (256 columns), so the GELU epilogue of
// chunk h overlaps G2(h-1) / G1(h+1) on the tensor pipe.  The kernel is bound by shared-memory bandwidth (operand reads +
// TMA fills, 128 B/clk/SM): keeping P out of shared memory takes 64 KB per chunk off that path.  W1 / W2 stream through a 4-stage ring
// of 32 KB TMA stages (1 MB of weights per tile, L2 resident).
//
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM alloc), warps 2..17 = epilogue (thread == row; the 4
// warps that share a TMEM lane quarter split the columns).
//
// kFused: the attention out-projection of the same transformer block (matcha transformer.py:274-297: to_out + residual,
// then norm3) runs as a PROLOGUE GEMM of the tile instead of as its own launch:
//   acc2 := hres + bo (tcgen05.st preload)  ->  acc2 += att . Wo^T   (K = 512: eight att k-blocks through the four 16 KB
//   slots of the X region, Wo k-blocks through the weight ring)  ->  epilogue warps read acc2 (the new residual stream h1,
//   which stays in tensor memory), normalise it and write X = LayerNorm(h1) as bf16 straight into the shared-memory X tile
//   in the 128B-swizzled K-major layout the MMA expects  ->  feed-forward as above, b2 added in the final drain.
// This removes the out-projection launch with its two HBM round trips per block: the bf16 X tile (written + read back)
// and the fp32 stream (read + rewritten) -- about 250 MB per block at the benchmarked size.
#include "gemm.h"
#include "ptx.cuh"
#include "epilogue.cuh"

#include <mutex>
#include <stdio.h>

namespace mossdec {

bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows);
bool encode_2d_ex(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_cols,
                  uint32_t box_rows, int swizzle64);

namespace {

#ifdef FFN_EXPERIMENT_NO_GELU   // timing experiment only (tests/): how much of E1 is the activation math
#define gelu2_bf16(x, y) pack_bf16x2(x, y)
#endif

constexpr int kC = 256;          // model width
constexpr int kHid = 1024;       // hidden width
constexpr int kChunk = 128;      // hidden columns per chunk
constexpr int kNChunk = kHid / kChunk;
constexpr int kCluster = 2;      // CTAs per cluster: every weight stage is fetched from L2 once and multicast to both
constexpr uint16_t kClusterMask = (1u << kCluster) - 1;
constexpr int kFStages = 4;   // L2 -> smem turnaround of a stage (both CTAs free it, refill, multicast) is ~1 700 clk against 512 clk of MMA
constexpr int kFStageBytes = 32768;
constexpr int kXBytes = kGemmBM * kC * 2;      // 64 KB
constexpr int kOutStageBytes = 2048;           // per epilogue warp: 32 rows x 32 bf16 (64B-swizzled), the box of the output TMA store
constexpr int kFfnSmem = kXBytes + kEpiWarps * kOutStageBytes + kFStages * kFStageBytes + 1024 /*align*/ + 1024 /*b2*/ + 256 /*barriers*/;
constexpr uint32_t kAcc2Col = 0, kAcc1Col = 256;

struct FfnParams {
    int32_t M, num_tiles;
    const float* b1;
    const float* b2;
    const float* bo;    // fused mode: bias of the attention out-projection
    int32_t ln;         // 1: out = (hres - mean) * rstd (affine folded into the consumer's weights), 0: out = bf16(hres)
    const int32_t* row_seq;
    float* hres;        // [M, 256] fp32 in the lane-major tiled layout (gemm.h tiled_f32_offset), read and written
    int32_t store_f32;
    float ln_eps;
    long long* trace;   // developer aid: clock64 stamps of CTA 0 (epilogue warp 0: slots 0..31, MMA thread: 32..47 per tile)
};

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int kOK = 512;                 // K of the fused out-projection (heads * head_dim)
constexpr int kOKB = kOK / kGemmBK;      // 8 k-blocks
constexpr int kASlots = kXBytes / kABytes;   // 4: att k-blocks cycle through the X region

// tmX: legacy = the normalised input X [M, 256]; fused = the attention output [M, 512].  tmWo: fused only.
template <bool kFused>
__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kThreads, 1)
ffn_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmWo,
              const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
              const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ FfnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays in the shared address space (STS / LDS)
    uint8_t* sX = smem;
    uint8_t* sStg = sX + kXBytes;                      // [16 warps][2 KB] bf16 output staging
    uint8_t* sRing = sStg + kEpiWarps * kOutStageBytes;
    float* sB2 = reinterpret_cast<float*>(sRing + kFStages * kFStageBytes);   // [256] second bias
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB2 + kC);
    uint64_t* full = bars;               // [kFStages]  TMA -> MMA
    uint64_t* empty = full + kFStages;   // [kFStages]  MMA -> TMA
    uint64_t* x_full = empty + kFStages; // TMA -> MMA
    uint64_t* x_free = x_full + 1;       // MMA -> TMA (all G1 of the tile done)
    uint64_t* a1_full = x_free + 1;      // [2] MMA -> epilogue
    uint64_t* p_ready = a1_full + 2;     // [2] epilogue -> MMA: P[b] written over the drained acc1[b]
    uint64_t* p_free = p_ready + 2;      // [2] (unused since P lives in TMEM: a1_full of G1(h + 2) implies G2(h) is done)
    uint64_t* acc2_full = p_free + 2;    // MMA -> epilogue
    uint64_t* a_full = acc2_full + 1;    // [kASlots] fused: TMA -> MMA, att k-block landed in slot
    uint64_t* a_empty = a_full + kASlots;   // [kASlots] fused: MMA -> TMA, slot may take k-block kb + 4
    uint64_t* acc2_init = a_empty + kASlots;   // fused: epilogue -> MMA, acc2 preloaded with hres + bo
    uint64_t* oproj_full = acc2_init + 1;   // fused: MMA -> epilogue, out-projection accumulated
    uint64_t* x_ready = oproj_full + 1;     // fused: epilogue -> MMA, X tile written to shared memory
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x_ready + 1);

    // (shuffle broadcasts mark values as warp-uniform for the compiler: descriptors and loop state of the TMA / MMA
    //  warps then live in uniform registers and every tcgen05.mma / TMA is a single instruction, not a per-lane loop)
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmX);
        if (kFused) tma_prefetch_desc(&tmWo);
        tma_prefetch_desc(&tmW1);
        tma_prefetch_desc(&tmW2);
        tma_prefetch_desc(&tmOut);
        for (int i = 0; i < kFStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], kCluster);   // one tcgen05.commit arrival from the MMA thread of every CTA of the cluster
        }
        mbar_init(x_full, 1);
        mbar_init(x_free, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&a1_full[i], 1);
            mbar_init(&p_ready[i], kEpiWarps);
            mbar_init(&p_free[i], 1);
        }
        mbar_init(acc2_full, 1);
        for (int i = 0; i < kASlots; ++i) {
            mbar_init(&a_full[i], 1);
            mbar_init(&a_empty[i], 1);
        }
        mbar_init(acc2_init, kEpiWarps);
        mbar_init(oproj_full, 1);
        mbar_init(x_ready, kEpiWarps);
        mbar_fence_init();
    }
    if (threadIdx.x >= 64 && threadIdx.x < 64 + kC) sB2[threadIdx.x - 64] = __ldg(p.b2 + threadIdx.x - 64);
    if (warp == 1) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    cluster_sync_all();   // peers' barriers are initialised before any multicast / remote arrive can reach them
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const int rank = __shfl_sync(0xffffffffu, (int)cluster_ctarank(), 0);
    pdl_wait();     // everything above overlapped the previous kernel's tail; its results are visible from here on
    pdl_launch();

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (whole warp, one elected lane issues)
        {
            int stage = 0;
            uint32_t phase = 0;
            auto next_stage = [&]() {
                if (++stage == kFStages) {
                    stage = 0;
                    phase ^= 1;
                }
            };
            auto load_g1 = [&](int h) {   // W1 rows [128 h, +128), all K = 256: two stages of two 64-wide k-blocks
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    if (elect_one()) {
                        mbar_expect_tx(&full[stage], kFStageBytes);
                        // this CTA fetches k-block 2j + rank and multicasts it; the peer supplies the other one
                        uint8_t* dst = sRing + stage * kFStageBytes + rank * (kFStageBytes / 2);
                        tma_load_2d_mc(dst, &tmW1, (2 * j + rank) * kGemmBK, h * kChunk, &full[stage], kClusterMask);
                    }
                    next_stage();
                }
            };
            auto load_g2 = [&](int h) {   // W2 all 256 rows, K columns [128 h, +128): two stages of one k-block
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    if (elect_one()) {
                        mbar_expect_tx(&full[stage], kFStageBytes);
                        // rows [128 rank, +128) of the 256-row k-block (8-row swizzle atoms are contiguous: offset 16 KB)
                        tma_load_2d_mc(sRing + stage * kFStageBytes + rank * (kFStageBytes / 2), &tmW2,
                                       h * kChunk + j * kGemmBK, rank * (kC / 2), &full[stage], kClusterMask);
                    }
                    next_stage();
                }
            };
            int it = 0;
            if ((int)blockIdx.x < p.num_tiles && lane == 0) l2_prefetch_bulk(p.hres + (size_t)blockIdx.x * kGemmBM * kC, kGemmBM * kC * 4);
            for (int tile = blockIdx.x; (tile & ~(kCluster - 1)) < p.num_tiles; tile += gridDim.x, ++it) {  // cluster-uniform trip count
                if (it > 0) mbar_wait(x_free, (uint32_t)((it - 1) & 1));
                if constexpr (kFused) {
                    // out-projection operands in consumption order: att k-block kb -> slot kb % 4 of the X region (the
                    // second four wait for the MMAs of the first four), Wo k-block kb -> the weight ring (multicast halves)
                    for (int kb = 0; kb < kOKB; ++kb) {
                        const int slot = kb & (kASlots - 1);
                        if (kb >= kASlots) mbar_wait(&a_empty[slot], (uint32_t)(it & 1));
                        if (elect_one()) {
                            mbar_expect_tx(&a_full[slot], kABytes);
                            tma_load_2d(sX + slot * kABytes, &tmX, kb * kGemmBK, tile * kGemmBM, &a_full[slot]);
                        }
                        mbar_wait(&empty[stage], phase ^ 1);
                        if (elect_one()) {
                            mbar_expect_tx(&full[stage], kFStageBytes);
                            tma_load_2d_mc(sRing + stage * kFStageBytes + rank * (kFStageBytes / 2), &tmWo, kb * kGemmBK,
                                           rank * (kC / 2), &full[stage], kClusterMask);
                        }
                        next_stage();
                    }
                } else if (elect_one()) {
                    mbar_expect_tx(x_full, kXBytes);
                    for (int kb = 0; kb < kC / kGemmBK; ++kb)
                        tma_load_2d(sX + kb * kABytes, &tmX, kb * kGemmBK, tile * kGemmBM, x_full);
                }
                // the residual rows of the next tile: pulled into L2 a whole tile ahead of their use
                if (tile + (int)gridDim.x < p.num_tiles && lane == 0)
                    l2_prefetch_bulk(p.hres + (size_t)(tile + gridDim.x) * kGemmBM * kC, kGemmBM * kC * 4);
                load_g1(0);
                load_g1(1);
                for (int h = 0; h < kNChunk; ++h) {   // same order as the MMA warp consumes the ring
                    load_g2(h);
                    if (h + 2 < kNChunk) load_g1(h + 2);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (whole warp, one elected lane issues)
        {
            const uint32_t idesc1 = umma_idesc_bf16(kGemmBM, kChunk);
            const uint32_t idesc2 = umma_idesc_bf16(kGemmBM, kC);
            int stage = 0;
            uint32_t phase = 0;
            auto next_stage = [&]() {
                if (++stage == kFStages) {
                    stage = 0;
                    phase ^= 1;
                }
            };
            const bool mtr = p.trace != nullptr && blockIdx.x == 0 && lane == 0;
            long long w_full = 0, w_ready = 0;
            auto g1 = [&](int b) {
                const uint32_t d = tmem_base + kAcc1Col + (uint32_t)b * kChunk;
                for (int j = 0; j < 2; ++j) {
                    const long long t0 = mtr ? clock64() : 0;
                    mbar_wait(&full[stage], phase);
                    if (mtr) w_full += clock64() - t0;
                    tc_fence_after();
                    if (elect_one()) {
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            const int kb = 2 * j + q;
                            const uint64_t da = umma_desc_sw128(smem_u32(sX + kb * kABytes));
                            const uint64_t db = umma_desc_sw128(smem_u32(sRing + stage * kFStageBytes + q * (kFStageBytes / 2)));
#pragma unroll
                            for (int k = 0; k < kGemmBK / 16; ++k)
                                umma_bf16(d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc1, (uint32_t)((kb | k) != 0));
                        }
                        umma_commit_mc(&empty[stage], kClusterMask);   // the stage is refilled by both CTAs: both must see it free
                        if (j == 1) umma_commit(&a1_full[b]);
                    }
                    next_stage();
                }
            };
            auto g2 = [&](int h) {
                const int b = h & 1;
                const uint32_t d = tmem_base + kAcc2Col;
                for (int j = 0; j < 2; ++j) {
                    const long long t0 = mtr ? clock64() : 0;
                    mbar_wait(&full[stage], phase);
                    if (mtr) w_full += clock64() - t0;
                    tc_fence_after();
                    if (elect_one()) {
                        // A = P[b] straight from tensor memory (bf16 pairs, written by E1 over the drained acc1[b]):
                        // K-step kk covers hidden [16 kk, +16) = words [8 (kk & 1), +8) of epilogue part kk >> 1
                        const uint32_t pa = tmem_base + kAcc1Col + (uint32_t)(b * kChunk);
                        const uint64_t db = umma_desc_sw128(smem_u32(sRing + stage * kFStageBytes));
#pragma unroll
                        for (int k = 0; k < kGemmBK / 16; ++k) {
                            const int kk = j * (kGemmBK / 16) + k;
                            umma_bf16_ts(d, pa + (uint32_t)((kk >> 1) * 32 + (kk & 1) * 8), db + (uint64_t)(2 * k), idesc2, 1u);  // acc2 was preloaded with hres + b2
                        }
                        umma_commit_mc(&empty[stage], kClusterMask);   // the stage is refilled by both CTAs: both must see it free
                    }
                    next_stage();
                }
            };
            int it = 0;
            uint32_t pr_phase[2] = {0, 0};
            for (int tile = blockIdx.x; (tile & ~(kCluster - 1)) < p.num_tiles; tile += gridDim.x, ++it) {  // cluster-uniform trip count
                if constexpr (kFused) {
                    mbar_wait(acc2_init, (uint32_t)(it & 1));   // every epilogue warp has drained acc2 of the previous tile
                    tc_fence_after();
                    for (int kb = 0; kb < kOKB; ++kb) {
                        const int slot = kb & (kASlots - 1);
                        mbar_wait(&a_full[slot], (uint32_t)(kb >= kASlots));   // two completions per tile and slot
                        mbar_wait(&full[stage], phase);
                        tc_fence_after();
                        if (elect_one()) {
                            const uint64_t da = umma_desc_sw128(smem_u32(sX + slot * kABytes));
                            const uint64_t db = umma_desc_sw128(smem_u32(sRing + stage * kFStageBytes));
#pragma unroll
                            for (int k = 0; k < kGemmBK / 16; ++k)
                                umma_bf16(tmem_base + kAcc2Col, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc2, (uint32_t)((kb | k) != 0));
                            umma_commit_mc(&empty[stage], kClusterMask);
                            if (kb < kASlots) umma_commit(&a_empty[slot]);
                            if (kb + 1 == kOKB) umma_commit(oproj_full);
                        }
                        next_stage();
                    }
                    mbar_wait(x_ready, (uint32_t)(it & 1));   // X = LayerNorm(h1) written to shared memory by the epilogue warps
                } else {
                    mbar_wait(x_full, (uint32_t)(it & 1));
                }
                tc_fence_after();
                if (mtr) p.trace[it * 48 + 32] = clock64();
                g1(0);
                g1(1);
                for (int h = 0; h < kNChunk; ++h) {
                    const int b = h & 1;
                    const long long t0 = mtr ? clock64() : 0;
                    mbar_wait(&p_ready[b], pr_phase[b]);   // P[b] written, acc1[b] drained (h == 0: acc2 preloaded too)
                    pr_phase[b] ^= 1;
                    tc_fence_after();
                    if (mtr) w_ready += clock64() - t0;
                    if (mtr) p.trace[it * 48 + 33 + h] = clock64();
                    g2(h);                                 // reads P[b] where acc1[b] was ...
                    if (h + 2 < kNChunk) g1(b);            // ... so the G1 that overwrites acc1[b] follows it on the (in-order) tensor pipe
                    if (h + 2 == kNChunk - 1 && elect_one()) umma_commit(x_free);  // last G1 of the tile issued: X may be reloaded
                }
                if (elect_one()) umma_commit(acc2_full);
                if (mtr) {
                    p.trace[it * 48 + 41] = clock64();
                    p.trace[it * 48 + 42] = w_full;
                    p.trace[it * 48 + 43] = w_ready;
                    w_full = w_ready = 0;
                }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue warps
        const int ew = warp - 2;
        const int part = ew >> 2;
        const int quarter = warp & 3;
        const int trow = quarter * 32 + lane;  // row of the tile owned by this thread
        // per-warp 4 KB staging tile (32 rows x 128 B, 128B-swizzled) for the bf16 TMA store
        uint8_t* stg = sStg + ew * kOutStageBytes;
        const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        uint32_t a1_phase[2] = {0, 0};
        int it = 0;
        const bool tr = p.trace != nullptr && blockIdx.x == 0 && ew == 0 && lane == 0;
#define FFN_STAMP(k) do { if (tr) p.trace[it * 48 + (k)] = clock64(); } while (0)
        for (int tile = blockIdx.x; (tile & ~(kCluster - 1)) < p.num_tiles; tile += gridDim.x, ++it) {  // cluster-uniform trip count
            FFN_STAMP(0);
            const int row = tile * kGemmBM + trow;
            const bool live = tile < p.num_tiles;   // the last cluster may carry a padding tile: it only keeps the weight ring in step
            const bool valid = live && (p.row_seq == nullptr || __ldg(p.row_seq + row) >= 0);

            if constexpr (!kFused)
            // ---- preload acc2 := hres + b2 (own row, own 64 columns): the MMAs then accumulate the residual stream
            // directly, and the L2 latency of these (coalesced, tiled-layout) loads hides behind G1(0) / G1(1).  This
            // warp drained the same TMEM lanes / columns in the previous tile; the p_ready[0] arrival of E1(0)
            // publishes the stores to the MMA thread.
            {
                const float* hrow = p.hres + tiled_f32_offset(row, part * 64);  // float4 j of the row: + j * 512 floats
                float4 t[16];   // all 16 loads of the thread in flight at once: one L2 round trip per tile, not four
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    t[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (live) t[j] = *reinterpret_cast<const float4*>(hrow + (size_t)j * 512);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float r[16];
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 bb = *reinterpret_cast<const float4*>(sB2 + part * 64 + q * 16 + i);   // broadcast LDS
                        const float4 tt = t[q * 4 + (i >> 2)];
                        r[i] = tt.x + bb.x; r[i + 1] = tt.y + bb.y; r[i + 2] = tt.z + bb.z; r[i + 3] = tt.w + bb.w;
                    }
                    tmem_st16(tlane + kAcc2Col + (uint32_t)(part * 64 + q * 16), r);
                }
                tmem_st_wait();
            }
            FFN_STAMP(1);
            if constexpr (kFused) {
                // acc2 is free: the out-projection MMAs start from zero at once and overlap the residual loads below; h1 = acc2 +
                // hres + bo is formed in registers and written back afterwards.  For every tile but the CTA's first the hand-over
                // was already made at the end of the previous tile's accumulator drain (below), so that this tile's
                // out-projection runs under the previous tile's LayerNorm and stores.
                if (it == 0) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc2_init);
                }
                const float* hrow = p.hres + tiled_f32_offset(row, part * 64);  // float4 j of the row: + j * 512 floats
                float4 hq[4], hn[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) hq[u] = live ? *reinterpret_cast<const float4*>(hrow + (size_t)u * 512) : make_float4(0.f, 0.f, 0.f, 0.f);
                // ---- out-projection done.  X = LayerNorm(h1) (affine folded into W1 / b1) -> bf16 -> shared memory
                mbar_wait(oproj_full, (uint32_t)(it & 1));
                tc_fence_after();
                float v[64];
                {
                    const uint32_t ta = tlane + kAcc2Col + (uint32_t)(part * 64);
                    tmem_ld16(ta, v);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        tmem_ld_wait();
                        if (q < 3) {
                            tmem_ld16(ta + (q + 1) * 16, v + (q + 1) * 16);
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                hn[u] = live ? *reinterpret_cast<const float4*>(hrow + (size_t)((q + 1) * 4 + u) * 512) : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                        float* w = v + q * 16;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const float4 bb = ldg4(p.bo + part * 64 + q * 16 + 4 * u);
                            w[4 * u] = valid ? w[4 * u] + hq[u].x + bb.x : 0.f;
                            w[4 * u + 1] = valid ? w[4 * u + 1] + hq[u].y + bb.y : 0.f;
                            w[4 * u + 2] = valid ? w[4 * u + 2] + hq[u].z + bb.z : 0.f;
                            w[4 * u + 3] = valid ? w[4 * u + 3] + hq[u].w + bb.w : 0.f;
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) hq[u] = hn[u];
                    }
                }
                if (lane == 0) tma_store_wait_read();   // the exchange slot is the head of the staging tile the last store read
                __syncwarp();
                const float2 lm = local_mean_m2<64>(v);
                reinterpret_cast<float2*>(stg)[lane] = lm;
                group_sync(quarter);
                float2 o[kEpiParts];
                float mean = 0.f;
#pragma unroll
                for (int q = 0; q < kEpiParts; ++q) {
                    o[q] = reinterpret_cast<const float2*>(sStg + ((q << 2) | (ew & 3)) * kOutStageBytes)[lane];
                    mean += o[q].x;
                }
                group_sync(quarter);
                mean *= (1.f / kEpiParts);
                float m2 = 0.f;
#pragma unroll
                for (int q = 0; q < kEpiParts; ++q) {
                    const float d = o[q].x - mean;
                    m2 += o[q].y + 64.f * d * d;
                }
                const float rstd = rsqrtf(m2 * (1.f / kC) + p.ln_eps);
                const float a = valid ? rstd : 0.f, bsh = valid ? -mean * rstd : 0.f;
                // this thread's 64 columns are exactly k-block `part` of the X tile: row trow, 8 chunks of 16 bytes, chunk
                // index XOR (row & 7) = the 128-byte swizzle TMA would have applied (conflict-free per quarter warp)
                uint8_t* xrow = sX + part * kABytes + trow * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float* w = v + 8 * j;
                    uint4 t;
                    t.x = pack_bf16x2(fmaf(w[0], a, bsh), fmaf(w[1], a, bsh));
                    t.y = pack_bf16x2(fmaf(w[2], a, bsh), fmaf(w[3], a, bsh));
                    t.z = pack_bf16x2(fmaf(w[4], a, bsh), fmaf(w[5], a, bsh));
                    t.w = pack_bf16x2(fmaf(w[6], a, bsh), fmaf(w[7], a, bsh));
                    *reinterpret_cast<uint4*>(xrow + ((j ^ (trow & 7)) << 4)) = t;
                }
                fence_proxy_async_smem();   // generic-proxy stores -> visible to the tensor core's async-proxy reads
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(x_ready);
                // h1 back into acc2 (own lanes / columns): the feed-forward MMAs accumulate onto it.  Off the critical path of
                // G1; the p_ready[0] arrival of E1(0) publishes the stores to the MMA thread.
#pragma unroll
                for (int q = 0; q < 4; ++q) tmem_st16(tlane + kAcc2Col + (uint32_t)(part * 64 + q * 16), v + q * 16);
                tmem_st_wait();
            }

            // ---- E1: hidden chunks -> GELU -> P[b].  TMEM -> register reads run at ~64 B/clk/SM (a 128 x 128 fp32 chunk
            // takes ~1000 clk to drain), so they are software-pipelined against the math: each thread's 32 columns
            // are read as two 16-column halves, and the load of the next half (or of the next chunk's first half)
            // is always in flight while the current half goes through the GELU.
            float va[16], vb[16];
            mbar_wait(&a1_full[0], a1_phase[0]);
            a1_phase[0] ^= 1;
            tc_fence_after();
            tmem_ld16(tlane + kAcc1Col + (uint32_t)(part * 32), va);
            bool have_va = true;
            for (int h = 0; h < kNChunk; ++h) {
                const int b = h & 1;
                if (!have_va) {   // the accumulator was not ready when the previous chunk looked: wait for it now
                    mbar_wait(&a1_full[b], a1_phase[b]);
                    a1_phase[b] ^= 1;
                    tc_fence_after();
                    tmem_ld16(tlane + kAcc1Col + (uint32_t)(b * kChunk + part * 32), va);
                }
                const float* bp = p.b1 + h * kChunk + part * 32;
                float4 bias[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) bias[i] = ldg4(bp + 4 * i);
                tmem_ld_wait();   // va of this chunk
                FFN_STAMP(2 + 2 * h);
                tmem_ld16(tlane + kAcc1Col + (uint32_t)(b * kChunk + part * 32 + 16), vb);
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    pk[2 * i] = gelu2_bf16(va[4 * i] + bias[i].x, va[4 * i + 1] + bias[i].y);
                    pk[2 * i + 1] = gelu2_bf16(va[4 * i + 2] + bias[i].z, va[4 * i + 3] + bias[i].w);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) bias[i] = ldg4(bp + 16 + 4 * i);
                tmem_ld_wait();   // vb of this chunk: acc1[b] is now fully drained by this thread
                // first half of the next chunk (other accumulator buffer) flies during the math below -- if G1 has already
                // delivered it.  Only probed: blocking here would put G2(h - 1) + G1(h + 1) into every chunk's critical path.
                have_va = false;
                if (h + 1 < kNChunk && __shfl_sync(0xffffffffu, (int)mbar_test(&a1_full[b ^ 1], a1_phase[b ^ 1]), 0)) {
                    a1_phase[b ^ 1] ^= 1;
                    tc_fence_after();
                    tmem_ld16(tlane + kAcc1Col + (uint32_t)((b ^ 1) * kChunk + part * 32), va);
                    have_va = true;
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    pk[8 + 2 * i] = gelu2_bf16(vb[4 * i] + bias[i].x, vb[4 * i + 1] + bias[i].y);
                    pk[8 + 2 * i + 1] = gelu2_bf16(vb[4 * i + 2] + bias[i].z, vb[4 * i + 3] + bias[i].w);
                }
                // P[b] goes back into tensor memory, over the first 16 of the 32 accumulator columns this thread has just
                // drained (no other thread reads them): G2 takes its A operand from there, no shared-memory round trip
                tmem_st16(tlane + kAcc1Col + (uint32_t)(b * kChunk + part * 32), reinterpret_cast<const float*>(pk));
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_ready[b]);
                FFN_STAMP(3 + 2 * h);
            }

            // ---- final: acc2 (= hres + b2 + FFN) -> registers -> fp32 stream (tiled, coalesced) + LayerNorm -> bf16
            mbar_wait(acc2_full, (uint32_t)(it & 1));
            tc_fence_after();
            FFN_STAMP(18);
            float v[64];
            {   // drain 16 columns at a time, the next load in flight while this one is stored
                const uint32_t ta = tlane + kAcc2Col + (uint32_t)(part * 64);
                tmem_ld16(ta, v);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    tmem_ld_wait();
                    if (q < 3) tmem_ld16(ta + (q + 1) * 16, v + (q + 1) * 16);
                    float* w = v + q * 16;
                    if constexpr (kFused) {   // b2 was not part of the preload here
#pragma unroll
                        for (int i = 0; i < 16; i += 4) {
                            const float4 bb = *reinterpret_cast<const float4*>(sB2 + part * 64 + q * 16 + i);   // broadcast LDS
                            w[i] += bb.x; w[i + 1] += bb.y; w[i + 2] += bb.z; w[i + 3] += bb.w;
                        }
                    }
                    if (!valid) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) w[i] = 0.f;
                    }
                    if (p.store_f32 && live) {
                        float* dst = p.hres + tiled_f32_offset(row, part * 64 + q * 16);
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            *reinterpret_cast<float4*>(dst + (size_t)u * 512) = make_float4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
                    }
                }
            }
            if constexpr (kFused) {
                // acc2 is drained into registers: hand it to the MMA thread now if this CTA has another tile (cluster-uniform
                // trip count as in the loop header), its out-projection GEMM then overlaps the LayerNorm + stores below
                if (((tile + (int)gridDim.x) & ~(kCluster - 1)) < p.num_tiles) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc2_init);
                }
            }
            if (lane == 0) tma_store_wait_read();   // the previous tile's last store has long finished reading this staging tile
            __syncwarp();
            if (p.ln) {   // (x - mean) * rstd; gamma / beta live in the consumer's weights
                // row statistics over the 4 warps that share this TMEM lane quarter (slot = head of the staging tile):
                // per-warp mean and centred sum of squares, merged with the parallel-variance formula (no E[x^2] - mean^2
                // cancellation when the residual stream carries a large mean)
                const float2 lm = local_mean_m2<64>(v);
                reinterpret_cast<float2*>(stg)[lane] = lm;
                group_sync(quarter);
                float2 o[kEpiParts];
                float mean = 0.f;
#pragma unroll
                for (int q = 0; q < kEpiParts; ++q) {
                    o[q] = reinterpret_cast<const float2*>(sStg + ((q << 2) | (ew & 3)) * kOutStageBytes)[lane];
                    mean += o[q].x;
                }
                group_sync(quarter);
                mean *= (1.f / kEpiParts);
                float m2 = 0.f;
#pragma unroll
                for (int q = 0; q < kEpiParts; ++q) {
                    const float d = o[q].x - mean;
                    m2 += o[q].y + 64.f * d * d;
                }
                const float rstd = rsqrtf(m2 * (1.f / kC) + p.ln_eps);
                const float a = valid ? rstd : 0.f, bsh = valid ? -mean * rstd : 0.f;
#pragma unroll
                for (int i = 0; i < 64; ++i) v[i] = fmaf(v[i], a, bsh);
            }
            // bf16 tile of this warp (32 rows x 64 columns) -> 64B-swizzled 2 KB staging -> two TMA stores of 32 columns
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                if (hh == 1) {
                    if (lane == 0) tma_store_wait_read();
                    __syncwarp();
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float* w = v + hh * 32 + u * 8;
                    uint4 t;
                    t.x = pack_bf16x2(w[0], w[1]);
                    t.y = pack_bf16x2(w[2], w[3]);
                    t.z = pack_bf16x2(w[4], w[5]);
                    t.w = pack_bf16x2(w[6], w[7]);
                    *reinterpret_cast<uint4*>(stg + lane * 64 + ((u ^ ((lane >> 1) & 3)) << 4)) = t;
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0 && live) {
                    tma_store_2d(&tmOut, stg, part * 64 + hh * 32, tile * kGemmBM + quarter * 32);
                    tma_store_commit();
                }
            }
            FFN_STAMP(19);
        }
        if (lane == 0) tma_store_wait_all();   // the global writes must be complete before the CTA exits
#undef FFN_STAMP
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // no CTA leaves while a peer may still multicast into it or arrive on its barriers
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

cudaError_t ffn_tc_init_device() {
    static std::atomic<uint64_t> done{0};
    static std::mutex mu;
    return once_per_device(done, mu, [] {
        cudaError_t e = cudaFuncSetAttribute(ffn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfnSmem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(ffn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfnSmem);
        return e;
    });
}

cudaError_t ffn_tc_launch(const FfnProblem& f, int num_sms, cudaStream_t stream) {
    const bool fused = f.att != nullptr;
    if (f.M <= 0 || (f.M % kGemmBM) != 0 || !f.w1 || !f.w2 || !f.b1 || !f.b2 || !f.hres || !f.out_bf16)
        return cudaErrorInvalidValue;
    if (fused ? (!f.wo || !f.bo) : !f.x) return cudaErrorInvalidValue;
    const cudaError_t attr_err = ffn_tc_init_device();
    if (attr_err != cudaSuccess) return attr_err;

    FfnParams p;
    p.M = f.M;
    p.num_tiles = f.M / kGemmBM;
    p.b1 = f.b1;
    p.b2 = f.b2;
    p.bo = f.bo;
    p.ln = f.ln;
    p.row_seq = f.row_seq;
    p.hres = f.hres;
    p.store_f32 = f.store_f32;
    p.ln_eps = f.ln_eps;
    p.trace = f.trace;

    CUtensorMap tX, tWo, tW1, tW2, tOut;
    if (fused) {
        if (!encode_2d(&tX, f.att, kOK, (uint64_t)f.M, kOK, kGemmBM)) return cudaErrorInvalidValue;
        if (!encode_2d(&tWo, f.wo, kOK, kC, kOK, kC / kCluster)) return cudaErrorInvalidValue;
    } else {
        if (!encode_2d(&tX, f.x, kC, (uint64_t)f.M, kC, kGemmBM)) return cudaErrorInvalidValue;
    }
    if (!encode_2d(&tW1, f.w1, kC, kHid, kC, kChunk)) return cudaErrorInvalidValue;
    if (!encode_2d(&tW2, f.w2, kHid, kC, kHid, kC / kCluster)) return cudaErrorInvalidValue;
    if (!encode_2d_ex(&tOut, f.out_bf16, kC, (uint64_t)f.M, (uint64_t)f.ldo, 32, 32, 1)) return cudaErrorInvalidValue;
    if (!fused) tWo = tW1;

    int grid = (p.num_tiles + kCluster - 1) / kCluster * kCluster;
    if (grid > num_sms) grid = num_sms / kCluster * kCluster;
    const cudaError_t le = fused ? launch_pdl(ffn_tc_kernel<true>, dim3(grid), dim3(kThreads), kFfnSmem, stream, tX, tWo, tW1, tW2, tOut, p)
                                 : launch_pdl(ffn_tc_kernel<false>, dim3(grid), dim3(kThreads), kFfnSmem, stream, tX, tWo, tW1, tW2, tOut, p);
    if (le != cudaSuccess) return le;
    return cudaGetLastError();
}

}  // namespace mossdec
