// Fused feed-forward of BasicTransformerBlock (matcha/models/components/transformer.py:299-314 with diffusers'
// FeedForward = Linear(256 -> 1024) + exact-erf GELU, Linear(1024 -> 256)) as ONE persistent tcgen05 kernel:
//
//   hres[r, :] += W2 . gelu(W1 . xn[r, :] + b1) + b2          (fp32 residual stream, in place)
//   out[r, :]   = LayerNorm_next(hres[r, :])  or  bf16(hres[r, :])          (operand of the next GEMM)
//
// The 1024-wide hidden activation never leaves the SM.  Per 128-row tile:
//   X tile (128 x 256 bf16, 64 KB) is loaded once by TMA and stays in shared memory;
//   for each of 8 hidden chunks of 128:  G1: acc1[b] = X . W1[chunk]^T        (tcgen05.mma M128 N128, K = 256)
//                                        E1: acc1[b] -> +b1 -> GELU -> bf16 -> P[b] in shared memory, already in the
//                                            128B-swizzled K-major layout the tensor core reads its A operand from
//                                        G2: acc2 += P[b] . W2[:, chunk]^T    (tcgen05.mma M128 N256, K = 128)
//   final: acc2 -> registers (TMEM released at once) -> + b2 + hres -> hres, LayerNorm -> bf16 out.
// acc1 is double buffered in TMEM (2 x 128 columns) next to acc2 (256 columns), P is double buffered in shared memory,
// so the GELU epilogue of chunk h overlaps G2(h-1) / G1(h+1) on the tensor pipe.  W1 / W2 stream through a 3-stage ring
// of 32 KB TMA stages (1 MB of weights per tile, L2 resident).
//
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM alloc), warps 2..17 = epilogue (thread == row; the 4
// warps that share a TMEM lane quarter split the columns).
#include "gemm.h"
#include "ptx.cuh"
#include "epilogue.cuh"

#include <mutex>
#include <stdio.h>

namespace mossdec {

bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows);

namespace {

constexpr int kC = 256;          // model width
constexpr int kHid = 1024;       // hidden width
constexpr int kChunk = 128;      // hidden columns per chunk
constexpr int kNChunk = kHid / kChunk;
constexpr int kFStages = 3;
constexpr int kFStageBytes = 32768;
constexpr int kXBytes = kGemmBM * kC * 2;      // 64 KB
constexpr int kPBytes = kGemmBM * kChunk * 2;  // 32 KB per buffer
constexpr int kFfnSmem = kXBytes + 2 * kPBytes + kFStages * kFStageBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t kAcc2Col = 0, kAcc1Col = 256;
static_assert(kEpiWarps * kStageBytes <= kPBytes, "final-epilogue staging tiles alias P[0]");

struct FfnParams {
    int32_t M, num_tiles;
    const float* b1;
    const float* b2;
    const float* ln_g;  // nullptr: out_bf16 = bf16(hres)
    const float* ln_b;
    const int32_t* row_seq;
    float* hres;        // [M, 256] fp32, read and written
    __nv_bfloat16* out_bf16;
    int32_t ldo;
    int32_t store_f32;
    float ln_eps;
};

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void epi_sync_all() { asm volatile("bar.sync 5, %0;" ::"n"(32 * kEpiWarps) : "memory"); }

__global__ void __launch_bounds__(kThreads, 1)
ffn_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
              const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ FfnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays in the shared address space (STS / LDS)
    uint8_t* sX = smem;
    uint8_t* sP = sX + kXBytes;
    uint8_t* sRing = sP + 2 * kPBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sRing + kFStages * kFStageBytes);
    uint64_t* full = bars;               // [kFStages]  TMA -> MMA
    uint64_t* empty = full + kFStages;   // [kFStages]  MMA -> TMA
    uint64_t* x_full = empty + kFStages; // TMA -> MMA
    uint64_t* x_free = x_full + 1;       // MMA -> TMA (all G1 of the tile done)
    uint64_t* a1_full = x_free + 1;      // [2] MMA -> epilogue
    uint64_t* p_ready = a1_full + 2;     // [2] epilogue -> MMA: P[b] written, acc1[b] drained
    uint64_t* p_free = p_ready + 2;      // [2] MMA -> epilogue: G2 finished reading P[b]
    uint64_t* acc2_full = p_free + 2;    // MMA -> epilogue
    uint64_t* acc2_free = acc2_full + 1; // epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc2_free + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmW1);
        tma_prefetch_desc(&tmW2);
        for (int i = 0; i < kFStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        mbar_init(x_full, 1);
        mbar_init(x_free, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&a1_full[i], 1);
            mbar_init(&p_ready[i], kEpiWarps);
            mbar_init(&p_free[i], 1);
        }
        mbar_init(acc2_full, 1);
        mbar_init(acc2_free, kEpiWarps);
        mbar_fence_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
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
                    mbar_expect_tx(&full[stage], kFStageBytes);
                    uint8_t* dst = sRing + stage * kFStageBytes;
                    tma_load_2d(dst, &tmW1, (2 * j) * kGemmBK, h * kChunk, &full[stage]);
                    tma_load_2d(dst + kFStageBytes / 2, &tmW1, (2 * j + 1) * kGemmBK, h * kChunk, &full[stage]);
                    next_stage();
                }
            };
            auto load_g2 = [&](int h) {   // W2 all 256 rows, K columns [128 h, +128): two stages of one k-block
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_expect_tx(&full[stage], kFStageBytes);
                    tma_load_2d(sRing + stage * kFStageBytes, &tmW2, h * kChunk + j * kGemmBK, 0, &full[stage]);
                    next_stage();
                }
            };
            int it = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
                if (it > 0) mbar_wait(x_free, (uint32_t)((it - 1) & 1));
                mbar_expect_tx(x_full, kXBytes);
                for (int kb = 0; kb < kC / kGemmBK; ++kb)
                    tma_load_2d(sX + kb * kABytes, &tmX, kb * kGemmBK, tile * kGemmBM, x_full);
                load_g1(0);
                load_g1(1);
                for (int h = 0; h < kNChunk; ++h) {
                    load_g2(h);
                    if (h + 2 < kNChunk) load_g1(h + 2);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
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
            auto g1 = [&](int b) {
                const uint32_t d = tmem_base + kAcc1Col + (uint32_t)b * kChunk;
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int kb = 2 * j + q;
                        const uint64_t da = umma_desc_sw128(smem_u32(sX + kb * kABytes));
                        const uint64_t db = umma_desc_sw128(smem_u32(sRing + stage * kFStageBytes + q * (kFStageBytes / 2)));
#pragma unroll
                        for (int k = 0; k < kGemmBK / 16; ++k)
                            umma_bf16(d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc1, (uint32_t)((kb | k) != 0));
                    }
                    umma_commit(&empty[stage]);
                    next_stage();
                }
                umma_commit(&a1_full[b]);
            };
            auto g2 = [&](int h) {
                const int b = h & 1;
                const uint32_t d = tmem_base + kAcc2Col;
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint64_t da = umma_desc_sw128(smem_u32(sP + b * kPBytes + j * kABytes));
                    const uint64_t db = umma_desc_sw128(smem_u32(sRing + stage * kFStageBytes));
#pragma unroll
                    for (int k = 0; k < kGemmBK / 16; ++k)
                        umma_bf16(d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc2, (uint32_t)((h | j | k) != 0));
                    umma_commit(&empty[stage]);
                    next_stage();
                }
                umma_commit(&p_free[b]);
            };
            int it = 0;
            uint32_t pr_phase[2] = {0, 0};
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
                mbar_wait(x_full, (uint32_t)(it & 1));
                tc_fence_after();
                g1(0);
                g1(1);
                for (int h = 0; h < kNChunk; ++h) {
                    const int b = h & 1;
                    mbar_wait(&p_ready[b], pr_phase[b]);
                    pr_phase[b] ^= 1;
                    if (h == 0 && it > 0) mbar_wait(acc2_free, (uint32_t)((it - 1) & 1));
                    tc_fence_after();
                    g2(h);
                    if (h + 2 < kNChunk) g1(b);
                    if (h + 2 == kNChunk - 1) umma_commit(x_free);  // last G1 of the tile issued: X may be reloaded
                }
                umma_commit(acc2_full);
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue warps
        const int ew = warp - 2;
        const int part = ew >> 2;
        const int quarter = warp & 3;
        const int trow = quarter * 32 + lane;  // row of the tile owned by this thread
        KParams kp;                            // only M is read by the staging helpers
        kp.M = p.M;
        EpiCtx c;
        c.p = &kp;
        c.lane = lane;
        c.ew = ew;
        c.quarter = quarter;
        c.stg = sP + ew * kStageBytes;  // aliases P[0]: only used while no G2 reads it (see the sync below)
        c.stg_all = sP;
        c.n0 = 0;
        const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        uint32_t a1_phase[2] = {0, 0};
        int it = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            c.row0 = tile * kGemmBM + quarter * 32;
            const int row = c.row0 + lane;
            c.seq = (p.row_seq != nullptr && row < p.M) ? __ldg(p.row_seq + row) : 0;
            c.valid = row < p.M && c.seq >= 0;

            // ---- E1: hidden chunks -> GELU -> P[b]
            for (int h = 0; h < kNChunk; ++h) {
                const int b = h & 1;
                mbar_wait(&a1_full[b], a1_phase[b]);
                a1_phase[b] ^= 1;
                if (h >= 2) mbar_wait(&p_free[b], (uint32_t)(((h - 2) >> 1) & 1));
                tc_fence_after();
                float v[32];
                const uint32_t ta = tlane + kAcc1Col + (uint32_t)(b * kChunk + part * 32);
                tmem_ld16(ta, v);
                tmem_ld16(ta + 16, v + 16);
                tmem_ld_wait();
                const float* bias = p.b1 + h * kChunk + part * 32;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 bb = ldg4(bias + i);
                    v[i] = fast_gelu_erf(v[i] + bb.x);
                    v[i + 1] = fast_gelu_erf(v[i + 1] + bb.y);
                    v[i + 2] = fast_gelu_erf(v[i + 2] + bb.z);
                    v[i + 3] = fast_gelu_erf(v[i + 3] + bb.w);
                }
                // K-major SW128 operand layout: k-block (64 hidden) = 128 rows x 128 B, 16-byte units XOR row % 8
                uint8_t* prow = sP + b * kPBytes + (part >> 1) * kABytes + trow * 128;
                const int ub = (part & 1) * 4;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    uint4 t;
                    t.x = pack_bf16x2(v[u * 8], v[u * 8 + 1]);
                    t.y = pack_bf16x2(v[u * 8 + 2], v[u * 8 + 3]);
                    t.z = pack_bf16x2(v[u * 8 + 4], v[u * 8 + 5]);
                    t.w = pack_bf16x2(v[u * 8 + 6], v[u * 8 + 7]);
                    *reinterpret_cast<uint4*>(prow + (((ub + u) ^ (trow & 7)) << 4)) = t;
                }
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_ready[b]);
            }

            // ---- final: acc2 -> registers, release TMEM, then bias + residual + LayerNorm from registers
            mbar_wait(acc2_full, (uint32_t)(it & 1));
            tc_fence_after();
            float v[64];
            {
                const uint32_t ta = tlane + kAcc2Col + (uint32_t)(part * 64);
                tmem_ld16(ta, v);
                tmem_ld16(ta + 16, v + 16);
                tmem_ld16(ta + 32, v + 32);
                tmem_ld16(ta + 48, v + 48);
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc2_free);

            float s = 0.f, ss = 0.f;
#pragma unroll
            for (int sc = 0; sc < 2; ++sc) {
                const int nb = part * 64 + sc * 32;
                float* w = v + sc * 32;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 bb = ldg4(p.b2 + nb + i);
                    w[i] += bb.x; w[i + 1] += bb.y; w[i + 2] += bb.z; w[i + 3] += bb.w;
                }
                add_rows_f32(c, p.hres, kC, nb, kC, w);
                if (!c.valid) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) w[i] = 0.f;
                }
                if (p.store_f32) store_rows_f32(c, p.hres, kC, nb, kC, w);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    s += w[i];
                    ss += w[i] * w[i];
                }
            }
            if (p.ln_g != nullptr) {
                const float2 t = group_exchange(c, make_float2(s, ss));
                const float mean = t.x * (1.f / kC);
                const float rstd = rsqrtf(fmaxf(t.y * (1.f / kC) - mean * mean, 0.f) + p.ln_eps);
#pragma unroll
                for (int i = 0; i < 64; i += 4) {
                    const int n = part * 64 + i;
                    const float4 g = ldg4(p.ln_g + n), bb = ldg4(p.ln_b + n);
                    v[i] = c.valid ? (v[i] - mean) * rstd * g.x + bb.x : 0.f;
                    v[i + 1] = c.valid ? (v[i + 1] - mean) * rstd * g.y + bb.y : 0.f;
                    v[i + 2] = c.valid ? (v[i + 2] - mean) * rstd * g.z + bb.z : 0.f;
                    v[i + 3] = c.valid ? (v[i + 3] - mean) * rstd * g.w + bb.w : 0.f;
                }
            }
            // (the per-warp staging tile holds 32 rows x 64 B: two passes of 32 columns)
            store_rows_bf16<32>(c, p.out_bf16, p.ldo, part * 64, kC, v);
            store_rows_bf16<32>(c, p.out_bf16, p.ldo, part * 64 + 32, kC, v + 32);
            // the staging tiles alias P[0]: every epilogue warp must be done with them before the next tile's E1
            epi_sync_all();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

cudaError_t ffn_tc_launch(const FfnProblem& f, int num_sms, cudaStream_t stream) {
    if (f.M <= 0 || (f.M % kGemmBM) != 0 || !f.x || !f.w1 || !f.w2 || !f.b1 || !f.b2 || !f.hres || !f.out_bf16)
        return cudaErrorInvalidValue;
    static std::once_flag attr_once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(attr_once, [] {
        attr_err = cudaFuncSetAttribute(ffn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfnSmem);
    });
    if (attr_err != cudaSuccess) return attr_err;

    FfnParams p;
    p.M = f.M;
    p.num_tiles = f.M / kGemmBM;
    p.b1 = f.b1;
    p.b2 = f.b2;
    p.ln_g = f.ln_g;
    p.ln_b = f.ln_b;
    p.row_seq = f.row_seq;
    p.hres = f.hres;
    p.out_bf16 = f.out_bf16;
    p.ldo = f.ldo;
    p.store_f32 = f.store_f32;
    p.ln_eps = f.ln_eps;

    CUtensorMap tX, tW1, tW2;
    if (!encode_2d(&tX, f.x, kC, (uint64_t)f.M, kC, kGemmBM)) return cudaErrorInvalidValue;
    if (!encode_2d(&tW1, f.w1, kC, kHid, kC, kChunk)) return cudaErrorInvalidValue;
    if (!encode_2d(&tW2, f.w2, kHid, kC, kHid, kC)) return cudaErrorInvalidValue;

    const int grid = p.num_tiles < num_sms ? p.num_tiles : num_sms;
    ffn_tc_kernel<<<grid, kThreads, kFfnSmem, stream>>>(tX, tW1, tW2, p);
    return cudaGetLastError();
}

}  // namespace mossdec
