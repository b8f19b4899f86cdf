// Self-attention of BasicTransformerBlock (diffusers Attention behind matcha transformer.py:266-271: softmax(q k^T / 8 +
// additive mask) v, 8 heads of 64) over packed rows, warp-specialised on tcgen05 / TMEM.  The additive mask of the
// reference ([B, T, T] fp32, cosyvoice/utils/mask.py:161-236 + common.py:160-168) is a predicate here: key j of a sequence
// is visible to query i iff j < len and (chunk == 0 or j < (i / chunk + 1) * chunk).
//   * a dedicated warp (warp 4) issues every TMA load and every tcgen05.mma, with mbarriers instead of block barriers;
//   * key tiles of 32 with a DOUBLE-BUFFERED score tile in tensor memory: S0 | S1 | O = 32 + 32 + 64 = 128 columns,
//     four CTAs per SM.  QK(j + 2) is issued right behind PV(j), so S_{j+1} is complete before the softmax warps finish
//     tile j: they never wait for an MMA issue or for MMA latency in steady state.
//   warps 0..3: thread == query row: tcgen05.ld S_j, predicate mask, online max (lazy rescaling of O), exp2, bf16 P_j
//               written over S_j with tcgen05.st, one mbarrier arrival per warp.
//   warp 4    : Q once, K/V tiles through a 4-stage ring (K and V of a tile share a stage and a barrier);
//               per tile: PV(j) (A = P_j from TMEM, B = V_j MN-major), QK(j + 2) into the S buffer PV(j) has just read.
#include <stdio.h>

#include "kernels.h"
#include "gemm.h"
#include "ptx.cuh"

namespace mossdec {

bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows);

namespace {

constexpr int kBQ = 128, kBK = 32, kD = 64, kNS = 4;
constexpr int kQBytes = kBQ * kD * 2;     // 16 KB
constexpr int kTileBytes = kBK * kD * 2;  // 4 KB: one K or V tile
constexpr int kStageBytes = 2 * kTileBytes;
constexpr int kSmem = kQBytes + kNS * kStageBytes + 1024 + 256;
constexpr int kThreadsWs = 160;

__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t smem_addr) {   // B operand MN-major (rows = keys, 64 dims contiguous)
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)(1024 >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.f;   // log2 units: P values stay below 2^8, sums in fp32

// Softmax of one [128 x NK] score tile (NK = 16 or 32), thread == query row.  `o_stable` is the barrier whose phase
// `o_phase` completes with PV(j - 1): O may only be rescaled once that MMA has finished.
// `bias` (kBias only): bf16 additive score term of this thread's query row, element kbase + i for key kbase + i (the
// relative-position term matrix_bd of the conformer encoder, already shifted to this row; engine_encoder.cu).
template <int NK, bool kBias>
__device__ __forceinline__ void softmax_tile(uint32_t t_s, uint32_t t_o, int kbase, int lim, int j, float& m2, float& l,
                                             uint64_t* o_stable, uint32_t o_phase, const unsigned short* bias) {
    float sc[NK];
    unsigned short braw[kBias ? NK : 1];
    if (kBias) {   // issued before the TMEM load so both latencies overlap
#pragma unroll
        for (int i = 0; i < NK; ++i) braw[i] = __ldg(bias + kbase + i);
    }
#pragma unroll
    for (int c = 0; c < NK / 16; ++c) tmem_ld16(t_s + c * 16, sc + c * 16);
    tmem_ld_wait();
    if (kBias) {
#pragma unroll
        for (int i = 0; i < NK; ++i) sc[i] += __uint_as_float((uint32_t)braw[i] << 16);
    }
    if (kbase + NK > lim) {
#pragma unroll
        for (int i = 0; i < NK; ++i)
            if (kbase + i >= lim) sc[i] = -INFINITY;
    }
    float mxa = fmaxf(sc[0], sc[1]), mxb = fmaxf(sc[2], sc[3]), mxc = fmaxf(sc[4], sc[5]), mxd = fmaxf(sc[6], sc[7]);
#pragma unroll
    for (int i = 8; i < NK; i += 8) {
        mxa = fmaxf(mxa, fmaxf(sc[i], sc[i + 1]));
        mxb = fmaxf(mxb, fmaxf(sc[i + 2], sc[i + 3]));
        mxc = fmaxf(mxc, fmaxf(sc[i + 4], sc[i + 5]));
        mxd = fmaxf(mxd, fmaxf(sc[i + 6], sc[i + 7]));
    }
    const float mx2 = fmaxf(fmaxf(mxa, mxb), fmaxf(mxc, mxd)) * kLog2e;
    const bool need = mx2 > m2 + kRescaleThreshold;   // also true for the first visible key (m2 == -inf)
    const float m2n = need ? mx2 : m2;
    const float factor = (m2 == -INFINITY) ? 0.f : exp2f(m2 - m2n);   // 1 when the maximum is kept
    const bool any_rescale = __any_sync(0xffffffffu, need && j > 0);
    const float nms = (m2n == -INFINITY) ? 0.f : -m2n;
    m2 = m2n;
    // exponent arguments and row sums as packed f32x2 (fma.rn.f32x2 / add.rn.f32x2): the softmax warps are bound by issue
    // slots, and this halves the FP32 instructions around every pair of MUFU.EX2
    uint64_t rs0 = 0ull, rs1 = 0ull;
    uint32_t pk[NK / 2];
    const uint64_t l2e = f2_pack(kLog2e, kLog2e), nms2 = f2_pack(nms, nms);
#pragma unroll
    for (int i = 0; i < NK; i += 4) {
        float x0, x1, x2, x3;
        f2_unpack(f2_fma(f2_pack(sc[i], sc[i + 1]), l2e, nms2), x0, x1);
        f2_unpack(f2_fma(f2_pack(sc[i + 2], sc[i + 3]), l2e, nms2), x2, x3);
        const float p0 = ex2_approx(x0), p1 = ex2_approx(x1), p2 = ex2_approx(x2), p3 = ex2_approx(x3);
        rs0 = f2_add(rs0, f2_pack(p0, p1));
        rs1 = f2_add(rs1, f2_pack(p2, p3));
        pk[i >> 1] = pack_bf16x2(p0, p1);
        pk[(i >> 1) + 1] = pack_bf16x2(p2, p3);
    }
    float rsa, rsb, rsc, rsd;
    f2_unpack(rs0, rsa, rsb);
    f2_unpack(rs1, rsc, rsd);
    l = fmaf(l, factor, (rsa + rsb) + (rsc + rsd));
    if (any_rescale) {   // warp-uniform; lanes that keep their maximum scale by 1
        mbar_wait(o_stable, o_phase);   // PV(j - 1) has finished accumulating into O
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float t[16];
            tmem_ld16(t_o + c * 16, t);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) t[i] *= factor;
            tmem_st16(t_o + c * 16, t);
        }
    }
    if (NK == 32) tmem_st16(t_s, reinterpret_cast<const float*>(pk));
    else tmem_st8(t_s, reinterpret_cast<const float*>(pk));
    tmem_st_wait();
}

// kPrefix: every sequence has `prefix_len` FROZEN keys in front of its live rows (streaming with block-causal masks: the
// first complete blocks of a stream are identical in every chunk, engine_solve.cu).  Their K | V rows come from a cache
// tensor (tmP: [rows, 2 * H * 64] = [k | v], sequence s at row prefix_row[s]); the live keys follow at absolute position
// prefix_len.  Key tiles never straddle the boundary: ceil(prefix_len / 32) prefix tiles (tail masked), then the live ones.
template <bool kBias, bool kPrefix>
__global__ void __launch_bounds__(kThreadsWs, 4)
attention_ws_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tmP,
                    __nv_bfloat16* __restrict__ out, const int32_t* __restrict__ seq_start,
                    const int32_t* __restrict__ seq_len, int heads, int chunk, const __nv_bfloat16* __restrict__ bias,
                    int64_t bias_head_stride, int bias_ld, int bias_tmax, int prefix_len,
                    const int32_t* __restrict__ prefix_row) {
    const int s = blockIdx.z, h = blockIdx.y;
    const int len = __shfl_sync(0xffffffffu, seq_len[s], 0);
    const int q0 = blockIdx.x * kBQ;
    if (q0 >= len) return;
    const int start = __shfl_sync(0xffffffffu, seq_start[s], 0);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sQ = smem;
    uint8_t* sKV = sQ + kQBytes;   // [kNS][K 4 KB | V 4 KB]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + kNS * kStageBytes);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;            // [kNS]  TMA -> MMA
    uint64_t* kv_empty = kv_full + kNS;      // [kNS]  PV(j) complete: stage free, O stable
    uint64_t* s_full = kv_empty + kNS;       // [2]    QK(j) complete
    uint64_t* p_ready = s_full + 2;          // [2]    softmax warps -> MMA: P_j stored
    uint64_t* o_full = p_ready + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int n_active = min(4, (min(len - q0, kBQ) + 31) >> 5);   // softmax warps that own at least one real query row

    if (tid == 0) {
        tma_prefetch_desc(&tm);
        mbar_init(q_full, 1);
        for (int i = 0; i < kNS; ++i) {
            mbar_init(&kv_full[i], 1);
            mbar_init(&kv_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&p_ready[i], (uint32_t)n_active);
        }
        mbar_init(o_full, 1);
        mbar_fence_init();
    }
    if (warp == 4) {
        tmem_alloc(tmem_slot, 128);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    // keys visible to this query tile (LIVE key index; absolute position = P + index)
    const int P = kPrefix ? prefix_len : 0;
    int kv_end = len;
    if (chunk > 0) kv_end = min(len, ((P + min(q0 + kBQ, len) - 1) / chunk + 1) * chunk - P);
    const int npt = (P + kBK - 1) / kBK;                 // prefix key tiles (the last one may be partly masked)
    const int nkt = npt + (kv_end + kBK - 1) / kBK;
    const int nk_last = ((kv_end - (nkt - npt - 1) * kBK) + 15) & ~15;   // 16 or 32: the last tile only computes its 16-key groups
    const int prow = kPrefix ? __shfl_sync(0xffffffffu, prefix_row[s], 0) : 0;

    if (warp == 4) {
        // ------------------------------------------------------------ TMA + MMA issue (whole warp, one elected lane issues)
        const int qcol = h * kD, kcol = (heads + h) * kD, vcol = (2 * heads + h) * kD;
        const uint32_t idesc_s = umma_idesc_bf16(kBQ, kBK);
        const uint32_t idesc_s_last = umma_idesc_bf16(kBQ, (uint32_t)nk_last);
        const uint32_t idesc_o = umma_idesc_bf16(kBQ, kD) | (1u << 16);  // B operand MN-major
        const uint64_t dq = umma_desc_sw128(smem_u32(sQ));
        auto load_kv = [&](int t) {   // tile t -> stage t % kNS
            uint8_t* st = sKV + (t % kNS) * kStageBytes;
            mbar_expect_tx(&kv_full[t % kNS], kStageBytes);
            if (kPrefix && t < npt) {   // frozen keys: cache rows, k at column h * 64, v at (H + h) * 64
                tma_load_2d(st, &tmP, h * kD, prow + t * kBK, &kv_full[t % kNS]);
                tma_load_2d(st + kTileBytes, &tmP, (heads + h) * kD, prow + t * kBK, &kv_full[t % kNS]);
            } else {
                tma_load_2d(st, &tm, kcol, start + (t - npt) * kBK, &kv_full[t % kNS]);
                tma_load_2d(st + kTileBytes, &tm, vcol, start + (t - npt) * kBK, &kv_full[t % kNS]);
            }
        };
        auto issue_qk = [&](int t) {   // S[t & 1] = Q K_t^T
            const uint64_t dk = umma_desc_sw128(smem_u32(sKV + (t % kNS) * kStageBytes));
            const uint32_t id = (t + 1 == nkt) ? idesc_s_last : idesc_s;
#pragma unroll
            for (int k = 0; k < kD / 16; ++k)
                umma_bf16(tmem + (uint32_t)((t & 1) * kBK), dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), id, (uint32_t)(k != 0));
            umma_commit(&s_full[t & 1]);
        };
        if (elect_one()) {
            mbar_expect_tx(q_full, kQBytes);
#pragma unroll
            for (int i = 0; i < kBQ / kBK; ++i) tma_load_2d(sQ + i * kTileBytes, &tm, qcol, start + q0 + i * kBK, q_full);
            for (int t = 0; t < kNS && t < nkt; ++t) load_kv(t);
        }
        mbar_wait(q_full, 0);
        mbar_wait(&kv_full[0], 0);
        tc_fence_after();
        if (elect_one()) issue_qk(0);
        if (nkt > 1) {
            mbar_wait(&kv_full[1], 0);
            tc_fence_after();
            if (elect_one()) issue_qk(1);
        }
        for (int j = 0; j < nkt; ++j) {
            const int b = j & 1, st = j % kNS;
            if (j >= 1 && j - 1 + kNS < nkt) {   // refill the stage PV(j - 1) has released (it finished long ago)
                mbar_wait(&kv_empty[(j - 1) % kNS], (uint32_t)(((j - 1) / kNS) & 1));
                if (elect_one()) load_kv(j - 1 + kNS);
            }
            mbar_wait(&p_ready[b], (uint32_t)((j >> 1) & 1));   // P_j stored over S_j (K_j / V_j landed before QK(j))
            tc_fence_after();
            const int nk = (j + 1 == nkt) ? nk_last : kBK;
            if (elect_one()) {
                const uint64_t dv = desc_mn_sw128(smem_u32(sKV + st * kStageBytes + kTileBytes));
                for (int k = 0; k < nk / 16; ++k)
                    umma_bf16_ts(tmem + 64, tmem + (uint32_t)(b * kBK + k * 8), dv + (uint64_t)(k * (2048 >> 4)), idesc_o, (uint32_t)((j | k) != 0));
                umma_commit(&kv_empty[st]);
                if (j + 1 == nkt) umma_commit(o_full);
            }
            if (j + 2 < nkt) {   // next-but-one scores into the buffer PV(j) reads: ordered behind it on the tensor pipe
                mbar_wait(&kv_full[(j + 2) % kNS], (uint32_t)(((j + 2) / kNS) & 1));
                tc_fence_after();
                if (elect_one()) issue_qk(j + 2);
            }
        }
    } else if (warp < n_active) {
        // ------------------------------------------------------------ softmax warps, thread == query row
        const int qi = q0 + tid;
        int lim = len;   // live key index limit of this query row
        if (chunk > 0) lim = min(len, ((P + min(qi, len - 1)) / chunk + 1) * chunk - P);
        const uint32_t t_lane = tmem + ((uint32_t)(warp * 32) << 16);
        const uint32_t t_o = t_lane + 64;
        float m2 = -INFINITY, l = 0.f;
        // relative-position bias row of this query: element j holds the term of key j (m = tmax - 1 - qi + j of the table)
        const unsigned short* brow = nullptr;
        if (kBias) {
            const int qc = min(qi, len - 1);   // padding lanes of the last tile read a valid row
            brow = reinterpret_cast<const unsigned short*>(bias) + (size_t)h * bias_head_stride +
                   (size_t)(start + qc) * bias_ld + (bias_tmax - 1 - qc);
        }
        for (int j = 0; j < nkt; ++j) {
            const int b = j & 1;
            mbar_wait(&s_full[b], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            uint64_t* o_stable = &kv_empty[(j + kNS - 1) % kNS];                      // completes with PV(j - 1)
            const uint32_t o_phase = (uint32_t)((((j > 0 ? j - 1 : 0)) / kNS) & 1);
            // prefix tiles: key index inside the prefix, valid below P; live tiles: live key index, valid below lim
            const int kb0 = (kPrefix && j < npt) ? j * kBK : (j - npt) * kBK;
            const int lm = (kPrefix && j < npt) ? P : lim;
            if (j + 1 == nkt && nk_last == 16) softmax_tile<16, kBias>(t_lane + (uint32_t)(b * kBK), t_o, kb0, lm, j, m2, l, o_stable, o_phase, brow);
            else softmax_tile<32, kBias>(t_lane + (uint32_t)(b * kBK), t_o, kb0, lm, j, m2, l, o_stable, o_phase, brow);
            tc_fence_before();
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&p_ready[b]);
        }
        mbar_wait(o_full, 0);
        tc_fence_after();
        const float inv = l > 0.f ? 1.f / l : 0.f;
        __nv_bfloat16* dst = out + (size_t)(start + qi) * (heads * kD) + h * kD;
#pragma unroll
        for (int c = 0; c < 4; ++c) {   // (tcgen05.ld is warp-collective: every lane takes part, only the store is predicated)
            float t[16];
            tmem_ld16(t_o + c * 16, t);
            tmem_ld_wait();
            if (qi < len) {
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    uint4 w;
                    w.x = pack_bf16x2(t[u * 8] * inv, t[u * 8 + 1] * inv);
                    w.y = pack_bf16x2(t[u * 8 + 2] * inv, t[u * 8 + 3] * inv);
                    w.z = pack_bf16x2(t[u * 8 + 4] * inv, t[u * 8 + 5] * inv);
                    w.w = pack_bf16x2(t[u * 8 + 6] * inv, t[u * 8 + 7] * inv);
                    *reinterpret_cast<uint4*>(dst + c * 16 + u * 8) = w;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 128);
}

}  // namespace

cudaError_t attention_ws_init_device() {
    static std::atomic<uint64_t> done{0};
    static std::mutex mu;
    return once_per_device(done, mu, [] {
        cudaError_t e = cudaFuncSetAttribute(attention_ws_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(attention_ws_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(attention_ws_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
        return e;
    });
}

cudaError_t launch_attention_ws(const __nv_bfloat16* qkv, int64_t rows, __nv_bfloat16* out, const int32_t* seq_start,
                                const int32_t* seq_len, int nseq, int max_len, int heads, int chunk, cudaStream_t st) {
    return launch_attention_ws_ex(qkv, rows, 3 * heads * kD, out, seq_start, seq_len, nseq, max_len, heads, chunk, nullptr, 0, 0,
                                  0, st);
}

cudaError_t launch_attention_ws_ex(const __nv_bfloat16* qkv, int64_t rows, int qkv_ld, __nv_bfloat16* out,
                                   const int32_t* seq_start, const int32_t* seq_len, int nseq, int max_len, int heads, int chunk,
                                   const __nv_bfloat16* bias, int64_t bias_head_stride, int bias_ld, int bias_tmax,
                                   cudaStream_t st) {
    const cudaError_t ie = attention_ws_init_device();
    if (ie != cudaSuccess) return ie;
    CUtensorMap tm;
    if (!encode_2d(&tm, qkv, (uint64_t)qkv_ld, (uint64_t)rows, (uint64_t)qkv_ld, kBK)) return cudaErrorInvalidValue;
    dim3 grid((max_len + kBQ - 1) / kBQ, heads, nseq);
    if (bias != nullptr)
        attention_ws_kernel<true, false><<<grid, kThreadsWs, kSmem, st>>>(tm, tm, out, seq_start, seq_len, heads, chunk, bias,
                                                                          bias_head_stride, bias_ld, bias_tmax, 0, nullptr);
    else
        attention_ws_kernel<false, false><<<grid, kThreadsWs, kSmem, st>>>(tm, tm, out, seq_start, seq_len, heads, chunk, nullptr,
                                                                           0, 0, 0, 0, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_attention_ws_prefix(const __nv_bfloat16* qkv, int64_t rows, __nv_bfloat16* out, const int32_t* seq_start,
                                       const int32_t* seq_len, int nseq, int max_len, int heads, int chunk,
                                       const __nv_bfloat16* prefix_kv, int64_t prefix_rows, int prefix_len,
                                       const int32_t* prefix_row, cudaStream_t st) {
    const cudaError_t ie = attention_ws_init_device();
    if (ie != cudaSuccess) return ie;
    if (prefix_len <= 0 || (chunk > 0 && (prefix_len % chunk) != 0)) return cudaErrorInvalidValue;
    CUtensorMap tm, tmP;
    if (!encode_2d(&tm, qkv, (uint64_t)3 * heads * kD, (uint64_t)rows, (uint64_t)3 * heads * kD, kBK)) return cudaErrorInvalidValue;
    if (!encode_2d(&tmP, prefix_kv, (uint64_t)2 * heads * kD, (uint64_t)prefix_rows, (uint64_t)2 * heads * kD, kBK))
        return cudaErrorInvalidValue;
    dim3 grid((max_len + kBQ - 1) / kBQ, heads, nseq);
    attention_ws_kernel<false, true><<<grid, kThreadsWs, kSmem, st>>>(tm, tmP, out, seq_start, seq_len, heads, chunk, nullptr, 0, 0,
                                                                      0, prefix_len, prefix_row);
    return cudaGetLastError();
}

}  // namespace mossdec
