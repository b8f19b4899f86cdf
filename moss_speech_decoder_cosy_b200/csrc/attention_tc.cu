// Self-attention of BasicTransformerBlock on the 5th-gen tensor cores (tcgen05 + TMEM), packed variable-length rows.
// Reference: matcha/models/components/transformer.py:262-274 + diffusers Attention (SDPA with an additive mask built by
// cosyvoice/flow/decoder.py:439-443 from cosyvoice/utils/mask.py:127-158).  Masked keys get an exact weight of 0 in the
// reference (bias -1e10 in fp32), so masking by predicate is equivalent; the [T, T] bias is never materialised.
//
// One CTA = 128 queries of one (sequence, head); keys/values stream through in tiles of 64:
//   S = Q K^T      tcgen05.mma  M128 N64 K64  (Q, K tiles K-major in smem via TMA, 128B swizzle)      -> TMEM cols 0..63
//   P = softmax    4 warps, thread == query row: tcgen05.ld, predicate mask, online max / exp2 / sum   -> tcgen05.st over S (bf16, A operand from TMEM)
//   O  += P V      tcgen05.mma  M128 N64 K64  (V tile as MN-major B operand: rows = keys, 64 dims contiguous) -> TMEM cols 64..127,
//                  accumulated in TMEM over all key tiles; rescaled (tcgen05.ld / .st) only when a row maximum grows by more
//                  than 2^8 (lazy rescaling), normalised and stored as bf16 at the end.
// Everything inside a CTA is sequential; 4 CTAs per SM overlap each other's TMA / MMA / softmax phases.  With head_dim
// 64 the kernel is bound by the exp2 throughput of the SFUs (16 / clk / SM), not by the tensor pipe.
#include <stdio.h>

#include "kernels.h"
#include "gemm.h"
#include "ptx.cuh"

namespace mossdec {

bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows);

namespace {

constexpr int kBQ = 128, kBK = 64, kD = 64;
constexpr int kQBytes = kBQ * kD * 2;   // 16 KB
constexpr int kKVBytes = kBK * kD * 2;  // 8 KB
constexpr int kSmem = kQBytes + 4 * kKVBytes + 1024 + 128;

// B operand stored MN-major (rows = K index, 64 N-elements = 128 bytes contiguous per row), 128B swizzle:
// 8 K-rows form one 1024-byte swizzle atom, consecutive atoms along K are SBO = 1024 bytes apart.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)(1024 >> 4) << 16;  // LBO: stride between 64-element blocks along N (single block here)
    d |= (uint64_t)(1024 >> 4) << 32;  // SBO: stride between 8-row groups along K
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// experiment switches (scripts/build_variant.sh); none is defined in the shipped build
#ifdef ATT_X_NOEXP
#define ATT_EX2(x) (x)
#else
#define ATT_EX2(x) ex2_approx(x)
#endif
#ifdef ATT_X_NOPACK
#define ATT_PACK(a, b) (__float_as_uint(a) ^ __float_as_uint(b))
#else
#define ATT_PACK(a, b) pack_bf16x2(a, b)
#endif
#ifdef ATT_TRACE   // clock64 timeline of one CTA's warp 0 over three key tiles, printed by the launcher (experiments only)
__device__ long long g_att_trace[64];
#define ATT_STAMP(k) do { if (trace_on && j >= 4 && j < 7) g_att_trace[(j - 4) * 12 + (k)] = clock64(); } while (0)
#else
#define ATT_STAMP(k) do { } while (0)
#endif
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.f;   // log2 units: P values stay below 2^8, sums in fp32

// Softmax of one [128 x NK] score tile, thread == query row.  Reads S from TMEM columns [0, NK), writes P (bf16, two keys
// per 32-bit column) over its first NK / 2 columns -- the A operand of the PV MMA, no shared-memory round trip -- and
// rescales O (TMEM columns 64..127) when some row of the warp moved its maximum.
template <int NK>
__device__ __forceinline__ void softmax_tile(uint32_t t_s, int kbase, int lim, int j, float& m2, float& l, bool trace_on) {
    float sc[NK];
#pragma unroll
#ifdef ATT_X_NOLD
    for (int i = 0; i < NK; ++i) sc[i] = l + (float)(i & 7);
#else
    for (int c = 0; c < NK / 16; ++c) tmem_ld16(t_s + c * 16, sc + c * 16);
    tmem_ld_wait();
#endif
    ATT_STAMP(1);
    if (kbase + NK > lim) {
#pragma unroll
        for (int i = 0; i < NK; ++i)
            if (kbase + i >= lim) sc[i] = -INFINITY;
    }
    // four independent chains (the serial 64-deep max / sum chains were a visible part of the per-tile latency)
    float mxa = fmaxf(sc[0], sc[1]), mxb = fmaxf(sc[2], sc[3]), mxc = fmaxf(sc[4], sc[5]), mxd = fmaxf(sc[6], sc[7]);
#pragma unroll
    for (int i = 8; i < NK; i += 8) {
        mxa = fmaxf(mxa, fmaxf(sc[i], sc[i + 1]));
        mxb = fmaxf(mxb, fmaxf(sc[i + 2], sc[i + 3]));
        mxc = fmaxf(mxc, fmaxf(sc[i + 4], sc[i + 5]));
        mxd = fmaxf(mxd, fmaxf(sc[i + 6], sc[i + 7]));
    }
    const float mx = fmaxf(fmaxf(mxa, mxb), fmaxf(mxc, mxd));
    const float mx2 = mx * kLog2e;
    const bool need = mx2 > m2 + kRescaleThreshold;   // also true for the first visible key (m2 == -inf)
    const float m2n = need ? mx2 : m2;
    const float factor = (m2 == -INFINITY) ? 0.f : exp2f(m2 - m2n);   // 1 when the maximum is kept
    const bool any_rescale = __any_sync(0xffffffffu, need && j > 0);
    const float nms = (m2n == -INFINITY) ? 0.f : -m2n;
    m2 = m2n;
    float rsa = 0.f, rsb = 0.f, rsc = 0.f, rsd = 0.f;
    uint32_t pk[NK / 2];
#pragma unroll
    for (int i = 0; i < NK; i += 8) {
        const float p0 = ATT_EX2(fmaf(sc[i], kLog2e, nms)), p1 = ATT_EX2(fmaf(sc[i + 1], kLog2e, nms));
        const float p2 = ATT_EX2(fmaf(sc[i + 2], kLog2e, nms)), p3 = ATT_EX2(fmaf(sc[i + 3], kLog2e, nms));
        const float p4 = ATT_EX2(fmaf(sc[i + 4], kLog2e, nms)), p5 = ATT_EX2(fmaf(sc[i + 5], kLog2e, nms));
        const float p6 = ATT_EX2(fmaf(sc[i + 6], kLog2e, nms)), p7 = ATT_EX2(fmaf(sc[i + 7], kLog2e, nms));
        rsa += p0 + p1; rsb += p2 + p3; rsc += p4 + p5; rsd += p6 + p7;
        pk[i >> 1] = ATT_PACK(p0, p1);
        pk[(i >> 1) + 1] = ATT_PACK(p2, p3);
        pk[(i >> 1) + 2] = ATT_PACK(p4, p5);
        pk[(i >> 1) + 3] = ATT_PACK(p6, p7);
    }
    l = fmaf(l, factor, (rsa + rsb) + (rsc + rsd));
    ATT_STAMP(2);
    if (any_rescale) {   // warp-uniform: tcgen05.ld / .st are warp-collective; lanes that keep their maximum scale by 1
        const uint32_t t_o = t_s + 64;
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
#pragma unroll
    for (int c = 0; c < NK / 32; ++c) tmem_st16(t_s + c * 16, reinterpret_cast<const float*>(pk + c * 16));
    if (NK % 32) tmem_st8(t_s + (NK / 32) * 16, reinterpret_cast<const float*>(pk + (NK / 32) * 16));
    tmem_st_wait();
    ATT_STAMP(3);
}

__global__ void __launch_bounds__(128, 4)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out,
                    const int32_t* __restrict__ seq_start, const int32_t* __restrict__ seq_len, int heads, int chunk) {
    const int s = blockIdx.z, h = blockIdx.y;
    // (shuffle broadcasts tell the compiler these are warp-uniform: loop counters and MMA descriptors stay in uniform registers)
    const int len = __shfl_sync(0xffffffffu, seq_len[s], 0);
    const int q0 = blockIdx.x * kBQ;
    if (q0 >= len) return;
    const int start = __shfl_sync(0xffffffffu, seq_start[s], 0);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays in the shared address space (STS / LDS)
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kQBytes;           // [2][8 KB]
    uint8_t* sV = sK + 2 * kKVBytes;      // [2][8 KB]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * kKVBytes);
    uint64_t* bar_q = bars;
    uint64_t* bar_k = bars + 1;  // [2]
    uint64_t* bar_v = bars + 3;  // [2]
    uint64_t* bar_s = bars + 5;
    uint64_t* bar_o = bars + 6;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 7);

    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    if (tid == 0) {
        tma_prefetch_desc(&tm);
        mbar_init(bar_q, 1);
        mbar_init(&bar_k[0], 1);
        mbar_init(&bar_k[1], 1);
        mbar_init(&bar_v[0], 1);
        mbar_init(&bar_v[1], 1);
        mbar_init(bar_s, 1);
        mbar_init(bar_o, 1);
        mbar_fence_init();
    }
    if (warp == 0) {
        tmem_alloc(tmem_slot, 128);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_wait();     // q | k | v of the preceding GEMM are visible from here on
    pdl_launch();

    // keys visible to this query tile
    int kv_end = len;
    if (chunk > 0) kv_end = min(len, ((min(q0 + kBQ, len) - 1) / chunk + 1) * chunk);
    const int nkt = (kv_end + kBK - 1) / kBK;
    const int qcol = h * kD, kcol = (heads + h) * kD, vcol = (2 * heads + h) * kD;
    // the last key tile only computes the 16-key groups that hold visible keys (T = 648: 8 of 64)
#ifdef ATT_NO_SHORT_TAIL
    const int nk_last = kBK;
#else
    const int nk_last = ((kv_end - (nkt - 1) * kBK) + 15) & ~15;
#endif
    const uint32_t idesc_s = umma_idesc_bf16(kBQ, kBK);
    const uint32_t idesc_s_last = umma_idesc_bf16(kBQ, (uint32_t)nk_last);
    const uint32_t idesc_o = umma_idesc_bf16(kBQ, kD) | (1u << 16);  // B operand MN-major

    if (warp == 0 && elect_one()) {
        mbar_expect_tx(bar_q, kQBytes);
        tma_load_2d(sQ, &tm, qcol, start + q0, bar_q);
        tma_load_2d(sQ + kQBytes / 2, &tm, qcol, start + q0 + 64, bar_q);
        mbar_expect_tx(&bar_k[0], kKVBytes);
        tma_load_2d(sK, &tm, kcol, start, &bar_k[0]);
        mbar_expect_tx(&bar_v[0], kKVBytes);
        tma_load_2d(sV, &tm, vcol, start, &bar_v[0]);
        if (nkt > 1) {
            mbar_expect_tx(&bar_k[1], kKVBytes);
            tma_load_2d(sK + kKVBytes, &tm, kcol, start + kBK, &bar_k[1]);
        }
        mbar_wait(bar_q, 0);
        mbar_wait(&bar_k[0], 0);
        tc_fence_after();
        const uint64_t dq = umma_desc_sw128(smem_u32(sQ));
        const uint64_t dk = umma_desc_sw128(smem_u32(sK));
#pragma unroll
        for (int k = 0; k < kD / 16; ++k) umma_bf16(tmem, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), nkt == 1 ? idesc_s_last : idesc_s, (uint32_t)(k != 0));
        umma_commit(bar_s);
    }

    const int qi = q0 + tid;  // this thread's query row
    int lim = len;
    if (chunk > 0) lim = min(len, (min(qi, len - 1) / chunk + 1) * chunk);
    const uint32_t t_s = tmem + ((uint32_t)(warp * 32) << 16);
    const uint32_t t_o = t_s + 64;

    // Running maximum in log2 units.  O accumulates in TMEM across all key tiles (tcgen05.mma accumulate); it is only
    // rescaled when some row's maximum grows by more than the threshold (lazy rescaling: exact, because numerator and
    // denominator use the same stale maximum).
    float m2 = -INFINITY, l = 0.f;
    // A warp whose 32 query rows all lie past the end of the sequence (the last query tile of T = 648 has 8 rows) only
    // takes part in the block barriers: its SFU / issue slots go to the other CTAs resident on the SM.
#ifdef ATT_TRACE
    const bool trace_on = tid == 0 && blockIdx.x == 2 && blockIdx.y == 3 && blockIdx.z == gridDim.z / 2;
#else
    const bool trace_on = false;
#endif
#ifdef ATT_NO_IDLE_SKIP
    const bool active = true;
#else
    const bool active = q0 + warp * 32 < len;
#endif

    const uint64_t dq = umma_desc_sw128(smem_u32(sQ));
    const uint64_t dk0 = umma_desc_sw128(smem_u32(sK)), dk1 = umma_desc_sw128(smem_u32(sK + kKVBytes));
    const uint64_t dv0 = umma_desc_mn_sw128(smem_u32(sV)), dv1 = umma_desc_mn_sw128(smem_u32(sV + kKVBytes));

    for (int j = 0; j < nkt; ++j) {
        const int b = j & 1;
        const int kbase = j * kBK;
        const int nk = (j + 1 == nkt) ? nk_last : kBK;   // keys of this tile that take part (multiple of 16)
        if (active) {
            mbar_wait(bar_s, (uint32_t)(j & 1));
            tc_fence_after();
            ATT_STAMP(0);
            // bar_s(j) implies that every earlier MMA, PV(j - 1) included, is complete (the tensor pipe runs in order and
            // a commit covers all prior MMAs): V buffer b^1 (tile j - 1) and K buffer b (tile j) are free, O is stable.
            if (warp == 0 && elect_one()) {
                if (j + 1 < nkt) {
                    mbar_expect_tx(&bar_v[b ^ 1], kKVBytes);
                    tma_load_2d(sV + (b ^ 1) * kKVBytes, &tm, vcol, start + (j + 1) * kBK, &bar_v[b ^ 1]);
                }
                if (j + 2 < nkt) {
                    mbar_expect_tx(&bar_k[b], kKVBytes);
                    tma_load_2d(sK + b * kKVBytes, &tm, kcol, start + (j + 2) * kBK, &bar_k[b]);
                }
            }
            if (nk == 64) softmax_tile<64>(t_s, kbase, lim, j, m2, l, trace_on);
            else if (nk == 16) softmax_tile<16>(t_s, kbase, lim, j, m2, l, trace_on);
            else if (nk == 32) softmax_tile<32>(t_s, kbase, lim, j, m2, l, trace_on);
            else softmax_tile<48>(t_s, kbase, lim, j, m2, l, trace_on);
        }
        tc_fence_before();
        __syncthreads();
        ATT_STAMP(4);
        if (warp == 0) {   // warp-uniform branch, one elected lane issues; descriptors were built before the loop
            const bool more = j + 1 < nkt;
            tc_fence_after();
            mbar_wait(&bar_v[b], (uint32_t)((j >> 1) & 1));
            if (more) mbar_wait(&bar_k[b ^ 1], (uint32_t)(((j + 1) >> 1) & 1));
            ATT_STAMP(7);
            if (elect_one()) {
                const uint64_t dv = b ? dv1 : dv0;
#ifndef ATT_X_NOPV
                if (nk == kBK) {
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k)
                        umma_bf16_ts(tmem + 64, tmem + (uint32_t)(k * 8), dv + (uint64_t)(k * (2048 >> 4)), idesc_o, (uint32_t)((j | k) != 0));
                } else {
                    for (int k = 0; k < nk / 16; ++k)
                        umma_bf16_ts(tmem + 64, tmem + (uint32_t)(k * 8), dv + (uint64_t)(k * (2048 >> 4)), idesc_o, (uint32_t)((j | k) != 0));
                }
#endif
                if (more) {   // next scores right behind it on the tensor pipe (every thread has read S_j)
                    const uint64_t dk = b ? dk0 : dk1;
                    const uint32_t id = (j + 2 == nkt) ? idesc_s_last : idesc_s;
#pragma unroll
                    for (int k = 0; k < kD / 16; ++k) umma_bf16(tmem, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), id, (uint32_t)(k != 0));
                    umma_commit(bar_s);
                } else {
                    umma_commit(bar_o);
                }
            }
            ATT_STAMP(6);
        }
    }

    if (active) {
        mbar_wait(bar_o, 0);
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
    if (warp == 0) tmem_dealloc(tmem, 128);
}

}  // namespace

cudaError_t launch_attention_tc(const __nv_bfloat16* qkv, int64_t rows, __nv_bfloat16* out, const int32_t* seq_start,
                                const int32_t* seq_len, int nseq, int max_len, int heads, int chunk, cudaStream_t st) {
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    CUtensorMap tm;
    if (!encode_2d(&tm, qkv, (uint64_t)3 * heads * kD, (uint64_t)rows, (uint64_t)3 * heads * kD, 64)) return cudaErrorInvalidValue;
    dim3 grid((max_len + kBQ - 1) / kBQ, heads, nseq);
    const cudaError_t le = launch_pdl(attention_tc_kernel, grid, dim3(128), kSmem, st, tm, out, seq_start, seq_len, heads, chunk);
    if (le != cudaSuccess) return le;
#ifdef ATT_TRACE
    static int n_launch = 0;
    if (++n_launch == 300) {
        long long tr[64];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(tr, g_att_trace, sizeof(tr));
        for (int t = 0; t < 3; ++t) {
            const long long* r = tr + t * 12;
            fprintf(stderr, "[att trace] tile %d: S ready +0 | ld done %lld | exp done %lld | P stored %lld | barrier %lld | K,V ready %lld | MMAs issued %lld | next S ready %lld\n",
                    t + 4, r[1] - r[0], r[2] - r[0], r[3] - r[0], r[4] - r[0], r[7] - r[0], r[6] - r[0],
                    t < 2 ? r[12] - r[0] : 0LL);
        }
    }
#endif
    return cudaGetLastError();
}

}  // namespace mossdec
