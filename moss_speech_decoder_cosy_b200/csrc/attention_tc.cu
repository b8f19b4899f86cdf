// Self-attention of BasicTransformerBlock on the 5th-gen tensor cores (tcgen05 + TMEM), packed variable-length rows.
// Reference: matcha/models/components/transformer.py:262-274 + diffusers Attention (SDPA with an additive mask built by
// cosyvoice/flow/decoder.py:439-443 from cosyvoice/utils/mask.py:127-158).  Masked keys get an exact weight of 0 in the
// reference (bias -1e10 in fp32), so masking by predicate is equivalent; the [T, T] bias is never materialised.
//
// One CTA = 128 queries of one (sequence, head); keys/values stream through in tiles of 64:
//   S = Q K^T      tcgen05.mma  M128 N64 K64  (Q, K tiles K-major in smem via TMA, 128B swizzle)      -> TMEM cols 0..63
//   P = softmax    4 warps, thread == query row: tcgen05.ld, predicate mask, online max / exp2 / sum   -> smem (bf16, A operand)
//   O_j = P V      tcgen05.mma  M128 N64 K64  (V tile as MN-major B operand: rows = keys, 64 dims contiguous) -> TMEM cols 64..127
//   O   = O * corr + O_j in registers (fp32), normalised and stored as bf16 at the end.
// Everything inside a CTA is sequential; 3 CTAs per SM overlap each other's TMA / MMA / softmax phases.  With head_dim
// 64 the kernel is bound by the exp2 throughput of the SFUs (16 / clk / SM), not by the tensor pipe.
#include "kernels.h"
#include "ptx.cuh"

namespace mossdec {

bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows);

namespace {

constexpr int kBQ = 128, kBK = 64, kD = 64;
constexpr int kQBytes = kBQ * kD * 2;   // 16 KB
constexpr int kKVBytes = kBK * kD * 2;  // 8 KB
constexpr int kPBytes = kBQ * kBK * 2;  // 16 KB
constexpr int kSmem = kQBytes + 4 * kKVBytes + kPBytes + 1024 + 128;

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
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__global__ void __launch_bounds__(128, 3)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out,
                    const int32_t* __restrict__ seq_start, const int32_t* __restrict__ seq_len, int heads, int chunk) {
    const int s = blockIdx.z, h = blockIdx.y;
    const int len = seq_len[s];
    const int q0 = blockIdx.x * kBQ;
    if (q0 >= len) return;
    const int start = seq_start[s];

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays in the shared address space (STS / LDS)
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kQBytes;           // [2][8 KB]
    uint8_t* sV = sK + 2 * kKVBytes;      // [2][8 KB]
    uint8_t* sP = sV + 2 * kKVBytes;      // 16 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kPBytes);
    uint64_t* bar_q = bars;
    uint64_t* bar_kv = bars + 1;  // [2]
    uint64_t* bar_s = bars + 3;
    uint64_t* bar_o = bars + 4;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        tma_prefetch_desc(&tm);
        mbar_init(bar_q, 1);
        mbar_init(&bar_kv[0], 1);
        mbar_init(&bar_kv[1], 1);
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
    const uint32_t tmem = *tmem_slot;

    // keys visible to this query tile
    int kv_end = len;
    if (chunk > 0) kv_end = min(len, ((min(q0 + kBQ, len) - 1) / chunk + 1) * chunk);
    const int nkt = (kv_end + kBK - 1) / kBK;
    const int qcol = h * kD, kcol = (heads + h) * kD, vcol = (2 * heads + h) * kD;

    if (tid == 0) {
        mbar_expect_tx(bar_q, kQBytes);
        tma_load_2d(sQ, &tm, qcol, start + q0, bar_q);
        tma_load_2d(sQ + kQBytes / 2, &tm, qcol, start + q0 + 64, bar_q);
        mbar_expect_tx(&bar_kv[0], 2 * kKVBytes);
        tma_load_2d(sK, &tm, kcol, start, &bar_kv[0]);
        tma_load_2d(sV, &tm, vcol, start, &bar_kv[0]);
    }

    const int qi = q0 + tid;  // this thread's query row
    int lim = len;
    if (chunk > 0) lim = min(len, (min(qi, len - 1) / chunk + 1) * chunk);
    const uint32_t t_s = tmem + ((uint32_t)(warp * 32) << 16);
    const uint32_t t_o = t_s + 64;
    const uint32_t idesc_s = umma_idesc_bf16(kBQ, kBK);
    const uint32_t idesc_o = umma_idesc_bf16(kBQ, kD) | (1u << 16);  // B operand MN-major
    constexpr float kLog2e = 1.4426950408889634f;

    float o[kD];
#pragma unroll
    for (int i = 0; i < kD; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;

    for (int j = 0; j < nkt; ++j) {
        const int b = j & 1;
        if (tid == 0) {
            if (j + 1 < nkt) {  // prefetch the next K/V tile (its buffer was released by the PV MMA of tile j-1)
                mbar_expect_tx(&bar_kv[b ^ 1], 2 * kKVBytes);
                tma_load_2d(sK + (b ^ 1) * kKVBytes, &tm, kcol, start + (j + 1) * kBK, &bar_kv[b ^ 1]);
                tma_load_2d(sV + (b ^ 1) * kKVBytes, &tm, vcol, start + (j + 1) * kBK, &bar_kv[b ^ 1]);
            }
            if (j == 0) mbar_wait(bar_q, 0);
            mbar_wait(&bar_kv[b], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            const uint64_t dq = umma_desc_sw128(smem_u32(sQ));
            const uint64_t dk = umma_desc_sw128(smem_u32(sK + b * kKVBytes));
#pragma unroll
            for (int k = 0; k < kD / 16; ++k) umma_bf16(tmem, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc_s, (uint32_t)(k != 0));
            umma_commit(bar_s);
        }
        mbar_wait(bar_s, (uint32_t)(j & 1));
        tc_fence_after();

        // ---- scores of this row: 64 keys
        float sc[kBK];
        tmem_ld16(t_s, sc);
        tmem_ld16(t_s + 16, sc + 16);
        tmem_ld16(t_s + 32, sc + 32);
        tmem_ld16(t_s + 48, sc + 48);
        tmem_ld_wait();
        const int kbase = j * kBK;
        float mx = m;
        if (kbase + kBK <= lim) {
#pragma unroll
            for (int i = 0; i < kBK; ++i) mx = fmaxf(mx, sc[i]);
        } else {
#pragma unroll
            for (int i = 0; i < kBK; ++i) {
                if (kbase + i >= lim) sc[i] = -INFINITY;
                mx = fmaxf(mx, sc[i]);
            }
        }
        const float ms = mx == -INFINITY ? 0.f : mx;
        const float corr = exp2f((m - ms) * kLog2e);
        m = mx;
        const float nms = -ms * kLog2e;
        float rs = 0.f;
        uint8_t* prow = sP + (tid >> 3) * 1024 + (tid & 7) * 128;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            float p[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                p[i] = exp2f(fmaf(sc[u * 8 + i], kLog2e, nms));
                rs += p[i];
            }
            uint4 t;
            t.x = pack_bf16x2(p[0], p[1]);
            t.y = pack_bf16x2(p[2], p[3]);
            t.z = pack_bf16x2(p[4], p[5]);
            t.w = pack_bf16x2(p[6], p[7]);
            *reinterpret_cast<uint4*>(prow + ((u ^ (tid & 7)) << 4)) = t;
        }
        l = l * corr + rs;
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint64_t dp = umma_desc_sw128(smem_u32(sP));
            const uint64_t dv = umma_desc_mn_sw128(smem_u32(sV + b * kKVBytes));
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k)
                umma_bf16(tmem + 64, dp + (uint64_t)(2 * k), dv + (uint64_t)(k * (2048 >> 4)), idesc_o, (uint32_t)(k != 0));
            umma_commit(bar_o);
        }
        mbar_wait(bar_o, (uint32_t)(j & 1));
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float t[16];
            tmem_ld16(t_o + c * 16, t);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[c * 16 + i] = fmaf(o[c * 16 + i], corr, t[i]);
        }
    }

    if (qi < len) {
        const float inv = l > 0.f ? 1.f / l : 0.f;
        __nv_bfloat16* dst = out + (size_t)(start + qi) * (heads * kD) + h * kD;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            uint4 t;
            t.x = pack_bf16x2(o[u * 8] * inv, o[u * 8 + 1] * inv);
            t.y = pack_bf16x2(o[u * 8 + 2] * inv, o[u * 8 + 3] * inv);
            t.z = pack_bf16x2(o[u * 8 + 4] * inv, o[u * 8 + 5] * inv);
            t.w = pack_bf16x2(o[u * 8 + 6] * inv, o[u * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(dst + u * 8) = t;
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
    attention_tc_kernel<<<grid, 128, kSmem, st>>>(tm, out, seq_start, seq_len, heads, chunk);
    return cudaGetLastError();
}

}  // namespace mossdec
