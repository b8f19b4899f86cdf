// Shared pieces of the tcgen05 kernels' epilogues (gemm_tc.cu, ffn_tc.cu): activation math, the per-warp swizzled
// staging tiles that turn thread==row register tiles into coalesced global accesses, and the row-group reductions.
#pragma once
#include "gemm.h"
#include "ptx.cuh"

namespace mossdec {
namespace {

// warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM alloc), warps 2..17 = epilogue.
// A warp may only touch TMEM lanes [32 * (warp % 4), +32), so the 16 epilogue warps form 4 groups of 4: the warps of
// a group own the same 32 rows of the tile and split its columns in four parts.  (The epilogue is ALU/latency bound
// at one row per thread; 4 resident warps per scheduler hide that latency, 2 did not -- profiles/r01_*.)
constexpr int kEpiWarps = 16;
constexpr int kEpiParts = kEpiWarps / 4;
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kABytes = kGemmBM * kGemmBK * 2;  // 16 KB
constexpr int kBBytesMax = 256 * kGemmBK * 2;   // 32 KB
constexpr int kStageBytes = 2048;               // per-epilogue-warp staging tile: 32 rows x 64 B
constexpr int kSmemBytes = kGemmStages * (kABytes + kBBytesMax) + kEpiWarps * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t kTmemCols = 512;

struct KParams {
    int32_t M, N, BN;
    int32_t tiles_n, num_tiles, total_kb, num_segs;
    int32_t group, num_units, a_stationary;   // work unit = `group` N tiles of one M tile; A fetched once per unit (gemm_tc.cu producer)
    int32_t b_stationary, tiles_m;            // 1: every CTA owns ONE N tile, its B tile is fetched once and stays in the ring; M tiles stream (gemm_tc.cu)
    int32_t fast_bf16;   // epilogue = bias + act1 + mask -> bf16 through TMA stores (gemm_tc.cu epi_fast)
    GemmSeg segs[kGemmMaxSegs];
    GemmEpilogue epi;
};

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ float fast_mish(float x) {
    // x * tanh(softplus(x)) with u = e^x: tanh(log(1 + u)) = (u^2 + 2u) / (u^2 + 2u + 2)
    // (ex2.approx / rcp.approx directly: __expf / __fdividef carry per-element range fix-up branches that serialise the
    //  unrolled epilogue -- measured 41 us of a 74 us resnet conv)
    const float u = ex2_approx(fminf(x, 20.f) * 1.4426950408889634f);
    const float n = u * (u + 2.f);
    return x > 20.f ? x : x * (n * rcp_approx(n + 2.f));
}
__device__ __forceinline__ float fast_gelu_erf(float x) {
    // exact-erf GELU with one SFU op:  x Phi(x) = max(x, 0) - |x| * 0.5 erfc(|x| / sqrt 2)  and
    // 0.5 erfc(a / sqrt 2) = 2^P(a) with P the weighted-minimax degree-6 fit of log2(erfc(a / sqrt 2)) - 1 on [0, 6]
    // (max |error| of the GELU value 4.8e-7 over all x, i.e. fp32 rounding level; fit: scripts/fit_gelu.py)
    const float a = fminf(fabsf(x), 6.f);
    float p = fmaf(3.291721077e-05f, a, -7.622798439e-04f);
    p = fmaf(p, a, 8.039220236e-03f);
    p = fmaf(p, a, -5.331587791e-02f);
    p = fmaf(p, a, -4.588711560e-01f);
    p = fmaf(p, a, -1.151156902e+00f);
    p = fmaf(p, a, -9.999995828e-01f);
    return fmaf(-a, ex2_approx(p), fmaxf(x, 0.f));
}
#ifdef FFN_X_NOEX2   // timing experiment only: the GELU without its SFU instruction
#define GELU_EX2(x) (x)
#else
#define GELU_EX2(x) ex2_approx(x)
#endif
// Two GELUs at once, packed to bf16x2: Horner and the final FMA run as FFMA2 (half the FMA-pipe instructions).  The result
// is rounded to bf16, so the fit is degree 4 here (max |error| of the GELU value 1.5e-5 = 13 % of half a bf16 ulp; the
// degree-6 form above is kept for fp32 outputs): two fewer instructions per pair in the issue-bound GELU epilogue.
__device__ __forceinline__ uint32_t gelu2_bf16(float x0, float x1) {
    const float a0 = fminf(fabsf(x0), 6.f), a1 = fminf(fabsf(x1), 6.f);
    const uint64_t a = f2_pack(a0, a1);
    uint64_t p = f2_fma(f2_pack(2.972227521e-03f, 2.972227521e-03f), a, f2_pack(-3.981290013e-02f, -3.981290013e-02f));
    p = f2_fma(p, a, f2_pack(-4.745878577e-01f, -4.745878577e-01f));
    p = f2_fma(p, a, f2_pack(-1.143665910e+00f, -1.143665910e+00f));
    p = f2_fma(p, a, f2_pack(-1.001076221e+00f, -1.001076221e+00f));
    float p0, p1;
    f2_unpack(p, p0, p1);
    const uint64_t r = f2_fma(f2_pack(-a0, -a1), f2_pack(GELU_EX2(p0), GELU_EX2(p1)), f2_pack(fmaxf(x0, 0.f), fmaxf(x1, 0.f)));
    float r0, r1;
    f2_unpack(r, r0, r1);
    return pack_bf16x2(r0, r1);
}
__device__ __forceinline__ float fast_sin(float x) {
    // two-term Cody-Waite reduction to [-pi, pi], then the SFU sine (abs error ~1e-6 for |x| < 1e4)
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(k, -6.2831854820251465f, x);
    r = fmaf(k, 1.7484555e-7f, r);
    return __sinf(r);
}

// activation over a register vector (the switch is evaluated once per vector, not per element)
template <int NV>
__device__ __forceinline__ void apply_act_vec(float* v, int act, float p, const float* alpha) {
    switch (act) {
        case ACT_MISH:
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = fast_mish(v[i]);
            break;
        case ACT_GELU:
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = fast_gelu_erf(v[i]);
            break;
        case ACT_ELU:
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = v[i] > 0.f ? v[i] : expm1f(v[i]);
            break;
        case ACT_LRELU:
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * p;
            break;
        case ACT_SNAKE:
#pragma unroll
            for (int i = 0; i < NV; i += 4) {
                const float4 a = ldg4(alpha + i);
                const float aa[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float s = fast_sin(v[i + j] * aa[j]);
                    v[i + j] = fmaf(rcp_approx(aa[j] + 1e-9f), s * s, v[i + j]);
                }
            }
            break;
        case ACT_SILU:
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = silu_f(v[i]);
            break;
        case ACT_ABS:
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = fabsf(v[i]);
            break;
        default: break;
    }
}

// ---- per-warp staging tile: 32 rows x RB bytes (RB = 128 or 64) held as 128-byte lines with the 16-byte units of a
// line XOR-swizzled by the line index, so both the thread==row side and the coalesced side are bank-conflict free.
template <int RB>
__device__ __forceinline__ uint32_t stage_off(int r, int u) {  // logical (row, 16-byte unit) -> byte offset
    if (RB == 128) return (uint32_t)(r * 128 + ((u ^ (r & 7)) << 4));
    const int line = r >> 1;
    return (uint32_t)(line * 128 + (((((r & 1) << 2) | u) ^ (line & 7)) << 4));
}
// coalesced side: pass k (0 .. RB/16 - 1), lane -> logical (row, unit); the 32 lanes of a pass cover 4 whole lines
template <int RB>
__device__ __forceinline__ void stage_coalesced(int k, int lane, int& r, int& u) {
    const int line = k * 4 + (lane >> 3);
    if (RB == 128) {
        r = line;
        u = (lane & 7) ^ (line & 7);
    } else {
        const int pos = (lane & 7) ^ (line & 7);
        r = line * 2 + (pos >> 2);
        u = pos & 3;
    }
}

struct EpiCtx {
    const KParams* p;
    uint8_t* stg;       // this warp's staging tile
    uint8_t* stg_all;   // staging tiles of all epilogue warps
    int ew;             // epilogue warp index
    uint32_t taddr;     // TMEM address of (lane quarter, accumulator)
    int lane, quarter, row0, n0, seq;
    bool valid;         // this thread's row exists and belongs to a sequence
};

__device__ __forceinline__ void group_sync(int quarter) {
    asm volatile("bar.sync %0, %1;" ::"r"(quarter + 1), "n"(32 * kEpiParts) : "memory");
}
// sum one float2 per row over the warps of a group (row statistics over the full tile width)
__device__ __forceinline__ float2 group_exchange(const EpiCtx& c, float2 mine) {
    reinterpret_cast<float2*>(c.stg)[c.lane] = mine;
    group_sync(c.quarter);
    float2 t = make_float2(0.f, 0.f);
#pragma unroll
    for (int part = 0; part < kEpiParts; ++part) {
        const float2 o = reinterpret_cast<const float2*>(c.stg_all + ((part << 2) | (c.ew & 3)) * kStageBytes)[c.lane];
        t.x += o.x;
        t.y += o.y;
    }
    group_sync(c.quarter);
    return t;
}

// LayerNorm statistics of a row whose columns are split over the kEpiParts warps of a group, without the
// E[x^2] - mean^2 cancellation: every warp contributes the mean and the centred sum of squares (M2) of ITS columns,
// the parts are merged with the parallel-variance formula  M2 = sum M2_p + n_p * sum (mean_p - mean)^2  (equal counts).
// Returns (mean, rstd).  A residual stream with a DC offset of many standard deviations keeps full fp32 accuracy.
__device__ __forceinline__ float2 group_mean_rstd(const EpiCtx& c, float mean_p, float m2_p, float n_p, float eps) {
    reinterpret_cast<float2*>(c.stg)[c.lane] = make_float2(mean_p, m2_p);
    group_sync(c.quarter);
    float2 o[kEpiParts];
    float mean = 0.f;
#pragma unroll
    for (int part = 0; part < kEpiParts; ++part) {
        o[part] = reinterpret_cast<const float2*>(c.stg_all + ((part << 2) | (c.ew & 3)) * kStageBytes)[c.lane];
        mean += o[part].x;
    }
    group_sync(c.quarter);
    mean *= (1.f / kEpiParts);
    float m2 = 0.f;
#pragma unroll
    for (int part = 0; part < kEpiParts; ++part) {
        const float d = o[part].x - mean;
        m2 += o[part].y + n_p * d * d;
    }
    return make_float2(mean, rsqrtf(m2 / (n_p * kEpiParts) + eps));
}
// mean and centred sum of squares of NV register values (two passes over registers)
template <int NV>
__device__ __forceinline__ float2 local_mean_m2(const float* v) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
        s0 += v[i]; s1 += v[i + 1]; s2 += v[i + 2]; s3 += v[i + 3];
    }
    const float mean = ((s0 + s1) + (s2 + s3)) * (1.f / NV);
    float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
        const float d0 = v[i] - mean, d1 = v[i + 1] - mean, d2 = v[i + 2] - mean, d3 = v[i + 3] - mean;
        q0 = fmaf(d0, d0, q0); q1 = fmaf(d1, d1, q1); q2 = fmaf(d2, d2, q2); q3 = fmaf(d3, d3, q3);
    }
    return make_float2(mean, (q0 + q1) + (q2 + q3));
}
// the same from shifted running sums  s = sum (x - K), ss = sum (x - K)^2  over n values (K = any value of the row)
__device__ __forceinline__ float2 shifted_mean_m2(float K, float s, float ss, float n) {
    return make_float2(K + s / n, fmaxf(ss - s * s / n, 0.f));
}

// coalesced global -> registers (thread == row) through the staging tile; fp32, 32 columns starting at column nb
// (two rounds of 16 columns = 64-byte rows)
__device__ __forceinline__ void load_rows_f32(const EpiCtx& c, const float* base, int ld, int nb, int nlim, float* out32) {
    const KParams& p = *c.p;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int r, u;
            stage_coalesced<64>(k, c.lane, r, u);
            const int grow = c.row0 + r, col = nb + h * 16 + u * 4;
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (grow < p.M && col < nlim) t = *reinterpret_cast<const float4*>(base + (size_t)grow * ld + col);
            *reinterpret_cast<float4*>(c.stg + k * 512 + c.lane * 16) = t;
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float4 t = *reinterpret_cast<const float4*>(c.stg + stage_off<64>(c.lane, u));
            float* o = out32 + h * 16 + u * 4;
            o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
        }
        __syncwarp();
    }
}
// tiled fp32 layout (gemm.h tiled_f32_offset): direct, coalesced thread == row access of 32 columns starting at nb
__device__ __forceinline__ void add_rows_f32_tiled(const float* base, int row, int nb, float* acc32) {
    const float* q = base + tiled_f32_offset(row, nb);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const float4 t = *reinterpret_cast<const float4*>(q + (size_t)u * 512);
        acc32[u * 4] += t.x; acc32[u * 4 + 1] += t.y; acc32[u * 4 + 2] += t.z; acc32[u * 4 + 3] += t.w;
    }
}
__device__ __forceinline__ void load_rows_f32_tiled(const float* base, int row, int nb, float* out32) {
    const float* q = base + tiled_f32_offset(row, nb);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const float4 t = *reinterpret_cast<const float4*>(q + (size_t)u * 512);
        out32[u * 4] = t.x; out32[u * 4 + 1] = t.y; out32[u * 4 + 2] = t.z; out32[u * 4 + 3] = t.w;
    }
}
__device__ __forceinline__ void store_rows_f32_tiled(float* base, int row, int nb, const float* v32) {
    float* q = base + tiled_f32_offset(row, nb);
#pragma unroll
    for (int u = 0; u < 8; ++u)
        *reinterpret_cast<float4*>(q + (size_t)u * 512) = make_float4(v32[u * 4], v32[u * 4 + 1], v32[u * 4 + 2], v32[u * 4 + 3]);
}
// same access pattern, but accumulates into acc32 (keeps the register footprint at the accumulator itself)
__device__ __forceinline__ void add_rows_f32(const EpiCtx& c, const float* base, int ld, int nb, int nlim, float* acc32) {
    const KParams& p = *c.p;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int r, u;
            stage_coalesced<64>(k, c.lane, r, u);
            const int grow = c.row0 + r, col = nb + h * 16 + u * 4;
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (grow < p.M && col < nlim) t = *reinterpret_cast<const float4*>(base + (size_t)grow * ld + col);
            *reinterpret_cast<float4*>(c.stg + k * 512 + c.lane * 16) = t;
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float4 t = *reinterpret_cast<const float4*>(c.stg + stage_off<64>(c.lane, u));
            float* o = acc32 + h * 16 + u * 4;
            o[0] += t.x; o[1] += t.y; o[2] += t.z; o[3] += t.w;
        }
        __syncwarp();
    }
}
__device__ __forceinline__ void store_rows_f32(const EpiCtx& c, float* base, int ld, int nb, int nlim, const float* v32) {
    const KParams& p = *c.p;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float* v = v32 + h * 16 + u * 4;
            *reinterpret_cast<float4*>(c.stg + stage_off<64>(c.lane, u)) = make_float4(v[0], v[1], v[2], v[3]);
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int r, u;
            stage_coalesced<64>(k, c.lane, r, u);
            const int grow = c.row0 + r, col = nb + h * 16 + u * 4;
            const float4 t = *reinterpret_cast<const float4*>(c.stg + k * 512 + c.lane * 16);
            if (grow < p.M && col < nlim) *reinterpret_cast<float4*>(base + (size_t)grow * ld + col) = t;
        }
        __syncwarp();
    }
}
// registers (thread == row, NV values) -> bf16 global, coalesced; RB = NV * 2 bytes per row
template <int NV>
__device__ __forceinline__ void store_rows_bf16(const EpiCtx& c, __nv_bfloat16* base, int ld, int nb, int nlim,
                                                const float* v) {
    constexpr int RB = NV * 2;
    const KParams& p = *c.p;
#pragma unroll
    for (int u = 0; u < NV / 8; ++u) {
        uint4 t;
        t.x = pack_bf16x2(v[u * 8], v[u * 8 + 1]);
        t.y = pack_bf16x2(v[u * 8 + 2], v[u * 8 + 3]);
        t.z = pack_bf16x2(v[u * 8 + 4], v[u * 8 + 5]);
        t.w = pack_bf16x2(v[u * 8 + 6], v[u * 8 + 7]);
        *reinterpret_cast<uint4*>(c.stg + stage_off<RB>(c.lane, u)) = t;
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < RB / 16; ++k) {
        int r, u;
        stage_coalesced<RB>(k, c.lane, r, u);
        const int grow = c.row0 + r, col = nb + u * 8;
        const uint4 t = *reinterpret_cast<const uint4*>(c.stg + k * 512 + c.lane * 16);
        if (grow < p.M && col < nlim) *reinterpret_cast<uint4*>(base + (size_t)grow * ld + col) = t;
    }
    __syncwarp();
}

}  // namespace
}  // namespace mossdec
