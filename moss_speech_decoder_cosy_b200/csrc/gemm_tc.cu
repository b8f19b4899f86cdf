// Persistent warp-specialised tcgen05 implicit-GEMM kernel (see gemm.h for the contract).
//
// Replaces, for the decode path, every dense contraction the reference runs through
// cuDNN/cuBLAS: CausalConv1d (cosyvoice/flow/decoder.py:36-62), the Linear layers of
// BasicTransformerBlock (matcha/models/components/transformer.py:262-314), and the HiFT
// Conv1d / ConvTranspose1d layers (cosyvoice/hifigan/generator.py:46-101,520-552).
#include "gemm.h"
#include "ptx.cuh"
#include "epilogue.cuh"

#include <mutex>
#include <stdlib.h>
#include <stdio.h>

namespace mossdec {

namespace {

// One pass over this warp's column range [cbeg, cend) (16-column chunks of the tile), W16 chunks at a time.
// W16 == 4 is the bf16-only fast path (no fp32 input / output); W16 == 2 handles everything.
template <int W16>
__device__ __forceinline__ void epi_main(const EpiCtx& c, int cbeg, int cend, float mean1, float rstd1, float& s2,
                                         float& ss2, float& k2) {
    constexpr int NV = W16 * 16;
    const KParams& p = *c.p;
    const GemmEpilogue& e = p.epi;
    for (int cc = cbeg; cc < cend; cc += W16) {
        const int nch = min(W16, cend - cc);
        const int nb = c.n0 + cc * 16;
        const int nlim = min(p.N, nb + nch * 16);  // columns of this super-chunk that exist and are ours
        float v[NV];
#pragma unroll
        for (int j = 0; j < W16; ++j) {
            if (j < nch) {
                tmem_ld16(c.taddr + (uint32_t)((cc + j) * 16), v + j * 16);
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[j * 16 + i] = 0.f;
            }
        }
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < W16; ++j) {
            if (j < nch) {
                float* w = v + j * 16;
                const int n = nb + j * 16;
                if (e.bias) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 b = ldg4(e.bias + n + i);
                        w[i] += b.x; w[i + 1] += b.y; w[i + 2] += b.z; w[i + 3] += b.w;
                    }
                }
                if (e.ln1_g) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 g = ldg4(e.ln1_g + n + i), b = ldg4(e.ln1_b + n + i);
                        w[i] = (w[i] - mean1) * rstd1 * g.x + b.x;
                        w[i + 1] = (w[i + 1] - mean1) * rstd1 * g.y + b.y;
                        w[i + 2] = (w[i + 2] - mean1) * rstd1 * g.z + b.z;
                        w[i + 3] = (w[i + 3] - mean1) * rstd1 * g.w + b.w;
                    }
                }
                apply_act_vec<16>(w, e.act1, e.act1_p, e.act1_alpha ? e.act1_alpha + n : nullptr);
                if (e.tbias) {
                    const float* tb = e.tbias + (size_t)(c.seq < 0 ? 0 : c.seq) * e.tbias_ld + n;
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 b = ldg4(tb + i);
                        w[i] += b.x; w[i + 1] += b.y; w[i + 2] += b.z; w[i + 3] += b.w;
                    }
                }
            }
        }
        if (W16 == 2) {
            if (e.resid) {
                if (e.f32_tiled) add_rows_f32_tiled(e.resid, c.row0 + c.lane, nb, v);
                else add_rows_f32(c, e.resid, e.ldr, nb, nlim, v);
            }
            if (e.accum_f32) {
                float r[32];
                load_rows_f32(c, e.out_f32, e.ldf, nb, nlim, r);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] += r[i];
            }
        }
        if (c.valid) {
            if (e.out_scale != 1.f) {
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] *= e.out_scale;
            }
        } else {
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = 0.f;
        }
        if (W16 == 2) {
            if (e.out_f32) {
                if (e.f32_tiled) store_rows_f32_tiled(e.out_f32, c.row0 + c.lane, nb, v);
                else store_rows_f32(c, e.out_f32, e.ldf, nb, nlim, v);
            }
            if (e.ln2_g || e.ln2_plain) {   // sums shifted by the first value of the row part
                if (cc == cbeg) k2 = v[0];
#pragma unroll
                for (int i = 0; i < NV; ++i) {
                    const float d = v[i] - k2;
                    s2 += d;
                    ss2 = fmaf(d, d, ss2);
                }
                continue;  // bf16 side is produced by the second pass
            }
        }
        if (e.out_bf16) {
            if (e.act2 != ACT_NONE) {
#pragma unroll
                for (int j = 0; j < W16; ++j)
                    if (j < nch)
                        apply_act_vec<16>(v + j * 16, e.act2, e.act2_p, e.act2_alpha ? e.act2_alpha + nb + j * 16 : nullptr);
                if (!c.valid) {
#pragma unroll
                    for (int i = 0; i < NV; ++i) v[i] = 0.f;
                }
            }
            store_rows_bf16<NV>(c, e.out_bf16, e.ldo, nb, nlim, v);
            if (e.out_bf16_lo) {
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] -= __bfloat162float(__float2bfloat16(v[i]));
                store_rows_bf16<NV>(c, e.out_bf16_lo, e.ldo, nb, nlim, v);
            }
        }
    }
}

// ---- fast path for "bias + activation + mask -> bf16" tiles (q|k|v projection, most vocoder convs): no LayerNorm, no
// fp32 traffic.  Each epilogue warp owns 32 rows x (BN / 4) columns; the TMEM drain of the second 32-column half is in
// flight while the first goes through the math, and the bf16 tile leaves through a 64B-swizzled 2 KB staging tile and
// one TMA store per half (no LDS / STG / predicates; TMA clips the M / N tails).
__device__ __forceinline__ void epi_fast_half(const EpiCtx& c, const CUtensorMap* tmOut, float* v, int n) {
    const GemmEpilogue& e = c.p->epi;
    if (e.bias) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 b = ldg4(e.bias + n + 4 * i);
            v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
        }
    }
    apply_act_vec<16>(v, e.act1, e.act1_p, e.act1_alpha ? e.act1_alpha + n : nullptr);
    apply_act_vec<16>(v + 16, e.act1, e.act1_p, e.act1_alpha ? e.act1_alpha + n + 16 : nullptr);
    if (!c.valid) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0.f;
    }
    // the previous TMA store of this warp must have finished reading the staging tile
    if (c.lane == 0) tma_store_wait_read();
    __syncwarp();
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        uint4 t;
        t.x = pack_bf16x2(v[u * 8], v[u * 8 + 1]);
        t.y = pack_bf16x2(v[u * 8 + 2], v[u * 8 + 3]);
        t.z = pack_bf16x2(v[u * 8 + 4], v[u * 8 + 5]);
        t.w = pack_bf16x2(v[u * 8 + 6], v[u * 8 + 7]);
        // 64B swizzle: 16-byte unit index XOR bits [7, 9) of the byte offset (= (row >> 1) & 3)
        *reinterpret_cast<uint4*>(c.stg + c.lane * 64 + ((u ^ ((c.lane >> 1) & 3)) << 4)) = t;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (c.lane == 0) {
        tma_store_2d(tmOut, c.stg, n, c.row0);
        tma_store_commit();
    }
}

// returns after the last TMEM read of the tile; the caller hands the accumulator back to the MMA warp
__device__ __forceinline__ void epi_fast(const EpiCtx& c, const CUtensorMap* tmOut, int part, uint64_t* tempty_bar) {
    const KParams& p = *c.p;
    const int wcols = p.BN >> 2;            // 64, 32 or 16 columns per warp
    const int col0 = part * wcols;          // inside the tile
    const int n = c.n0 + col0;
    if (wcols == 16) {   // BN = 64 (vocoder C = 64 stage): one 16-column chunk, 32 rows x 32 B staging tile without swizzle
        const GemmEpilogue& e = p.epi;
        float v[16];
        tmem_ld16(c.taddr + (uint32_t)col0, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (c.lane == 0) mbar_arrive(tempty_bar);
        if (e.bias) {
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 b = ldg4(e.bias + n + i);
                v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
            }
        }
        apply_act_vec<16>(v, e.act1, e.act1_p, e.act1_alpha ? e.act1_alpha + n : nullptr);
        if (!c.valid) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        if (c.lane == 0) tma_store_wait_read();
        __syncwarp();
        uint4 t0, t1;
        t0.x = pack_bf16x2(v[0], v[1]); t0.y = pack_bf16x2(v[2], v[3]); t0.z = pack_bf16x2(v[4], v[5]); t0.w = pack_bf16x2(v[6], v[7]);
        t1.x = pack_bf16x2(v[8], v[9]); t1.y = pack_bf16x2(v[10], v[11]); t1.z = pack_bf16x2(v[12], v[13]); t1.w = pack_bf16x2(v[14], v[15]);
        *reinterpret_cast<uint4*>(c.stg + c.lane * 32) = t0;
        *reinterpret_cast<uint4*>(c.stg + c.lane * 32 + 16) = t1;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (c.lane == 0) {
            tma_store_2d(tmOut, c.stg, n, c.row0);
            tma_store_commit();
        }
        return;
    }
    float va[32], vb[32];
    tmem_ld16(c.taddr + (uint32_t)col0, va);
    tmem_ld16(c.taddr + (uint32_t)(col0 + 16), va + 16);
    tmem_ld_wait();
    if (wcols == 64) {
        tmem_ld16(c.taddr + (uint32_t)(col0 + 32), vb);
        tmem_ld16(c.taddr + (uint32_t)(col0 + 48), vb + 16);
        epi_fast_half(c, tmOut, va, n);
        tmem_ld_wait();
    }
    tc_fence_before();
    __syncwarp();
    if (c.lane == 0) mbar_arrive(tempty_bar);
    if (wcols == 64) epi_fast_half(c, tmOut, vb, n + 32);
    else epi_fast_half(c, tmOut, va, n);
}

// ---- "bias + fp32 residual (tiled layout) -> fp32 stream + LayerNorm (affine folded away) -> bf16" tiles: the attention
// out-projection.  N == BN == 256; each warp owns 32 rows x 64 columns.  The residual loads are issued BEFORE the wait for
// the accumulator (they do not depend on the MMA), so their latency hides behind the tile's MMAs; the TMEM drain is
// pipelined 16 columns at a time into the same registers; the bf16 tile leaves through two TMA stores.
__device__ __forceinline__ void epi_resln(const EpiCtx& c, const CUtensorMap* tmOut, int part, uint32_t tmem_acc,
                                          uint64_t* tfull_bar, uint32_t tfull_phase, uint64_t* tempty_bar) {
    const KParams& p = *c.p;
    const GemmEpilogue& e = p.epi;
    const int row = c.row0 + c.lane;
    const int n = part * 64;
    float v[64];
    {
        const float* q = e.resid + tiled_f32_offset(row, n);
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const float4 t = *reinterpret_cast<const float4*>(q + (size_t)u * 512);
            const float4 b = e.bias ? ldg4(e.bias + n + 4 * u) : make_float4(0.f, 0.f, 0.f, 0.f);
            v[4 * u] = t.x + b.x; v[4 * u + 1] = t.y + b.y; v[4 * u + 2] = t.z + b.z; v[4 * u + 3] = t.w + b.w;
        }
    }
    mbar_wait(tfull_bar, tfull_phase);
    tc_fence_after();
    const uint32_t ta = tmem_acc + (uint32_t)n;
    float t0[16], t1[16];
    tmem_ld16(ta, t0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float* t = (q & 1) ? t1 : t0;
        tmem_ld_wait();
        if (q < 3) tmem_ld16(ta + (q + 1) * 16, (q & 1) ? t0 : t1);
        if (q == 3) {   // accumulator fully read: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (c.lane == 0) mbar_arrive(tempty_bar);
        }
        float* w = v + q * 16;
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i] = c.valid ? w[i] + t[i] : 0.f;
        float* dst = e.out_f32 + tiled_f32_offset(row, n + q * 16);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            *reinterpret_cast<float4*>(dst + (size_t)u * 512) = make_float4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
    }
    if (c.lane == 0) tma_store_wait_read();   // the exchange slot is the head of the staging tile the last store read
    __syncwarp();
    // centred statistics (no E[x^2] - mean^2 cancellation when the residual stream carries a large mean)
    const float2 lm = local_mean_m2<64>(v);
    const float2 mr = group_mean_rstd(c, lm.x, lm.y, 64.f, e.ln_eps);
    const float mean = mr.x, rstd = mr.y;
    const float a = c.valid ? rstd : 0.f, bsh = c.valid ? -mean * rstd : 0.f;
#pragma unroll
    for (int hlf = 0; hlf < 2; ++hlf) {
        float* w = v + hlf * 32;
        if (c.lane == 0) tma_store_wait_read();
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint4 t;
            t.x = pack_bf16x2(fmaf(w[u * 8], a, bsh), fmaf(w[u * 8 + 1], a, bsh));
            t.y = pack_bf16x2(fmaf(w[u * 8 + 2], a, bsh), fmaf(w[u * 8 + 3], a, bsh));
            t.z = pack_bf16x2(fmaf(w[u * 8 + 4], a, bsh), fmaf(w[u * 8 + 5], a, bsh));
            t.w = pack_bf16x2(fmaf(w[u * 8 + 6], a, bsh), fmaf(w[u * 8 + 7], a, bsh));
            *reinterpret_cast<uint4*>(c.stg + c.lane * 64 + ((u ^ ((c.lane >> 1) & 3)) << 4)) = t;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (c.lane == 0) {
            tma_store_2d(tmOut, c.stg, n + hlf * 32, c.row0);
            tma_store_commit();
        }
    }
}

// ---- single-drain variant of the resnet conv epilogue below (used for block1, where it measured 40 us against 51 us for
// the two-pass form; block2 is the other way round, 77 us against 74 us):
//   kRes == false:  mask(mish(LayerNorm(acc + bias))) + time bias                       -> bf16            (block1)
//   kRes == true :  mask(mish(LayerNorm(acc + bias))) + fp32 residual -> fp32 stream, plain LayerNorm -> bf16 (block2)
// The accumulator is read ONCE (pipelined 16-column drains) into 64 registers per thread and handed back to the MMA warp
// before any of the math, so LayerNorm + Mish + the second LayerNorm run behind the next tile's MMAs.  bias comes from
// shared memory (sBias, filled once per CTA), gamma / beta and the addend are prefetched one 16-column group ahead.
template <bool kRes>
__device__ __forceinline__ void epi_lnact_single(const EpiCtx& c, const CUtensorMap* tmOut, int part, uint32_t tmem_acc,
                                          uint64_t* tfull_bar, uint32_t tfull_phase, uint64_t* tempty_bar,
                                          const float* sBias) {
    const KParams& p = *c.p;
    const GemmEpilogue& e = p.epi;
    const int row = c.row0 + c.lane;
    const int n = part * 64;
    float v[64];
    mbar_wait(tfull_bar, tfull_phase);
    tc_fence_after();
    const uint32_t ta = tmem_acc + (uint32_t)n;
    {
        float t0[16], t1[16];
        tmem_ld16(ta, t0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float* t = (q & 1) ? t1 : t0;
            tmem_ld_wait();
            if (q < 3) tmem_ld16(ta + (q + 1) * 16, (q & 1) ? t0 : t1);
            if (q == 3) {   // accumulator fully read: hand it back to the MMA warp
                tc_fence_before();
                __syncwarp();
                if (c.lane == 0) mbar_arrive(tempty_bar);
            }
            float* w = v + q * 16;
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 b = *reinterpret_cast<const float4*>(sBias + n + q * 16 + i);   // broadcast LDS
                w[i] = t[i] + b.x; w[i + 1] = t[i + 1] + b.y; w[i + 2] = t[i + 2] + b.z; w[i + 3] = t[i + 3] + b.w;
            }
        }
    }
    if (c.lane == 0) tma_store_wait_read();   // the exchange slot is the head of the staging tile the last store read
    __syncwarp();
    const float2 lm = local_mean_m2<64>(v);
    const float2 mr = group_mean_rstd(c, lm.x, lm.y, 64.f, e.ln_eps);
    const float mean = mr.x, rstd = mr.y;

    // addend of 16 columns: fp32 residual (lane-major tiled layout) or the time bias row of this row's sequence
    const float* addp = kRes ? e.resid + tiled_f32_offset(row, n)
                             : (e.tbias ? e.tbias + (size_t)(c.seq < 0 ? 0 : c.seq) * e.tbias_ld + n : nullptr);
    // groups of 4 columns, the three parameter vectors of the next group in flight while this one goes through the math
    auto load4 = [&](int u, float4& a, float4& g, float4& bt) {
        if (kRes) a = *reinterpret_cast<const float4*>(addp + (size_t)u * 512);
        else a = addp ? ldg4(addp + 4 * u) : make_float4(0.f, 0.f, 0.f, 0.f);
        g = ldg4(e.ln1_g + n + 4 * u);
        bt = ldg4(e.ln1_b + n + 4 * u);
    };
    float4 an, gn, bn;
    load4(0, an, gn, bn);
    const float nmr = -mean * rstd;
    const float vmask = c.valid ? 1.f : 0.f;
#pragma unroll
    for (int u = 0; u < 16; ++u) {
        const float4 a = an, g = gn, bt = bn;
        if (u < 15) load4(u + 1, an, gn, bn);
        float* w = v + 4 * u;
        const float gx[4] = {g.x, g.y, g.z, g.w}, bx[4] = {bt.x, bt.y, bt.z, bt.w}, ax[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float y = fmaf(fmaf(w[k], rstd, nmr), gx[k], bx[k]);
            // (masking by multiplication: a per-element "valid ? f(y) : 0" becomes a branch around every Mish, which
            //  serialises the unrolled loop; all inputs are finite, padding rows included)
            w[k] = (fast_mish(y) + ax[k]) * vmask;
        }
        if (kRes)
            *reinterpret_cast<float4*>(e.out_f32 + tiled_f32_offset(row, n) + (size_t)u * 512) = make_float4(w[0], w[1], w[2], w[3]);
    }
    float a = 1.f, bsh = 0.f;
    if (kRes) {
        const float2 lm2 = local_mean_m2<64>(v);
        const float2 mr2 = group_mean_rstd(c, lm2.x, lm2.y, 64.f, e.ln_eps);
        a = c.valid ? mr2.y : 0.f;
        bsh = c.valid ? -mr2.x * mr2.y : 0.f;
    }
#pragma unroll
    for (int hlf = 0; hlf < 2; ++hlf) {
        float* w = v + hlf * 32;
        if (hlf == 1) {
            if (c.lane == 0) tma_store_wait_read();
            __syncwarp();
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint4 t;
            t.x = pack_bf16x2(fmaf(w[u * 8], a, bsh), fmaf(w[u * 8 + 1], a, bsh));
            t.y = pack_bf16x2(fmaf(w[u * 8 + 2], a, bsh), fmaf(w[u * 8 + 3], a, bsh));
            t.z = pack_bf16x2(fmaf(w[u * 8 + 4], a, bsh), fmaf(w[u * 8 + 5], a, bsh));
            t.w = pack_bf16x2(fmaf(w[u * 8 + 6], a, bsh), fmaf(w[u * 8 + 7], a, bsh));
            *reinterpret_cast<uint4*>(c.stg + c.lane * 64 + ((u ^ ((c.lane >> 1) & 3)) << 4)) = t;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (c.lane == 0) {
            tma_store_2d(tmOut, c.stg, n + hlf * 32, c.row0);
            tma_store_commit();
        }
    }
}

// ---- the two conv blocks of a resnet (cosyvoice decoder.py:65-78, matcha decoder.py:49-61), N == BN == 256:
//   kRes == false:  mask(mish(LayerNorm(acc + bias)) + time bias)                       -> bf16            (block1)
//   kRes == true :  mask(mish(LayerNorm(acc + bias)) + fp32 residual) -> fp32 stream, plain LayerNorm -> bf16 (block2)
// Each warp owns 32 rows x 64 columns.  The addend (residual rows in the tiled layout, or the time-bias row of the
// sequence) is loaded into the thread's 64 result registers BEFORE the wait for the accumulator, so its L2 latency hides
// behind the tile's MMAs; the accumulator is then drained twice with pipelined 16-column tcgen05.ld (statistics, then
// normalise + Mish + add) -- TMEM reads are cheap next to K = 768 of MMAs, registers are not.  bias comes from shared
// memory (sBias, filled once per CTA); all masking is by multiplication (a per-element select turns into a branch around
// every Mish and serialises the unrolled loop).
template <bool kRes>
__device__ __forceinline__ void epi_lnact(const EpiCtx& c, const CUtensorMap* tmOut, int part, uint32_t tmem_acc,
                                          uint64_t* tfull_bar, uint32_t tfull_phase, uint64_t* tempty_bar,
                                          const float* sBias) {
    const KParams& p = *c.p;
    const GemmEpilogue& e = p.epi;
    const int row = c.row0 + c.lane;
    const int n = part * 64;
    float v[64];
    if (kRes) {
        const float* q = e.resid + tiled_f32_offset(row, n);
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const float4 t = *reinterpret_cast<const float4*>(q + (size_t)u * 512);
            v[4 * u] = t.x; v[4 * u + 1] = t.y; v[4 * u + 2] = t.z; v[4 * u + 3] = t.w;
        }
    } else if (e.tbias) {
        const float* q = e.tbias + (size_t)(c.seq < 0 ? 0 : c.seq) * e.tbias_ld + n;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const float4 t = ldg4(q + 4 * u);
            v[4 * u] = t.x; v[4 * u + 1] = t.y; v[4 * u + 2] = t.z; v[4 * u + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = 0.f;
    }
    mbar_wait(tfull_bar, tfull_phase);
    tc_fence_after();
    const uint32_t ta = tmem_acc + (uint32_t)n;
    float t0[16], t1[16];
    float s = 0.f, ss = 0.f, K = 0.f;
    tmem_ld16(ta, t0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {   // pass 1: row statistics of acc + bias, sums shifted by the row's first value
        float* t = (q & 1) ? t1 : t0;
        tmem_ld_wait();
        if (q < 3) tmem_ld16(ta + (q + 1) * 16, (q & 1) ? t0 : t1);
        if (q == 0) K = t[0] + sBias[n];
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            const float4 b = *reinterpret_cast<const float4*>(sBias + n + q * 16 + i);   // broadcast LDS
            const float x0 = t[i] + b.x - K, x1 = t[i + 1] + b.y - K, x2 = t[i + 2] + b.z - K, x3 = t[i + 3] + b.w - K;
            s += (x0 + x1) + (x2 + x3);
            ss = fmaf(x0, x0, fmaf(x1, x1, fmaf(x2, x2, fmaf(x3, x3, ss))));
        }
    }
    tmem_ld16(ta, t0);   // first group of pass 2 flies during the exchange
    if (c.lane == 0) tma_store_wait_read();   // the exchange slot is the head of the staging tile the last store read
    __syncwarp();
    const float2 lm = shifted_mean_m2(K, s, ss, 64.f);
    const float2 mr = group_mean_rstd(c, lm.x, lm.y, 64.f, e.ln_eps);
    const float mean = mr.x, rstd = mr.y;
    const float vmask = c.valid ? 1.f : 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {   // pass 2: LayerNorm + Mish + addend
        float* t = (q & 1) ? t1 : t0;
        tmem_ld_wait();
        if (q < 3) tmem_ld16(ta + (q + 1) * 16, (q & 1) ? t0 : t1);
        if (q == 3) {   // accumulator fully read: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (c.lane == 0) mbar_arrive(tempty_bar);
        }
        float* w = v + q * 16;
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            const float4 b = *reinterpret_cast<const float4*>(sBias + n + q * 16 + i);
            const float4 g = ldg4(e.ln1_g + n + q * 16 + i), bt = ldg4(e.ln1_b + n + q * 16 + i);
            const float bx[4] = {b.x, b.y, b.z, b.w}, gx[4] = {g.x, g.y, g.z, g.w}, tx[4] = {bt.x, bt.y, bt.z, bt.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float y = fmaf((t[i + k] + bx[k] - mean) * rstd, gx[k], tx[k]);
                w[i + k] = (fast_mish(y) + w[i + k]) * vmask;
            }
        }
        if (kRes) {
            float* dst = e.out_f32 + tiled_f32_offset(row, n + q * 16);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                *reinterpret_cast<float4*>(dst + (size_t)u * 512) = make_float4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
        }
    }
    float a = 1.f, bsh = 0.f;
    if (kRes) {
        const float2 lm2 = local_mean_m2<64>(v);
        const float2 mr2 = group_mean_rstd(c, lm2.x, lm2.y, 64.f, e.ln_eps);
        a = c.valid ? mr2.y : 0.f;
        bsh = c.valid ? -mr2.x * mr2.y : 0.f;
    }
#pragma unroll
    for (int hlf = 0; hlf < 2; ++hlf) {
        float* w = v + hlf * 32;
        if (hlf == 1) {
            if (c.lane == 0) tma_store_wait_read();
            __syncwarp();
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint4 t;
            t.x = pack_bf16x2(fmaf(w[u * 8], a, bsh), fmaf(w[u * 8 + 1], a, bsh));
            t.y = pack_bf16x2(fmaf(w[u * 8 + 2], a, bsh), fmaf(w[u * 8 + 3], a, bsh));
            t.z = pack_bf16x2(fmaf(w[u * 8 + 4], a, bsh), fmaf(w[u * 8 + 5], a, bsh));
            t.w = pack_bf16x2(fmaf(w[u * 8 + 6], a, bsh), fmaf(w[u * 8 + 7], a, bsh));
            *reinterpret_cast<uint4*>(c.stg + c.lane * 64 + ((u ^ ((c.lane >> 1) & 3)) << 4)) = t;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (c.lane == 0) {
            tma_store_2d(tmOut, c.stg, n + hlf * 32, c.row0);
            tma_store_commit();
        }
    }
}

// ---- "bias + fp32 residual (+ running sum) -> fp32 stream, Snake / leaky-ReLU of it -> bf16" tiles: the second conv of the
// vocoder ResBlocks (cosyvoice/hifigan/generator.py:94-101) at C = 64 / 128, N == BN.  In the general epilogue these tiles
// cost ~570 instructions per warp for 16 elements per thread (index arithmetic, predicates, staged row I/O): the kernel
// ran at 4 % of the tensor pipe whatever K was.  Here a warp owns 32 rows x 16 (BN = 64) or 2 x 16 (BN = 128) columns;
// per 16-column sub-tile the residual (and, when accumulating, the running sum) arrives by TMA in a 64B-swizzled 2 KB
// tile -- requested BEFORE the wait for the accumulator -- the result overwrites it in place and leaves by a TMA store,
// the bf16 side by a second one.  No per-element predicates: TMA clips the M tail, masked rows are zeroed by multiplication.
// The fp32 tiles live in the part of the B stage slots that a BN <= 128 problem does not use (r1), the bf16 tile in the
// warp's staging tile.
__device__ __forceinline__ void epi_vocres(const EpiCtx& c, const CUtensorMap* tmRes, const CUtensorMap* tmF32,
                                           const CUtensorMap* tmOut, int part, uint64_t* tfull_bar, uint32_t accphase,
                                           uint64_t* tempty_bar, uint64_t* rbar, uint32_t& rphase, uint8_t* r1) {
    const KParams& p = *c.p;
    const GemmEpilogue& e = p.epi;
    const int wcols = p.BN >> 2;   // 16 or 32 columns per warp
    const int nsub = wcols >> 4;
    const uint32_t swz = (uint32_t)((c.lane >> 1) & 3);   // 64B swizzle: 16-byte unit index XOR bits [7, 9) of the byte offset
    const float keep = c.valid ? e.out_scale : 0.f;
    for (int sub = 0; sub < nsub; ++sub) {
        const int col = part * wcols + sub * 16;   // column inside the tile == global column (N == BN)
        if (c.lane == 0) {
            tma_store_wait_read();   // the previous stores of this warp have finished reading the staging tiles
            mbar_expect_tx(rbar, e.accum_f32 ? 4096u : 2048u);
            tma_load_2d(r1, tmRes, col, c.row0, rbar);
            if (e.accum_f32) tma_load_2d(r1 + 2048, tmF32, col, c.row0, rbar);
        }
        if (sub == 0) {
            mbar_wait(tfull_bar, accphase);
            tc_fence_after();
        }
        float v[16];
        tmem_ld16(c.taddr + (uint32_t)col, v);
        tmem_ld_wait();
        if (sub + 1 == nsub) {   // last TMEM read of the tile: the accumulator goes back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (c.lane == 0) mbar_arrive(tempty_bar);
        }
        if (e.bias) {
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 b = ldg4(e.bias + col + i);
                v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
            }
        }
        mbar_wait(rbar, rphase);
        rphase ^= 1;
        uint8_t* rrow = r1 + c.lane * 64;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            float4* q = reinterpret_cast<float4*>(rrow + ((u ^ swz) << 4));
            float4 t = *q;
            if (e.accum_f32) {
                const float4 o = *reinterpret_cast<const float4*>(rrow + 2048 + ((u ^ swz) << 4));
                t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w;
            }
            v[4 * u] = (v[4 * u] + t.x) * keep;
            v[4 * u + 1] = (v[4 * u + 1] + t.y) * keep;
            v[4 * u + 2] = (v[4 * u + 2] + t.z) * keep;
            v[4 * u + 3] = (v[4 * u + 3] + t.w) * keep;
            *q = make_float4(v[4 * u], v[4 * u + 1], v[4 * u + 2], v[4 * u + 3]);   // in place: the thread's own row
        }
        if (e.out_bf16) {
            if (e.act2 != ACT_NONE) {
                apply_act_vec<16>(v, e.act2, e.act2_p, e.act2_alpha ? e.act2_alpha + col : nullptr);
                if (!c.valid) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = 0.f;
                }
            }
            uint4 t0, t1;
            t0.x = pack_bf16x2(v[0], v[1]); t0.y = pack_bf16x2(v[2], v[3]); t0.z = pack_bf16x2(v[4], v[5]); t0.w = pack_bf16x2(v[6], v[7]);
            t1.x = pack_bf16x2(v[8], v[9]); t1.y = pack_bf16x2(v[10], v[11]); t1.z = pack_bf16x2(v[12], v[13]); t1.w = pack_bf16x2(v[14], v[15]);
            *reinterpret_cast<uint4*>(c.stg + c.lane * 32) = t0;   // 32 rows x 32 B, no swizzle
            *reinterpret_cast<uint4*>(c.stg + c.lane * 32 + 16) = t1;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (c.lane == 0) {
            tma_store_2d(tmF32, r1, col, c.row0);
            if (e.out_bf16) tma_store_2d(tmOut, c.stg, col, c.row0);
            tma_store_commit();
        }
    }
}

// The it-th tile of this CTA.  Two schedules:
//   units (default): work unit = `group` consecutive N tiles of one M tile, units strided over the CTAs; with a_stationary
//     the A tile of a unit is fetched once (keep_a for its following N tiles);
//   b_stationary: K fits the ring exactly and there are few N tiles -- CTA b owns N tile b % tiles_n for the whole launch,
//     fetches its B tile ONCE (it stays in the B halves of the ring's stages) and streams the M tiles
//     b / tiles_n, + gridDim.x / tiles_n, ...: what an SM takes in per tile drops from A / group + B to A alone
//     (q | k | v projection: 149 KB -> 64 KB per 128 x 256 tile, against ~42 B/clk that an SM can take in).
struct TileIter {   // incremental: no division beyond the one the unit schedule always had
    int m0, n0;
    bool keep_a, keep_b;
    int unit, ju, m_tile, per_n;
    __device__ __forceinline__ void init(const KParams& p) {
        unit = (int)blockIdx.x;
        ju = -1;
        keep_a = keep_b = false;
        if (p.b_stationary) {
            per_n = (int)gridDim.x / p.tiles_n;            // CTAs per N tile (the grid is a multiple of tiles_n)
            const int owner = (int)blockIdx.x / p.tiles_n;
            n0 = ((int)blockIdx.x - owner * p.tiles_n) * p.BN;
            m_tile = owner - per_n;
        }
    }
    __device__ __forceinline__ bool next(const KParams& p) {
        if (p.b_stationary) {
            keep_b = ju >= 0;
            ju = 0;
            m_tile += per_n;
            m0 = m_tile * kGemmBM;
            return m_tile < p.tiles_m;
        }
        if (++ju == p.group) {
            ju = 0;
            unit += (int)gridDim.x;
        }
        const int tile = unit * p.group + ju;
        if (unit >= p.num_units || tile >= p.num_tiles) return false;
        const int mt = tile / p.tiles_n;
        m0 = mt * kGemmBM;
        n0 = (tile - mt * p.tiles_n) * p.BN;
        keep_a = p.a_stationary && ju > 0;
        return true;
    }
};

// kMode: 0 = general epilogue, 1 = epi_fast (bias + act -> bf16), 2 = epi_resln (residual + LayerNorm),
//        3 / 4 = epi_lnact (LayerNorm + Mish resnet blocks without / with residual + second LayerNorm),
//        5 = epi_vocres (fp32 residual / running sum through TMA: vocoder ResBlock convs at C = 64 / 128)
template <int kMode>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
               const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmF32,
               const __grid_constant__ KParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays in the shared address space (STS / LDS)
    uint8_t* sA = smem;
    uint8_t* sB = smem + kGemmStages * kABytes;
    uint8_t* sStage = smem + kGemmStages * (kABytes + kBBytesMax);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sStage + kEpiWarps * kStageBytes);
    uint64_t* full = bars;                     // [stages]
    uint64_t* empty = bars + kGemmStages;      // [stages]
    uint64_t* tfull = bars + 2 * kGemmStages;  // [2]
    uint64_t* tempty = tfull + 2;              // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    float* sBias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [256], modes 3 / 4 only

    // (shuffle broadcasts mark values as warp-uniform for the compiler: descriptors and loop state of the TMA / MMA
    //  warps then live in uniform registers and every tcgen05.mma / TMA is a single instruction, not a per-lane loop)
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmB);
        if (kMode != 0) tma_prefetch_desc(&tmOut);
        for (int i = 0; i < kGemmStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], kEpiWarps);
        }
        if (kMode == 5) {
            tma_prefetch_desc(&tmRes);
            tma_prefetch_desc(&tmF32);
            for (int i = 0; i < kEpiWarps; ++i) mbar_init(&bars[16 + i], 1);   // one transaction barrier per epilogue warp
        }
        mbar_fence_init();
    }
    if ((kMode == 3 || kMode == 4) && threadIdx.x >= 64 && threadIdx.x < 64 + 256) sBias[threadIdx.x - 64] = __ldg(p.epi.bias + threadIdx.x - 64);
    if (warp == 1) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_wait();     // everything above overlapped the previous kernel's tail; its results are visible from here on
    pdl_launch();

    const uint32_t b_bytes = (uint32_t)p.BN * kGemmBK * 2;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (whole warp, one elected lane issues)
        {
            int stage = 0;
            uint32_t phase = 0;
            // Work unit = `group` consecutive N tiles of one M tile.  When the whole K extent fits the ring (total_kb ==
            // stages, one segment) the A tile of a unit is fetched once and stays in its stages; only B is refilled for
            // the following N tiles (these GEMMs are bound by the bytes delivered into the SM, ~42 B/clk/SM chip-wide).
            TileIter tr;
            for (tr.init(p); tr.next(p);) {
                const bool keep_a = tr.keep_a, keep_b = tr.keep_b;
                const int m0 = tr.m0, n0 = tr.n0;
                int kb = 0;
                for (int s = 0; s < p.num_segs; ++s) {
                    const GemmSeg sg = p.segs[s];
                    const CUtensorMap* tmA = sg.map ? &tmA1 : &tmA0;
                    for (int j = 0; j < sg.nkb; ++j, ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        if (elect_one()) {
                            mbar_expect_tx(&full[stage], (keep_a ? 0u : (uint32_t)kABytes) + (keep_b ? 0u : b_bytes));
                            if (!keep_a) tma_load_2d(sA + stage * kABytes, tmA, sg.col0 + j * kGemmBK, m0 + sg.row_shift, &full[stage]);
                            if (!keep_b) tma_load_2d(sB + stage * kBBytesMax, &tmB, kb * kGemmBK, n0, &full[stage]);
                        }
                        if (++stage == kGemmStages) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (whole warp, one elected lane issues)
        {
            const uint32_t idesc = umma_idesc_bf16(kGemmBM, (uint32_t)p.BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t accphase = 0;
            TileIter tr;
            for (tr.init(p); tr.next(p);) {
                mbar_wait(&tempty[acc], accphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256u;
                for (int kb = 0; kb < p.total_kb; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint64_t da = umma_desc_sw128(smem_u32(sA + stage * kABytes));
                        const uint64_t db = umma_desc_sw128(smem_u32(sB + stage * kBBytesMax));
#pragma unroll
                        for (int k = 0; k < kGemmBK / 16; ++k)
                            umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
                        umma_commit(&empty[stage]);
                        if (kb + 1 == p.total_kb) umma_commit(&tfull[acc]);
                    }
                    if (++stage == kGemmStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                acc ^= 1;
                if (acc == 0) accphase ^= 1;
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue (thread == row, 4 warps per row group)
        const GemmEpilogue& e = p.epi;
        const int ew = warp - 2;
        const int part = ew >> 2;
        EpiCtx c;
        c.p = &p;
        c.lane = lane;
        c.ew = ew;
        c.quarter = warp & 3;  // TMEM lane quarter this warp may access
        c.stg = sStage + ew * kStageBytes;
        c.stg_all = sStage;
        const int nchunks = p.BN / 16;
        const int cbase = nchunks / kEpiParts, crem = nchunks % kEpiParts;
        const int cbeg = part * cbase + min(part, crem);
        const int cend = cbeg + cbase + (part < crem ? 1 : 0);
        int acc = 0;
        uint32_t accphase = 0;
        [[maybe_unused]] uint32_t vr_phase = 0;
        TileIter tr;
        for (tr.init(p); tr.next(p);) {
            const int m0 = tr.m0;
            c.n0 = tr.n0;
            c.row0 = m0 + c.quarter * 32;
            const int row = c.row0 + lane;
            const bool inb = row < p.M;
            c.seq = 0;
            if (e.row_seq != nullptr && inb) c.seq = __ldg(e.row_seq + row);
            c.valid = inb && c.seq >= 0;

            c.taddr = tmem_base + ((uint32_t)(c.quarter * 32) << 16) + (uint32_t)acc * 256u;
            if constexpr (kMode == 5) {
                // fp32 staging of this warp: 4 KB behind the B tile of stage slot (ew & 3) (BN <= 128 leaves >= 16 KB of each slot free)
                epi_vocres(c, &tmRes, &tmF32, &tmOut, part, &tfull[acc], accphase, &tempty[acc], &bars[16 + ew], vr_phase,
                           sB + (ew & 3) * kBBytesMax + p.BN * (kGemmBK * 2) + (ew >> 2) * 4096);
                acc ^= 1;
                if (acc == 0) accphase ^= 1;
                continue;
            }
            if constexpr (kMode == 2) {
                epi_resln(c, &tmOut, part, c.taddr, &tfull[acc], accphase, &tempty[acc]);
                acc ^= 1;
                if (acc == 0) accphase ^= 1;
                continue;
            }
            if constexpr (kMode == 3 || kMode == 4) {
                if constexpr (kMode == 3) epi_lnact_single<false>(c, &tmOut, part, c.taddr, &tfull[acc], accphase, &tempty[acc], sBias);
                else epi_lnact<true>(c, &tmOut, part, c.taddr, &tfull[acc], accphase, &tempty[acc], sBias);
                acc ^= 1;
                if (acc == 0) accphase ^= 1;
                continue;
            }
            mbar_wait(&tfull[acc], accphase);
            tc_fence_after();

            if constexpr (kMode == 1) {
                epi_fast(c, &tmOut, part, &tempty[acc]);
                acc ^= 1;
                if (acc == 0) accphase ^= 1;
                continue;
            } else {
            float mean1 = 0.f, rstd1 = 1.f;
            // (LayerNorm in this general mode needs BN % 64 == 0 so that the four warps of a group own equal column counts)
            if (e.ln1_g != nullptr) {  // LayerNorm over the whole row (N == BN): statistics of acc + bias, shifted sums
                float s = 0.f, ss = 0.f, K = 0.f;
                for (int cc = cbeg; cc < cend; ++cc) {
                    float v[16];
                    tmem_ld16(c.taddr + (uint32_t)(cc * 16), v);
                    tmem_ld_wait();
                    if (cc == cbeg) K = v[0] + (e.bias ? __ldg(e.bias + c.n0 + cc * 16) : 0.f);
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (e.bias) b = ldg4(e.bias + c.n0 + cc * 16 + i);
                        const float x0 = v[i] + b.x - K, x1 = v[i + 1] + b.y - K, x2 = v[i + 2] + b.z - K, x3 = v[i + 3] + b.w - K;
                        s += (x0 + x1) + (x2 + x3);
                        ss += (x0 * x0 + x1 * x1) + (x2 * x2 + x3 * x3);
                    }
                }
                const float np = (float)((cend - cbeg) * 16);
                const float2 lm = shifted_mean_m2(K, s, ss, np);
                const float2 mr = group_mean_rstd(c, lm.x, lm.y, np, e.ln_eps);
                mean1 = mr.x;
                rstd1 = mr.y;
            }
            float s2 = 0.f, ss2 = 0.f, k2 = 0.f;
            epi_main<2>(c, cbeg, cend, mean1, rstd1, s2, ss2, k2);

            // all TMEM reads of this tile are done: hand the accumulator back before the (TMEM-free) LN2 pass
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
            acc ^= 1;
            if (acc == 0) accphase ^= 1;

            if (e.ln2_g != nullptr || e.ln2_plain) {  // second LayerNorm on the fp32 value just stored -> bf16 operand of the next GEMM
                const float np2 = (float)((cend - cbeg) * 16);
                const float2 lm2 = shifted_mean_m2(k2, s2, ss2, np2);
                const float2 mr2 = group_mean_rstd(c, lm2.x, lm2.y, np2, e.ln_eps);
                const float mean2 = mr2.x, rstd2 = mr2.y;
                for (int cc = cbeg; cc < cend; cc += 2) {
                    const int nb = c.n0 + cc * 16;
                    const int nlim = min(p.N, nb + min(2, cend - cc) * 16);
                    float v[32];
                    if (e.f32_tiled) load_rows_f32_tiled(e.out_f32, c.row0 + c.lane, nb, v);
                    else load_rows_f32(c, e.out_f32, e.ldf, nb, nlim, v);
                    if (e.ln2_plain) {
                        const float a = c.valid ? rstd2 : 0.f, b = c.valid ? -mean2 * rstd2 : 0.f;
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = fmaf(v[i], a, b);
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            const float4 g = ldg4(e.ln2_g + nb + i), b = ldg4(e.ln2_b + nb + i);
                            v[i] = c.valid ? (v[i] - mean2) * rstd2 * g.x + b.x : 0.f;
                            v[i + 1] = c.valid ? (v[i + 1] - mean2) * rstd2 * g.y + b.y : 0.f;
                            v[i + 2] = c.valid ? (v[i + 2] - mean2) * rstd2 * g.z + b.z : 0.f;
                            v[i + 3] = c.valid ? (v[i + 3] - mean2) * rstd2 * g.w + b.w : 0.f;
                        }
                    }
                    store_rows_bf16<32>(c, e.out_bf16, e.ldo, nb, nlim, v);
                }
            }
            }  // kMode == 0
        }
    }

    if (kMode != 0 && warp >= 2 && lane == 0) tma_store_wait_all();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// 2-D bf16 tensor map: dims {cols, rows}, row stride ld elements, box {64, box_rows}, 128B swizzle.
bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)kGemmBK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr, "[mossdec] cuTensorMapEncodeTiled failed: %d (ptr=%p cols=%llu rows=%llu ld=%llu box=%u)\n",
                (int)r, ptr, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld, box_rows);
        return false;
    }
    return true;
}

// general form: box {box_cols, box_rows}; swizzle 0 = 128B, 1 = 64B
bool encode_2d_ex(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_cols,
                  uint32_t box_rows, int swizzle64) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr, "[mossdec] cuTensorMapEncodeTiled (store map) failed: %d (ptr=%p cols=%llu rows=%llu ld=%llu)\n",
                (int)r, ptr, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld);
        return false;
    }
    return true;
}

// any element type: dtype 0 = bf16, 1 = fp32; swizzle 0 = none, 1 = 64B, 2 = 128B (the inner box must fit the swizzle span)
bool encode_2d_any(CUtensorMap* m, const void* ptr, int dtype, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_cols,
                   uint32_t box_rows, int swizzle) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return false;
    const uint64_t esz = dtype ? 4 : 2;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * esz};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, dtype ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims,
                    strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_128B : (swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE),
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr, "[mossdec] cuTensorMapEncodeTiled (%s map) failed: %d (ptr=%p cols=%llu rows=%llu ld=%llu)\n",
                dtype ? "fp32" : "bf16", (int)r, ptr, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld);
        return false;
    }
    return true;
}

cudaError_t gemm_tc_init_device() {
    static std::atomic<uint64_t> done{0};
    static std::mutex mu;
    return once_per_device(done, mu, [] {
        if (!get_encode_fn()) return cudaErrorNotSupported;
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes + 1024);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes + 1024);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        return e;
    });
}

cudaError_t gemm_tc_launch(const GemmProblem& g, int num_sms, cudaStream_t stream) {
    if (g.M <= 0 || g.N <= 0 || g.BN < 16 || g.BN > 256 || (g.BN % 16) != 0 || g.num_segs < 1 ||
        g.num_segs > kGemmMaxSegs)
        return cudaErrorInvalidValue;
    if (g.epi.ln1_g && g.BN != g.N) return cudaErrorInvalidValue;
    if ((g.epi.ln2_g || g.epi.ln2_plain) && (g.BN != g.N || g.epi.accum_f32)) return cudaErrorInvalidValue;
    if ((g.epi.ln1_g || g.epi.ln2_g || g.epi.ln2_plain) && (g.BN % 64) != 0) return cudaErrorInvalidValue;  // equal column counts per epilogue warp
    if (g.epi.f32_tiled && (g.N != 256 || g.BN != 256 || g.epi.accum_f32 || (g.epi.resid && g.epi.ldr != 256) ||
                            (g.epi.out_f32 && g.epi.ldf != 256)))
        return cudaErrorInvalidValue;

    const cudaError_t attr_err = gemm_tc_init_device();
    if (attr_err != cudaSuccess) return attr_err;

    KParams kp;
    kp.M = g.M;
    kp.N = g.N;
    kp.BN = g.BN;
    kp.tiles_n = (g.N + g.BN - 1) / g.BN;
    const int tiles_m = (g.M + kGemmBM - 1) / kGemmBM;
    kp.num_tiles = tiles_m * kp.tiles_n;
    kp.group = 1;
    kp.a_stationary = 0;
    kp.b_stationary = 0;
    kp.tiles_m = tiles_m;
    kp.num_segs = g.num_segs;
    kp.total_kb = 0;
    for (int s = 0; s < g.num_segs; ++s) {
        kp.segs[s] = g.segs[s];
        kp.total_kb += g.segs[s].nkb;
    }
    kp.epi = g.epi;

    CUtensorMap tA0, tA1, tB;
    if (!encode_2d(&tA0, g.a[0].ptr, (uint64_t)g.a[0].cols, (uint64_t)g.a[0].rows, (uint64_t)g.a[0].ld, kGemmBM))
        return cudaErrorInvalidValue;
    if (g.a[1].ptr) {
        if (!encode_2d(&tA1, g.a[1].ptr, (uint64_t)g.a[1].cols, (uint64_t)g.a[1].rows, (uint64_t)g.a[1].ld, kGemmBM))
            return cudaErrorInvalidValue;
    } else {
        tA1 = tA0;
    }
    const uint64_t K = (uint64_t)kp.total_kb * kGemmBK;
    if (!encode_2d(&tB, g.w, K, (uint64_t)g.N, g.w_ld > 0 ? (uint64_t)g.w_ld : K, (uint32_t)g.BN)) return cudaErrorInvalidValue;

    // "bias + activation + mask -> bf16" epilogues leave through TMA stores (epi_fast)
    const GemmEpilogue& e = g.epi;
    kp.fast_bf16 = e.out_bf16 != nullptr && !e.out_f32 && !e.resid && !e.ln1_g && !e.ln2_g && !e.ln2_plain && !e.tbias &&
                   !e.out_bf16_lo && !e.accum_f32 && e.act2 == ACT_NONE && e.out_scale == 1.f &&
                   (g.BN == 256 || g.BN == 128 || g.BN == 64) && (g.N % g.BN) == 0 && (e.ldo % 8) == 0 &&
                   (reinterpret_cast<uintptr_t>(e.out_bf16) & 15) == 0;
    // residual + LayerNorm epilogue of the attention out-projection (epi_resln)
    const bool resln = e.out_bf16 != nullptr && e.out_f32 != nullptr && e.resid != nullptr && e.f32_tiled && e.ln2_plain &&
                       !e.ln1_g && !e.ln2_g && !e.tbias && !e.out_bf16_lo && !e.accum_f32 && e.act1 == ACT_NONE &&
                       e.act2 == ACT_NONE && e.out_scale == 1.f && g.BN == 256 && g.N == 256 && (g.M % kGemmBM) == 0 &&
                       (e.ldo % 8) == 0 && (reinterpret_cast<uintptr_t>(e.out_bf16) & 15) == 0;
    // LayerNorm + Mish conv blocks of the resnets (epi_lnact): block1 (+ time bias) and block2 (+ residual + next LayerNorm)
    const bool ln_common = e.out_bf16 != nullptr && e.ln1_g && e.ln1_b && e.bias && e.act1 == ACT_MISH && !e.ln2_g &&
                           !e.out_bf16_lo && !e.accum_f32 && e.act2 == ACT_NONE && e.out_scale == 1.f && g.BN == 256 &&
                           g.N == 256 && (g.M % kGemmBM) == 0 && (e.ldo % 8) == 0 &&
                           (reinterpret_cast<uintptr_t>(e.out_bf16) & 15) == 0;
    static const bool force_general = getenv("MOSSDEC_GEMM_GENERAL") != nullptr;   // A/B aid: every epilogue through mode 0
    const bool lnact_a = !force_general && ln_common && !e.out_f32 && !e.resid && !e.ln2_plain && (!e.tbias || (e.tbias_ld % 4) == 0);
    const bool lnact_b = !force_general && ln_common && e.out_f32 != nullptr && e.resid != nullptr && e.f32_tiled && e.ln2_plain && !e.tbias;
    CUtensorMap tOut = tB;
    if (kp.fast_bf16 && g.BN == 64) {   // 16 columns per warp: 32-byte rows, no swizzle
        if (!encode_2d_any(&tOut, e.out_bf16, 0, (uint64_t)g.N, (uint64_t)g.M, (uint64_t)e.ldo, 16, 32, 0)) return cudaErrorInvalidValue;
    } else if ((kp.fast_bf16 || resln || lnact_a || lnact_b) &&
               !encode_2d_ex(&tOut, e.out_bf16, (uint64_t)g.N, (uint64_t)g.M, (uint64_t)e.ldo, 32, 32, 1)) {
        return cudaErrorInvalidValue;
    }
    // fp32 residual (+ running sum) convs of the vocoder ResBlocks at C = 64 / 128 (epi_vocres); MOSSDEC_GEMM_VOCRES=0: general epilogue
    static const bool vocres_on = [] {
        const char* v = getenv("MOSSDEC_GEMM_VOCRES");
        return !(v && v[0] == '0');
    }();
    const bool vocres = vocres_on && !force_general && e.resid != nullptr && e.out_f32 != nullptr && !e.f32_tiled && !e.ln1_g && !e.ln2_g &&
                        !e.ln2_plain && !e.tbias && !e.out_bf16_lo && e.act1 == ACT_NONE && g.N == g.BN && (g.BN == 64 || g.BN == 128) &&
                        e.ldr == g.N && e.ldf == g.N && (!e.out_bf16 || e.ldo == g.N) &&
                        (reinterpret_cast<uintptr_t>(e.resid) & 15) == 0 && (reinterpret_cast<uintptr_t>(e.out_f32) & 15) == 0 &&
                        (!e.out_bf16 || (reinterpret_cast<uintptr_t>(e.out_bf16) & 15) == 0);
    CUtensorMap tRes = tB, tF32 = tB;
    if (vocres) {
        if (!encode_2d_any(&tRes, e.resid, 1, (uint64_t)g.N, (uint64_t)g.M, (uint64_t)e.ldr, 16, 32, 1) ||
            !encode_2d_any(&tF32, e.out_f32, 1, (uint64_t)g.N, (uint64_t)g.M, (uint64_t)e.ldf, 16, 32, 1) ||
            (e.out_bf16 && !encode_2d_any(&tOut, e.out_bf16, 0, (uint64_t)g.N, (uint64_t)g.M, (uint64_t)e.ldo, 16, 32, 0)))
            return cudaErrorInvalidValue;
    }

    // A-stationary units (see the producer): K fits the ring exactly, several N tiles per M tile, enough units to fill the GPU
    if (kp.total_kb == kGemmStages && g.num_segs == 1 && kp.tiles_n > 1) {
        const int grp = (kp.tiles_n % 3) == 0 ? 3 : ((kp.tiles_n % 2) == 0 ? 2 : 1);
        if (grp > 1 && kp.num_tiles / grp >= 2 * num_sms) {
            kp.group = grp;
            kp.a_stationary = 1;
        }
    }
    kp.num_units = (kp.num_tiles + kp.group - 1) / kp.group;
    int grid = kp.num_units < num_sms ? kp.num_units : num_sms;
    // B-stationary schedule (see TileIter): same condition on K, few N tiles, several M tiles per CTA.  MOSSDEC_GEMM_BSTAT=0: off
    static const bool bstat_on = [] {
        const char* v = getenv("MOSSDEC_GEMM_BSTAT");
        return !(v && v[0] == '0');
    }();
    if (bstat_on && kp.total_kb == kGemmStages && g.num_segs == 1 && kp.tiles_n > 1 && kp.tiles_n <= 8 &&
        tiles_m >= 2 * (num_sms / kp.tiles_n)) {
        kp.b_stationary = 1;
        kp.a_stationary = 0;
        kp.group = 1;
        kp.num_units = kp.num_tiles;
        grid = (num_sms / kp.tiles_n) * kp.tiles_n;
    }
    static const bool trace = getenv("MOSSDEC_GEMM_TRACE") != nullptr;   // developer aid: the first launches' geometry on stderr
    static int traced = 0;
    if (trace && traced < 40) {
        ++traced;
        fprintf(stderr, "[mossdec] gemm M=%d N=%d BN=%d K=%d segs=%d group=%d a_stationary=%d b_stationary=%d grid=%d\n", g.M, g.N, g.BN, (int)K,
                g.num_segs, kp.group, kp.a_stationary, kp.b_stationary, grid);
    }
    cudaError_t le;
    if (lnact_a) le = launch_pdl(gemm_tc_kernel<3>, dim3(grid), dim3(kThreads), kSmemBytes + 1024, stream, tA0, tA1, tB, tOut, tRes, tF32, kp);
    else if (lnact_b) le = launch_pdl(gemm_tc_kernel<4>, dim3(grid), dim3(kThreads), kSmemBytes + 1024, stream, tA0, tA1, tB, tOut, tRes, tF32, kp);
    else if (resln) le = launch_pdl(gemm_tc_kernel<2>, dim3(grid), dim3(kThreads), kSmemBytes, stream, tA0, tA1, tB, tOut, tRes, tF32, kp);
    else if (kp.fast_bf16) le = launch_pdl(gemm_tc_kernel<1>, dim3(grid), dim3(kThreads), kSmemBytes, stream, tA0, tA1, tB, tOut, tRes, tF32, kp);
    else if (vocres) le = launch_pdl(gemm_tc_kernel<5>, dim3(grid), dim3(kThreads), kSmemBytes, stream, tA0, tA1, tB, tOut, tRes, tF32, kp);
    else le = launch_pdl(gemm_tc_kernel<0>, dim3(grid), dim3(kThreads), kSmemBytes, stream, tA0, tA1, tB, tOut, tRes, tF32, kp);
    return le != cudaSuccess ? le : cudaGetLastError();
}

bool pdl_enabled() {
    static const bool on = [] {
        const char* v = getenv("MOSSDEC_PDL");
        return v && v[0] == '1';   // opt-in: measured neutral on one stream and 3 % slower with the two-stream CFG halves
                                   // (early-scheduled CTAs wait on SMs the other stream's kernel could be using)
    }();
    return on;
}

}  // namespace mossdec
