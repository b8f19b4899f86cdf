// Building blocks of the non-causal cosyvoice1 / Matcha estimator on the same tcgen05
// implicit-GEMM core as the live causal path:
//
//   cv1_resnet_block  <- matcha ResnetBlock1D / Block1D   third_party/Matcha-TTS/matcha/models/components/decoder.py:32-61
//                        (Conv1d k3 pad 1 -> GroupNorm(8) -> Mish, time-MLP bias, 1x1 residual conv)
//   cv1_block1d       <- matcha Block1D alone (the decoder's final_block)                                 decoder.py:32-43
//   cv1_downsample    <- Downsample1D  Conv1d(C, C, 3, stride 2, pad 1)            decoder.py:64-70
//   cv1_upsample      <- Upsample1D    ConvTranspose1d(C, C, 4, stride 2, pad 1)   decoder.py:120-158
//
// Dense [B, C, T] fp32 tensors at the boundary (the reference's layout); inside, activations are channels-last bf16
// rows with zero guard rows between utterances, so a conv tap is a row-shifted K-slab of one matrix (gemm.h).
//   * symmetric k3: taps at row shifts -1 / 0 / +1.
//   * stride 2: utterance b occupies 2 * S_out input rows starting with one zero row, and the A operand is the
//     overlapping-window view (row stride 2C, 3C columns): view row b * S_out + m = [x[2m-1] | x[2m] | x[2m+1]], one
//     K = 3C segment, no strided gather.
//   * ConvTranspose k4 s2 p1: polyphase form, one GEMM with N = 2C whose output row q holds frames 2q and 2q + 1.
//   * GroupNorm needs statistics over (32 channels x T frames) of an utterance, which no thread == row epilogue can
//     see: the conv GEMM leaves fp32 through its general epilogue, gn_stats_kernel reduces (fp32 per thread over <= 64
//     rows, warp == group shuffle, one fp64 atomic pair per warp) and gn_apply_kernel normalises + Mish + mask (+ time
//     bias, + residual).  Both are HBM-bound: 1 KB read per row for the statistics, 1 KB read + 0.5-2 KB written per
//     row for the apply.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>

#include <string>

#include "engine.h"

namespace mossdec {

namespace {

constexpr int kC = 256;       // block width (channels=[256, 256]); GroupNorm(8) -> 32 channels == one warp per group
constexpr int kGroups = 8;

__device__ __forceinline__ float mish_exact(float x) {
    // torch.nn.Mish: x * tanh(softplus(x)), softplus threshold 20
    const float sp = x > 20.f ? x : log1pf(expf(x));
    return x * tanhf(sp);
}

// Same form as the GEMM epilogue's Mish (epilogue.cuh fast_mish): u = e^x, tanh(log(1 + u)) = (u^2 + 2u) / (u^2 + 2u + 2); one
// ex2.approx + one rcp.approx per element, no range fix-up branches (the libm form made the apply kernels SFU / issue bound:
// 83 us for 21 M elements, profiles/r01_cv1_launches_summary.txt).
__device__ __forceinline__ float mish_fast(float x) {
    float u, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(u) : "f"(fminf(x, 20.f) * 1.4426950408889634f));
    const float n = u * (u + 2.f);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(n + 2.f));
    return x > 20.f ? x : x * (n * r);
}

// [B, C, T] fp32 (x mask) -> channels-last bf16 rows; row of (b, t) = b * S + g + t.  32 x 32 smem transpose.
__global__ void cv1_pack_kernel(const float* __restrict__ x, const float* __restrict__ mask, __nv_bfloat16* __restrict__ out,
                                int C, int Cpad, int T, int S, int g) {
    // tile of 64 channels x 32 frames; block (32, 8); Cpad % 64 == 0.  Stores are bf16x2: 128 B per warp and row.
    __shared__ float tile[64][33];
    const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 64;
    const int tr = t0 + threadIdx.x;
    const float m = (mask && tr < T) ? mask[(size_t)b * T + tr] : 1.f;
    for (int j = threadIdx.y; j < 64; j += 8) {
        const int c = c0 + j;
        tile[j][threadIdx.x] = (c < C && tr < T) ? x[((size_t)b * C + c) * T + tr] * m : 0.f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += 8) {
        const int t = t0 + j, c = c0 + 2 * threadIdx.x;
        if (t < T)
            *reinterpret_cast<__nv_bfloat162*>(out + ((size_t)b * S + g + t) * Cpad + c) =
                __floats2bfloat162_rn(tile[2 * threadIdx.x][j], tile[2 * threadIdx.x + 1][j]);
    }
}

// rows fp32 [*, ld] -> [B, C, T_out] fp32 with T_out = sfac * Tq: out[b, c, t] = rows[b * S + g + t / sfac, (t % sfac) * C + c]
__global__ void cv1_unpack_kernel(const float* __restrict__ rows, float* __restrict__ out, int C, int ld, int Tq, int sfac,
                                  int S, int g, int T_out) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, q0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int r = 0; r < sfac; ++r) {
        for (int j = threadIdx.y; j < 32; j += blockDim.y) {
            const int q = q0 + j, c = c0 + threadIdx.x;
            tile[j][threadIdx.x] = (q < Tq && c < C) ? rows[((size_t)b * S + g + q) * ld + (size_t)r * C + c] : 0.f;
        }
        __syncthreads();
        for (int j = threadIdx.y; j < 32; j += blockDim.y) {
            const int c = c0 + j, q = q0 + threadIdx.x, t = q * sfac + r;
            if (c < C && q < Tq && t < T_out) out[((size_t)b * C + c) * T_out + t] = tile[threadIdx.x][j];
        }
        __syncthreads();
    }
}

__global__ void cv1_rowseq_kernel(int32_t* __restrict__ row_seq, int Mp, int B, int S, int g, int T) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= Mp) return;
    const int b = r / S, t = r - b * S - g;
    row_seq[r] = (b < B && t >= 0 && t < T) ? b : -1;
}

// tbias[b, n] = bias[n] + sum_k w[n, k] * mish(temb[b, k])     (ResnetBlock1D.mlp, decoder.py:49,58)
// grid (n_out / 32, B), 256 threads: mish(temb[b]) once per block into smem, then 4 outputs per warp with coalesced weight rows.
__global__ void __launch_bounds__(256) cv1_time_mlp_kernel(const float* __restrict__ temb, const float* __restrict__ w,
                                                            const float* __restrict__ bias, float* __restrict__ tbias, int dim,
                                                            int n_out) {
    extern __shared__ float tm[];
    const int b = blockIdx.y, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < dim; k += 256) tm[k] = mish_exact(temb[(size_t)b * dim + k]);
    __syncthreads();
    for (int i = 0; i < 4; ++i) {
        const int n = blockIdx.x * 32 + wid * 4 + i;
        if (n >= n_out) break;
        const float* wr = w + (size_t)n * dim;
        float acc = 0.f;
        for (int k = lane; k < dim; k += 32) acc = fmaf(wr[k], tm[k], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) tbias[(size_t)b * n_out + n] = acc + bias[n];
    }
}

// GroupNorm statistics of h [rows, 256] fp32: stats[b, g] = {sum, sum of squares} over the T valid rows of utterance b.
// grid (ceil(T / 64), B), 256 threads: thread == channel (coalesced 1 KB rows), warp == group.
__global__ void __launch_bounds__(kC) gn_stats_kernel(const float* __restrict__ h, double* __restrict__ stats, int T, int S, int g) {
    const int b = blockIdx.y, t0 = blockIdx.x * 64, t1 = min(T, t0 + 64);
    const int c = threadIdx.x;
    const float* p = h + ((size_t)b * S + g + t0) * kC + c;
    float s = 0.f, ss = 0.f;
    for (int t = t0; t < t1; ++t, p += kC) {
        const float v = *p;
        s += v;
        ss = fmaf(v, v, ss);
    }
    double ds = s, dss = ss;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ds += __shfl_xor_sync(0xffffffffu, ds, o);
        dss += __shfl_xor_sync(0xffffffffu, dss, o);
    }
    if ((c & 31) == 0) {
        double* d = stats + ((size_t)b * kGroups + (c >> 5)) * 2;
        atomicAdd(d, ds);
        atomicAdd(d + 1, dss);
    }
}

struct GnCoef {
    float a[4], sh[4];
};
// per-thread affine of 4 consecutive channels c..c+3 (one group): y = h * a + sh
__device__ __forceinline__ GnCoef gn_coef(const double* __restrict__ stats, const float* __restrict__ gamma,
                                          const float* __restrict__ beta, int b, int c, int T, float eps) {
    const double* d = stats + ((size_t)b * kGroups + (c >> 5)) * 2;
    const double n = 32.0 * (double)T;
    const double mean = d[0] / n;
    double var = d[1] / n - mean * mean;
    if (var < 0.0) var = 0.0;
    const float rstd = (float)(1.0 / sqrt(var + (double)eps)), fm = (float)mean;
    const float4 gm = *reinterpret_cast<const float4*>(gamma + c), bt = *reinterpret_cast<const float4*>(beta + c);
    GnCoef k;
    k.a[0] = rstd * gm.x; k.a[1] = rstd * gm.y; k.a[2] = rstd * gm.z; k.a[3] = rstd * gm.w;
    k.sh[0] = bt.x - fm * k.a[0]; k.sh[1] = bt.y - fm * k.a[1]; k.sh[2] = bt.z - fm * k.a[2]; k.sh[3] = bt.w - fm * k.a[3];
    return k;
}

// v = mish(GroupNorm(h)) * mask  [+ tbias[b, c], result * mask: the input of block2]  [+ resid]
// -> bf16 channels-last (next conv operand) and / or fp32 channels-last.  grid (ceil(T / 32), B), 256 threads: a thread owns 4
// consecutive channels (16 B loads, 8 B bf16 stores), 64 threads per row, 4 rows per pass.
__global__ void __launch_bounds__(kC) gn_apply_kernel(const float* __restrict__ h, const double* __restrict__ stats,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       const float* __restrict__ mask, const float* __restrict__ tbias,
                                                       const float* __restrict__ resid, __nv_bfloat16* __restrict__ out_bf16,
                                                       float* __restrict__ out_f32, int T, int S, int g, float eps) {
    const int b = blockIdx.y, c = (threadIdx.x & 63) * 4, r = threadIdx.x >> 6;
    const GnCoef k = gn_coef(stats, gamma, beta, b, c, T, eps);
    float tb[4] = {0.f, 0.f, 0.f, 0.f};
    if (tbias) {
        const float4 t4 = *reinterpret_cast<const float4*>(tbias + (size_t)b * kC + c);
        tb[0] = t4.x; tb[1] = t4.y; tb[2] = t4.z; tb[3] = t4.w;
    }
    const int t0 = blockIdx.x * 32, t1 = min(T, t0 + 32);
    for (int t = t0 + r; t < t1; t += 4) {
        const size_t idx = ((size_t)b * S + g + t) * kC + c;
        const float m = mask ? mask[(size_t)b * T + t] : 1.f;
        const float4 h4 = *reinterpret_cast<const float4*>(h + idx);
        float v[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[i] = mish_fast(fmaf(v[i], k.a[i], k.sh[i])) * m;
            if (tbias) v[i] = (v[i] + tb[i]) * m;
        }
        if (resid) {
            const float4 r4 = *reinterpret_cast<const float4*>(resid + idx);
            v[0] += r4.x; v[1] += r4.y; v[2] += r4.z; v[3] += r4.w;
        }
        if (out_bf16) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(out_bf16 + idx) = pk;
        }
        if (out_f32) *reinterpret_cast<float4*>(out_f32 + idx) = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// Last stage of a ResnetBlock1D: out[b, c, t] = mish(GroupNorm(h)) * mask + resid, written straight in the reference's
// channel-first layout through a 32-frame smem tile (no fp32 channels-last round trip).  grid (ceil(T / 32), B), 256 threads.
// Tile column of channel c = (c & 3) * 64 + (c >> 2): conflict-free for the channel-quad writes and for the frame-major reads.
__global__ void __launch_bounds__(kC) gn_apply_bct_kernel(const float* __restrict__ h, const double* __restrict__ stats,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           const float* __restrict__ mask, const float* __restrict__ resid,
                                                           float* __restrict__ out, int T, int S, int g, float eps) {
    __shared__ float tile[32][kC + 1];
    const int b = blockIdx.y, q = threadIdx.x & 63, c = q * 4, r = threadIdx.x >> 6;
    const GnCoef k = gn_coef(stats, gamma, beta, b, c, T, eps);
    const int t0 = blockIdx.x * 32, nt = min(32, T - t0);
    for (int j = r; j < nt; j += 4) {
        const size_t idx = ((size_t)b * S + g + t0 + j) * kC + c;
        const float m = mask ? mask[(size_t)b * T + t0 + j] : 1.f;
        const float4 h4 = *reinterpret_cast<const float4*>(h + idx);
        float v[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = mish_fast(fmaf(v[i], k.a[i], k.sh[i])) * m;
        if (resid) {
            const float4 r4 = *reinterpret_cast<const float4*>(resid + idx);
            v[0] += r4.x; v[1] += r4.y; v[2] += r4.z; v[3] += r4.w;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) tile[j][i * 64 + q] = v[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane < nt) {
        for (int cc = w * 32; cc < w * 32 + 32; ++cc)
            out[((size_t)b * kC + cc) * T + t0 + lane] = tile[lane][(cc & 3) * 64 + (cc >> 2)];
    }
}

// ---- glue of the assembled two-level decoder (cv1_estimator below), all on dense [B, C, T] fp32 tensors
// xin [B, 4 * 80, T] = x | mu | spks (broadcast over t) | cond      (decoder.py:230-236 pack)
__global__ void cv1_cat4_kernel(const float* __restrict__ x, const float* __restrict__ mu, const float* __restrict__ spks,
                                const float* __restrict__ cond, float* __restrict__ out, int B, int F, int T) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)B * 4 * F * T) return;
    const int t = (int)(idx % T), c = (int)((idx / T) % (4 * F)), b = (int)(idx / ((int64_t)T * 4 * F));
    const int part = c / F, cc = c % F;
    const size_t o = ((size_t)b * F + cc) * T + t;
    float v = 0.f;
    if (part == 0) v = x[o];
    else if (part == 1) v = mu[o];
    else if (part == 2) v = spks ? spks[(size_t)b * F + cc] : 0.f;
    else v = cond ? cond[o] : 0.f;
    out[idx] = v;
}
// out [B, 2C, T] = a[:, :, :T] (a has Ta >= T frames) | skip [B, C, T]      (decoder.py:272 pack([x[:, :, :T], skip]))
__global__ void cv1_cat2_kernel(const float* __restrict__ a, int Ta, const float* __restrict__ skip, float* __restrict__ out,
                                int B, int C, int T) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)B * 2 * C * T) return;
    const int t = (int)(idx % T), c = (int)((idx / T) % (2 * C)), b = (int)(idx / ((int64_t)T * 2 * C));
    out[idx] = c < C ? a[((size_t)b * C + c) * Ta + t] : skip[((size_t)b * C + (c - C)) * T + t];
}
// m1[b, t] = m0[b, 2 t]      (decoder.py:252 masks.append(mask_down[:, :, ::2]))
__global__ void cv1_mask_stride2_kernel(const float* __restrict__ m0, float* __restrict__ m1, int B, int T, int T1) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * T1) return;
    m1[idx] = m0[(size_t)(idx / T1) * T + 2 * (idx % T1)];
}
__global__ void cv1_mask_mul_kernel(float* __restrict__ x, const float* __restrict__ mask, int B, int C, int T) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)B * C * T) return;
    x[idx] *= mask[(size_t)(idx / ((int64_t)C * T)) * T + idx % T];
}

// Entry of a transformer group: dense [B, 256, T] fp32 -> the packed-row fp32 residual stream (lane-major tiled layout,
// gemm.h) + plain LayerNorm -> bf16 operand of the first q|k|v GEMM (norm1's affine is folded into its weights).
// grid (ceil(T / 32), B), 256 threads: 32-frame x 256-channel tile through shared memory, then one warp per row.
__global__ void __launch_bounds__(256) cv1_tf_pack_kernel(const float* __restrict__ x, int T, const int32_t* __restrict__ seq_start,
                                                           const int32_t* __restrict__ seq_len, float* __restrict__ hres,
                                                           __nv_bfloat16* __restrict__ xn, float eps) {
    __shared__ float tile[32][kC + 1];
    const int b = blockIdx.y, t0 = blockIdx.x * 32, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int len = seq_len[b], start = seq_start[b];
    for (int c = w; c < kC; c += 8) tile[lane][c] = (t0 + lane < T) ? x[((size_t)b * kC + c) * T + t0 + lane] : 0.f;
    __syncthreads();
    for (int j = w * 4; j < w * 4 + 4; ++j) {
        const int t = t0 + j;
        if (t >= len) continue;
        float v[8], s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            v[k] = tile[j][lane + 32 * k];
            s += v[k];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mean = s * (1.f / kC);
        float q = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) q = fmaf(v[k] - mean, v[k] - mean, q);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
        const float rstd = rsqrtf(q * (1.f / kC) + eps);
        const int row = start + t;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int c = lane + 32 * k;
            hres[tiled_f32_offset(row, c)] = v[k];
            xn[(size_t)row * kC + c] = __float2bfloat16((v[k] - mean) * rstd);
        }
    }
}
// Exit of a transformer group: tiled fp32 stream -> dense [B, 256, T] fp32 (zeros on padded frames)
__global__ void __launch_bounds__(256) cv1_tf_unpack_kernel(const float* __restrict__ hres, int T, const int32_t* __restrict__ seq_start,
                                                             const int32_t* __restrict__ seq_len, float* __restrict__ out) {
    __shared__ float tile[32][kC + 1];
    const int b = blockIdx.y, t0 = blockIdx.x * 32, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int len = seq_len[b], start = seq_start[b];
    for (int j = w * 4; j < w * 4 + 4; ++j) {
        const int t = t0 + j;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int c = lane + 32 * k;
            tile[j][c] = t < len ? hres[tiled_f32_offset(start + t, c)] : 0.f;
        }
    }
    __syncthreads();
    for (int c = w; c < kC; c += 8)
        if (t0 + lane < T) out[((size_t)b * kC + c) * T + t0 + lane] = tile[lane][c];
}

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

// ResnetBlock1D of the non-causal decoder (decoder.py:46-61): out = block2(block1(x, mask) + mlp(temb), mask) + res_conv(x * mask)
int Engine::cv1_resnet_block(const char* prefix, int B, int T, int Cin, const float* x, const float* mask, const float* temb,
                             int time_dim, float* out, cudaStream_t st) {
    DeviceGuard guard_(device_);
    if (B <= 0 || T <= 0 || Cin <= 0) return fail("cv1_resnet_block: bad shape");
    const std::string P(prefix);
    missing_.clear();
    const __nv_bfloat16 *w1 = W<__nv_bfloat16>(P + ".c1.w"), *w2 = W<__nv_bfloat16>(P + ".c2.w"), *wr = W<__nv_bfloat16>(P + ".res.w");
    const float *b1 = W<float>(P + ".c1.b"), *b2 = W<float>(P + ".c2.b"), *br = W<float>(P + ".res.b");
    const float *g1 = W<float>(P + ".gn1.g"), *be1 = W<float>(P + ".gn1.b"), *g2 = W<float>(P + ".gn2.g"), *be2 = W<float>(P + ".gn2.b");
    const float *wm = W<float>(P + ".mlp.w"), *bm = W<float>(P + ".mlp.b");
    if (!missing_.empty()) return fail("cv1_resnet_block: missing tensors: " + missing_);
    const DevTensor* tw1 = find(P + ".c1.w");
    const int cin_pad = (Cin + kGemmBK - 1) / kGemmBK * kGemmBK;
    if (tw1->shape.size() != 2 || tw1->shape[0] != kC || tw1->shape[1] != 3 * cin_pad)
        return fail("cv1_resnet_block: " + P + ".c1.w must be [256, 3 * ceil64(Cin)]");

    const int S = T + 2, g = 1;
    const int64_t rows = (int64_t)B * S;
    const int Mp = (int)((rows + kGemmBM - 1) / kGemmBM * kGemmBM);
    // workspace: xa bf16 [Mp, cin_pad] | a1 bf16 [Mp, 256] | h f32 [Mp, 256] | res f32 [Mp, 256] | row_seq | tbias | stats
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t o_xa = take((size_t)(Mp + 2) * cin_pad * 2), o_a1 = take((size_t)(Mp + 2) * kC * 2), o_h = take((size_t)Mp * kC * 4),
                 o_res = take((size_t)Mp * kC * 4), o_rs = take((size_t)Mp * 4),
                 o_tb = take((size_t)B * kC * 4), o_st = take((size_t)B * kGroups * 2 * 8 * 2);
    ws_cv1_.reserve(off);
    char* base = ws_cv1_.as<char>();
    if (!base) return fail("cv1_resnet_block: workspace allocation failed");
    auto* xa = reinterpret_cast<__nv_bfloat16*>(base + o_xa);
    auto* a1 = reinterpret_cast<__nv_bfloat16*>(base + o_a1);
    float *h = reinterpret_cast<float*>(base + o_h), *res = reinterpret_cast<float*>(base + o_res);
    auto* row_seq = reinterpret_cast<int32_t*>(base + o_rs);
    float* tbias = reinterpret_cast<float*>(base + o_tb);
    double* stats = reinterpret_cast<double*>(base + o_st);

    cudaMemsetAsync(xa, 0, (size_t)(Mp + 2) * cin_pad * 2, st);
    cudaMemsetAsync(a1, 0, (size_t)(Mp + 2) * kC * 2, st);
    cudaMemsetAsync(stats, 0, (size_t)B * kGroups * 2 * 8 * 2, st);
    cv1_rowseq_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(row_seq, Mp, B, S, g, T);
    cv1_pack_kernel<<<dim3((T + 31) / 32, cin_pad / 64, B), dim3(32, 8), 0, st>>>(x, mask, xa, Cin, cin_pad, T, S, g);
    launches_ += 2;
    if (temb) {
        cv1_time_mlp_kernel<<<dim3(kC / 32, B), 256, (size_t)time_dim * 4, st>>>(temb, wm, bm, tbias, time_dim, kC);
        ++launches_;
    }

    auto conv3 = [&](const __nv_bfloat16* A, int cp, const __nv_bfloat16* w, const float* bias, float* dst) {
        GemmProblem p;
        p.a[0].ptr = A; p.a[0].rows = Mp; p.a[0].cols = cp; p.a[0].ld = cp;
        p.num_segs = 3;
        for (int j = 0; j < 3; ++j) p.segs[j] = {0, j - 1, 0, cp / kGemmBK};
        p.M = Mp; p.N = kC; p.BN = kC; p.w = w;
        p.epi.bias = bias; p.epi.row_seq = row_seq;
        p.epi.out_f32 = dst; p.epi.ldf = kC;
        return gemm(p, st);
    };
    const dim3 gs((T + 63) / 64, B), ga((T + 31) / 32, B);
    // block1: conv -> GroupNorm -> Mish -> mask, + time bias; a1 = block2's masked input
    if (conv3(xa, cin_pad, w1, b1, h)) return -1;
    gn_stats_kernel<<<gs, kC, 0, st>>>(h, stats, T, S, g);
    gn_apply_kernel<<<ga, kC, 0, st>>>(h, stats, g1, be1, mask, temb ? tbias : nullptr, nullptr, a1, nullptr, T, S, g, 1e-5f);
    // residual 1x1 conv on x * mask
    {
        GemmProblem p;
        p.a[0].ptr = xa; p.a[0].rows = Mp; p.a[0].cols = cin_pad; p.a[0].ld = cin_pad;
        p.num_segs = 1;
        p.segs[0] = {0, 0, 0, cin_pad / kGemmBK};
        p.M = Mp; p.N = kC; p.BN = kC; p.w = wr;
        p.epi.bias = br; p.epi.row_seq = row_seq;
        p.epi.out_f32 = res; p.epi.ldf = kC;
        if (gemm(p, st)) return -1;
    }
    // block2 + residual
    if (conv3(a1, kC, w2, b2, h)) return -1;
    double* stats2 = stats + (size_t)B * kGroups * 2;
    gn_stats_kernel<<<gs, kC, 0, st>>>(h, stats2, T, S, g);
    gn_apply_bct_kernel<<<dim3((T + 31) / 32, B), kC, 0, st>>>(h, stats2, g2, be2, mask, res, out, T, S, g, 1e-5f);
    launches_ += 4;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("cv1_resnet_block: ") + cudaGetErrorString(e));
    return 0;
}

// Block1D of the non-causal decoder (decoder.py:32-43; also its `final_block`): out = mish(GroupNorm(conv3(x * mask))) * mask
int Engine::cv1_block1d(const char* prefix, int B, int T, int Cin, const float* x, const float* mask, float* out, cudaStream_t st) {
    DeviceGuard guard_(device_);
    if (B <= 0 || T <= 0 || Cin <= 0) return fail("cv1_block1d: bad shape");
    const std::string P(prefix);
    missing_.clear();
    const __nv_bfloat16* w = W<__nv_bfloat16>(P + ".c.w");
    const float *bias = W<float>(P + ".c.b"), *gg = W<float>(P + ".gn.g"), *gb = W<float>(P + ".gn.b");
    if (!missing_.empty()) return fail("cv1_block1d: missing tensors: " + missing_);
    const int cin_pad = (Cin + kGemmBK - 1) / kGemmBK * kGemmBK;
    const DevTensor* tw = find(P + ".c.w");
    if (tw->shape.size() != 2 || tw->shape[0] != kC || tw->shape[1] != 3 * cin_pad)
        return fail("cv1_block1d: " + P + ".c.w must be [256, 3 * ceil64(Cin)]");
    const int S = T + 2, g = 1;
    const int64_t rows = (int64_t)B * S;
    const int Mp = (int)((rows + kGemmBM - 1) / kGemmBM * kGemmBM);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t o_xa = take((size_t)(Mp + 2) * cin_pad * 2), o_h = take((size_t)Mp * kC * 4), o_rs = take((size_t)Mp * 4),
                 o_st = take((size_t)B * kGroups * 2 * 8);
    ws_cv1_.reserve(off);
    char* base = ws_cv1_.as<char>();
    if (!base) return fail("cv1_block1d: workspace allocation failed");
    auto* xa = reinterpret_cast<__nv_bfloat16*>(base + o_xa);
    float* h = reinterpret_cast<float*>(base + o_h);
    auto* row_seq = reinterpret_cast<int32_t*>(base + o_rs);
    double* stats = reinterpret_cast<double*>(base + o_st);
    cudaMemsetAsync(xa, 0, (size_t)(Mp + 2) * cin_pad * 2, st);
    cudaMemsetAsync(stats, 0, (size_t)B * kGroups * 2 * 8, st);
    cv1_rowseq_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(row_seq, Mp, B, S, g, T);
    cv1_pack_kernel<<<dim3((T + 31) / 32, cin_pad / 64, B), dim3(32, 8), 0, st>>>(x, mask, xa, Cin, cin_pad, T, S, g);
    GemmProblem p;
    p.a[0].ptr = xa; p.a[0].rows = Mp; p.a[0].cols = cin_pad; p.a[0].ld = cin_pad;
    p.num_segs = 3;
    for (int j = 0; j < 3; ++j) p.segs[j] = {0, j - 1, 0, cin_pad / kGemmBK};
    p.M = Mp; p.N = kC; p.BN = kC; p.w = w;
    p.epi.bias = bias; p.epi.row_seq = row_seq;
    p.epi.out_f32 = h; p.epi.ldf = kC;
    if (gemm(p, st)) return -1;
    gn_stats_kernel<<<dim3((T + 63) / 64, B), kC, 0, st>>>(h, stats, T, S, g);
    gn_apply_bct_kernel<<<dim3((T + 31) / 32, B), kC, 0, st>>>(h, stats, gg, gb, mask, nullptr, out, T, S, g, 1e-5f);
    launches_ += 4;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("cv1_block1d: ") + cudaGetErrorString(e));
    return 0;
}

// Downsample1D (decoder.py:64-70): Conv1d(C, C, 3, stride 2, padding 1); out [B, C, (T - 1) / 2 + 1]
int Engine::cv1_downsample(const char* prefix, int B, int T, int C, const float* x, float* out, cudaStream_t st) {
    DeviceGuard guard_(device_);
    if (B <= 0 || T <= 0 || C <= 0 || (C % kGemmBK) != 0 || C > 256) return fail("cv1_downsample: C must be a multiple of 64, <= 256");
    const std::string P(prefix);
    missing_.clear();
    const __nv_bfloat16* w = W<__nv_bfloat16>(P + ".w");
    const float* bias = W<float>(P + ".b");
    if (!missing_.empty()) return fail("cv1_downsample: missing tensors: " + missing_);
    const int T_out = (T - 1) / 2 + 1;
    const int S_out = T_out + 1, S_in = 2 * S_out;   // input row of (b, t) = b * S_in + 1 + t, the rest zero
    const int64_t rows = (int64_t)B * S_out;
    const int Mp = (int)((rows + kGemmBM - 1) / kGemmBM * kGemmBM);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t in_bytes = (size_t)(2 * Mp + 2) * C * 2;
    const size_t o_xa = take(in_bytes), o_y = take((size_t)Mp * C * 4), o_rs = take((size_t)Mp * 4);
    ws_cv1_.reserve(off);
    char* base = ws_cv1_.as<char>();
    if (!base) return fail("cv1_downsample: workspace allocation failed");
    auto* xa = reinterpret_cast<__nv_bfloat16*>(base + o_xa);
    float* y = reinterpret_cast<float*>(base + o_y);
    auto* row_seq = reinterpret_cast<int32_t*>(base + o_rs);
    cudaMemsetAsync(xa, 0, in_bytes, st);
    cv1_rowseq_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(row_seq, Mp, B, S_out, 0, T_out);
    cv1_pack_kernel<<<dim3((T + 31) / 32, C / 64, B), dim3(32, 8), 0, st>>>(x, nullptr, xa, C, C, T, S_in, 1);
    launches_ += 2;
    GemmProblem p;
    p.a[0].ptr = xa; p.a[0].rows = Mp; p.a[0].cols = 3 * C; p.a[0].ld = 2 * C;   // overlapping windows [x[2m-1] | x[2m] | x[2m+1]]
    p.num_segs = 1;
    p.segs[0] = {0, 0, 0, 3 * C / kGemmBK};
    p.M = Mp; p.N = C; p.BN = C; p.w = w;
    p.epi.bias = bias; p.epi.row_seq = row_seq;
    p.epi.out_f32 = y; p.epi.ldf = C;
    if (gemm(p, st)) return -1;
    cv1_unpack_kernel<<<dim3((T_out + 31) / 32, (C + 31) / 32, B), dim3(32, 8), 0, st>>>(y, out, C, C, T_out, 1, S_out, 0, T_out);
    ++launches_;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("cv1_downsample: ") + cudaGetErrorString(e));
    return 0;
}

// Upsample1D (decoder.py:120-158, use_conv_transpose): ConvTranspose1d(C, C, 4, stride 2, padding 1); out [B, C, 2 T]
int Engine::cv1_upsample(const char* prefix, int B, int T, int C, const float* x, float* out, cudaStream_t st) {
    DeviceGuard guard_(device_);
    if (B <= 0 || T <= 0 || C <= 0 || (C % kGemmBK) != 0 || C > 256) return fail("cv1_upsample: C must be a multiple of 64, <= 256");
    const std::string P(prefix);
    missing_.clear();
    const __nv_bfloat16* w = W<__nv_bfloat16>(P + ".w");   // polyphase [2C, 3C] (weights.py convtranspose_polyphase)
    const float* bias2 = W<float>(P + ".b");               // [2C]: the bias repeated per phase
    if (!missing_.empty()) return fail("cv1_upsample: missing tensors: " + missing_);
    const int S = T + 2, g = 1;
    const int64_t rows = (int64_t)B * S;
    const int Mp = (int)((rows + kGemmBM - 1) / kGemmBM * kGemmBM);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t in_bytes = (size_t)(Mp + 2) * C * 2;
    const size_t o_xa = take(in_bytes), o_y = take((size_t)Mp * 2 * C * 4), o_rs = take((size_t)Mp * 4);
    ws_cv1_.reserve(off);
    char* base = ws_cv1_.as<char>();
    if (!base) return fail("cv1_upsample: workspace allocation failed");
    auto* xa = reinterpret_cast<__nv_bfloat16*>(base + o_xa);
    float* y = reinterpret_cast<float*>(base + o_y);
    auto* row_seq = reinterpret_cast<int32_t*>(base + o_rs);
    cudaMemsetAsync(xa, 0, in_bytes, st);
    cv1_rowseq_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(row_seq, Mp, B, S, g, T);
    cv1_pack_kernel<<<dim3((T + 31) / 32, C / 64, B), dim3(32, 8), 0, st>>>(x, nullptr, xa, C, C, T, S, g);
    launches_ += 2;
    GemmProblem p;
    p.a[0].ptr = xa; p.a[0].rows = Mp; p.a[0].cols = C; p.a[0].ld = C;
    p.num_segs = 3;
    for (int j = 0; j < 3; ++j) p.segs[j] = {0, j - 1, 0, C / kGemmBK};
    p.M = Mp; p.N = 2 * C; p.BN = C; p.w = w;
    p.epi.bias = bias2; p.epi.row_seq = row_seq;
    p.epi.out_f32 = y; p.epi.ldf = 2 * C;
    if (gemm(p, st)) return -1;
    cv1_unpack_kernel<<<dim3((T + 31) / 32, (C + 31) / 32, B), dim3(32, 8), 0, st>>>(y, out, C, 2 * C, T, 2, S, g, 2 * T);
    ++launches_;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("cv1_upsample: ") + cudaGetErrorString(e));
    return 0;
}

// Plain Conv1d of the decoder with "same" padding (k = 3: the `downsample` / `upsample` slots of the last level,
// decoder.py:137-139,182-186; k = 1: final_proj, :188): out [B, Cout, T] = conv(x * mask_in) (* mask_out)
int Engine::cv1_conv(const char* prefix, int B, int T, int Cin, int Cout, int k, const float* x, const float* mask_in,
                     const float* mask_out, float* out, cudaStream_t st) {
    DeviceGuard guard_(device_);
    if (B <= 0 || T <= 0 || (k != 1 && k != 3) || Cout > 256 || (Cout % 16) != 0) return fail("cv1_conv: bad shape");
    const std::string P(prefix);
    missing_.clear();
    const __nv_bfloat16* w = W<__nv_bfloat16>(P + ".w");
    const float* bias = W<float>(P + ".b");
    if (!missing_.empty()) return fail("cv1_conv: missing tensors: " + missing_);
    const int cin_pad = (Cin + kGemmBK - 1) / kGemmBK * kGemmBK;
    const int S = T + 2, g = 1;
    const int Mp = (int)(((int64_t)B * S + kGemmBM - 1) / kGemmBM * kGemmBM);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t o_xa = take((size_t)(Mp + 2) * cin_pad * 2), o_y = take((size_t)Mp * Cout * 4), o_rs = take((size_t)Mp * 4);
    ws_cv1_.reserve(off);
    char* base = ws_cv1_.as<char>();
    if (!base) return fail("cv1_conv: workspace allocation failed");
    auto* xa = reinterpret_cast<__nv_bfloat16*>(base + o_xa);
    float* y = reinterpret_cast<float*>(base + o_y);
    auto* row_seq = reinterpret_cast<int32_t*>(base + o_rs);
    cudaMemsetAsync(xa, 0, (size_t)(Mp + 2) * cin_pad * 2, st);
    cv1_rowseq_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(row_seq, Mp, B, S, g, T);
    cv1_pack_kernel<<<dim3((T + 31) / 32, cin_pad / 64, B), dim3(32, 8), 0, st>>>(x, mask_in, xa, Cin, cin_pad, T, S, g);
    GemmProblem p;
    p.a[0].ptr = xa; p.a[0].rows = Mp; p.a[0].cols = cin_pad; p.a[0].ld = cin_pad;
    p.num_segs = k;
    for (int j = 0; j < k; ++j) p.segs[j] = {0, j - k / 2, 0, cin_pad / kGemmBK};
    p.M = Mp; p.N = Cout; p.BN = Cout; p.w = w;
    p.epi.bias = bias; p.epi.row_seq = row_seq;
    p.epi.out_f32 = y; p.epi.ldf = Cout;
    if (gemm(p, st)) return -1;
    cv1_unpack_kernel<<<dim3((T + 31) / 32, (Cout + 31) / 32, B), dim3(32, 8), 0, st>>>(y, out, Cout, Cout, T, 1, S, g, T);
    launches_ += 3;
    if (mask_out) {
        const int64_t n = (int64_t)B * Cout * T;
        cv1_mask_mul_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, mask_out, B, Cout, T);
        ++launches_;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("cv1_conv: ") + cudaGetErrorString(e));
    return 0;
}

// n_blocks BasicTransformerBlocks of one decoder level (decoder.py:238-249: rearrange, key-padding attention bias, blocks,
// rearrange back) on the kernels of the live path: q|k|v GEMM, attention_ws_kernel with chunk = 0 (non-causal: the
// key-padding predicate only), fused out-projection + feed-forward.  lens: HOST valid frames per utterance.
int Engine::cv1_transformer_group(const char* prefix, int n_blocks, int B, int T, const float* x, const int32_t* lens, float* out,
                                  cudaStream_t st) {
    DeviceGuard guard_(device_);
    const std::string P(prefix);
    const int C = cfg_.channels, H = cfg_.heads * cfg_.head_dim;
    if (build_flow_layout(B, lens, /*cfg_double=*/false, st, /*dense_T=*/T)) return -1;
    const int M = flow_M_;
    cv1_tf_pack_kernel<<<dim3((T + 31) / 32, B), 256, 0, st>>>(x, T, layout_.seq_start, layout_.seq_len, hres_, xn_, 1e-5f);
    ++launches_;
    missing_.clear();
    for (int j = 0; j < n_blocks; ++j) {
        char buf[32];
        snprintf(buf, sizeof buf, ".%d.", j);
        const std::string t = P + buf;
        const __nv_bfloat16 *wqkv = W<__nv_bfloat16>(t + "qkv.w"), *wo = W<__nv_bfloat16>(t + "o.w"),
                            *w1 = W<__nv_bfloat16>(t + "ff1.w"), *w2 = W<__nv_bfloat16>(t + "ff2.w");
        const float *bqkv = W<float>(t + "qkv.b"), *bo = W<float>(t + "o.b"), *b1 = W<float>(t + "ff1.b"), *b2 = W<float>(t + "ff2.b");
        if (!missing_.empty()) return fail("cv1_transformer_group: missing tensors: " + missing_);
        {
            GemmProblem p;
            p.a[0].ptr = xn_; p.a[0].rows = M; p.a[0].cols = C; p.a[0].ld = C;
            p.num_segs = 1;
            p.segs[0] = {0, 0, 0, C / kGemmBK};
            p.M = M; p.N = 3 * H; p.BN = 256; p.w = wqkv;
            p.epi.bias = bqkv;
            p.epi.row_seq = layout_.row_seq;
            p.epi.out_bf16 = qkv_; p.epi.ldo = 3 * H;
            if (gemm(p, st)) return -1;
        }
        cudaError_t ae = launch_attention_ws(qkv_, M, att_, layout_.seq_start, layout_.seq_len, B, flow_max_len_, cfg_.heads, 0, st);
        if (ae != cudaSuccess) return fail(std::string("cv1 attention: ") + cudaGetErrorString(ae));
        ++launches_;
        FfnProblem f;
        f.M = M;
        f.att = att_; f.wo = wo; f.bo = bo;
        f.w1 = w1; f.b1 = b1; f.w2 = w2; f.b2 = b2;
        f.hres = hres_;
        f.row_seq = layout_.row_seq;
        f.ln = j + 1 < n_blocks;   // norm1 of the next block (affine folded into its q|k|v weights)
        f.out_bf16 = xn_; f.ldo = C;
        f.store_f32 = 1;
        if (ffn(f, st)) return -1;
    }
    cv1_tf_unpack_kernel<<<dim3((T + 31) / 32, B), 256, 0, st>>>(hres_, T, layout_.seq_start, layout_.seq_len, out);
    ++launches_;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("cv1_transformer_group: ") + cudaGetErrorString(e));
    return 0;
}

// The assembled non-causal decoder: ConditionalDecoder.forward of cosyvoice (flow/decoder.py:206-291; cosyvoice1/flow/decoder.py
// :151-222 is the same graph) with channels = [256, 256]: two resolutions (T and ceil(T / 2)), skip concatenation, GroupNorm
// resnets, non-causal transformer groups, stride-2 down / ConvTranspose up.  Chained from the blocks above on dense
// [B, C, T] fp32 tensors -- every op packs / unpacks its own rows (correct, not fast: see DESIGN.md section 7).
// Tensors: "<p>.time.{w1,b1,w2,b2}", "<p>.{d0,d1,m<i>,u0,u1}.res.*" (pack_cv1_resnet), "<p>.<level>.tf.<j>.*" (q|k|v, o, ff1,
// ff2 as in the live path), "<p>.d0.down.*", "<p>.d1.conv.*", "<p>.u0.up.*", "<p>.u1.conv.*", "<p>.final.*", "<p>.proj.*".
int Engine::cv1_estimator(const char* prefix, int n_blocks, int n_mid, int B, int T, const float* x, const float* mask, const float* mu,
                          const float* t, const float* spks, const float* cond, float* out, cudaStream_t st) {
    DeviceGuard guard_(device_);
    if (B <= 0 || T < 2 || n_blocks <= 0 || n_mid < 0) return fail("cv1_estimator: bad shape");
    const std::string P(prefix);
    const int F = cfg_.n_feats, C = kC, TD = cfg_.time_dim, T1 = (T - 1) / 2 + 1;
    // valid lengths (prefix masks only), read back once: this entry point mirrors a synchronous module call
    std::vector<int32_t> lens0(B), lens1(B);
    {
        if (mask_lens_cap_ < B + 1) {
            if (h_mask_lens_) cudaFreeHost(h_mask_lens_);
            if (d_mask_lens_) cudaFree(d_mask_lens_);
            h_mask_lens_ = nullptr;
            d_mask_lens_ = nullptr;
            mask_lens_cap_ = 0;
            if (cudaMallocHost(reinterpret_cast<void**>(&h_mask_lens_), (size_t)(B + 1) * 4) != cudaSuccess ||
                cudaMalloc(reinterpret_cast<void**>(&d_mask_lens_), (size_t)(B + 1) * 4) != cudaSuccess)
                return fail("cv1_estimator: allocation failed");
            mask_lens_cap_ = B + 1;
        }
        cudaMemsetAsync(d_mask_lens_ + B, 0, 4, st);
        launch_mask_lens(mask, B, T, d_mask_lens_, d_mask_lens_ + B, st);
        cudaMemcpyAsync(h_mask_lens_, d_mask_lens_, (size_t)(B + 1) * 4, cudaMemcpyDeviceToHost, st);
        if (cudaStreamSynchronize(st) != cudaSuccess) return fail("cv1_estimator: sync failed");
        if (h_mask_lens_[B] != 0) return fail("cv1_estimator: mask must be a prefix (padding) mask");
        for (int b = 0; b < B; ++b) {
            lens0[b] = h_mask_lens_[b];
            if (lens0[b] <= 0) return fail("cv1_estimator: a sequence is fully masked");
            lens1[b] = (lens0[b] + 1) / 2;   // mask[:, :, ::2]
        }
        ++launches_;
    }
    // dense intermediates
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
    const size_t nT = (size_t)B * C * T, nT1 = (size_t)B * C * T1;
    const size_t o_xin = take((size_t)B * 4 * F * T * 4), o_temb = take((size_t)B * TD * 4), o_m1 = take((size_t)B * T1 * 4),
                 o_cat = take(2 * nT * 4), o_a = take(nT * 4), o_b = take(nT * 4), o_s0 = take(nT * 4), o_s1 = take(nT1 * 4),
                 o_up = take((size_t)B * C * 2 * T1 * 4);
    ws_cv1e_.reserve(off);
    char* base = ws_cv1e_.as<char>();
    if (!base) return fail("cv1_estimator: workspace allocation failed");
    float *xin = (float*)(base + o_xin), *temb = (float*)(base + o_temb), *m1 = (float*)(base + o_m1), *cat = (float*)(base + o_cat),
          *A = (float*)(base + o_a), *Bf = (float*)(base + o_b), *skip0 = (float*)(base + o_s0), *skip1 = (float*)(base + o_s1),
          *up = (float*)(base + o_up);
    missing_.clear();
    const float *tw1 = W<float>(P + ".time.w1"), *tb1 = W<float>(P + ".time.b1"), *tw2 = W<float>(P + ".time.w2"), *tb2 = W<float>(P + ".time.b2");
    if (!missing_.empty()) return fail("cv1_estimator: missing tensors: " + missing_);
    launch_time_embed(t, B, 4 * F, TD, tw1, tb1, tw2, tb2, temb, st, /*apply_mish=*/0);   // raw: ResnetBlock1D.mlp applies Mish
    {
        const int64_t n = (int64_t)B * 4 * F * T;
        cv1_cat4_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, mu, spks, cond, xin, B, F, T);
        cv1_mask_stride2_kernel<<<(B * T1 + 255) / 256, 256, 0, st>>>(mask, m1, B, T, T1);
    }
    launches_ += 3;
    auto level = [&](const std::string& name, int Cin, int Tl, const float* in, const float* m, const int32_t* lens, float* dst) -> int {
        if (cv1_resnet_block((P + "." + name + ".res").c_str(), B, Tl, Cin, in, m, temb, TD, A, st)) return -1;
        return cv1_transformer_group((P + "." + name + ".tf").c_str(), n_blocks, B, Tl, A, lens, dst, st);
    };
    auto cat2 = [&](const float* a, int Ta, const float* skip, int Tl) {
        const int64_t n = (int64_t)B * 2 * C * Tl;
        cv1_cat2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a, Ta, skip, cat, B, C, Tl);
        ++launches_;
    };
    // down path
    if (level("d0", 4 * F, T, xin, mask, lens0.data(), skip0)) return -1;
    if (cv1_downsample((P + ".d0.down").c_str(), B, T, C, skip0, Bf, st)) return -1;             // x * mask: padded frames are already zero
    if (level("d1", C, T1, Bf, m1, lens1.data(), skip1)) return -1;
    if (cv1_conv((P + ".d1.conv").c_str(), B, T1, C, C, 3, skip1, m1, nullptr, Bf, st)) return -1;
    // mid
    for (int i = 0; i < n_mid; ++i) {
        char nm[16];
        snprintf(nm, sizeof nm, "m%d", i);
        if (level(nm, C, T1, Bf, m1, lens1.data(), Bf)) return -1;
    }
    // up path
    cat2(Bf, T1, skip1, T1);
    if (level("u0", 2 * C, T1, cat, m1, lens1.data(), Bf)) return -1;
    if (cv1_upsample((P + ".u0.up").c_str(), B, T1, C, Bf, up, st)) return -1;
    cat2(up, 2 * T1, skip0, T);
    if (level("u1", 2 * C, T, cat, mask, lens0.data(), Bf)) return -1;
    if (cv1_conv((P + ".u1.conv").c_str(), B, T, C, C, 3, Bf, mask, nullptr, A, st)) return -1;
    if (cv1_block1d((P + ".final").c_str(), B, T, C, A, mask, Bf, st)) return -1;
    return cv1_conv((P + ".proj").c_str(), B, T, C, F, 1, Bf, mask, mask, out, st);
}

}  // namespace mossdec
