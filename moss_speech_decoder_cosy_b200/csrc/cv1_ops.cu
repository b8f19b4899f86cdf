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

}  // namespace mossdec
