// Bandwidth-class kernels of the token -> mu front end (CausalMaskedDiffWithXvec.inference prologue, flow.py:250-270, and
// UpsampleConformerEncoder, cosyvoice/transformer/upsample_encoder.py:244-307).  The contractions run on gemm_tc_kernel and
// attention_ws_kernel<bias>; this file holds what is left: the embedding gather, LayerNorm over 512 channels (one warp
// per row; the GEMM epilogues' fused LayerNorm needs the row inside one 256-column tile), the sinusoidal relative-position
// table, the speaker projection, the prompt-mel scatter and the packed-rows -> ragged gather.
#include "kernels.h"
#include "ptx.cuh"

namespace mossdec {

namespace {

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// out[row, :] = emb[clamp(token, 0), :] for rows that hold a token (incl. look-ahead context), 0 on guard rows
// (flow.py:255-256: input_embedding(clamp(token, min=0)) * mask).  64 threads per row, 8 bf16 each.
__global__ void enc_gather_kernel(EncLayout L, const int64_t* __restrict__ tokens, const __nv_bfloat16* __restrict__ emb,
                                  int vocab, __nv_bfloat16* __restrict__ out) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int row = (int)(idx >> 6), u = (int)(idx & 63);
    if (row >= L.rows) return;
    const int s = L.row_seq_all[row];
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (s >= 0) {
        int64_t tok = tokens[L.tok_off[s] + (row - L.seq_start[s])];
        tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);
        v = __ldg(reinterpret_cast<const uint4*>(emb + (size_t)tok * 512) + u);
    }
    reinterpret_cast<uint4*>(out + (size_t)row * 512)[u] = v;
}

// LayerNorm over 512 channels, one warp per row (16 values per lane, two passes over registers):
//   y = ((x - mean) * rstd * gamma + beta) * scale on rows with row_seq >= 0, 0 elsewhere; out_f32 and / or out_bf16.
__global__ void layernorm512_kernel(const float* __restrict__ x, int rows, const float* __restrict__ gamma,
                                    const float* __restrict__ beta, float eps, float scale,
                                    const int32_t* __restrict__ row_seq, float* __restrict__ out_f32,
                                    __nv_bfloat16* __restrict__ out_bf16) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= rows) return;
    const bool valid = row_seq == nullptr || row_seq[row] >= 0;
    float v[16];
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * 512);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 t = xr[lane + 32 * i];
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += v[i];
    const float mean = warp_sum_f(s) * (1.f / 512.f);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float d = v[i] - mean;
        q = fmaf(d, d, q);
    }
    const float rstd = rsqrtf(warp_sum_f(q) * (1.f / 512.f) + eps);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = (lane + 32 * i) * 4;
        const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c)), b = __ldg(reinterpret_cast<const float4*>(beta + c));
        float4 y;
        y.x = valid ? ((v[4 * i] - mean) * rstd * g.x + b.x) * scale : 0.f;
        y.y = valid ? ((v[4 * i + 1] - mean) * rstd * g.y + b.y) * scale : 0.f;
        y.z = valid ? ((v[4 * i + 2] - mean) * rstd * g.z + b.z) * scale : 0.f;
        y.w = valid ? ((v[4 * i + 3] - mean) * rstd * g.w + b.w) * scale : 0.f;
        if (out_f32) *reinterpret_cast<float4*>(out_f32 + (size_t)row * 512 + c) = y;
        if (out_bf16) {
            uint2 p;
            p.x = pack_bf16x2(y.x, y.y);
            p.y = pack_bf16x2(y.z, y.w);
            *reinterpret_cast<uint2*>(out_bf16 + (size_t)row * 512 + c) = p;
        }
    }
}

// EspnetRelPositionalEncoding table (embedding.py:224-252): row m holds the sinusoid of relative position tmax - 1 - m
// (sin on even, cos on odd channels), m < 2 tmax - 1; zero rows up to `rows` (padding read by masked keys only).
__global__ void rel_pos_table_kernel(int tmax, int rows, float* __restrict__ pe) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * 256) return;
    const int m = idx >> 8, k = idx & 255;
    float sv = 0.f, cv = 0.f;
    if (m < 2 * tmax - 1) {
        const float div = expf((float)(2 * k) * -(9.210340371976184f / 512.f));   // exp(arange(0, d, 2) * -(ln 1e4 / d))
        const float a = (float)(tmax - 1 - m) * div;
        sv = sinf(a);
        cv = cosf(a);
    }
    pe[(size_t)m * 512 + 2 * k] = sv;
    pe[(size_t)m * 512 + 2 * k + 1] = cv;
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ x, int64_t n, __nv_bfloat16* __restrict__ y) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const float4 t = *reinterpret_cast<const float4*>(x + i);
    uint2 p;
    p.x = pack_bf16x2(t.x, t.y);
    p.y = pack_bf16x2(t.z, t.w);
    *reinterpret_cast<uint2*>(y + i) = p;
}

// spks[b, :] = W . (e / max(||e||, 1e-12)) + bias  (flow.py:250-251: F.normalize(dim=1) then Linear(192 -> 80)).  One block per b.
__global__ void spk_project_kernel(const float* __restrict__ emb, int dim_in, const float* __restrict__ w,
                                   const float* __restrict__ bias, int dim_out, float* __restrict__ out) {
    extern __shared__ float se[];
    __shared__ float s_norm;
    const int b = blockIdx.x;
    float acc = 0.f;
    for (int i = threadIdx.x; i < dim_in; i += blockDim.x) {
        const float v = emb[(size_t)b * dim_in + i];
        se[i] = v;
        acc = fmaf(v, v, acc);
    }
    acc = warp_sum_f(acc);
    if (threadIdx.x == 0) s_norm = 0.f;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_norm, acc);
    __syncthreads();
    const float inv = 1.f / fmaxf(sqrtf(s_norm), 1e-12f);
    for (int o = threadIdx.x; o < dim_out; o += blockDim.x) {
        float a = 0.f;
        for (int i = 0; i < dim_in; ++i) a = fmaf(w[(size_t)o * dim_in + i], se[i] * inv, a);
        out[(size_t)b * dim_out + o] = a + bias[o];
    }
}

// cond (ragged, time-major [T_b, 80]) = prompt_feat in the first F_b frames, zero after (flow.py:268-270);
// mu (ragged) = rows [seq_start * 4 + t] of the packed mel-rate matrix.  One thread per (frame, 4 channels).
__global__ void enc_finish_kernel(EncLayout L, const float* __restrict__ mu_rows, const float* __restrict__ prompt_feat,
                                  float* __restrict__ mu, float* __restrict__ cond) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int row = (int)(idx / 20), c = (int)(idx % 20) * 4;   // mel-rate packed row
    if (row >= L.rows * 4) return;
    const int s = L.row_seq_main[row >> 2];
    if (s < 0) return;
    const int t = row - 4 * L.seq_start[s];
    const int64_t o = (L.mel_off[s] + t) * 80 + c;
    *reinterpret_cast<float4*>(mu + o) = *reinterpret_cast<const float4*>(mu_rows + (size_t)row * 80 + c);
    float4 cv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < L.prompt_frames[s]) cv = *reinterpret_cast<const float4*>(prompt_feat + (L.feat_off[s] + t) * 80 + c);
    *reinterpret_cast<float4*>(cond + o) = cv;
}

}  // namespace

void launch_enc_gather(const EncLayout& L, const int64_t* tokens, const __nv_bfloat16* emb, int vocab, __nv_bfloat16* out,
                       cudaStream_t st) {
    const int64_t n = (int64_t)L.rows * 64;
    enc_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(L, tokens, emb, vocab, out);
}
void launch_layernorm512(const float* x, int rows, const float* gamma, const float* beta, float eps, float scale,
                         const int32_t* row_seq, float* out_f32, __nv_bfloat16* out_bf16, cudaStream_t st) {
    layernorm512_kernel<<<(rows * 32 + 255) / 256, 256, 0, st>>>(x, rows, gamma, beta, eps, scale, row_seq, out_f32, out_bf16);
}
void launch_rel_pos_table(int tmax, int rows, float* pe, cudaStream_t st) {
    rel_pos_table_kernel<<<(rows * 256 + 255) / 256, 256, 0, st>>>(tmax, rows, pe);
}
void launch_f32_to_bf16(const float* x, int64_t n, __nv_bfloat16* y, cudaStream_t st) {
    f32_to_bf16_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, st>>>(x, n, y);
}
void launch_spk_project(const float* emb, int B, int dim_in, const float* w, const float* bias, int dim_out, float* out,
                        cudaStream_t st) {
    spk_project_kernel<<<B, 128, dim_in * sizeof(float), st>>>(emb, dim_in, w, bias, dim_out, out);
}
void launch_enc_finish(const EncLayout& L, const float* mu_rows, const float* prompt_feat, float* mu, float* cond,
                       cudaStream_t st) {
    const int64_t n = (int64_t)L.rows * 4 * 20;
    enc_finish_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(L, mu_rows, prompt_feat, mu, cond);
}

}  // namespace mossdec
