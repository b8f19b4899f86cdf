// Bandwidth-class kernels of the flow-matching estimator / Euler solver.
//   time embedding   : matcha/models/components/decoder.py:14-29 (SinusoidalPosEmb), :73-117 (TimestepEmbedding)
//   resnet time bias : matcha/models/components/decoder.py:49 (Mish -> Linear), decoder.py:57
//   input packing    : cosyvoice/flow/decoder.py:425-431 (pack x, mu, spks, cond -> 320 channels)
//   CFG + Euler      : cosyvoice/flow/flow_matching.py:106-125
#include "kernels.h"
#include "ptx.cuh"

namespace mossdec {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One block per distinct timestep value.  temb[s, :] = lin2(silu(lin1(sinusoid(1000 t_s)))).
__global__ void time_embed_kernel(const float* __restrict__ t, int dim_in, int dim_h, const float* __restrict__ w1,
                                  const float* __restrict__ b1, const float* __restrict__ w2,
                                  const float* __restrict__ b2, float* __restrict__ temb_mish, int apply_mish) {
    extern __shared__ float sm[];
    float* pe = sm;           // [dim_in]
    float* h1 = sm + dim_in;  // [dim_h]
    const int s = blockIdx.x;
    const float tv = t[s];
    const int half = dim_in / 2;
    const float step = logf(10000.f) / (float)(half - 1);
    for (int i = threadIdx.x; i < half; i += blockDim.x) {
        float e = expf((float)i * -step);
        float a = 1000.f * tv * e;
        pe[i] = sinf(a);
        pe[half + i] = cosf(a);
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int o = warp; o < dim_h; o += nw) {
        const float* w = w1 + (size_t)o * dim_in;
        float acc = 0.f;
        for (int i = lane; i < dim_in; i += 32) acc += w[i] * pe[i];
        acc = warp_sum(acc);
        if (lane == 0) h1[o] = silu_f(acc + b1[o]);
    }
    __syncthreads();
    for (int o = warp; o < dim_h; o += nw) {
        const float* w = w2 + (size_t)o * dim_h;
        float acc = 0.f;
        for (int i = lane; i < dim_h; i += 32) acc += w[i] * h1[i];
        acc = warp_sum(acc);
        // every consumer applies Mish first (ResnetBlock1D.mlp = Sequential(Mish, Linear)), so store mish(temb)
        // (apply_mish == 0: the raw embedding, for consumers that apply their own activation -- cv1_ops.cu)
        if (lane == 0) temb_mish[(size_t)s * dim_h + o] = apply_mish ? mish_f(acc + b2[o]) : acc + b2[o];
    }
}

// tbias[s, o] = W[o, :] . mish_temb[s, :] + b[o] for the concatenated [n_out, dim_h] weight of all resnet blocks.
__global__ void time_bias_kernel(const float* __restrict__ temb_mish, int dim_h, const float* __restrict__ w,
                                 const float* __restrict__ b, int n_out, float* __restrict__ tbias) {
    const int s = blockIdx.y;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_out) return;
    const float* wr = w + (size_t)warp * dim_h;
    const float* x = temb_mish + (size_t)s * dim_h;
    float acc = 0.f;
    for (int i = lane; i < dim_h; i += 32) acc += wr[i] * x[i];
    acc = warp_sum(acc);
    if (lane == 0) tbias[(size_t)s * n_out + warp] = acc + b[warp];
}

// Static part of the estimator input (mu | spks | cond -> channels 80..319) and the Euler state x (fp32 rows),
// from the reference's channel-first tensors.  One thread per (row, channel) of the conditional half.
__global__ void flow_prepare_kernel(FlowLayout L, const float* __restrict__ mu, const float* __restrict__ spks,
                                    const float* __restrict__ cond, const float* __restrict__ noise,
                                    int noise_bstride, float temperature, __nv_bfloat16* __restrict__ xin,
                                    float* __restrict__ xstate) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int C = L.n_feats;
    if (idx >= L.rows_half * C) return;
    const int row = idx / C, c = idx % C;
    const int s = L.row_seq[row];
    const int ld = 4 * C;
    __nv_bfloat16 z = __float2bfloat16(0.f);
    __nv_bfloat16* xc = xin + (size_t)row * ld;                   // conditional half
    __nv_bfloat16* xu = xin + (size_t)(row + L.rows_half) * ld;   // unconditional half (mu, spks, cond = 0)
    xu[C + c] = z; xu[2 * C + c] = z; xu[3 * C + c] = z;
    if (s < 0) {
        xc[c] = z; xc[C + c] = z; xc[2 * C + c] = z; xc[3 * C + c] = z; xu[c] = z;
        xstate[(size_t)row * C + c] = 0.f;
        return;
    }
    const int t = row - L.seq_start[s] + L.frame0;   // frame inside the utterance's ragged block
    const int T = L.seq_len[s] + L.frame0;
    const size_t off = L.in_offset[s] + (L.time_major ? (size_t)t * C + c : (size_t)c * T + t);  // ragged [C, T] or [T, C]
    xc[C + c] = __float2bfloat16(mu[off]);
    xc[2 * C + c] = __float2bfloat16(spks[(size_t)s * C + c]);
    xc[3 * C + c] = __float2bfloat16(cond[off]);
    const float x0 = noise[(size_t)s * noise_bstride + (size_t)c * L.noise_cstride + t] * temperature;
    xstate[(size_t)row * C + c] = x0;
    const __nv_bfloat16 xb = __float2bfloat16(x0);
    xc[c] = xb;
    xu[c] = xb;
}

// x += dt * ((1 + r) * d_cond - r * d_uncond); refresh the bf16 x slab of both halves of the estimator input.
__global__ void euler_cfg_kernel(FlowLayout L, const float* __restrict__ d, float dt, float cfg_rate,
                                 float* __restrict__ xstate, __nv_bfloat16* __restrict__ xin) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per 4 channels
    const int C = L.n_feats, C4 = C / 4;
    if (idx >= L.rows_half * C4) return;
    const int row = idx / C4, c = (idx % C4) * 4;
    if (L.row_seq[row] < 0) return;
    const float4 dc = *reinterpret_cast<const float4*>(d + (size_t)row * C + c);
    const float4 du = *reinterpret_cast<const float4*>(d + (size_t)(row + L.rows_half) * C + c);
    float4 x = *reinterpret_cast<float4*>(xstate + (size_t)row * C + c);
    const float a = 1.f + cfg_rate;
    x.x += dt * (a * dc.x - cfg_rate * du.x);
    x.y += dt * (a * dc.y - cfg_rate * du.y);
    x.z += dt * (a * dc.z - cfg_rate * du.z);
    x.w += dt * (a * dc.w - cfg_rate * du.w);
    *reinterpret_cast<float4*>(xstate + (size_t)row * C + c) = x;
    uint2 p;
    p.x = pack_bf16x2(x.x, x.y);
    p.y = pack_bf16x2(x.z, x.w);
    const int ld = 4 * C;
    *reinterpret_cast<uint2*>(xin + (size_t)row * ld + c) = p;
    *reinterpret_cast<uint2*>(xin + (size_t)(row + L.rows_half) * ld + c) = p;
}

// rows [rows, C] fp32 -> channel-first ragged [B, C, T] fp32 (optionally skipping the first `skip[s]` frames).
__global__ void rows_to_bct_kernel(FlowLayout L, const float* __restrict__ rows, int row_offset, float* __restrict__ out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int C = L.n_feats;
    if (idx >= L.rows_half * C) return;
    const int row = idx / C, c = idx % C;
    const int s = L.row_seq[row];
    if (s < 0) return;
    const int t = row - L.seq_start[s] + L.frame0;
    const int cs = L.dense_T > 0 ? L.dense_T : L.seq_len[s] + L.frame0;   // channel stride of the channel-first form
    out[L.in_offset[s] + (L.time_major ? (size_t)t * C + c : (size_t)c * cs + t)] = rows[(size_t)(row + row_offset) * C + c];
}

// generic estimator entry: x/mu/cond [nseq, C, T] + spks [nseq, C] -> bf16 rows [M, 4C] (all sequences explicit)
__global__ void est_pack_kernel(FlowLayout L, int nseq_rows, const float* __restrict__ x, const float* __restrict__ mu,
                                const float* __restrict__ spks, const float* __restrict__ cond,
                                const float* __restrict__ mask, __nv_bfloat16* __restrict__ xin) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int C = L.n_feats;
    if (idx >= nseq_rows * C) return;
    const int row = idx / C, c = idx % C;
    const int s = L.row_seq[row];
    __nv_bfloat16* xr = xin + (size_t)row * 4 * C;
    if (s < 0) {
        __nv_bfloat16 z = __float2bfloat16(0.f);
        xr[c] = z; xr[C + c] = z; xr[2 * C + c] = z; xr[3 * C + c] = z;
        return;
    }
    const int t = row - L.seq_start[s];
    const int T = L.dense_T > 0 ? L.dense_T : L.seq_len[s];
    const size_t off = L.in_offset[s] + (L.time_major ? (size_t)t * C + c : (size_t)c * T + t);
    (void)mask;   // prefix masks are turned into sequence lengths by the caller (mask_lens_kernel); masked frames have no row
    xr[c] = __float2bfloat16(x[off]);
    xr[C + c] = __float2bfloat16(mu[off]);
    xr[2 * C + c] = __float2bfloat16(spks ? spks[(size_t)s * C + c] : 0.f);
    xr[3 * C + c] = __float2bfloat16(cond ? cond[off] : 0.f);
}

// mask [B, 1, T] fp32 -> lens[b] = number of non-zero entries; bad[0] != 0 if some mask is not of the prefix form
// 1 ... 1 0 ... 0 (the only form `make_pad_mask` produces, cosyvoice/utils/mask.py:239-265).  One block per sequence.
__global__ void mask_lens_kernel(const float* __restrict__ mask, int T, int32_t* __restrict__ lens, int32_t* __restrict__ bad) {
    __shared__ int s_cnt, s_last;
    if (threadIdx.x == 0) { s_cnt = 0; s_last = 0; }
    __syncthreads();
    const float* m = mask + (size_t)blockIdx.x * T;
    int cnt = 0, last = 0;
    for (int t = threadIdx.x; t < T; t += blockDim.x)
        if (m[t] != 0.f) { ++cnt; last = t + 1; }
    atomicAdd(&s_cnt, cnt);
    atomicMax(&s_last, last);
    __syncthreads();
    if (threadIdx.x == 0) {
        lens[blockIdx.x] = s_cnt;
        if (s_cnt != s_last) atomicExch(bad, 1);
    }
}

__global__ void prefix_kv_save_kernel(const __nv_bfloat16* __restrict__ qkv, int hd, const int32_t* __restrict__ seq_start,
                                      const int32_t* __restrict__ slot, int prefix, __nv_bfloat16* __restrict__ slab) {
    const int s = blockIdx.y, per_row = 2 * hd / 8;   // uint4 per cached row
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= prefix * per_row) return;
    const int r = idx / per_row, j = idx % per_row;
    const uint4 v = *reinterpret_cast<const uint4*>(qkv + (size_t)(seq_start[s] + r) * 3 * hd + hd + j * 8);
    *reinterpret_cast<uint4*>(slab + ((size_t)slot[s] * prefix + r) * 2 * hd + j * 8) = v;
}

__global__ void prefix_halo_kernel(int save, __nv_bfloat16* __restrict__ buf, int64_t ld, int width,
                                   const int32_t* __restrict__ seq_start, int row_delta, const int32_t* __restrict__ slot,
                                   __nv_bfloat16* __restrict__ cache) {
    const int s = blockIdx.x, per_row = width / 8;
    for (int idx = threadIdx.x; idx < 2 * per_row; idx += blockDim.x) {
        const int r = idx / per_row, j = idx % per_row;
        uint4* m = reinterpret_cast<uint4*>(buf + (size_t)(seq_start[s] + row_delta + r) * ld + j * 8);
        uint4* c = reinterpret_cast<uint4*>(cache + ((size_t)slot[s] * 2 + r) * 512 + j * 8);
        if (save) *c = *m;
        else *m = *c;
    }
}

}  // namespace

void launch_prefix_kv_save(const __nv_bfloat16* qkv, int hd, const int32_t* seq_start, const int32_t* slot, int nseq, int prefix,
                           __nv_bfloat16* slab, cudaStream_t st) {
    dim3 grid((prefix * (2 * hd / 8) + 255) / 256, nseq);
    prefix_kv_save_kernel<<<grid, 256, 0, st>>>(qkv, hd, seq_start, slot, prefix, slab);
}
void launch_prefix_halo(int save, __nv_bfloat16* buf, int64_t ld, int width, const int32_t* seq_start, int row_delta,
                        const int32_t* slot, int nseq, __nv_bfloat16* cache, cudaStream_t st) {
    prefix_halo_kernel<<<nseq, 128, 0, st>>>(save, buf, ld, width, seq_start, row_delta, slot, cache);
}

void launch_mask_lens(const float* mask, int B, int T, int32_t* lens, int32_t* bad, cudaStream_t st) {
    mask_lens_kernel<<<B, 256, 0, st>>>(mask, T, lens, bad);
}

void launch_time_embed(const float* t, int nseq, int dim_in, int dim_h, const float* w1, const float* b1,
                       const float* w2, const float* b2, float* temb_mish, cudaStream_t st, int apply_mish) {
    time_embed_kernel<<<nseq, 512, (dim_in + dim_h) * sizeof(float), st>>>(t, dim_in, dim_h, w1, b1, w2, b2, temb_mish, apply_mish);
}
void launch_time_bias(const float* temb_mish, int nseq, int dim_h, const float* w, const float* b, int n_out,
                      float* tbias, cudaStream_t st) {
    dim3 grid((n_out * 32 + 255) / 256, nseq);
    time_bias_kernel<<<grid, 256, 0, st>>>(temb_mish, dim_h, w, b, n_out, tbias);
}
void launch_flow_prepare(const FlowLayout& L, const float* mu, const float* spks, const float* cond,
                         const float* noise, int noise_bstride, float temperature, __nv_bfloat16* xin, float* xstate,
                         cudaStream_t st) {
    const int n = L.rows_half * L.n_feats;
    flow_prepare_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, mu, spks, cond, noise, noise_bstride, temperature, xin, xstate);
}
void launch_euler_cfg(const FlowLayout& L, const float* d, float dt, float cfg_rate, float* xstate,
                      __nv_bfloat16* xin, cudaStream_t st) {
    const int n = L.rows_half * (L.n_feats / 4);
    euler_cfg_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, d, dt, cfg_rate, xstate, xin);
}
void launch_rows_to_bct(const FlowLayout& L, const float* rows, int row_offset, float* out, cudaStream_t st) {
    const int n = L.rows_half * L.n_feats;
    rows_to_bct_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, rows, row_offset, out);
}
void launch_est_pack(const FlowLayout& L, int nrows, const float* x, const float* mu, const float* spks,
                     const float* cond, const float* mask, __nv_bfloat16* xin, cudaStream_t st) {
    const int n = nrows * L.n_feats;
    est_pack_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, nrows, x, mu, spks, cond, mask, xin);
}

}  // namespace mossdec
