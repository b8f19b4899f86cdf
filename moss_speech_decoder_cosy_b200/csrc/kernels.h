// Launch prototypes of the non-GEMM kernels (flow_kernels.cu, attention.cu, hift_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace mossdec {

// Packed-row layout of a batch of sequences (device pointers).  Rows of one sequence are contiguous; sequences are
// separated by >= guard zero rows; the unconditional CFG half repeats the layout `rows_half` rows further down.
struct FlowLayout {
    int32_t n_feats;          // 80
    int32_t rows_half;        // rows of one CFG half (multiple of 128)
    int32_t noise_cstride;    // channel stride of the noise bank
    const int32_t* row_seq;   // [2 * rows_half] sequence id or -1
    const int32_t* seq_start; // [nseq] first row
    const int32_t* seq_len;   // [nseq] frames
    const int64_t* in_offset; // [nseq] element offset of the sequence inside ragged [B, C, T_b] tensors
};

void launch_time_embed(const float* t, int nseq, int dim_in, int dim_h, const float* w1, const float* b1,
                       const float* w2, const float* b2, float* temb_mish, cudaStream_t st);
void launch_time_bias(const float* temb_mish, int nseq, int dim_h, const float* w, const float* b, int n_out,
                      float* tbias, cudaStream_t st);
void launch_flow_prepare(const FlowLayout& L, const float* mu, const float* spks, const float* cond,
                         const float* noise, int noise_bstride, float temperature, __nv_bfloat16* xin, float* xstate,
                         cudaStream_t st);
void launch_euler_cfg(const FlowLayout& L, const float* d, float dt, float cfg_rate, float* xstate,
                      __nv_bfloat16* xin, cudaStream_t st);
void launch_rows_to_bct(const FlowLayout& L, const float* rows, int row_offset, float* out, cudaStream_t st);
void launch_est_pack(const FlowLayout& L, int nrows, const float* x, const float* mu, const float* spks,
                     const float* cond, const float* mask, __nv_bfloat16* xin, cudaStream_t st);

// Self-attention over packed rows.  qkv: bf16 [M, 3*H*64] = [q | k | v], q pre-scaled by 1/sqrt(64).
// Key j of sequence s is visible to query i iff j < len_s and (chunk == 0 or j < (i / chunk + 1) * chunk)
// (cosyvoice/utils/mask.py:127-158 + key padding).  out: bf16 [M, H*64].
void launch_attention(const __nv_bfloat16* qkv, __nv_bfloat16* out, const int32_t* seq_start, const int32_t* seq_len,
                      int nseq, int max_len, int heads, int chunk, cudaStream_t st);

}  // namespace mossdec
