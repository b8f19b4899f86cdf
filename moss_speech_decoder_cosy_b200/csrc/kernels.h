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
    int32_t time_major;       // ragged tensors are [T_b, C] (1) instead of the reference's [C, T_b] (0)
    int32_t dense_T;          // > 0: tensors are dense [nseq, C, dense_T] (channel stride dense_T); 0: ragged
    int32_t frame0;           // packed row t of a sequence is frame frame0 + t of its ragged block (frozen-prefix reuse: the
                              // first frame0 frames of every utterance have no rows); the block holds frame0 + seq_len frames
    const int32_t* row_seq;   // [2 * rows_half] sequence id or -1
    const int32_t* seq_start; // [nseq] first row
    const int32_t* seq_len;   // [nseq] frames
    const int64_t* in_offset; // [nseq] element offset of the sequence inside ragged [B, C, T_b] tensors
};

void launch_time_embed(const float* t, int nseq, int dim_in, int dim_h, const float* w1, const float* b1,
                       const float* w2, const float* b2, float* temb_mish, cudaStream_t st, int apply_mish = 1);
void launch_time_bias(const float* temb_mish, int nseq, int dim_h, const float* w, const float* b, int n_out,
                      float* tbias, cudaStream_t st);
void launch_flow_prepare(const FlowLayout& L, const float* mu, const float* spks, const float* cond,
                         const float* noise, int noise_bstride, float temperature, __nv_bfloat16* xin, float* xstate,
                         cudaStream_t st);
void launch_euler_cfg(const FlowLayout& L, const float* d, float dt, float cfg_rate, float* xstate,
                      __nv_bfloat16* xin, cudaStream_t st);
void launch_rows_to_bct(const FlowLayout& L, const float* rows, int row_offset, float* out, cudaStream_t st);
// Frozen-prefix cache of a streaming solve (engine_solve.cu): copies between the packed activation matrices and the cache.
//   kv:   rows [seq_start[s], + prefix) of qkv [*, 3 * HD] (columns HD .. 3 HD = k | v)  ->  slab rows [slot[s] * prefix, + prefix)
//   halo: the two rows in front of row seq_start[s] + row_delta + 2 of a conv input (width channels, row stride ld)
//         <-> cache[slot[s]][2][512]; save = 1 copies matrix -> cache, 0 cache -> matrix
void launch_prefix_kv_save(const __nv_bfloat16* qkv, int hd, const int32_t* seq_start, const int32_t* slot, int nseq, int prefix,
                           __nv_bfloat16* slab, cudaStream_t st);
void launch_prefix_halo(int save, __nv_bfloat16* buf, int64_t ld, int width, const int32_t* seq_start, int row_delta,
                        const int32_t* slot, int nseq, __nv_bfloat16* cache, cudaStream_t st);
void launch_mask_lens(const float* mask, int B, int T, int32_t* lens, int32_t* bad, cudaStream_t st);
void launch_est_pack(const FlowLayout& L, int nrows, const float* x, const float* mu, const float* spks,
                     const float* cond, const float* mask, __nv_bfloat16* xin, cudaStream_t st);

// Self-attention over packed rows.  qkv: bf16 [M, 3*H*64] = [q | k | v], q pre-scaled by 1/sqrt(64).
// Key j of sequence s is visible to query i iff j < len_s and (chunk == 0 or j < (i / chunk + 1) * chunk)
// (cosyvoice/utils/mask.py:127-158 + key padding).  out: bf16 [M, H*64].
// warp-specialised tcgen05 / TMEM kernel (attention_ws.cu): dedicated TMA / MMA issue warp, 32-key tiles, double-buffered
// scores in TMEM; qkv has `rows` rows of 3*H*64 bf16.
cudaError_t launch_attention_ws(const __nv_bfloat16* qkv, int64_t rows, __nv_bfloat16* out, const int32_t* seq_start,
                                const int32_t* seq_len, int nseq, int max_len, int heads, int chunk, cudaStream_t st);
// general form: qkv rows of `qkv_ld` bf16 ([q | k | v | ...], q / k / v of head h at columns h*64, (H+h)*64, (2H+h)*64), output
// rows of H*64; optional additive score bias (bf16) bias[h][row][bias_tmax - 1 - t + j] for query position t and key j
// (the shifted relative-position term of the conformer encoder, attention.py:223-244, 317-324), NULL = none.
cudaError_t launch_attention_ws_ex(const __nv_bfloat16* qkv, int64_t rows, int qkv_ld, __nv_bfloat16* out,
                                   const int32_t* seq_start, const int32_t* seq_len, int nseq, int max_len, int heads, int chunk,
                                   const __nv_bfloat16* bias, int64_t bias_head_stride, int bias_ld, int bias_tmax,
                                   cudaStream_t st);

// Attention over live rows whose sequences each have `prefix_len` frozen keys in front (absolute key positions
// 0 .. prefix_len - 1; the live rows start at position prefix_len).  prefix_kv: bf16 [prefix_rows, 2 * H * 64] = [k | v];
// sequence s's prefix starts at row prefix_row[s] (device array).  chunk must divide prefix_len.
cudaError_t launch_attention_ws_prefix(const __nv_bfloat16* qkv, int64_t rows, __nv_bfloat16* out, const int32_t* seq_start,
                                       const int32_t* seq_len, int nseq, int max_len, int heads, int chunk,
                                       const __nv_bfloat16* prefix_kv, int64_t prefix_rows, int prefix_len,
                                       const int32_t* prefix_row, cudaStream_t st);

// ------------------------------------------------------------------------------------------------ encoder front end
// Packed token-rate layout of a group of utterances (device pointers): sequence b owns token rows
// [seq_start[b], seq_start[b] + n_tok[b]) (prompt + new tokens + look-ahead context), >= 4 zero guard rows between
// sequences.  The mel-rate matrices use the same axis scaled by 4 (the x4 up-conv is a polyphase GEMM whose
// [rows, 4 * 512] output IS the [4 * rows, 512] mel-rate matrix).
struct EncLayout {
    int32_t rows;                  // token-rate rows incl. guards, multiple of 128 (mel-rate rows = 4 * rows)
    int32_t nseq;
    const int32_t* row_seq_all;    // [rows]      sequence id of rows that hold a token (context included), -1 on guard rows
    const int32_t* row_seq_main;   // [rows]      the same, but -1 on the look-ahead context rows
    const int32_t* row_seq_mel;    // [4 * rows]  mel-rate rows of the main tokens
    const int32_t* seq_start;      // [nseq] first token row
    const int32_t* seq_len;        // [nseq] main tokens (attention length at token rate)
    const int32_t* seq_start4;     // [nseq] 4 * seq_start
    const int32_t* seq_len4;       // [nseq] 4 * seq_len
    const int32_t* prompt_frames;  // [nseq] prompt mel frames F_b
    const int64_t* tok_off;        // [nseq] offset into the ragged token array
    const int64_t* mel_off;        // [nseq] FRAME offset into the ragged mu / cond outputs
    const int64_t* feat_off;       // [nseq] FRAME offset into the ragged prompt_feat input
};
void launch_enc_gather(const EncLayout& L, const int64_t* tokens, const __nv_bfloat16* emb, int vocab, __nv_bfloat16* out,
                       cudaStream_t st);
void launch_layernorm512(const float* x, int rows, const float* gamma, const float* beta, float eps, float scale,
                         const int32_t* row_seq, float* out_f32, __nv_bfloat16* out_bf16, cudaStream_t st);
void launch_rel_pos_table(int tmax, int rows, float* pe, cudaStream_t st);
void launch_f32_to_bf16(const float* x, int64_t n, __nv_bfloat16* y, cudaStream_t st);
void launch_spk_project(const float* emb, int B, int dim_in, const float* w, const float* bias, int dim_out, float* out,
                        cudaStream_t st);
void launch_enc_finish(const EncLayout& L, const float* mu_rows, const float* prompt_feat, float* mu, float* cond,
                       cudaStream_t st);

// ------------------------------------------------------------------------------------------------ HiFT
// Packed layout of one group of mel segments: sequence b owns frame rows [seq_start[b], seq_start[b] + seq_len[b]);
// at least 4 zero guard frames separate sequences.  Stage rows are frame rows x 8 / 40 / 120 (the last stage has one
// extra row per sequence: ReflectionPad1d((1, 0)), generator.py:529-530).
struct HiftLayout {
    int32_t nseq;
    int32_t rows_f;            // frame rows incl. guards, multiple of 128
    int32_t time_major;        // mel input is ragged [T_b, 80] (1) instead of [80, T_b] (0)
    const int32_t* seq_start;  // [nseq]
    const int32_t* seq_len;    // [nseq]
    const int64_t* mel_off;    // [nseq] element offset inside the ragged [80, T_b] input
    const int64_t* wav_off;    // [nseq] sample offset inside the ragged [480 T_b] output
};

void launch_hift_row_seq(const HiftLayout& L, int scale, int extra, int rows, int32_t* out, cudaStream_t st);
void launch_hift_pack_mel(const HiftLayout& L, const int32_t* row_seq, const float* mel, __nv_bfloat16* out,
                          float* out32, cudaStream_t st);
// fp32 f0 predictor (f0_fp32.cu): Conv1d(k3, pad 1) + ELU over packed rows, weights Wt [3 * Cin, N] (k = tap * Cin + c)
cudaError_t launch_conv3_elu_f32(const float* in, int M, int Cin, const float* wt, int N, const float* bias,
                                 const int32_t* row_seq, float* out, cudaStream_t st);
// generic fp32 GEMM on the CUDA cores: out [M, N] = view [M, K] . Wt [K, N]; row m of the view starts at in + m * ld_in
cudaError_t launch_gemm_f32(const float* in, int M, int64_t ld_in, int K, const float* wt, int N, float* out, cudaStream_t st);
cudaError_t launch_f0_classifier(const float* h, int M, int C, const float* w, const float* b, const int32_t* row_seq,
                                 float* f0, int f0_ld, cudaStream_t st);
void launch_hift_phase(const HiftLayout& L, const float* f0, int f0_ld, float sample_rate, float upsample, int n_harm,
                       float* p480, cudaStream_t st);
void launch_hift_source(const HiftLayout& L, int max_len, const float* f0, int f0_ld, const float* p480,
                        const float* noise, const float* lin_w, const float* lin_b, const float* cache,
                        int cache_stride, const int32_t* cache_len, int n_harm, float sine_amp, float noise_std,
                        float voiced_thr, int up, uint64_t seed, float* src_out, cudaStream_t st);
void launch_hift_stft(const HiftLayout& L, const int32_t* row_seq2, int rows2, const float* src, __nv_bfloat16* out,
                      cudaStream_t st);
void launch_hift_fuse_snake(const HiftLayout& L, const int32_t* row_seq, int scale, int reflect, int64_t rows, int C,
                            const float* u, const float* si, const float* a0, const float* a1, const float* a2,
                            float* x, __nv_bfloat16* o0, __nv_bfloat16* o1, __nv_bfloat16* o2, cudaStream_t st);
void launch_hift_istft(const HiftLayout& L, int max_len, const float* post, int post_ld, float limit, float* wav,
                       cudaStream_t st);
void launch_fade(float* speech, const float* prev, const float* w, int n, cudaStream_t st);


}  // namespace mossdec
