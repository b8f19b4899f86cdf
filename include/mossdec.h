/* mossdec -- C ABI of the B200-native token->waveform decode path of MOSS-Speech-decoder-cosy.
 *
 * The reference is pure Python/PyTorch; the one FFI-shaped boundary it already has for this path is the TensorRT
 * estimator contract in GLM_modules/cosyvoice/flow/flow_matching.py:129-156 (raw data_ptr()s of contiguous fp32
 * tensors, enqueue on a CUDA stream).  Every entry point below replaces a reference call with the same tensor
 * layouts, so a maintainer can bind it with ctypes:
 *
 *   mossdec_estimator   <- ConditionalCFM.forward_estimator          flow_matching.py:129-156 / decoder.py:405-494
 *   mossdec_flow_solve  <- CausalConditionalCFM.forward + solve_euler flow_matching.py:206-230, 76-127
 *   mossdec_hift        <- HiFTGenerator.inference                    hifigan/generator.py:570-582
 *   mossdec_token2mel   <- CausalMaskedDiffWithXvec.inference         cosyvoice/flow/flow.py:237-283 (encoder + solve)
 *   mossdec_mel_spectrogram <- feat_extractor (matcha mel_spectrogram) third_party/Matcha-TTS/matcha/utils/audio.py:45-82
 *   mossdec_set_tensor  <- nn.Module.load_state_dict                  flow_inference.py:59-61
 *
 * All data pointers are DEVICE pointers owned by the caller unless stated otherwise; `stream` is a cudaStream_t
 * passed as void*.  Calls only enqueue work: the per-call layout arrays are staged in pinned host memory and uploaded
 * with async copies on `stream`, so back-to-back calls on one (also non-blocking) stream are ordered by the stream.
 * Host synchronisation happens only in mossdec_set_tensor, in mossdec_estimator when a mask is given, in the debug /
 * profile readers, and when a workspace has to grow (see mossdec_reserve).  Return value: 0 = OK, negative = error (message via
 * mossdec_last_error).  A handle must not be used from two threads at once.
 */
#ifndef MOSSDEC_H_
#define MOSSDEC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mossdec_handle mossdec_handle;

enum { MOSSDEC_F32 = 0, MOSSDEC_BF16 = 1, MOSSDEC_I32 = 2 };

typedef struct mossdec_config {
    int32_t n_feats;      /* 80   mel bins                                   flow.py:154            */
    int32_t channels;     /* 256  estimator width (channels=[256])           decoder.py:294         */
    int32_t heads;        /* 8                                                                    */
    int32_t head_dim;     /* 64                                                                   */
    int32_t n_blocks;     /* 4    transformer blocks per group                                    */
    int32_t n_mid;        /* 12   mid groups                                                      */
    int32_t noise_len;    /* 15000 frames of the fixed noise bank             flow_matching.py:203 */
    int32_t hift_base;    /* 512                                              generator.py:400     */
    int32_t n_harmonics;  /* 8 (+1 fundamental)                                                   */
    float sample_rate;    /* 24000                                                                */
    float nsf_alpha, nsf_sigma, voiced_threshold, lrelu_slope, audio_limit;
} mossdec_config;

/* Fill `cfg` with the pinned CosyVoice2-style configuration. */
void mossdec_default_config(mossdec_config* cfg);

int mossdec_create(const mossdec_config* cfg, mossdec_handle** out);
void mossdec_destroy(mossdec_handle* h);
const char* mossdec_last_error(const mossdec_handle* h);

/* Upload one engine-ready tensor (HOST pointer; copied).  Names/layouts: moss_speech_decoder_cosy_b200/weights.py. */
int mossdec_set_tensor(mossdec_handle* h, const char* name, const void* host_data, int32_t dtype, int32_t ndim,
                       const int64_t* shape);
/* Check that every tensor the engine needs is present. */
int mossdec_finalize(mossdec_handle* h);

/* One estimator evaluation, dense batch.  x, mu, cond, out: [B2, 80, T] fp32; t: [B2]; spks: [B2, 80] (spks / cond may be
 * NULL = zeros).  mask: [B2, 1, T] fp32 key-padding mask as the reference passes it (decoder.py:405-494), NULL = all
 * valid.  It must be a prefix mask (1...1 0...0 per sequence, what make_pad_mask produces); masked frames are excluded
 * from attention and the convolutions' receptive field sees zeros there, the output is zero on them (output * mask).  A
 * mask of another form is rejected with an error.  With a mask the call synchronises `stream` once (the lengths are read
 * back), like the TensorRT path it mirrors (flow_matching.py:153-154).  `out` may alias `x`: the TensorRT contract binds
 * the output onto x (flow_matching.py:143-151).  chunk = 0 -> full attention, > 0 -> block-causal mask with `chunk` mel
 * frames (decoder.py:439-440, streaming=True). */
int mossdec_estimator(mossdec_handle* h, int32_t B2, int32_t T, const float* x, const float* mask, const float* mu,
                      const float* t, const float* spks, const float* cond, float* out, int32_t chunk, void* stream);

/* flags bit 0 (MOSSDEC_TIME_MAJOR): the ragged mel-like tensors (mu, cond, mel_out of mossdec_flow_solve; mel of
 * mossdec_hift) are [lens[b], 80] blocks instead of the reference's channel-first [80, lens[b]]. */
enum { MOSSDEC_TIME_MAJOR = 1 };

/* n_steps Euler steps with classifier-free guidance for B utterances of lens[b] frames (HOST array).
 * mu, cond, mel_out: ragged [80, lens[b]] blocks back to back (for equal lengths this is [B, 80, T]); spks [B, 80].
 * noise: NULL -> the bank uploaded as tensor "flow.rand_noise" [80, noise_len]; else a device bank [80, noise_len_arg]
 * shared by all utterances (z = bank[:, :T] * temperature, flow_matching.py:225). */
int mossdec_flow_solve(mossdec_handle* h, int32_t B, const int32_t* lens, const float* mu, const float* spks,
                       const float* cond, const float* noise, int32_t noise_len_arg, int32_t n_steps, float cfg_rate,
                       float temperature, int32_t chunk, float* mel_out, int32_t flags, void* stream);

/* HiFT vocoder for B mel segments of lens[b] frames (HOST array): mel ragged [80, lens[b]] back to back ->
 * wav_out ragged [480 * lens[b]], source_out ragged [480 * lens[b]] (may be NULL).
 * cache_source / cache_len: optional per-utterance source prefix (generator.py:579-580); cache_source is
 * [B, cache_stride] with cache_len[b] valid samples (HOST array) -- NULL = none.
 * nsf_noise: optional injected N(0,1) draws, ragged [480 * lens[b], 9] (sample-major, torch.randn_like order of
 * generator.py:334); NULL -> counter-based Philox on device seeded with `seed`. */
int mossdec_hift(mossdec_handle* h, int32_t B, const int32_t* lens, const float* mel, const float* cache_source,
                 int32_t cache_stride, const int32_t* cache_len, const float* nsf_noise, float* wav_out,
                 float* source_out, uint64_t seed, int32_t flags, void* stream);

/* mossdec_hift over the GENERATED part of flow results: the mel block of utterance b holds mel_skip[b] + lens[b] frames
 * (time-major [frames, 80], what mossdec_token2mel writes); the first mel_skip[b] (prompt) frames are passed over
 * (flow.py:281 feat[:, :, mel_len1:]).  lens / mel_skip: HOST arrays.  No source cache (non-streaming call). */
int mossdec_hift_skip(mossdec_handle* h, int32_t B, const int32_t* lens, const int32_t* mel_skip, const float* mel,
                      const float* nsf_noise, float* wav_out, float* source_out, uint64_t seed, void* stream);

/* CausalMaskedDiffWithXvec.inference (flow.py:237-283) for B independent utterances (the reference asserts B == 1):
 * speaker projection (L2-normalise + Linear 192 -> 80), token embedding, UpsampleConformerEncoder, encoder_proj, prompt
 * condition, then the CFG Euler solve -- all on this library's kernels (row N1).
 *   n_tok[b] (HOST): tokens of utterance b = prompt tokens + new tokens (+ ctx_len look-ahead tokens at the end when the
 *     call is not final, flow.py:262-263: ctx_len = pre_lookahead_len = 3, else 0);
 *   tokens: device int64, the utterances' token ids back to back (negative ids are clamped to 0 like flow.py:256);
 *   prompt_frames[b] (HOST) / prompt_feat: device fp32 ragged [prompt_frames[b], 80] prompt mels back to back;
 *   embedding: device fp32 [B, 192];
 *   chunk_tokens: 0 = full attention (streaming = False), else the encoder's static_chunk_size in tokens;
 *   est_chunk_frames: the estimator's block-causal chunk in mel frames (0 = full attention);
 *   mel_out: device fp32 ragged time-major [4 * (n_tok[b] - ctx_len), 80] blocks = `feat` of flow.py:272-280 (prompt
 *     frames included; strip them with mossdec_hift_skip or by offset). */
int mossdec_token2mel(mossdec_handle* h, int32_t B, const int32_t* n_tok, const int64_t* tokens, const int32_t* prompt_frames,
                      const float* prompt_feat, const float* embedding, int32_t ctx_len, int32_t chunk_tokens,
                      int32_t est_chunk_frames, int32_t n_steps, float cfg_rate, float* mel_out, void* stream);

/* ---- Frozen-prefix cache for streaming chunk calls.
 * The reference recomputes the flow over prompt + token window for every streamed chunk (flow_inference.py:188-218).  With
 * `stream=True` the estimator's attention is block-causal (chunks of est_chunk_frames) and its convolutions are causal,
 * so the first whole chunks that lie inside the prompt -- the frozen prefix -- are identical in every chunk call of a
 * stream.  mossdec_token2mel_prefix with mode 1 (record) runs the full solve and stores, per Euler step and CFG branch, the
 * K | V rows of the prefix for every transformer block and the two-row left context of every causal conv in cache slot
 * slots[b]; with mode 2 (reuse) only the live frames behind the prefix are computed (the frozen keys come from the cache
 * inside the attention kernel, with the same key tiling as in the recording call: the live frames are bit-identical to
 * a recording call on the same inputs).  Exact, also for sliding `max_token_len` windows.  mode 0 = mossdec_token2mel.
 *   mossdec_prefix_reserve: allocate the cache (n_slots streams x prefix_frames frames x n_steps; drops its contents).
 *     prefix_frames must be a multiple of est_chunk_frames and of 4 * chunk_tokens.
 *   slots / tags: HOST arrays [B]; tags[b] != 0 identifies the frozen inputs of the stream (prompt tokens, prompt mel,
 *     speaker embedding); a reuse call whose tag differs from the recorded one fails with an error instead of drifting.
 *   Requirements (checked): streaming call (chunk_tokens > 0, est_chunk_frames > 0), prompt_frames[b] >= prefix_frames and
 *     prompt tokens >= prefix_frames / 4 + 3, every utterance longer than the prefix, same n_steps / cfg_rate / chunk
 *     as recorded.  In reuse mode the first prefix_frames frames of every mel_out block are not written. */
int mossdec_prefix_reserve(mossdec_handle* h, int32_t n_slots, int32_t prefix_frames, int32_t n_steps);
int mossdec_token2mel_prefix(mossdec_handle* h, int32_t B, const int32_t* n_tok, const int64_t* tokens,
                             const int32_t* prompt_frames, const float* prompt_feat, const float* embedding, int32_t ctx_len,
                             int32_t chunk_tokens, int32_t est_chunk_frames, int32_t n_steps, float cfg_rate, int32_t mode,
                             const int32_t* slots, const uint64_t* tags, float* mel_out, void* stream);

/* Streaming glue (flow_inference.py:35-43,151-156): speech[:n] = speech[:n]*w[:n] + prev_tail[:n]*w[n:], fp32, device. */
int mossdec_fade_in_out(mossdec_handle* h, float* speech, const float* prev_tail, const float* window2n, int32_t n,
                        void* stream);

/* Prompt mel front end = the reference's `feat_extractor`, matcha mel_spectrogram
 * (GLM_modules/third_party/Matcha-TTS/matcha/utils/audio.py:45-82) with n_fft = win = 1920, hop 480, 80 mels, 24 kHz,
 * fmin 0, fmax 8000, center = False, called by GLM4Encoder._extract_speech_feat (whisper_encoder_decoder.py:191-195).
 * wav: device fp32 [n_samples] in [-1, 1]; mel_out: device fp32 [80, T] with T = n_samples / 480 (floor; the caller
 * allocates 80 * (n_samples / 480) floats).  Needs the tensors "mel.dft" and "mel.fb" (mel_frontend.py).  Returns T,
 * negative on error. */
int32_t mossdec_mel_spectrogram(mossdec_handle* h, const float* wav, int64_t n_samples, float* mel_out, void* stream);

/* Allocation-free variant of mossdec_mel_spectrogram.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same kernels and the same result as mossdec_mel_spectrogram, but the workspace
 * (padded signal, spectrum, magnitude, mel energies) is a file-scope `static __device__` array inside libmossdec.so instead
 * of a growable cudaMalloc buffer, so no call -- the first one included -- calls cudaMalloc, cudaFree or an allocating
 * library routine.
 *   limit : T = n_samples / 480 <= MOSSDEC_MEL_STATIC_MAX_FRAMES (512 frames = 10.24 s at 24 kHz, n_samples <= 246239); longer
 *           inputs are rejected with an error (negative return, text via mossdec_last_error) and must use
 *           mossdec_mel_spectrogram.
 *   cost  : MOSSDEC_MEL_STATIC_SCRATCH_BYTES (7.2 MiB) of device memory on every GPU that loads the library, held until the
 *           process exits, whether or not this entry point is ever called.
 *   races : every call on a device shares the one array (calls through different handles too).  Calls that overlap on the
 *           GPU (different streams without an ordering between them) race; calls on one stream are ordered by the stream.
 *           The array keeps the previous call's contents: the entry point rewrites or clears everything it later reads.
 *   set-up: the first call on a device resolves the array's device address (cudaGetSymbolAddress, with the handle's device
 *           current); like any first launch it loads the module, so make one warm-up call before starting a capture.
 * wav, mel_out: device pointers owned by the caller as above (mel_out: 80 * (n_samples / 480) floats).  Returns T. */
#define MOSSDEC_MEL_STATIC_MAX_FRAMES 512
#define MOSSDEC_MEL_STATIC_SCRATCH_BYTES 7544832
int32_t mossdec_mel_spectrogram_static(mossdec_handle* h, const float* wav, int64_t n_samples, float* mel_out, void* stream);

/* ---- Secondary configuration: building blocks of the NON-causal cosyvoice1 / Matcha
 * estimator (GLM_modules/cosyvoice1 uses third_party/Matcha-TTS/matcha/models/components/decoder.py).  Dense
 * [B, C, T] fp32 tensors exactly as the reference modules take them; block width 256, GroupNorm(8).  Weights are the
 * tensors "<prefix>.*" uploaded with mossdec_set_tensor (layouts: weights.py pack_cv1_resnet / _downsample / _upsample).
 *
 *   mossdec_cv1_resnet_block <- ResnetBlock1D.forward(x, mask, time_emb)   matcha decoder.py:46-61 (Block1D :32-43)
 *       x [B, Cin, T], mask [B, 1, T], time_emb [B, time_dim] (NULL = no time bias) -> out [B, 256, T]
 *   mossdec_cv1_block1d      <- Block1D.forward(x, mask) (also the decoder's final_block)   matcha decoder.py:32-43
 *       x [B, Cin, T], mask [B, 1, T] -> out [B, 256, T];  tensors "<prefix>.c.w / .c.b / .gn.g / .gn.b"
 *   mossdec_cv1_downsample   <- Downsample1D.forward   Conv1d(C, C, 3, 2, 1)            matcha decoder.py:64-70
 *       x [B, C, T] -> out [B, C, (T - 1) / 2 + 1]
 *   mossdec_cv1_upsample     <- Upsample1D.forward     ConvTranspose1d(C, C, 4, 2, 1)   matcha decoder.py:120-158
 *       x [B, C, T] -> out [B, C, 2 T]
 */
int mossdec_cv1_resnet_block(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t Cin, const float* x,
                             const float* mask, const float* time_emb, int32_t time_dim, float* out, void* stream);
int mossdec_cv1_block1d(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t Cin, const float* x, const float* mask,
                        float* out, void* stream);
int mossdec_cv1_downsample(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t C, const float* x, float* out,
                           void* stream);
int mossdec_cv1_upsample(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t C, const float* x, float* out,
                         void* stream);

/* The assembled non-causal decoder of the secondary configuration: ConditionalDecoder.forward(x, mask, mu, t, spks, cond)
 * (GLM_modules/cosyvoice/flow/decoder.py:206-291; cosyvoice1/flow/decoder.py:151-222 is the same graph) with
 * channels = [256, 256], n_blocks transformer blocks per level (8 heads x 64, GELU), n_mid mid blocks.  Same tensor
 * contract as mossdec_estimator: x, mu, cond, out [B, 80, T] fp32, mask [B, 1, T] (prefix masks), t [B], spks [B, 80]
 * (spks / cond may be NULL).  Weights: tensors "<prefix>.*" (weights.py pack_cv1_decoder).  Synchronises `stream` once
 * (mask lengths).  Chained from the building blocks above, each on its own packed rows: correct, not tuned. */
int mossdec_cv1_estimator(mossdec_handle* h, const char* prefix, int32_t n_blocks, int32_t n_mid, int32_t B, int32_t T, const float* x,
                          const float* mask, const float* mu, const float* t, const float* spks, const float* cond, float* out,
                          void* stream);

/* Test hook: copy up to max_floats of a named fp32 intermediate of the last call ("hift.f0", "hift.p480",
 * "hift.post", "hift.x"; "enc.mu", "enc.cond", "enc.spks" after mossdec_token2mel) to HOST memory; synchronises the device.  Returns floats copied, negative on error. */
int mossdec_debug_read(mossdec_handle* h, const char* name, float* host_dst, int64_t max_floats);

/* Per-launch device timing (CUDA events on the launch stream) of the three tensor-core kernels.
 * mossdec_profile_read synchronises and returns, since the last read: out9 = {gemm ms, gemm launches, gemm executed
 * FLOPs (2 M N K incl. tile padding), attention ms, attention launches, 0, fused-FFN ms, fused-FFN launches, FFN FLOPs}. */
void mossdec_profile(mossdec_handle* h, int32_t on);
int mossdec_profile_read(mossdec_handle* h, double* out9);
/* Bandwidth-class kernels ("euler_cfg_kernel", "flow_prepare_kernel", "hift_source_kernel", "hift_stft_kernel",
 * "hift_fuse_snake_kernel", "hift_istft_kernel") are timed the same way while profiling is on; after a
 * mossdec_profile_read, out3 = {device ms, launches, ALGORITHMIC bytes moved (SURVEY.md 8d figures)} of `name`. */
int mossdec_profile_named(mossdec_handle* h, const char* name, double* out3);

/* Pre-size the device workspaces so that the first decode does not allocate: solves of up to `flow_rows` mel frames in
 * total over up to `flow_seqs` utterances, vocoder groups of up to `hift_frames` frames.  (Workspaces otherwise grow on
 * demand; growing synchronises the device.) */
int mossdec_reserve(mossdec_handle* h, int64_t flow_rows, int32_t flow_seqs, int64_t hift_frames);

/* The Euler loop of mossdec_flow_solve is captured into a CUDA graph per packed-row geometry and replayed (the
 * reference's analogue: the enqueue-once TensorRT execution context, flow_matching.py:133-156).  mode 0 = plain
 * launches, 1 = capture a geometry the second time it is seen (default), 2 = capture on first sight.  Also settable
 * with the environment variable MOSSDEC_GRAPH at handle creation. */
void mossdec_set_graph_mode(mossdec_handle* h, int32_t mode);
int64_t mossdec_graph_replays(const mossdec_handle* h);

/* CUDA device ordinal the handle was created on.  Every call runs on that device regardless of the caller's current
 * device (and restores the caller's device before returning). */
int32_t mossdec_device(const mossdec_handle* h);

/* Frames per vocoder group (bounds the stage-2 working set; default 8192). */
void mossdec_set_hift_group_frames(mossdec_handle* h, int64_t frames);

/* Number of kernels this handle has enqueued since creation / last reset (bench.py reports it as gpu_launches). */
int64_t mossdec_launch_count(const mossdec_handle* h);
void mossdec_reset_launch_count(mossdec_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* MOSSDEC_H_ */
