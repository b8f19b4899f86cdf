// extern "C" surface of libmossdec.so (include/mossdec.h).
#include "mossdec.h"

#include "engine.h"

using mossdec::Engine;
using mossdec::EngineConfig;

struct mossdec_handle {
    Engine* eng;
    std::string create_err;
};

extern "C" {

void mossdec_default_config(mossdec_config* c) {
    c->n_feats = 80; c->channels = 256; c->heads = 8; c->head_dim = 64; c->n_blocks = 4; c->n_mid = 12;
    c->noise_len = 15000; c->hift_base = 512; c->n_harmonics = 8; c->sample_rate = 24000.f;
    c->nsf_alpha = 0.1f; c->nsf_sigma = 0.003f; c->voiced_threshold = 10.f; c->lrelu_slope = 0.1f;
    c->audio_limit = 0.99f;
}

int mossdec_create(const mossdec_config* c, mossdec_handle** out) {
    if (!c || !out) return -1;
    if (c->head_dim != 64 || c->channels != 256 || c->n_feats != 80 || c->heads * c->head_dim != 512) return -2;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return -3;  // no CPU fallback: fail loudly
    EngineConfig e;
    e.n_feats = c->n_feats; e.channels = c->channels; e.heads = c->heads; e.head_dim = c->head_dim;
    e.n_blocks = c->n_blocks; e.n_mid = c->n_mid; e.time_dim = 4 * c->channels; e.in_channels = 4 * c->n_feats;
    e.noise_len = c->noise_len; e.hift_base = c->hift_base; e.n_harm = c->n_harmonics + 1;
    e.sample_rate = c->sample_rate; e.sine_amp = c->nsf_alpha; e.noise_std = c->nsf_sigma;
    e.voiced_thr = c->voiced_threshold; e.lrelu = c->lrelu_slope; e.audio_limit = c->audio_limit;
    mossdec_handle* h = new mossdec_handle;
    h->eng = new Engine(e);
    *out = h;
    return 0;
}

void mossdec_destroy(mossdec_handle* h) {
    if (!h) return;
    delete h->eng;
    delete h;
}

const char* mossdec_last_error(const mossdec_handle* h) { return h ? h->eng->last_error() : "null handle"; }

int mossdec_set_tensor(mossdec_handle* h, const char* name, const void* host_data, int32_t dtype, int32_t ndim,
                       const int64_t* shape) {
    return h->eng->set_tensor(name, host_data, dtype, ndim, shape);
}
int mossdec_finalize(mossdec_handle* h) { return h->eng->finalize(); }

int mossdec_estimator(mossdec_handle* h, int32_t B2, int32_t T, const float* x, const float* mask, const float* mu,
                      const float* t, const float* spks, const float* cond, float* out, int32_t chunk, void* stream) {
    return h->eng->estimator(B2, T, x, mask, mu, t, spks, cond, out, chunk, (cudaStream_t)stream);
}

int mossdec_flow_solve(mossdec_handle* h, int32_t B, const int32_t* lens, const float* mu, const float* spks,
                       const float* cond, const float* noise, int32_t noise_len_arg, int32_t n_steps, float cfg_rate,
                       float temperature, int32_t chunk, float* mel_out, int32_t flags, void* stream) {
    return h->eng->flow_solve(B, lens, mu, spks, cond, noise, noise_len_arg, n_steps, cfg_rate, temperature, chunk,
                              mel_out, flags, (cudaStream_t)stream);
}

int mossdec_hift(mossdec_handle* h, int32_t B, const int32_t* lens, const float* mel, const float* cache_source,
                 int32_t cache_stride, const int32_t* cache_len, const float* nsf_noise, float* wav_out,
                 float* source_out, uint64_t seed, int32_t flags, void* stream) {
    return h->eng->hift(B, lens, mel, cache_source, cache_stride, cache_len, nsf_noise, wav_out, source_out, seed, flags,
                        (cudaStream_t)stream);
}

int mossdec_hift_skip(mossdec_handle* h, int32_t B, const int32_t* lens, const int32_t* mel_skip, const float* mel,
                      const float* nsf_noise, float* wav_out, float* source_out, uint64_t seed, void* stream) {
    return h->eng->hift(B, lens, mel, nullptr, 0, nullptr, nsf_noise, wav_out, source_out, seed, MOSSDEC_TIME_MAJOR,
                        (cudaStream_t)stream, mel_skip);
}

int mossdec_token2mel(mossdec_handle* h, int32_t B, const int32_t* n_tok, const int64_t* tokens, const int32_t* prompt_frames,
                      const float* prompt_feat, const float* embedding, int32_t ctx_len, int32_t chunk_tokens,
                      int32_t est_chunk_frames, int32_t n_steps, float cfg_rate, float* mel_out, void* stream) {
    return h->eng->token2mel(B, n_tok, tokens, prompt_frames, prompt_feat, embedding, ctx_len, chunk_tokens, est_chunk_frames,
                             n_steps, cfg_rate, mel_out, (cudaStream_t)stream);
}

int mossdec_prefix_reserve(mossdec_handle* h, int32_t n_slots, int32_t prefix_frames, int32_t n_steps) {
    return h->eng->prefix_reserve(n_slots, prefix_frames, n_steps);
}

int mossdec_token2mel_prefix(mossdec_handle* h, int32_t B, const int32_t* n_tok, const int64_t* tokens,
                             const int32_t* prompt_frames, const float* prompt_feat, const float* embedding, int32_t ctx_len,
                             int32_t chunk_tokens, int32_t est_chunk_frames, int32_t n_steps, float cfg_rate, int32_t mode,
                             const int32_t* slots, const uint64_t* tags, float* mel_out, void* stream) {
    Engine::PrefixArgs px;
    px.mode = mode;
    px.slots = slots;
    px.tags = tags;
    return h->eng->token2mel(B, n_tok, tokens, prompt_frames, prompt_feat, embedding, ctx_len, chunk_tokens, est_chunk_frames,
                             n_steps, cfg_rate, mel_out, (cudaStream_t)stream, &px);
}

int mossdec_fade_in_out(mossdec_handle* h, float* speech, const float* prev_tail, const float* window2n, int32_t n,
                        void* stream) {
    return h->eng->fade_in_out(speech, prev_tail, window2n, n, (cudaStream_t)stream);
}

int32_t mossdec_mel_spectrogram(mossdec_handle* h, const float* wav, int64_t n_samples, float* mel_out, void* stream) {
    return h->eng->mel_spectrogram(wav, n_samples, mel_out, (cudaStream_t)stream);
}

int32_t mossdec_mel_spectrogram_static(mossdec_handle* h, const float* wav, int64_t n_samples, float* mel_out, void* stream) {
    return h->eng->mel_spectrogram_static(wav, n_samples, mel_out, (cudaStream_t)stream);
}

int mossdec_cv1_resnet_block(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t Cin, const float* x,
                             const float* mask, const float* time_emb, int32_t time_dim, float* out, void* stream) {
    return h->eng->cv1_resnet_block(prefix, B, T, Cin, x, mask, time_emb, time_dim, out, (cudaStream_t)stream);
}
int mossdec_cv1_block1d(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t Cin, const float* x, const float* mask,
                        float* out, void* stream) {
    return h->eng->cv1_block1d(prefix, B, T, Cin, x, mask, out, (cudaStream_t)stream);
}
int mossdec_cv1_downsample(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t C, const float* x, float* out,
                           void* stream) {
    return h->eng->cv1_downsample(prefix, B, T, C, x, out, (cudaStream_t)stream);
}
int mossdec_cv1_upsample(mossdec_handle* h, const char* prefix, int32_t B, int32_t T, int32_t C, const float* x, float* out,
                         void* stream) {
    return h->eng->cv1_upsample(prefix, B, T, C, x, out, (cudaStream_t)stream);
}

int mossdec_cv1_estimator(mossdec_handle* h, const char* prefix, int32_t n_blocks, int32_t n_mid, int32_t B, int32_t T, const float* x,
                          const float* mask, const float* mu, const float* t, const float* spks, const float* cond, float* out,
                          void* stream) {
    return h->eng->cv1_estimator(prefix, n_blocks, n_mid, B, T, x, mask, mu, t, spks, cond, out, (cudaStream_t)stream);
}

int mossdec_debug_read(mossdec_handle* h, const char* name, float* host_dst, int64_t max_floats) {
    return h->eng->debug_read(name, host_dst, (size_t)max_floats * 4);
}

void mossdec_profile(mossdec_handle* h, int32_t on) { h->eng->set_profile(on != 0); }
int mossdec_profile_read(mossdec_handle* h, double* out9) { return h->eng->profile_read(out9); }
int mossdec_profile_named(mossdec_handle* h, const char* name, double* out3) { return h->eng->profile_named(name, out3); }
void mossdec_set_hift_group_frames(mossdec_handle* h, int64_t frames) { h->eng->set_hift_group_frames(frames); }

int mossdec_reserve(mossdec_handle* h, int64_t flow_rows, int32_t flow_seqs, int64_t hift_frames) {
    return h->eng->reserve(flow_rows, flow_seqs, hift_frames);
}
void mossdec_set_graph_mode(mossdec_handle* h, int32_t mode) { h->eng->set_graph_mode(mode); }
int64_t mossdec_graph_replays(const mossdec_handle* h) { return h->eng->graph_replays(); }
int32_t mossdec_device(const mossdec_handle* h) { return h->eng->device(); }

int64_t mossdec_launch_count(const mossdec_handle* h) { return h->eng->launches(); }
void mossdec_reset_launch_count(mossdec_handle* h) { h->eng->reset_launches(); }

}  // extern "C"
