// Decode engine behind the C-ABI (include/mossdec.h): owns device weights + workspace and enqueues the kernel
// sequences of the flow-matching solver (cosyvoice/flow/flow_matching.py:76-127 over decoder.py:405-494) and of the
// HiFT vocoder (cosyvoice/hifigan/generator.py:520-582).  No torch types; raw device pointers and streams only.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "gemm.h"
#include "kernels.h"

namespace mossdec {

// Every entry point of an Engine runs with the engine's device current and restores the caller's device afterwards.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        if (switched) cudaSetDevice(prev);
    }
};

struct DevTensor {
    void* ptr = nullptr;
    size_t bytes = 0;
    int dtype = 0;  // 0 = f32, 1 = bf16, 2 = i32
    std::vector<int64_t> shape;
};

struct EngineConfig {
    int n_feats = 80;
    int channels = 256;
    int heads = 8;
    int head_dim = 64;
    int n_blocks = 4;       // transformer blocks per group
    int n_mid = 12;         // mid groups
    int time_dim = 1024;
    int in_channels = 320;  // x | mu | spks | cond
    int noise_len = 15000;
    // HiFT
    int hift_base = 512;
    int n_harm = 9;         // nb_harmonics + 1
    float sample_rate = 24000.f;
    float sine_amp = 0.1f, noise_std = 0.003f, voiced_thr = 10.f, lrelu = 0.1f, audio_limit = 0.99f;
};

// Pinned host staging for the small per-call layout arrays (sequence starts / lengths / row -> sequence maps): the
// arrays are built in a pinned slot and uploaded with cudaMemcpyAsync on the caller's stream, so an entry point never
// blocks the host and the copy is ordered behind the kernels of an earlier call on the same stream that still read the
// previous layout.  A slot is reused only after the copy that read it has completed (event per slot).
class PinnedStager {
   public:
    ~PinnedStager();
    void* acquire(size_t bytes);   // host pointer to a pinned slot of >= bytes (nullptr on allocation failure)
    void commit(cudaStream_t st);  // call after the cudaMemcpyAsync(s) reading the slot returned by the last acquire()
   private:
    static constexpr int kSlots = 4;
    struct Slot {
        void* p = nullptr;
        size_t cap = 0;
        cudaEvent_t ev = nullptr;
        bool pending = false;
    } slots_[kSlots];
    int cur_ = -1;
};

class DeviceBuffer {
   public:
    ~DeviceBuffer() { release(); }
    // grows (never shrinks); returns true when the allocation moved.  Growing synchronises the device (cudaFree of the old
    // block + a completed memset of the new one), so it is safe with work in flight; steady state never allocates.
    bool reserve(size_t bytes);
    void release();
    template <typename T>
    T* as() const { return reinterpret_cast<T*>(ptr_); }
    size_t size() const { return bytes_; }

   private:
    void* ptr_ = nullptr;
    size_t bytes_ = 0;
};

class Engine {
   public:
    explicit Engine(const EngineConfig& cfg);
    ~Engine();

    // Frozen-prefix cache for streaming solves (engine_solve.cu).  With block-causal masks the
    // first `prefix_frames` frames of a stream (whole estimator chunks inside the prompt) are identical in every chunk of
    // that stream: their per-layer K | V and the two-row left context of every causal conv are recorded once
    // (mode 1) and later chunks only compute the live frames behind them (mode 2).
    int prefix_reserve(int n_slots, int prefix_frames, int n_steps);
    struct PrefixArgs {
        int mode = 0;                    // 0 = plain, 1 = record (full solve that also fills the cache), 2 = reuse
        const int32_t* slots = nullptr;  // HOST [B] cache slot of every utterance
        const uint64_t* tags = nullptr;  // HOST [B] caller's identity of the frozen inputs (checked on reuse)
    };
    int set_tensor(const char* name, const void* host, int dtype, int ndim, const int64_t* shape);
    int finalize();

    // 1:1 with the reference's TensorRT estimator contract (flow_matching.py:133-156): dense [B2, 80, T] inputs.
    int estimator(int B2, int T, const float* x, const float* mask, const float* mu, const float* t,
                  const float* spks, const float* cond, float* out, int chunk, cudaStream_t st);
    // Whole Euler solve with CFG batch doubling for B ragged utterances (lens[b] frames each, tensors [C, T_b] back
    // to back).  noise == nullptr -> built-in bank uploaded as tensor "flow.rand_noise".
    int flow_solve(int B, const int32_t* lens, const float* mu, const float* spks, const float* cond,
                   const float* noise, int noise_len_arg, int n_steps, float cfg_rate, float temperature, int chunk, float* mel_out,
                   int flags, cudaStream_t st, const PrefixArgs* px = nullptr);
    // HiFT vocoder for B ragged mel segments ([80, T_b] back to back) -> wav [480 T_b] back to back.
    int hift(int B, const int32_t* lens, const float* mel, const float* cache_source, int cache_stride,
             const int32_t* cache_len, const float* nsf_noise, float* wav_out, float* source_out, uint64_t seed,
             int flags, cudaStream_t st, const int32_t* mel_skip = nullptr);
    // per-launch device timing of the GEMM (0), attention (1) and fused-FFN (2) kernels with CUDA events on the launch stream
    void set_profile(bool on) { profile_ = on; }
    int profile_read(double* out9);
    // {ms, launches, algorithmic bytes} of a named bandwidth kernel accumulated by profile_read() calls; clears the entry
    int profile_named(const char* name, double* out3);
    // test hook: copy a named intermediate of the last call to the host; returns floats copied or -1
    int debug_read(const char* name, void* host_dst, size_t bytes);
    int fade_in_out(float* speech, const float* prev_tail, const float* window2n, int n, cudaStream_t st);
    // CausalMaskedDiffWithXvec.inference for B utterances (flow.py:237-283; engine_encoder.cu): tokens = ragged device
    // int64 (prompt + new tokens + `ctx` look-ahead tokens per utterance, n_tok[b] each, HOST counts), prompt_feat =
    // ragged device [F_b, 80], embedding device [B, 192] -> mel_out ragged time-major [4 (n_tok[b] - ctx), 80].
    int token2mel(int B, const int32_t* n_tok, const int64_t* tokens, const int32_t* prompt_frames, const float* prompt_feat,
                  const float* embedding, int ctx, int chunk_tokens, int est_chunk_frames, int n_steps, float cfg_rate,
                  float* mel_out, cudaStream_t st, const PrefixArgs* px = nullptr);
    // matcha mel_spectrogram (n_fft 1920, hop 480, 80 mels, 24 kHz, fmax 8000, center False; mel_kernels.cu):
    // wav device [L] -> mel_out device [80, T], T = (L - 480) / 480 + 1.  Returns T, negative on error.
    int mel_spectrogram(const float* wav, int64_t L, float* mel_out, cudaStream_t st);
    // the same on a file-scope static device scratch (<= 512 frames): no cudaMalloc / cudaFree on any call (mel_kernels.cu)
    int mel_spectrogram_static(const float* wav, int64_t L, float* mel_out, cudaStream_t st);
    // Building blocks of the non-causal cosyvoice1 / Matcha estimator (cv1_ops.cu): dense
    // [B, C, T] fp32 in / out like the reference modules, weights looked up as "<prefix>.*" (weights.py pack_cv1_*).
    int cv1_resnet_block(const char* prefix, int B, int T, int Cin, const float* x, const float* mask, const float* temb,
                         int time_dim, float* out, cudaStream_t st);
    int cv1_block1d(const char* prefix, int B, int T, int Cin, const float* x, const float* mask, float* out, cudaStream_t st);
    int cv1_downsample(const char* prefix, int B, int T, int C, const float* x, float* out, cudaStream_t st);
    int cv1_upsample(const char* prefix, int B, int T, int C, const float* x, float* out, cudaStream_t st);
    // the assembled two-level non-causal decoder (ConditionalDecoder.forward, flow/decoder.py:206-291) and its remaining parts
    int cv1_conv(const char* prefix, int B, int T, int Cin, int Cout, int k, const float* x, const float* mask_in,
                 const float* mask_out, float* out, cudaStream_t st);
    int cv1_transformer_group(const char* prefix, int n_blocks, int B, int T, const float* x, const int32_t* lens, float* out,
                              cudaStream_t st);
    int cv1_estimator(const char* prefix, int n_blocks, int n_mid, int B, int T, const float* x, const float* mask, const float* mu,
                      const float* t, const float* spks, const float* cond, float* out, cudaStream_t st);

    const char* last_error() const { return err_.c_str(); }
    long launches() const { return launches_; }
    void reset_launches() { launches_ = 0; }
    void set_hift_group_frames(int64_t f) { hift_group_frames_ = f > 0 ? f : 8192; }
    int num_sms() const { return num_sms_; }
    int device() const { return device_; }
    // Pre-size the workspaces for solves of up to `flow_rows` packed rows per CFG half over `flow_seqs` sequences and
    // vocoder groups of up to `hift_frames` frames, so that the first real call does not allocate.
    int reserve(int64_t flow_rows, int flow_seqs, int64_t hift_frames);
    // CUDA-graph replay of the Euler loop of flow_solve (MOSSDEC_GRAPH=0 disables): 0 = off, 1 = capture a shape the
    // second time it is seen (default), 2 = capture on first sight
    void set_graph_mode(int m) { graph_mode_ = m; }
    long graph_replays() const { return graph_replays_; }

   private:
    friend struct FlowRunner;
    friend struct HiftRunner;
    int fail(const std::string& m) {
        err_ = m;
        return -1;
    }
    const DevTensor* find(const std::string& name);
    template <typename T>
    const T* W(const std::string& name) {
        const DevTensor* t = find(name);
        return t ? reinterpret_cast<const T*>(t->ptr) : nullptr;
    }
    int gemm(const GemmProblem& p, cudaStream_t st);
    int ffn(const FfnProblem& p, cudaStream_t st);

    // flow internals
    struct FlowPlan;
    int build_flow_layout(int nseq_half, const int32_t* lens, bool cfg_double, cudaStream_t st, int dense_T = 0, int frame0 = 0);
    int reserve_flow(size_t M, int nseq);
    int reserve_hift(int64_t rows_f, int nseq);
    int run_euler_steps(int n_steps, int chunk, float cfg_rate, const float* dtv, cudaStream_t st, int px_mode = 0);
    // row0 > 0: run on the rows [row0, row0 + M) of every workspace matrix (the second CFG half on its own stream)
    // per-call view of the prefix cache for one (Euler step, CFG branch); nullptr = plain
    struct PrefixCtx {
        int mode, prefix, step, branch;
    };
    int estimator_forward(int M, int nseq, int max_len, const float* tbias, int tbias_ld, int chunk, cudaStream_t st,
                          int row0 = 0, const PrefixCtx* pc = nullptr);
    int prefix_halo(const PrefixCtx* pc, int conv_id, __nv_bfloat16* buf, int64_t ld, int width, int nseq, cudaStream_t st);
    int time_bias(const float* t_dev, int nt, float* temb, float* tbias, cudaStream_t st);
    int finalize_hift();
    int finalize_encoder();
    int encode_group(int b0, int b1, const int32_t* n_tok, const int64_t* tok_off, const int64_t* mel_off,
                     const int32_t* prompt_frames, const int64_t* feat_off, const int64_t* tokens, const float* prompt_feat,
                     int ctx, int chunk_tokens, float* mu, float* cond, cudaStream_t st);
    int encoder_layer(int l, int R, const int32_t* row_seq, const int32_t* seq_start, const int32_t* seq_len, int nseq,
                      int max_len, int chunk, const __nv_bfloat16* P, int npad, int tmax, bool emit_bf16, cudaStream_t st);
    int encoder_pos_tables(int l0, int l1, int tmax, int npad, __nv_bfloat16* P, cudaStream_t st);
    int hift_group(int b0, int b1, const int32_t* lens, int64_t mel_off0, int64_t wav_off0, const float* mel,
                   const float* cache_source, int cache_stride, const int32_t* cache_len, const float* nsf_noise,
                   float* wav_out, float* source_out, uint64_t seed, int flags, cudaStream_t st);

    EngineConfig cfg_;
    int device_ = 0;                   // the device this engine lives on; every entry point runs under it (DeviceGuard)
    int num_sms_ = 148;
    bool finalized_ = false;
    bool fuse_oproj_ = true;           // attention out-projection fused into ffn_tc_kernel (MOSSDEC_FUSE_OPROJ=0: separate GEMM)
    bool dual_stream_ = true;          // the two CFG halves of a solve run on two streams (MOSSDEC_DUAL=0 disables)
    cudaStream_t st2_ = nullptr;
    cudaStream_t cap_st_ = nullptr;    // capture origin stream (the caller's stream may be the legacy stream, which cannot capture)
    cudaEvent_t ev_fork_ = nullptr, ev_join_ = nullptr;
    PinnedStager stager_;
    int32_t* h_mask_lens_ = nullptr;   // pinned [max B2 + 1]: sequence lengths + "not a prefix mask" flag read back by estimator()
    int32_t* d_mask_lens_ = nullptr;
    int mask_lens_cap_ = 0;

    // ---- CUDA graphs of the Euler loop (engine_solve.cu)
    struct GraphKey {
        int rows_half, nseq_half, max_len, chunk, n_steps, dual;
        uint32_t cfg_bits;
        const void *ws, *lay, *tb, *px, *pxi;
        bool operator<(const GraphKey& o) const;
    };
    struct GraphEntry {
        cudaGraphExec_t exec = nullptr;
        long launches = 0;
        long seen = 0;
        uint64_t last_use = 0;
    };
    std::map<GraphKey, GraphEntry> graphs_;
    int graph_mode_ = 1;
    uint64_t graph_clock_ = 0;
    long graph_replays_ = 0;
    void drop_graphs();
    // ---- frozen-prefix cache (engine_solve.cu)
    DeviceBuffer ws_prefix_, ws_prefix_idx_;
    int px_slots_ = 0, px_frames_ = 0, px_steps_ = 0;
    int px_chunk_ = 0;
    uint32_t px_cfg_bits_ = 0;
    std::vector<uint64_t> px_tags_;          // host: what every slot holds (0 = empty)
    int32_t *px_d_slot_ = nullptr, *px_d_prow_ = nullptr, *px_d_len_pref_ = nullptr, *px_d_start_live_ = nullptr,
            *px_d_len_live_ = nullptr;      // device index arrays of the current call ([nseq_half] each)
    __nv_bfloat16* px_kv(int step, int branch, int block) const;
    __nv_bfloat16* px_halo(int step, int branch, int conv) const;
    size_t px_kv_slab_elems() const { return ((size_t)px_slots_ * px_frames_ + 64) * 2 * cfg_.heads * cfg_.head_dim; }
    int solve_tbias_steps_ = 0;        // ws_time_solve_ holds the time-embedding biases of an n-step cosine schedule
    std::unordered_map<std::string, DevTensor> tensors_;
    std::string err_;
    std::string missing_;
    long launches_ = 0;

    // ---- flow workspace
    DeviceBuffer ws_flow_, ws_layout_, ws_time_, ws_time_solve_;
    std::vector<int32_t> h_row_seq_, h_seq_start_, h_seq_len_;
    std::vector<int64_t> h_in_offset_;
    FlowLayout layout_{};
    int flow_M_ = 0, flow_rows_half_ = 0, flow_nseq_ = 0, flow_max_len_ = 0;
    __nv_bfloat16 *xin_ = nullptr, *xn_ = nullptr, *h1_ = nullptr, *qkv_ = nullptr, *att_ = nullptr, *ff_ = nullptr,
                  *cat_ = nullptr;
    float *xstate_ = nullptr, *hres_ = nullptr, *rres_ = nullptr, *dout_ = nullptr;
    float *temb_ = nullptr, *tbias_ = nullptr, *tvals_ = nullptr;

    // ---- cosyvoice1 building-block workspace (cv1_ops.cu)
    DeviceBuffer ws_cv1_, ws_cv1e_;
    DeviceBuffer ws_mel_;   // prompt mel front end (mel_kernels.cu)

    // ---- encoder front end (engine_encoder.cu)
    DeviceBuffer ws_enc_, ws_enc_layout_, ws_mu_;
    bool has_encoder_ = false;
    std::string enc_missing_;
    float *enc_x32_ = nullptr, *enc_pe_ = nullptr, *enc_ptmp_ = nullptr;
    __nv_bfloat16 *enc_xn_ = nullptr, *enc_qkv_ = nullptr, *enc_att_ = nullptr, *enc_ffh_ = nullptr, *enc_bias_ = nullptr;
    int enc_tmax_cached_ = -1;
    const void *enc_ptok_cached_ = nullptr, *enc_pmel_cached_ = nullptr;

    // ---- hift workspace
    DeviceBuffer ws_hift_, ws_hift_layout_;
    bool has_hift_ = false;
    std::map<std::string, std::pair<const void*, size_t>> taps_;
    struct ProfRec { cudaEvent_t a, b; int cls; double flops; uint64_t key; };
    std::map<uint64_t, std::pair<double, long>> prof_by_key_;
    bool profile_ = false;
    std::vector<ProfRec> prof_;
    std::vector<cudaEvent_t> prof_pool_;
    cudaEvent_t prof_event();
    void prof_begin(int cls, double flops, cudaStream_t st, uint64_t key = 0);
    void prof_end(cudaStream_t st);
    // bandwidth-class kernels: event pair around the launch + its ALGORITHMIC bytes, accumulated per kernel name
    struct BwRec { cudaEvent_t a, b; const char* name; double bytes; };
    struct BwAcc { double ms = 0, n = 0, bytes = 0; };
    std::vector<BwRec> bw_prof_;
    std::map<std::string, BwAcc> bw_acc_;
    void bw_begin(const char* name, double bytes, cudaStream_t st) {
        if (!profile_) return;
        BwRec r{prof_event(), prof_event(), name, bytes};
        cudaEventRecord(r.a, st);
        bw_prof_.push_back(r);
    }
    void bw_end(cudaStream_t st) {
        if (profile_) cudaEventRecord(bw_prof_.back().b, st);
    }
    const int32_t* hift_mel_skip_ = nullptr;   // (valid during a hift() call) leading frames of every mel block to skip
    int64_t hift_group_frames_ = 8192;  // frames per HiFT group (bounds the stage-2 working set)
};

}  // namespace mossdec
