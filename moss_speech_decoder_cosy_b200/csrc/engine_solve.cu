// Engine entry points of the flow path: single estimator call (reference TensorRT contract) and the full
// classifier-free-guidance Euler solve (cosyvoice/flow/flow_matching.py:76-127, 206-230).
//
// The Euler loop of a solve (n_steps x {both CFG halves of the estimator on two streams, CFG mix + Euler update}, about
// 580 launches per step) depends on the call only through the packed-row geometry, so it is captured ONCE per geometry
// into a CUDA graph and replayed: one cudaGraphLaunch instead of ~5 800 launches, each preceded on the host by 3-4
// tensor-map encodes.  The reference's analogue is the enqueue-once TensorRT context (flow_matching.py:133-156).
// What changes from call to call (caller pointers, sequence layout arrays) stays outside the graph: the layout arrays
// live at fixed device addresses and are refreshed by an async copy, flow_prepare / rows_to_bct are plain launches.
#include "engine.h"

#include <math.h>
#include <string.h>
#include <tuple>

namespace mossdec {

#define MD_CUDA(x)                                                                          \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) return fail(std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

int Engine::estimator(int B2, int T, const float* x, const float* mask, const float* mu, const float* t,
                      const float* spks, const float* cond, float* out, int chunk, cudaStream_t st) {
    DeviceGuard g(device_);
    if (!finalized_) return fail("weights not finalized");
    if (B2 <= 0 || T <= 0) return fail("estimator: bad shape");
    std::vector<int32_t> lens(B2, T);
    bool ragged = false;
    if (mask != nullptr) {
        // The reference multiplies by `mask` after every layer and turns it into the attention bias
        // (decoder.py:433-494); its masks are prefix masks (make_pad_mask).  Here a masked frame has no row at all:
        // read the lengths back (this entry point synchronises anyway, like the TensorRT path it mirrors,
        // flow_matching.py:153-154) and reject anything that is not a prefix mask.
        if (mask_lens_cap_ < B2 + 1) {
            if (h_mask_lens_) cudaFreeHost(h_mask_lens_);
            if (d_mask_lens_) cudaFree(d_mask_lens_);
            h_mask_lens_ = nullptr;
            d_mask_lens_ = nullptr;
            mask_lens_cap_ = 0;
            MD_CUDA(cudaMallocHost(reinterpret_cast<void**>(&h_mask_lens_), (size_t)(B2 + 1) * 4));
            MD_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_mask_lens_), (size_t)(B2 + 1) * 4));
            mask_lens_cap_ = B2 + 1;
        }
        MD_CUDA(cudaMemsetAsync(d_mask_lens_ + B2, 0, 4, st));
        launch_mask_lens(mask, B2, T, d_mask_lens_, d_mask_lens_ + B2, st);
        ++launches_;
        MD_CUDA(cudaMemcpyAsync(h_mask_lens_, d_mask_lens_, (size_t)(B2 + 1) * 4, cudaMemcpyDeviceToHost, st));
        MD_CUDA(cudaStreamSynchronize(st));
        if (h_mask_lens_[B2] != 0) return fail("estimator: mask must be a prefix (padding) mask: 1...1 0...0 per sequence");
        for (int b = 0; b < B2; ++b) {
            lens[b] = h_mask_lens_[b];
            if (lens[b] <= 0) return fail("estimator: a sequence is fully masked");
            ragged |= lens[b] != T;
        }
    }
    if (build_flow_layout(B2, lens.data(), /*cfg_double=*/false, st, /*dense_T=*/T)) return -1;
    const int M = flow_M_;
    launch_est_pack(layout_, M, x, mu, spks, cond, mask, xin_, st);
    ++launches_;
    const int ngroups = cfg_.n_mid + 2;
    const size_t tbytes = ((size_t)B2 * cfg_.time_dim + (size_t)B2 * ngroups * cfg_.channels + 64) * 4;
    ws_time_.reserve(tbytes);
    if (!ws_time_.as<void>()) return fail("cudaMalloc(time workspace) failed");
    float* temb = ws_time_.as<float>();
    float* tbias = temb + (size_t)B2 * cfg_.time_dim;
    if (time_bias(t, B2, temb, tbias, st)) return -1;
    if (estimator_forward(M, B2, flow_max_len_, tbias, ngroups * cfg_.channels, chunk, st)) return -1;
    // the input tensors are fully consumed by est_pack, so `out` may alias `x` (the TensorRT contract binds the output
    // onto x, flow_matching.py:143-151); masked frames of the output are zero like `output * mask` (decoder.py:494)
    if (ragged) MD_CUDA(cudaMemsetAsync(out, 0, (size_t)B2 * cfg_.n_feats * T * 4, st));
    launch_rows_to_bct(layout_, dout_, 0, out, st);
    ++launches_;
    MD_CUDA(cudaGetLastError());
    return 0;
}

// t_span = 1 - cos(linspace(0, 1, n + 1) * pi / 2) in fp32, like torch (flow_matching.py:227-229)
static void cosine_t_span(int n, std::vector<float>& ts) {
    ts.resize(n + 1);
    const float step = 1.0f / (float)n;
    for (int i = 0; i <= n; ++i) {
        // torch.linspace fills the first half forwards and the second half backwards from `end`
        const float v = (i < (n + 1) / 2) ? step * (float)i : 1.0f - step * (float)(n - i);
        ts[i] = 1.0f - cosf(v * 0.5f * 3.14159265358979323846f);
    }
}

bool Engine::GraphKey::operator<(const GraphKey& o) const {
    return std::tie(rows_half, nseq_half, max_len, chunk, n_steps, dual, cfg_bits, ws, lay, tb, px, pxi) <
           std::tie(o.rows_half, o.nseq_half, o.max_len, o.chunk, o.n_steps, o.dual, o.cfg_bits, o.ws, o.lay, o.tb, o.px, o.pxi);
}

void Engine::drop_graphs() {
    for (auto& kv : graphs_)
        if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    graphs_.clear();
}

// The Euler loop: every step runs the estimator on both CFG halves and applies the CFG mix + Euler update.
int Engine::run_euler_steps(int n_steps, int chunk, float cfg_rate, const float* dtv, cudaStream_t st, int px_mode) {
    const int ngroups = cfg_.n_mid + 2;
    for (int s = 0; s < n_steps; ++s) {
        // tbias_ld = 0: every sequence reads the row of this step
        const float* tb = tbias_ + (size_t)s * ngroups * cfg_.channels;
        PrefixCtx pc0{px_mode, px_frames_, s, 0}, pc1{px_mode, px_frames_, s, 1};
        const bool two = dual_stream_ && !profile_;
        if (two || px_mode != 0) {
            // The two CFG halves are independent until the Euler update: each runs its kernel chain on its own
            // stream, so the partial last wave of one half's persistent kernels (654 tiles on 148 SMs = 4.4 waves)
            // and its launch gaps are filled by the other half's kernels.  (Profiling runs serially so that the
            // per-kernel event times are not stretched by co-running kernels; with a prefix cache the halves always
            // run as two chains because each has its own cache slabs.)
            cudaStream_t s2 = two ? st2_ : st;
            if (two) {
                MD_CUDA(cudaEventRecord(ev_fork_, st));
                MD_CUDA(cudaStreamWaitEvent(st2_, ev_fork_, 0));
            }
            if (estimator_forward(flow_rows_half_, flow_nseq_ / 2, flow_max_len_, tb, 0, chunk, st, 0, px_mode ? &pc0 : nullptr)) return -1;
            if (estimator_forward(flow_rows_half_, flow_nseq_ / 2, flow_max_len_, tb, 0, chunk, s2, flow_rows_half_, px_mode ? &pc1 : nullptr)) return -1;
            if (two) {
                MD_CUDA(cudaEventRecord(ev_join_, st2_));
                MD_CUDA(cudaStreamWaitEvent(st, ev_join_, 0));
            }
        } else if (estimator_forward(flow_M_, flow_nseq_, flow_max_len_, tb, 0, chunk, st)) {
            return -1;
        }
        // read d_cond, d_uncond, x + write x (4 x 80 fp32 = 1 280 B per frame) + bf16 x slab of both halves
        bw_begin("euler_cfg_kernel", (double)flow_rows_half_ * (4.0 * 80 * 4 + 2.0 * 80 * 2), st);
        launch_euler_cfg(layout_, dout_, dtv[s], cfg_rate, xstate_, xin_, st);
        bw_end(st);
        ++launches_;
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------ frozen-prefix cache
// Streaming (`stream=True`) solves use block-causal attention with chunks of `chunk` frames and causal convolutions, so
// frame i of the estimator depends on frames < (i / chunk + 1) * chunk only.  A stream's flow input always starts with
// the same prompt (tokens, prompt mel, speaker, and the fixed noise bank by position), so the first whole chunks that lie
// inside the prompt -- the frozen prefix -- are IDENTICAL in every chunk call of that stream: the reference recomputes
// them every time (flow_inference.py:188-218 re-runs the flow over prompt + window).  Here their per-layer K | V rows and
// the two-row left context of every causal conv are recorded once per stream (mode 1) for every (Euler step, CFG branch);
// later chunk calls (mode 2) pack only the live frames, take the frozen keys from the cache inside the attention kernel
// (same key tiling as in the recording call, so the live frames come out bit-identical) and fill the conv halos from it.
// Exact for any window that slides behind the prefix, including `max_token_len` windows; not an approximation.
int Engine::prefix_reserve(int n_slots, int prefix_frames, int n_steps) {
    DeviceGuard g(device_);
    if (n_slots <= 0 || prefix_frames < 2 || n_steps <= 0) return fail("prefix_reserve: bad arguments");
    const int ngroups = cfg_.n_mid + 2, nblk = ngroups * cfg_.n_blocks, nconv = 2 * ngroups + 3;
    px_slots_ = n_slots;
    px_frames_ = prefix_frames;
    px_steps_ = n_steps;
    const size_t kv = (size_t)n_steps * 2 * nblk * px_kv_slab_elems();
    const size_t halo = (size_t)n_steps * 2 * nconv * ((size_t)n_slots * 2 * 512);
    cudaDeviceSynchronize();
    drop_graphs();
    ws_prefix_.release();
    ws_prefix_.reserve((kv + halo) * 2 + 1024);
    if (!ws_prefix_.as<void>()) {
        px_slots_ = 0;
        return fail("cudaMalloc(prefix cache) failed");
    }
    px_tags_.assign(n_slots, 0);
    return 0;
}

__nv_bfloat16* Engine::px_kv(int step, int branch, int block) const {
    const int nblk = (cfg_.n_mid + 2) * cfg_.n_blocks;
    return ws_prefix_.as<__nv_bfloat16>() + ((size_t)(step * 2 + branch) * nblk + block) * px_kv_slab_elems();
}
__nv_bfloat16* Engine::px_halo(int step, int branch, int conv) const {
    const int ngroups = cfg_.n_mid + 2, nblk = ngroups * cfg_.n_blocks, nconv = 2 * ngroups + 3;
    const size_t kv = (size_t)px_steps_ * 2 * nblk * px_kv_slab_elems();
    return ws_prefix_.as<__nv_bfloat16>() + kv + ((size_t)(step * 2 + branch) * nconv + conv) * ((size_t)px_slots_ * 2 * 512);
}

int Engine::flow_solve(int B, const int32_t* lens, const float* mu, const float* spks, const float* cond,
                       const float* noise, int noise_len_arg, int n_steps, float cfg_rate, float temperature, int chunk,
                       float* mel_out, int flags, cudaStream_t st, const PrefixArgs* px) {
    DeviceGuard g(device_);
    if (!finalized_) return fail("weights not finalized");
    if (B <= 0 || n_steps <= 0) return fail("flow_solve: bad arguments");
    const int px_mode = px ? px->mode : 0;
    uint32_t cfg_bits = 0;
    memcpy(&cfg_bits, &cfg_rate, 4);
    std::vector<int32_t> live_lens;
    if (px_mode != 0) {
        if (px_slots_ <= 0) return fail("prefix cache not reserved (mossdec_prefix_reserve)");
        if (px_mode < 0 || px_mode > 2 || !px->slots || !px->tags) return fail("flow_solve: bad prefix arguments");
        if (chunk <= 0 || (px_frames_ % chunk) != 0) return fail("the frozen prefix must be a whole number of attention chunks");
        if (n_steps != px_steps_) return fail("flow_solve: n_steps differs from the prefix cache");
        if (temperature != 1.0f || noise != nullptr) return fail("prefix reuse needs the built-in noise bank at temperature 1");
        std::vector<char> used(px_slots_, 0);
        for (int b = 0; b < B; ++b) {
            const int sl = px->slots[b];
            if (sl < 0 || sl >= px_slots_ || used[sl]) return fail("flow_solve: bad or repeated prefix cache slot");
            used[sl] = 1;
            if (lens[b] <= px_frames_) return fail("an utterance is not longer than the frozen prefix");
            if (px->tags[b] == 0) return fail("prefix tags must be non-zero");
        }
        if (px_mode == 2) {
            if (chunk != px_chunk_ || cfg_bits != px_cfg_bits_) return fail("prefix cache was recorded with another chunk / cfg rate");
            for (int b = 0; b < B; ++b)
                if (px_tags_[px->slots[b]] != px->tags[b])
                    return fail("prefix cache slot holds another stream (tag mismatch): record it first");
            live_lens.resize(B);
            for (int b = 0; b < B; ++b) live_lens[b] = lens[b] - px_frames_;
        } else {
            // contents recorded under other solver settings are invalid from here on
            if (chunk != px_chunk_ || cfg_bits != px_cfg_bits_) std::fill(px_tags_.begin(), px_tags_.end(), 0);
            px_chunk_ = chunk;
            px_cfg_bits_ = cfg_bits;
            for (int b = 0; b < B; ++b) px_tags_[px->slots[b]] = px->tags[b];
        }
    }
    const int noise_bstride = 0;  // one bank shared by all utterances, like the reference
    int noise_len = noise_len_arg;
    if (noise == nullptr) {
        noise = W<float>("flow.rand_noise");
        if (!noise) return fail("flow.rand_noise not loaded and no noise given");
        noise_len = cfg_.noise_len;
    }
    for (int b = 0; b < B; ++b)
        if (lens[b] > noise_len) return fail("sequence longer than the noise bank");
    if (build_flow_layout(B, px_mode == 2 ? live_lens.data() : lens, /*cfg_double=*/true, st, 0, px_mode == 2 ? px_frames_ : 0)) return -1;
    layout_.noise_cstride = noise_len;
    layout_.time_major = flags & 1;
    if (px_mode != 0) {   // device index arrays of this call: slot, first cache row, and the record-mode sub-layouts
        ws_prefix_idx_.reserve((size_t)B * 5 * 4 + 256);
        if (!ws_prefix_idx_.as<void>()) return fail("cudaMalloc(prefix index) failed");
        int32_t* hp = static_cast<int32_t*>(stager_.acquire((size_t)B * 5 * 4));
        if (!hp) return fail("cudaMallocHost(staging) failed");
        int r = 0;
        for (int b = 0; b < B; ++b) {
            r += 4;   // kGuard rows in front of every sequence (build_flow_layout)
            hp[b] = px->slots[b];
            hp[B + b] = px->slots[b] * px_frames_;
            hp[2 * B + b] = px_frames_;
            hp[3 * B + b] = r + px_frames_;            // record mode: first live row
            hp[4 * B + b] = lens[b] - px_frames_;      // live frames
            r += px_mode == 2 ? live_lens[b] : lens[b];
        }
        int32_t* d = ws_prefix_idx_.as<int32_t>();
        MD_CUDA(cudaMemcpyAsync(d, hp, (size_t)B * 5 * 4, cudaMemcpyHostToDevice, st));
        stager_.commit(st);
        px_d_slot_ = d;
        px_d_prow_ = d + B;
        px_d_len_pref_ = d + 2 * B;
        px_d_start_live_ = d + 3 * B;
        px_d_len_live_ = d + 4 * B;
    }
    // algorithmic bytes per row of the conditional half: read mu, cond, noise (3 x 80 fp32), write xstate (80 fp32) and
    // the 320-channel bf16 input rows of both CFG halves
    bw_begin("flow_prepare_kernel", (double)flow_rows_half_ * (4.0 * 80 * 4 + 2.0 * 320 * 2), st);
    launch_flow_prepare(layout_, mu, spks, cond, noise, noise_bstride, temperature, xin_, xstate_, st);
    bw_end(st);
    ++launches_;

    std::vector<float> ts;
    cosine_t_span(n_steps, ts);
    // t accumulates in fp32 exactly like the reference loop (t = t + dt; dt = t_span[step + 1] - t)
    std::vector<float> tv(n_steps), dtv(n_steps);
    float t = ts[0], dt = ts[1] - ts[0];
    for (int s = 1; s <= n_steps; ++s) {
        tv[s - 1] = t;
        dtv[s - 1] = dt;
        t = t + dt;
        if (s < n_steps) dt = ts[s + 1] - t;
    }
    // All time-embedding biases of the solve at once: the timestep is shared by every sequence of the batch
    // (t_in[:] = t, flow_matching.py:110), so one row per step.  They depend on n_steps only: computed once and kept.
    const int ngroups = cfg_.n_mid + 2;
    const size_t need = ((size_t)n_steps * cfg_.time_dim + (size_t)n_steps * ngroups * cfg_.channels + n_steps + 64) * 4;
    if (ws_time_solve_.reserve(need)) solve_tbias_steps_ = 0;
    if (!ws_time_solve_.as<void>()) return fail("cudaMalloc(time workspace) failed");
    temb_ = ws_time_solve_.as<float>();
    tbias_ = temb_ + (size_t)n_steps * cfg_.time_dim;
    tvals_ = tbias_ + (size_t)n_steps * ngroups * cfg_.channels;
    if (solve_tbias_steps_ != n_steps) {
        float* hp = static_cast<float*>(stager_.acquire((size_t)n_steps * 4));
        if (!hp) return fail("cudaMallocHost(staging) failed");
        memcpy(hp, tv.data(), (size_t)n_steps * 4);
        MD_CUDA(cudaMemcpyAsync(tvals_, hp, (size_t)n_steps * 4, cudaMemcpyHostToDevice, st));
        stager_.commit(st);
        if (time_bias(tvals_, n_steps, temb_, tbias_, st)) return -1;
        solve_tbias_steps_ = n_steps;
    }

    bool done = false;
    if (graph_mode_ > 0 && !profile_ && px_mode != 1) {   // (a recording call happens once per stream: plain launches)
        GraphKey key{flow_rows_half_, B, flow_max_len_, chunk, n_steps, (dual_stream_ ? 1 : 0) | (px_mode << 1) | (px_frames_ << 4), cfg_bits,
                     ws_flow_.as<void>(), ws_layout_.as<void>(), ws_time_solve_.as<void>(),
                     px_mode ? ws_prefix_.as<void>() : nullptr, px_mode ? ws_prefix_idx_.as<void>() : nullptr};
        GraphEntry& ge = graphs_[key];
        ++ge.seen;
        ge.last_use = ++graph_clock_;
        if (ge.exec == nullptr && ge.seen >= (graph_mode_ >= 2 ? 1 : 2)) {
            // capture on an internal stream (the caller's may be the legacy stream, which cannot capture)
            const long l0 = launches_;
            cudaGraph_t graph = nullptr;
            if (cudaStreamBeginCapture(cap_st_, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
                const int rc = run_euler_steps(n_steps, chunk, cfg_rate, dtv.data(), cap_st_, px_mode);
                const cudaError_t ee = cudaStreamEndCapture(cap_st_, &graph);
                if (rc == 0 && ee == cudaSuccess && graph != nullptr &&
                    cudaGraphInstantiate(&ge.exec, graph, 0) == cudaSuccess) {
                    ge.launches = launches_ - l0;
                } else {
                    ge.exec = nullptr;
                    graph_mode_ = 0;   // capture is not possible here: fall back to plain launches for good
                    cudaGetLastError();
                }
                if (graph) cudaGraphDestroy(graph);
            } else {
                graph_mode_ = 0;
                cudaGetLastError();
            }
            launches_ = l0;
            if (graphs_.size() > 24) {   // bound the cache: drop the least recently used executable
                auto victim = graphs_.end();
                for (auto it = graphs_.begin(); it != graphs_.end(); ++it)
                    if (it->second.exec && &it->second != &ge && (victim == graphs_.end() || it->second.last_use < victim->second.last_use))
                        victim = it;
                if (victim != graphs_.end()) {
                    cudaGraphExecDestroy(victim->second.exec);
                    graphs_.erase(victim);
                }
            }
        }
        if (ge.exec != nullptr) {
            MD_CUDA(cudaGraphLaunch(ge.exec, st));
            launches_ += ge.launches;
            ++graph_replays_;
            done = true;
        }
    }
    if (!done && run_euler_steps(n_steps, chunk, cfg_rate, dtv.data(), st, px_mode)) return -1;
    launch_rows_to_bct(layout_, xstate_, 0, mel_out, st);
    ++launches_;
    MD_CUDA(cudaGetLastError());
    return 0;
}

int Engine::reserve(int64_t flow_rows, int flow_seqs, int64_t hift_frames) {
    DeviceGuard g(device_);
    if (flow_rows > 0) {
        const size_t rows_half = ((size_t)flow_rows + 4 * (size_t)(flow_seqs + 1) + kGemmBM - 1) / kGemmBM * kGemmBM;
        if (reserve_flow(2 * rows_half, 2 * (flow_seqs > 0 ? flow_seqs : 1))) return -1;
    }
    if (hift_frames > 0 && reserve_hift(hift_frames, flow_seqs > 0 ? flow_seqs : 1)) return -1;
    return 0;
}

}  // namespace mossdec
