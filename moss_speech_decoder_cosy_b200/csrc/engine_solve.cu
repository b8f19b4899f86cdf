// Engine entry points of the flow path: single estimator call (reference TensorRT contract) and the full
// classifier-free-guidance Euler solve (cosyvoice/flow/flow_matching.py:76-127, 206-230).
#include "engine.h"

#include <math.h>

namespace mossdec {

#define MD_CUDA(x)                                                                          \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) return fail(std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

int Engine::estimator(int B2, int T, const float* x, const float* mask, const float* mu, const float* t,
                      const float* spks, const float* cond, float* out, int chunk, cudaStream_t st) {
    if (!finalized_) return fail("weights not finalized");
    if (B2 <= 0 || T <= 0) return fail("estimator: bad shape");
    std::vector<int32_t> lens(B2, T);
    if (build_flow_layout(B2, lens.data(), /*cfg_double=*/false)) return -1;
    layout_.time_major = 0;
    const int M = flow_M_;
    launch_est_pack(layout_, M, x, mu, spks, cond, mask, xin_, st);
    ++launches_;
    if (time_bias(t, B2, st)) return -1;
    const int ngroups = cfg_.n_mid + 2;
    if (estimator_forward(M, B2, T, tbias_, ngroups * cfg_.channels, chunk, st)) return -1;
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

int Engine::flow_solve(int B, const int32_t* lens, const float* mu, const float* spks, const float* cond,
                       const float* noise, int noise_len_arg, int n_steps, float cfg_rate, float temperature, int chunk,
                       float* mel_out, int flags, cudaStream_t st) {
    if (!finalized_) return fail("weights not finalized");
    if (B <= 0 || n_steps <= 0) return fail("flow_solve: bad arguments");
    const int noise_bstride = 0;  // one bank shared by all utterances, like the reference
    int noise_len = noise_len_arg;
    if (noise == nullptr) {
        noise = W<float>("flow.rand_noise");
        if (!noise) return fail("flow.rand_noise not loaded and no noise given");
        noise_len = cfg_.noise_len;
    }
    for (int b = 0; b < B; ++b)
        if (lens[b] > noise_len) return fail("sequence longer than the noise bank");
    if (build_flow_layout(B, lens, /*cfg_double=*/true)) return -1;
    layout_.noise_cstride = noise_len;
    layout_.time_major = flags & 1;
    const int M = flow_M_;
    launch_flow_prepare(layout_, mu, spks, cond, noise, noise_bstride, temperature, xin_, xstate_, st);
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
    // all time-embedding biases of the solve at once: the timestep is shared by every sequence of the batch
    // (t_in[:] = t, flow_matching.py:110), so one row per step
    const int ngroups = cfg_.n_mid + 2;
    const size_t need = ((size_t)n_steps * cfg_.time_dim + (size_t)n_steps * ngroups * cfg_.channels + n_steps + 64) * 4;
    ws_time_.reserve(need);
    if (!ws_time_.as<void>()) return fail("cudaMalloc(time workspace) failed");
    temb_ = ws_time_.as<float>();
    tbias_ = temb_ + (size_t)n_steps * cfg_.time_dim;
    tvals_ = tbias_ + (size_t)n_steps * ngroups * cfg_.channels;
    MD_CUDA(cudaMemcpyAsync(tvals_, tv.data(), (size_t)n_steps * 4, cudaMemcpyHostToDevice, st));
    if (time_bias(tvals_, n_steps, st)) return -1;

    for (int s = 0; s < n_steps; ++s) {
        // tbias_ld = 0: every sequence reads the row of this step
        const float* tb = tbias_ + (size_t)s * ngroups * cfg_.channels;
        if (dual_stream_ && !profile_) {
            // The two CFG halves are independent until the Euler update: each runs its kernel chain on its own
            // stream, so the partial last wave of one half's persistent kernels (654 tiles on 148 SMs = 4.4 waves)
            // and its launch gaps are filled by the other half's kernels.  (Profiling runs serially so that the
            // per-kernel event times are not stretched by co-running kernels.)
            MD_CUDA(cudaEventRecord(ev_fork_, st));
            MD_CUDA(cudaStreamWaitEvent(st2_, ev_fork_, 0));
            if (estimator_forward(flow_rows_half_, flow_nseq_ / 2, flow_max_len_, tb, 0, chunk, st, 0)) return -1;
            if (estimator_forward(flow_rows_half_, flow_nseq_ / 2, flow_max_len_, tb, 0, chunk, st2_, flow_rows_half_)) return -1;
            MD_CUDA(cudaEventRecord(ev_join_, st2_));
            MD_CUDA(cudaStreamWaitEvent(st, ev_join_, 0));
        } else if (estimator_forward(M, flow_nseq_, flow_max_len_, tb, 0, chunk, st)) {
            return -1;
        }
        launch_euler_cfg(layout_, dout_, dtv[s], cfg_rate, xstate_, xin_, st);
        ++launches_;
    }
    launch_rows_to_bct(layout_, xstate_, 0, mel_out, st);
    ++launches_;
    MD_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace mossdec
