// HiFT vocoder sequence (cosyvoice/hifigan/generator.py:520-582, f0_predictor.py:55-60) as implicit-GEMM launches on
// the tcgen05 core plus the bandwidth kernels of hift_kernels.cu.
//
// Layout: the mel segments of a group are packed on one frame axis with 4 zero guard frames between them; the three
// upsampling stages use the same axis scaled by 8 / 40 / 120, so
//   * a non-causal (dilated) Conv1d is k row-shifted K-slabs of one channels-last matrix, the guard rows supplying the
//     zero padding of every sequence;
//   * ConvTranspose1d(stride s) is ONE GEMM with N = s * Cout whose [rows, s * Cout] output is the [s * rows, Cout]
//     matrix of the next stage (polyphase form, <= 3 input taps per phase);
//   * the strided source_downs convs read k consecutive 24-wide STFT rows = one contiguous K-slab per output row
//     (overlapping-window tensor map, row stride = stride * 24 elements).
#include "engine.h"

#include <algorithm>
#include <stdio.h>

namespace mossdec {

#define MD_CUDA(x)                                                                          \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) return fail(std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

namespace {
constexpr int kGuardF = 4;        // guard frames between sequences (32 / 160 / 480 rows at the three stages)
constexpr int kStftLd = 24;
constexpr int kScale[3] = {8, 40, 120};
constexpr int kUp[3] = {8, 5, 3};
constexpr int kResK[3] = {3, 7, 11};
constexpr int kSrcResK[3] = {7, 7, 11};
constexpr int kDil[3] = {1, 3, 5};

constexpr size_t kHiftSlack = 256;  // rows of slack after every matrix (TMA boxes / shifted windows stay inside)
constexpr int kHiftBufs = 20;

// Offsets of the vocoder workspace buffers for a group of `rows_f` frame rows; returns the total size.
size_t hift_carve(int64_t rows_f, int base, size_t* offs) {
    size_t need = 0;
    int k = 0;
    auto take = [&](size_t bytes) {
        offs[k++] = need;
        need += (bytes + 1023) & ~size_t(1023);
    };
    const size_t slack = kHiftSlack;
    const size_t Fr = (size_t)rows_f + slack;
    const size_t M2 = (size_t)rows_f * 120;
    const size_t S = (M2 + slack) * 64;  // elements of one stage matrix (largest: stage 2; M0*256 and M1*128 are smaller)
    take(Fr * 128 * 2); take(Fr * 80 * 4);                                       // melb (bf16, 128 wide), mel32 (fp32 rows)
    take(Fr * base * 4); take(Fr * base * 4);                                    // fa, fb: fp32 activations of the f0 predictor
    take(Fr * 4 * 4); take(Fr * 16 * 4); take(((size_t)rows_f * 480 + 1024) * 4);     // f0, p480, src
    take((M2 + 2 * slack) * kStftLd * 2); take((M2 + slack) * 32 * 4);            // stft, post
    for (int i = 0; i < 5; ++i) take(S * 4);                                      // u, si, x, xr, xs
    for (int i = 0; i < 6; ++i) take(S * 2);                                      // a0, a1, a2, h, lr, sb
    return need;
}

std::string N(const char* fmt, int a = 0, int b = 0, int c = 0) {
    char buf[128];
    snprintf(buf, sizeof buf, fmt, a, b, c);
    return buf;
}
}  // namespace

struct HiftRunner {
    Engine& e;
    cudaStream_t st;
    HiftLayout L{};
    int rows_f = 0, max_len = 0;
    int64_t M[3] = {0, 0, 0};
    int32_t *rs_f = nullptr, *rs[3] = {nullptr, nullptr, nullptr};
    // frame level
    __nv_bfloat16 *melb = nullptr, *stft = nullptr;
    float *mel32 = nullptr, *fa = nullptr, *fb = nullptr;   // fp32 rows of the f0 predictor (f0_fp32.cu)
    float *f0 = nullptr, *p480 = nullptr, *src = nullptr, *post = nullptr;
    // stage buffers (sized for the largest stage)
    float *u = nullptr, *si = nullptr, *x = nullptr, *xr = nullptr, *xs = nullptr;
    __nv_bfloat16 *a[3] = {nullptr, nullptr, nullptr}, *h = nullptr, *lr = nullptr, *sb = nullptr;

    HiftRunner(Engine& eng, cudaStream_t s) : e(eng), st(s) {}

    void conv_taps(GemmProblem& p, const __nv_bfloat16* A, int64_t rows, int cin_pad, int k, int dil) {
        p.a[0].ptr = A;
        p.a[0].rows = rows;
        p.a[0].cols = cin_pad;
        p.a[0].ld = cin_pad;
        p.num_segs = k;
        for (int j = 0; j < k; ++j) p.segs[j] = {0, (j - (k - 1) / 2) * dil, 0, cin_pad / kGemmBK};
        p.M = (int32_t)rows;
    }

    // ResBlock (generator.py:94-101) of C channels, kernel k at a stage with `rows` rows.
    //   x_in: fp32 input, a_in: bf16 Snake_{a1[0]}(x_in).  mode 0: result -> out_f32 (plain), also emits nothing else.
    //   The last conv adds the residual and writes per `fin`:
    //     fin.accum / fin.scale / fin.out_f32 / fin.act2 + fin.out_bf16
    struct Final {
        float* out_f32 = nullptr;
        int accum = 0;
        float scale = 1.f;
        __nv_bfloat16* out_bf16 = nullptr;
        int act2 = ACT_NONE;
        float act2_p = 0.f;
    };
    int resblock(const std::string& w, int C, int k, int64_t rows, const int32_t* row_seq, const float* x_in,
                 const __nv_bfloat16* a_in, float* run, __nv_bfloat16* act_tmp, const Final& fin) {
        const float* xcur = x_in;
        const __nv_bfloat16* acur = a_in;
        for (int l = 0; l < 3; ++l) {
            {   // xt = convs1[l](snake1(x)); emit snake2(xt)
                GemmProblem p;
                conv_taps(p, acur, rows, C, k, kDil[l]);
                p.w = e.W<__nv_bfloat16>(w + N(".c1.%d.w", l));
                p.N = C; p.BN = C;
                p.epi.bias = e.W<float>(w + N(".c1.%d.b", l));
                p.epi.act1 = ACT_SNAKE;
                p.epi.act1_alpha = e.W<float>(w + N(".a2.%d", l));
                p.epi.row_seq = row_seq;
                p.epi.out_bf16 = h; p.epi.ldo = C;
                if (e.gemm(p, st)) return -1;
            }
            {   // x = convs2[l](.) + x
                GemmProblem p;
                conv_taps(p, h, rows, C, k, 1);
                p.w = e.W<__nv_bfloat16>(w + N(".c2.%d.w", l));
                p.N = C; p.BN = C;
                p.epi.bias = e.W<float>(w + N(".c2.%d.b", l));
                p.epi.row_seq = row_seq;
                p.epi.resid = xcur; p.epi.ldr = C;
                if (l < 2) {
                    p.epi.out_f32 = run; p.epi.ldf = C;
                    p.epi.act2 = ACT_SNAKE;
                    p.epi.act2_alpha = e.W<float>(w + N(".a1.%d", l + 1));
                    p.epi.out_bf16 = act_tmp; p.epi.ldo = C;
                } else {
                    p.epi.out_f32 = fin.out_f32; p.epi.ldf = C;
                    p.epi.accum_f32 = fin.accum;
                    p.epi.out_scale = fin.scale;
                    p.epi.out_bf16 = fin.out_bf16; p.epi.ldo = C;
                    p.epi.act2 = fin.act2;
                    p.epi.act2_p = fin.act2_p;
                }
                if (e.gemm(p, st)) return -1;
            }
            xcur = run;
            acur = act_tmp;
        }
        return 0;
    }
};

int Engine::finalize_hift() {
    missing_.clear();
    for (int i = 0; i < 5; ++i) {
        find(N("hift.f0.c%d.wt", i));
        find(N("hift.f0.c%d.b", i));
    }
    const char* once[] = {"hift.f0.cls.w", "hift.f0.cls.b", "hift.src.w", "hift.src.b", "hift.pre.w", "hift.pre.b",
                          "hift.post.w", "hift.post.b"};
    for (const char* n : once) find(n);
    auto rb = [&](const std::string& p) {
        for (int l = 0; l < 3; ++l)
            for (const char* n : {".c1.%d.w", ".c1.%d.b", ".c2.%d.w", ".c2.%d.b", ".a1.%d", ".a2.%d"}) find(p + N(n, l));
    };
    for (int i = 0; i < 3; ++i) {
        for (const char* n : {"hift.up%d.w", "hift.up%d.b", "hift.sd%d.w", "hift.sd%d.b"}) find(N(n, i));
        rb(N("hift.srb%d", i));
        for (int r = 0; r < 3; ++r) rb(N("hift.rb%d.%d", i, r));
    }
    has_hift_ = missing_.empty();
    return 0;
}

int Engine::reserve_hift(int64_t rows_f, int nseq) {
    rows_f = (rows_f + kGemmBM - 1) / kGemmBM * kGemmBM;
    const size_t n_rs = (size_t)rows_f * (1 + 8 + 40 + 120);
    ws_hift_layout_.reserve((size_t)nseq * 28 + n_rs * 4 + 1024);
    if (!ws_hift_layout_.as<void>()) return fail("cudaMalloc(hift layout) failed");
    size_t offs[kHiftBufs];
    ws_hift_.reserve(hift_carve(rows_f, cfg_.hift_base, offs));
    if (!ws_hift_.as<void>()) return fail("cudaMalloc(hift workspace) failed");
    return 0;
}

int Engine::hift(int B, const int32_t* lens, const float* mel, const float* cache_source, int cache_stride,
                 const int32_t* cache_len, const float* nsf_noise, float* wav_out, float* source_out, uint64_t seed,
                 int flags, cudaStream_t st, const int32_t* mel_skip) {
    DeviceGuard g(device_);
    if (!finalized_) return fail("weights not finalized");
    if (!has_hift_) return fail("HiFT tensors missing: " + missing_);
    if (B <= 0) return fail("hift: bad batch");
    for (int b = 0; b < B; ++b)
        if (lens[b] <= 0) return fail("hift: sequence length must be positive");
    if (cfg_.hift_base != 512) return fail("hift: only base_channels = 512 is built");
    if (mel_skip != nullptr && !(flags & 1)) return fail("hift: mel_skip needs the time-major mel layout");
    hift_mel_skip_ = mel_skip;
    // split the batch into groups of <= hift_group_frames_ frames so the stage-2 working set stays bounded
    int b0 = 0;
    int64_t mel_off = 0, wav_off = 0;
    while (b0 < B) {
        int b1 = b0;
        int64_t frames = 0;
        while (b1 < B && (b1 == b0 || frames + lens[b1] <= hift_group_frames_)) frames += lens[b1++];
        if (hift_group(b0, b1, lens, mel_off, wav_off, mel, cache_source, cache_stride, cache_len, nsf_noise, wav_out,
                       source_out, seed, flags, st))
            return -1;
        for (int b = b0; b < b1; ++b) {
            mel_off += (int64_t)(lens[b] + (mel_skip ? mel_skip[b] : 0)) * cfg_.n_feats;
            wav_off += (int64_t)lens[b] * 480;
        }
        b0 = b1;
    }
    MD_CUDA(cudaGetLastError());
    return 0;
}

int Engine::hift_group(int b0, int b1, const int32_t* lens, int64_t mel_off0, int64_t wav_off0, const float* mel,
                       const float* cache_source, int cache_stride, const int32_t* cache_len, const float* nsf_noise,
                       float* wav_out, float* source_out, uint64_t seed, int flags, cudaStream_t st) {
    const int nseq = b1 - b0;
    HiftRunner R(*this, st);
    // ---------------------------------------------------------------- host layout (built in a pinned slot, one async upload)
    int rows = 0, max_len = 0;
    for (int s = 0; s < nseq; ++s) {
        rows += kGuardF + lens[b0 + s];
        max_len = std::max(max_len, lens[b0 + s]);
    }
    rows += kGuardF;
    const int rows_f = (rows + kGemmBM - 1) / kGemmBM * kGemmBM;
    R.rows_f = rows_f;
    R.max_len = max_len;
    for (int i = 0; i < 3; ++i) R.M[i] = (int64_t)rows_f * kScale[i];
    if (R.M[2] > 0x7fffff00LL) return fail("hift: group too large");
    if (reserve_hift(rows_f, nseq)) return -1;

    // layout arrays: [mel_off | wav_off] int64, [start | len | cache_len] int32, then row_seq of 4 levels (device-built)
    const size_t head_bytes = (size_t)nseq * 28;
    uint8_t* hp = static_cast<uint8_t*>(stager_.acquire(head_bytes));
    if (!hp) return fail("cudaMallocHost(hift layout staging) failed");
    {
        int64_t* h_mel = reinterpret_cast<int64_t*>(hp);
        int64_t* h_wav = h_mel + nseq;
        int32_t* h_start = reinterpret_cast<int32_t*>(h_wav + nseq);
        int32_t* h_len = h_start + nseq;
        int32_t* h_clen = h_len + nseq;
        int r = 0;
        int64_t mo = mel_off0, wo = wav_off0;
        for (int s = 0; s < nseq; ++s) {
            r += kGuardF;
            const int skip = hift_mel_skip_ ? hift_mel_skip_[b0 + s] : 0;   // leading (prompt) frames of the block to pass over
            h_start[s] = r;
            h_len[s] = lens[b0 + s];
            h_mel[s] = mo + (int64_t)skip * cfg_.n_feats;
            h_wav[s] = wo;
            mo += (int64_t)(h_len[s] + skip) * cfg_.n_feats;
            wo += (int64_t)h_len[s] * 480;
            r += h_len[s];
            h_clen[s] = cache_len ? cache_len[b0 + s] : 0;
        }
    }
    uint8_t* lp = ws_hift_layout_.as<uint8_t>();
    int64_t* d_mel = reinterpret_cast<int64_t*>(lp);
    int64_t* d_wav = d_mel + nseq;
    int32_t* d_start = reinterpret_cast<int32_t*>(d_wav + nseq);
    int32_t* d_len = d_start + nseq;
    int32_t* d_clen = d_len + nseq;
    int32_t* d_rs = reinterpret_cast<int32_t*>(lp + ((head_bytes + 255) & ~size_t(255)));
    MD_CUDA(cudaMemcpyAsync(lp, hp, head_bytes, cudaMemcpyHostToDevice, st));
    stager_.commit(st);
    HiftLayout& L = R.L;
    L.nseq = nseq;
    L.rows_f = rows_f;
    L.time_major = flags & 1;
    L.seq_start = d_start;
    L.seq_len = d_len;
    L.mel_off = d_mel;
    L.wav_off = d_wav;
    R.rs_f = d_rs;
    R.rs[0] = R.rs_f + rows_f;
    R.rs[1] = R.rs[0] + R.M[0];
    R.rs[2] = R.rs[1] + R.M[1];
    launch_hift_row_seq(L, 1, 0, rows_f, R.rs_f, st);
    for (int i = 0; i < 3; ++i) launch_hift_row_seq(L, kScale[i], i == 2 ? 1 : 0, (int)R.M[i], R.rs[i], st);
    launches_ += 4;

    // ---------------------------------------------------------------- workspace
    const int base = cfg_.hift_base;
    const size_t slack = kHiftSlack;
    size_t offs[kHiftBufs];
    hift_carve(rows_f, base, offs);
    const size_t o_melb = offs[0], o_mel32 = offs[1], o_fa = offs[2], o_fb = offs[3],
                 o_f0 = offs[4], o_p480 = offs[5], o_src = offs[6], o_stft = offs[7], o_post = offs[8], o_u = offs[9],
                 o_si = offs[10], o_x = offs[11], o_xr = offs[12], o_xs = offs[13], o_a0 = offs[14], o_a1 = offs[15],
                 o_a2 = offs[16], o_h = offs[17], o_lr = offs[18], o_sb = offs[19];
    uint8_t* wb = ws_hift_.as<uint8_t>();
    R.melb = (__nv_bfloat16*)(wb + o_melb);
    R.mel32 = (float*)(wb + o_mel32);
    R.fa = (float*)(wb + o_fa);
    R.fb = (float*)(wb + o_fb);
    R.f0 = (float*)(wb + o_f0);
    R.p480 = (float*)(wb + o_p480);
    R.src = (float*)(wb + o_src);
    R.stft = (__nv_bfloat16*)(wb + o_stft) + slack * kStftLd;  // leading slack: windows start up to 7 rows early
    R.post = (float*)(wb + o_post);
    R.u = (float*)(wb + o_u); R.si = (float*)(wb + o_si); R.x = (float*)(wb + o_x); R.xr = (float*)(wb + o_xr);
    R.xs = (float*)(wb + o_xs);
    R.a[0] = (__nv_bfloat16*)(wb + o_a0); R.a[1] = (__nv_bfloat16*)(wb + o_a1); R.a[2] = (__nv_bfloat16*)(wb + o_a2);
    R.h = (__nv_bfloat16*)(wb + o_h); R.lr = (__nv_bfloat16*)(wb + o_lr); R.sb = (__nv_bfloat16*)(wb + o_sb);

    // ---------------------------------------------------------------- f0 predictor (f0_predictor.py:55-60)
    // fp32 on the CUDA cores with round-to-nearest FMAs (f0_fp32.cu explains why the tensor cores are not used here)
    launch_hift_pack_mel(L, R.rs_f, mel, R.melb, R.mel32, st);
    ++launches_;
    {
        const float* in = R.mel32;
        int cin = cfg_.n_feats;
        float* outs[2] = {R.fa, R.fb};
        for (int i = 0; i < 5; ++i) {
            cudaError_t ce = launch_conv3_elu_f32(in, rows_f, cin, W<float>(N("hift.f0.c%d.wt", i)), base,
                                                  W<float>(N("hift.f0.c%d.b", i)), R.rs_f, outs[i & 1], st);
            if (ce != cudaSuccess) return fail(std::string("launch_conv3_elu_f32: ") + cudaGetErrorString(ce));
            ++launches_;
            in = outs[i & 1];
            cin = base;
        }
        cudaError_t ce = launch_f0_classifier(in, rows_f, base, W<float>("hift.f0.cls.w"), W<float>("hift.f0.cls.b"), R.rs_f,
                                              R.f0, 4, st);
        if (ce != cudaSuccess) return fail(std::string("launch_f0_classifier: ") + cudaGetErrorString(ce));
        ++launches_;
    }
    // ---------------------------------------------------------------- NSF source + STFT (generator.py:574-580, 521-522)
    float* src = source_out ? source_out : R.src - wav_off0;  // ragged [480 T_b] signal addressed by wav_off
    double frames = 0;
    for (int s = 0; s < nseq; ++s) frames += lens[b0 + s];
    launch_hift_phase(L, R.f0, 4, cfg_.sample_rate, 480.f, cfg_.n_harm, R.p480, st);
    // per frame: 480 samples written (fp32) + 480 x 9 injected noise draws read (none with the device Philox stream)
    bw_begin("hift_source_kernel", frames * (480.0 * 4 + (nsf_noise ? 480.0 * 9 * 4 : 0.0)), st);
    launch_hift_source(L, max_len, R.f0, 4, R.p480, nsf_noise, W<float>("hift.src.w"), W<float>("hift.src.b"),
                       cache_source ? cache_source + (size_t)b0 * cache_stride : nullptr, cache_stride, d_clen,
                       cfg_.n_harm, cfg_.sine_amp, cfg_.noise_std, cfg_.voiced_thr, 480, seed, src, st);
    bw_end(st);
    // per stage-2 row (120 per frame): 4 new samples read, 24 bf16 written
    bw_begin("hift_stft_kernel", (double)R.M[2] * (4.0 * 4 + 24.0 * 2), st);
    launch_hift_stft(L, R.rs[2], (int)R.M[2], src, R.stft, st);
    bw_end(st);
    launches_ += 3;

    // ---------------------------------------------------------------- conv_pre (generator.py:524) -> leaky_relu
    {
        GemmProblem p;
        R.conv_taps(p, R.melb, rows_f, 128, 7, 1);
        p.w = W<__nv_bfloat16>("hift.pre.w");
        p.N = base; p.BN = 256;
        p.epi.bias = W<float>("hift.pre.b");
        p.epi.act1 = ACT_LRELU; p.epi.act1_p = cfg_.lrelu;
        p.epi.row_seq = R.rs_f;
        p.epi.out_bf16 = R.lr; p.epi.ldo = base;
        if (gemm(p, st)) return -1;
    }
    // ---------------------------------------------------------------- upsampling stages (generator.py:525-543)
    int cin = base;
    int64_t rows_in = rows_f;
    const int32_t* rs_in = R.rs_f;
    for (int i = 0; i < 3; ++i) {
        const int C = cin / 2, s = kUp[i];
        const int64_t Mi = R.M[i];
        const int reflect = i == 2 ? 1 : 0;
        {   // ConvTranspose1d as one polyphase GEMM: [rows_in, s * C] == [Mi, C] (shifted one row down on the last stage)
            GemmProblem p;
            p.a[0].ptr = R.lr; p.a[0].rows = rows_in; p.a[0].cols = cin; p.a[0].ld = cin;
            p.num_segs = 3;
            for (int d = 0; d < 3; ++d) p.segs[d] = {0, d - 1, 0, cin / kGemmBK};
            p.M = (int32_t)rows_in;
            p.w = W<__nv_bfloat16>(N("hift.up%d.w", i));
            p.N = s * C; p.BN = (s * C) % 256 == 0 ? 256 : (s * C) % 192 == 0 ? 192 : 128;
            p.epi.bias = W<float>(N("hift.up%d.b", i));
            p.epi.row_seq = rs_in;
            p.epi.out_f32 = R.u + (reflect ? C : 0); p.epi.ldf = s * C;
            if (gemm(p, st)) return -1;
        }
        {   // source_downs[i] on the STFT rows (generator.py:533), opening Snake of the source ResBlock on the bf16 copy
            GemmProblem p;
            const int k = i == 0 ? 30 : i == 1 ? 6 : 1, stride = i == 0 ? 15 : i == 1 ? 3 : 1, pad = i == 0 ? 7 : i == 1 ? 1 : 0;
            const int kb = (k * kStftLd + kGemmBK - 1) / kGemmBK;
            p.a[0].ptr = R.stft - (size_t)pad * kStftLd;
            p.a[0].rows = Mi;
            p.a[0].cols = std::max(k * kStftLd, kGemmBK);
            p.a[0].ld = (int64_t)stride * kStftLd;
            p.num_segs = 1;
            p.segs[0] = {0, 0, 0, kb};
            p.M = (int32_t)Mi;
            p.w = W<__nv_bfloat16>(N("hift.sd%d.w", i));
            p.N = C; p.BN = C;
            p.epi.bias = W<float>(N("hift.sd%d.b", i));
            p.epi.row_seq = R.rs[i];
            p.epi.out_f32 = R.si; p.epi.ldf = C;
            p.epi.act2 = ACT_SNAKE; p.epi.act2_alpha = W<float>(N("hift.srb%d.a1.0", i));
            p.epi.out_bf16 = R.sb; p.epi.ldo = C;
            if (gemm(p, st)) return -1;
        }
        {   // source ResBlock: si = source_resblocks[i](si)
            HiftRunner::Final fin;
            fin.out_f32 = R.si;
            if (R.resblock(N("hift.srb%d", i), C, kSrcResK[i], Mi, R.rs[i], R.si, R.sb, R.si, R.sb, fin)) return -1;
        }
        // x = ups(x) (+ reflection pad) + si; Snake operands of the three ResBlocks
        // per element: read u, si (fp32), write x (fp32) + three bf16 Snake operands
        bw_begin("hift_fuse_snake_kernel", (double)Mi * C * (3.0 * 4 + 3.0 * 2), st);
        launch_hift_fuse_snake(L, R.rs[i], kScale[i], reflect, Mi, C, R.u, R.si, W<float>(N("hift.rb%d.0.a1.0", i)),
                               W<float>(N("hift.rb%d.1.a1.0", i)), W<float>(N("hift.rb%d.2.a1.0", i)), R.x, R.a[0],
                               R.a[1], R.a[2], st);
        bw_end(st);
        ++launches_;
        for (int r = 0; r < 3; ++r) {   // xs = sum_r resblocks[3 i + r](x); x = xs / 3; next operand = leaky_relu(x)
            HiftRunner::Final fin;
            fin.out_f32 = R.xs;
            fin.accum = r > 0;
            if (r == 2) {
                fin.scale = 1.f / 3.f;
                fin.out_bf16 = R.lr;
                fin.act2 = ACT_LRELU;
                fin.act2_p = i == 2 ? 0.01f : cfg_.lrelu;  // F.leaky_relu default slope before conv_post (generator.py:545)
            }
            if (R.resblock(N("hift.rb%d.%d", i, r), C, kResK[r], Mi, R.rs[i], R.x, R.a[r], R.xr, R.a[r], fin)) return -1;
        }
        cin = C;
        rows_in = Mi;
        rs_in = R.rs[i];
    }
    {   // conv_post (generator.py:546) -> [M2, 32] fp32 (18 used)
        GemmProblem p;
        R.conv_taps(p, R.lr, R.M[2], 64, 7, 1);
        p.w = W<__nv_bfloat16>("hift.post.w");
        p.N = 18; p.BN = 32;
        p.epi.bias = W<float>("hift.post.b");
        p.epi.row_seq = R.rs[2];
        p.epi.out_f32 = R.post; p.epi.ldf = 32;
        if (gemm(p, st)) return -1;
    }
    // per frame: 120 rows x 32 fp32 of conv_post output read (18 used, 32-wide rows), 480 samples written
    bw_begin("hift_istft_kernel", frames * (120.0 * 32 * 4 + 480.0 * 4), st);
    launch_hift_istft(L, max_len, R.post, 32, cfg_.audio_limit, wav_out, st);
    bw_end(st);
    ++launches_;
    taps_["hift.f0"] = {R.f0, (size_t)rows_f * 4 * 4};          // [rows_f, 4] fp32, column 0; first sequence at row 4
    taps_["hift.p480"] = {R.p480, (size_t)rows_f * 16 * 4};
    taps_["hift.post"] = {R.post, (size_t)R.M[2] * 32 * 4};
    taps_["hift.x"] = {R.xs, (size_t)R.M[2] * 64 * 4};
    return 0;
}

int Engine::debug_read(const char* name, void* host_dst, size_t bytes) {
    DeviceGuard g(device_);
    auto it = taps_.find(name);
    if (it == taps_.end()) return fail(std::string("no such tap: ") + name);
    if (bytes > it->second.second) bytes = it->second.second;
    MD_CUDA(cudaDeviceSynchronize());
    MD_CUDA(cudaMemcpy(host_dst, it->second.first, bytes, cudaMemcpyDeviceToHost));
    return (int)(bytes >> 2);
}

int Engine::fade_in_out(float* speech, const float* prev_tail, const float* window2n, int n, cudaStream_t st) {
    DeviceGuard g(device_);
    if (n <= 0) return 0;
    launch_fade(speech, prev_tail, window2n, n, st);
    ++launches_;
    MD_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace mossdec
