// Token -> mu front end on the repo's own kernels:
//   CausalMaskedDiffWithXvec.inference prologue          cosyvoice/flow/flow.py:250-270
//   UpsampleConformerEncoder.forward                     cosyvoice/transformer/upsample_encoder.py:244-307
//     LinearNoSubsampling + EspnetRelPositionalEncoding  subsampling.py:85-93, embedding.py:224-302
//     PreLookaheadLayer                                  upsample_encoder.py:81-98
//     ConformerEncoderLayer (pre-norm, no macaron/conv)  encoder_layer.py:196-236
//     RelPositionMultiHeadedAttention                    attention.py:215-330
//     Upsample1D (nearest x4, left pad 8, conv k9)       upsample_encoder.py:37-64
// batched over a ragged group of utterances (the reference asserts batch 1, flow.py:248).
//
// Layout: one packed token-rate row axis with 4 zero guard rows between utterances; the mel-rate stage uses the same
// axis scaled by 4.  Every Linear / Conv1d is a gemm_tc_kernel launch (tcgen05):
//   * PreLookahead conv1 (k4, looks 3 tokens ahead: the context tokens are real rows, the guard rows behind them the zero
//     pad) and conv2 (k3 causal) are row-shifted K slabs;
//   * the x4 nearest up-sampling + conv k9 is ONE polyphase GEMM with N = 4 * 512 over three token rows
//     (out[4q + r] = sum_d (sum_{j: (r + j) / 4 == d} W_j) x[q - 2 + d]), whose [rows, 2048] output is the [4 rows, 512]
//     mel-rate matrix;
//   * q | k | v is one GEMM with N = 2048 = [(q + pos_bias_u) / 8 | k | v | (q + pos_bias_v) / 8]: the two biased query
//     copies of attention.py:305-307 come out of the bias vector, 1 / sqrt(d_k) is folded into the weights;
//   * relative-position scores (matrix_bd, attention.py:317-324): per head one K = 64 GEMM of the second query copy
//     against P = linear_pos(pos_emb) gives raw[row, m] for every table row m; rel_shift is then only an index:
//     bd[i, j] = raw[i, tmax - 1 - i + j], read by attention_ws_kernel<bias> as an additive score term.  P is computed
//     in fp32 (f0_fp32.cu GEMM) from the fp32 sinusoid table and kept per (layer, tmax).
// LayerNorm over 512 channels, embedding gather, speaker projection and the final ragged gather: enc_kernels.cu.
#include "engine.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>

namespace mossdec {

#define MD_CUDA(x)                                                                          \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) return fail(std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

namespace {
constexpr int kD = 512, kH = 8, kFF = 2048, kQkv = 2048, kEncGuard = 4, kTokLayers = 6, kMelLayers = 4;
std::string EN(const char* fmt, int a = 0) {
    char buf[96];
    snprintf(buf, sizeof buf, fmt, a);
    return buf;
}
int round_up(int x, int m) { return (x + m - 1) / m * m; }
}  // namespace

int Engine::finalize_encoder() {
    missing_.clear();
    const char* once[] = {"enc.emb", "enc.embed0.w", "enc.embed0.b", "enc.embed0.ln.g", "enc.embed0.ln.b", "enc.embed1.w",
                          "enc.embed1.b", "enc.embed1.ln.g", "enc.embed1.ln.b", "enc.pl.c1.w", "enc.pl.c1.b", "enc.pl.c2.w",
                          "enc.pl.c2.b", "enc.up.w", "enc.up.b", "enc.after.g", "enc.after.b", "enc.proj.w", "enc.proj.b",
                          "spk.w", "spk.b"};
    for (const char* n : once) find(n);
    const char* per[] = {"n1.g", "n1.b", "n2.g", "n2.b", "qkv.w", "qkv.b", "pos.wt", "o.w", "o.b", "ff1.w", "ff1.b", "ff2.w", "ff2.b"};
    for (int l = 0; l < kTokLayers + kMelLayers; ++l)
        for (const char* n : per) find(EN("enc.L%d.", l) + n);
    has_encoder_ = missing_.empty();
    enc_missing_ = missing_;
    return 0;
}

// One conformer layer on `R` packed rows (enc_x32_: fp32 residual stream, in place).
int Engine::encoder_layer(int l, int R, const int32_t* row_seq, const int32_t* seq_start, const int32_t* seq_len, int nseq,
                          int max_len, int chunk, const __nv_bfloat16* P, int npad, int tmax, bool emit_bf16, cudaStream_t st) {
    const std::string p = EN("enc.L%d.", l);
    launch_layernorm512(enc_x32_, R, W<float>(p + "n1.g"), W<float>(p + "n1.b"), 1e-12f, 1.f, row_seq, nullptr, enc_xn_, st);
    ++launches_;
    auto dense = [&](GemmProblem& g, const __nv_bfloat16* a, int cols, int ld, int col0, int k) {
        g.a[0].ptr = a; g.a[0].rows = R; g.a[0].cols = cols; g.a[0].ld = ld;
        g.num_segs = 1;
        g.segs[0] = {0, 0, col0, k / kGemmBK};
        g.M = R;
    };
    {   // [(q + u) / 8 | k | v | (q + v) / 8]
        GemmProblem g;
        dense(g, enc_xn_, kD, kD, 0, kD);
        g.w = W<__nv_bfloat16>(p + "qkv.w");
        g.N = kQkv; g.BN = 256;
        g.epi.bias = W<float>(p + "qkv.b");
        g.epi.row_seq = row_seq;
        g.epi.out_bf16 = enc_qkv_; g.epi.ldo = kQkv;
        if (gemm(g, st)) return -1;
    }
    for (int h = 0; h < kH; ++h) {   // raw[row, m] = q_v[row, h] . P[m, h]   (matrix_bd before rel_shift)
        GemmProblem g;
        dense(g, enc_qkv_, kQkv, kQkv, 3 * kD + h * 64, 64);
        g.w = P + h * 64;
        g.w_ld = kD;
        g.N = npad; g.BN = (npad % 256) == 0 ? 256 : 128;
        g.epi.out_bf16 = enc_bias_ + (size_t)h * R * npad; g.epi.ldo = npad;
        if (gemm(g, st)) return -1;
    }
    {
        if (profile_) prof_begin(1, 0.0, st);
        cudaError_t ae = launch_attention_ws_ex(enc_qkv_, R, kQkv, enc_att_, seq_start, seq_len, nseq, max_len, kH, chunk,
                                                enc_bias_, (int64_t)R * npad, npad, tmax, st);
        if (profile_) prof_end(st);
        if (ae != cudaSuccess) return fail(std::string("encoder attention: ") + cudaGetErrorString(ae));
        ++launches_;
    }
    {   // linear_out + residual
        GemmProblem g;
        dense(g, enc_att_, kD, kD, 0, kD);
        g.w = W<__nv_bfloat16>(p + "o.w");
        g.N = kD; g.BN = 256;
        g.epi.bias = W<float>(p + "o.b");
        g.epi.row_seq = row_seq;
        g.epi.resid = enc_x32_; g.epi.ldr = kD;
        g.epi.out_f32 = enc_x32_; g.epi.ldf = kD;
        if (gemm(g, st)) return -1;
    }
    launch_layernorm512(enc_x32_, R, W<float>(p + "n2.g"), W<float>(p + "n2.b"), 1e-12f, 1.f, row_seq, nullptr, enc_xn_, st);
    ++launches_;
    {   // w_1 + SiLU
        GemmProblem g;
        dense(g, enc_xn_, kD, kD, 0, kD);
        g.w = W<__nv_bfloat16>(p + "ff1.w");
        g.N = kFF; g.BN = 256;
        g.epi.bias = W<float>(p + "ff1.b");
        g.epi.act1 = ACT_SILU;
        g.epi.row_seq = row_seq;
        g.epi.out_bf16 = enc_ffh_; g.epi.ldo = kFF;
        if (gemm(g, st)) return -1;
    }
    {   // w_2 + residual
        GemmProblem g;
        dense(g, enc_ffh_, kFF, kFF, 0, kFF);
        g.w = W<__nv_bfloat16>(p + "ff2.w");
        g.N = kD; g.BN = 256;
        g.epi.bias = W<float>(p + "ff2.b");
        g.epi.row_seq = row_seq;
        g.epi.resid = enc_x32_; g.epi.ldr = kD;
        g.epi.out_f32 = enc_x32_; g.epi.ldf = kD;
        if (emit_bf16) {
            g.epi.out_bf16 = enc_xn_; g.epi.ldo = kD;   // operand of the up-conv
        }
        if (gemm(g, st)) return -1;
    }
    return 0;
}

// P[l] = bf16(linear_pos_l(pe(tmax))) for the layers [l0, l1): fp32 sinusoid table -> fp32 GEMM -> bf16, [npad, 512] each.
int Engine::encoder_pos_tables(int l0, int l1, int tmax, int npad, __nv_bfloat16* P, cudaStream_t st) {
    launch_rel_pos_table(tmax, npad, enc_pe_, st);
    ++launches_;
    for (int l = l0; l < l1; ++l) {
        cudaError_t e = launch_gemm_f32(enc_pe_, npad, kD, kD, W<float>(EN("enc.L%d.pos.wt", l)), kD, enc_ptmp_, st);
        if (e != cudaSuccess) return fail(std::string("pos table gemm: ") + cudaGetErrorString(e));
        launch_f32_to_bf16(enc_ptmp_, (int64_t)npad * kD, P + (size_t)(l - l0) * npad * kD, st);
        launches_ += 2;
    }
    return 0;
}

// Encoder for the utterances [b0, b1): writes mu and cond (ragged, time-major [T_b, 80]) at the frame offsets mel_off.
int Engine::encode_group(int b0, int b1, const int32_t* n_tok, const int64_t* tok_off, const int64_t* mel_off,
                         const int32_t* prompt_frames, const int64_t* feat_off, const int64_t* tokens,
                         const float* prompt_feat, int ctx, int chunk_tokens, float* mu, float* cond, cudaStream_t st) {
    const int nseq = b1 - b0;
    int rows = 0, tmax = 0;
    for (int s = 0; s < nseq; ++s) {
        rows += kEncGuard + n_tok[b0 + s];
        tmax = std::max(tmax, n_tok[b0 + s] - ctx);
    }
    rows += kEncGuard;
    const int R = round_up(rows, kGemmBM), R4 = 4 * R;
    const int tmax4 = 4 * tmax;
    const int npad = round_up(2 * tmax + 32, 128), npad4 = round_up(2 * tmax4 + 32, 128);

    // ---------------------------------------------------------------- layout (pinned staging, one async upload)
    const size_t o_i64 = 0, n_i64 = (size_t)nseq * 3;
    const size_t o_i32 = n_i64 * 8, n_i32 = (size_t)R * 2 + (size_t)R4 + (size_t)nseq * 5;
    const size_t lay_bytes = o_i32 + n_i32 * 4;
    ws_enc_layout_.reserve(lay_bytes + 256);
    if (!ws_enc_layout_.as<void>()) return fail("cudaMalloc(encoder layout) failed");
    uint8_t* hp = static_cast<uint8_t*>(stager_.acquire(lay_bytes));
    if (!hp) return fail("cudaMallocHost(encoder layout staging) failed");
    int64_t* h_tok_off = reinterpret_cast<int64_t*>(hp + o_i64);
    int64_t* h_mel_off = h_tok_off + nseq;
    int64_t* h_feat_off = h_mel_off + nseq;
    int32_t* h_rs_all = reinterpret_cast<int32_t*>(hp + o_i32);
    int32_t* h_rs_main = h_rs_all + R;
    int32_t* h_rs_mel = h_rs_main + R;
    int32_t* h_start = h_rs_mel + R4;
    int32_t* h_len = h_start + nseq;
    int32_t* h_start4 = h_len + nseq;
    int32_t* h_len4 = h_start4 + nseq;
    int32_t* h_pf = h_len4 + nseq;
    for (int i = 0; i < R; ++i) h_rs_all[i] = h_rs_main[i] = -1;
    for (int i = 0; i < R4; ++i) h_rs_mel[i] = -1;
    int r = 0;
    for (int s = 0; s < nseq; ++s) {
        const int L = n_tok[b0 + s], T = L - ctx;
        r += kEncGuard;
        h_start[s] = r;
        h_len[s] = T;
        h_start4[s] = 4 * r;
        h_len4[s] = 4 * T;
        h_pf[s] = prompt_frames[b0 + s];
        h_tok_off[s] = tok_off[b0 + s];
        h_mel_off[s] = mel_off[b0 + s];
        h_feat_off[s] = feat_off[b0 + s];
        for (int t = 0; t < L; ++t) h_rs_all[r + t] = s;
        for (int t = 0; t < T; ++t) h_rs_main[r + t] = s;
        for (int t = 0; t < 4 * T; ++t) h_rs_mel[4 * r + t] = s;
        r += L;
    }
    uint8_t* lp = ws_enc_layout_.as<uint8_t>();
    MD_CUDA(cudaMemcpyAsync(lp, hp, lay_bytes, cudaMemcpyHostToDevice, st));
    stager_.commit(st);
    EncLayout EL{};
    EL.rows = R;
    EL.nseq = nseq;
    EL.tok_off = reinterpret_cast<int64_t*>(lp + o_i64);
    EL.mel_off = EL.tok_off + nseq;
    EL.feat_off = EL.mel_off + nseq;
    EL.row_seq_all = reinterpret_cast<int32_t*>(lp + o_i32);
    EL.row_seq_main = EL.row_seq_all + R;
    EL.row_seq_mel = EL.row_seq_main + R;
    EL.seq_start = EL.row_seq_mel + R4;
    EL.seq_len = EL.seq_start + nseq;
    EL.seq_start4 = EL.seq_len + nseq;
    EL.seq_len4 = EL.seq_start4 + nseq;
    EL.prompt_frames = EL.seq_len4 + nseq;

    // ---------------------------------------------------------------- workspace (sized for the mel-rate stage)
    size_t need = 0;
    auto take = [&](size_t bytes) {
        size_t o = need;
        need += (bytes + 1023) & ~size_t(1023);
        return o;
    };
    const size_t Rr = (size_t)R4 + kGemmBM;
    const size_t o_xe = take(Rr * kD * 2), o_tmp = take(Rr * kD * 4), o_x32 = take(Rr * kD * 4), o_xn = take(Rr * kD * 2),
                 o_qkv = take(Rr * kQkv * 2), o_att = take(Rr * kD * 2), o_ffh = take(Rr * kFF * 2),
                 o_bias = take((size_t)kH * Rr * npad4 * 2), o_pe = take((size_t)npad4 * kD * 4),
                 o_ptmp = take((size_t)npad4 * kD * 4), o_mu = take(Rr * 80 * 4);
    const size_t o_ptok = take((size_t)kTokLayers * npad * kD * 2 + 4096), o_pmel = take((size_t)kMelLayers * npad4 * kD * 2 + 4096);
    if (ws_enc_.reserve(need)) enc_tmax_cached_ = -1;
    if (!ws_enc_.as<void>()) return fail("cudaMalloc(encoder workspace) failed");
    uint8_t* wb = ws_enc_.as<uint8_t>();
    __nv_bfloat16* xe = (__nv_bfloat16*)(wb + o_xe);
    float* tmp32 = (float*)(wb + o_tmp);
    enc_x32_ = (float*)(wb + o_x32);
    enc_xn_ = (__nv_bfloat16*)(wb + o_xn);
    enc_qkv_ = (__nv_bfloat16*)(wb + o_qkv);
    enc_att_ = (__nv_bfloat16*)(wb + o_att);
    enc_ffh_ = (__nv_bfloat16*)(wb + o_ffh);
    enc_bias_ = (__nv_bfloat16*)(wb + o_bias);
    enc_pe_ = (float*)(wb + o_pe);
    enc_ptmp_ = (float*)(wb + o_ptmp);
    float* mu_rows = (float*)(wb + o_mu);
    __nv_bfloat16* p_tok = (__nv_bfloat16*)(wb + o_ptok);
    __nv_bfloat16* p_mel = (__nv_bfloat16*)(wb + o_pmel);
    // (the tables depend on tmax only; the carve-out moves with R, so they are rebuilt when either changes)
    if (enc_tmax_cached_ != tmax || enc_ptok_cached_ != p_tok || enc_pmel_cached_ != p_mel) {
        if (encoder_pos_tables(0, kTokLayers, tmax, npad, p_tok, st)) return -1;
        if (encoder_pos_tables(kTokLayers, kTokLayers + kMelLayers, tmax4, npad4, p_mel, st)) return -1;
        enc_tmax_cached_ = tmax;
        enc_ptok_cached_ = p_tok;
        enc_pmel_cached_ = p_mel;
    }

    // ---------------------------------------------------------------- token-rate stage
    launch_enc_gather(EL, tokens, W<__nv_bfloat16>("enc.emb"), (int)find("enc.emb")->shape[0], xe, st);
    ++launches_;
    auto embed = [&](int which, int rows_n, const int32_t* rs) -> int {   // Linear + LayerNorm(1e-5) + * sqrt(d)
        GemmProblem g;
        g.a[0].ptr = xe; g.a[0].rows = rows_n; g.a[0].cols = kD; g.a[0].ld = kD;
        g.num_segs = 1;
        g.segs[0] = {0, 0, 0, kD / kGemmBK};
        g.M = rows_n;
        g.w = W<__nv_bfloat16>(EN("enc.embed%d.w", which));
        g.N = kD; g.BN = 256;
        g.epi.bias = W<float>(EN("enc.embed%d.b", which));
        g.epi.out_f32 = tmp32; g.epi.ldf = kD;
        if (gemm(g, st)) return -1;
        launch_layernorm512(tmp32, rows_n, W<float>(EN("enc.embed%d.ln.g", which)), W<float>(EN("enc.embed%d.ln.b", which)),
                            1e-5f, sqrtf((float)kD), rs, enc_x32_, enc_xn_, st);
        ++launches_;
        return 0;
    };
    if (embed(0, R, EL.row_seq_all)) return -1;
    {   // PreLookaheadLayer conv1: k = 4 over rows t .. t + 3 (context rows, then zero guard rows), leaky_relu
        GemmProblem g;
        g.a[0].ptr = enc_xn_; g.a[0].rows = R; g.a[0].cols = kD; g.a[0].ld = kD;
        g.num_segs = 4;
        for (int j = 0; j < 4; ++j) g.segs[j] = {0, j, 0, kD / kGemmBK};
        g.M = R;
        g.w = W<__nv_bfloat16>("enc.pl.c1.w");
        g.N = kD; g.BN = 256;
        g.epi.bias = W<float>("enc.pl.c1.b");
        g.epi.act1 = ACT_LRELU; g.epi.act1_p = 0.01f;
        g.epi.row_seq = EL.row_seq_main;
        g.epi.out_bf16 = enc_att_; g.epi.ldo = kD;
        if (gemm(g, st)) return -1;
    }
    {   // conv2: k = 3 causal (rows t - 2 .. t) + residual `inputs`
        GemmProblem g;
        g.a[0].ptr = enc_att_; g.a[0].rows = R; g.a[0].cols = kD; g.a[0].ld = kD;
        g.num_segs = 3;
        for (int j = 0; j < 3; ++j) g.segs[j] = {0, j - 2, 0, kD / kGemmBK};
        g.M = R;
        g.w = W<__nv_bfloat16>("enc.pl.c2.w");
        g.N = kD; g.BN = 256;
        g.epi.bias = W<float>("enc.pl.c2.b");
        g.epi.row_seq = EL.row_seq_main;
        g.epi.resid = enc_x32_; g.epi.ldr = kD;
        g.epi.out_f32 = enc_x32_; g.epi.ldf = kD;
        if (gemm(g, st)) return -1;
    }
    for (int l = 0; l < kTokLayers; ++l)
        if (encoder_layer(l, R, EL.row_seq_main, EL.seq_start, EL.seq_len, nseq, tmax, chunk_tokens,
                          p_tok + (size_t)l * npad * kD, npad, tmax, l + 1 == kTokLayers, st))
            return -1;
    // ---------------------------------------------------------------- x4 up-sampling + conv k9 as one polyphase GEMM
    {
        GemmProblem g;
        g.a[0].ptr = enc_xn_; g.a[0].rows = R; g.a[0].cols = kD; g.a[0].ld = kD;
        g.num_segs = 3;
        for (int d = 0; d < 3; ++d) g.segs[d] = {0, d - 2, 0, kD / kGemmBK};
        g.M = R;
        g.w = W<__nv_bfloat16>("enc.up.w");
        g.N = 4 * kD; g.BN = 256;
        g.epi.bias = W<float>("enc.up.b");
        g.epi.row_seq = EL.row_seq_main;
        g.epi.out_bf16 = xe; g.epi.ldo = 4 * kD;   // [R, 2048] == [4 R, 512]
        if (gemm(g, st)) return -1;
    }
    // ---------------------------------------------------------------- mel-rate stage
    if (embed(1, R4, EL.row_seq_mel)) return -1;
    for (int l = 0; l < kMelLayers; ++l)
        if (encoder_layer(kTokLayers + l, R4, EL.row_seq_mel, EL.seq_start4, EL.seq_len4, nseq, tmax4, chunk_tokens * 4,
                          p_mel + (size_t)l * npad4 * kD, npad4, tmax4, false, st))
            return -1;
    launch_layernorm512(enc_x32_, R4, W<float>("enc.after.g"), W<float>("enc.after.b"), 1e-5f, 1.f, EL.row_seq_mel, nullptr,
                        enc_xn_, st);
    ++launches_;
    {   // encoder_proj 512 -> 80 (flow.py:265)
        GemmProblem g;
        g.a[0].ptr = enc_xn_; g.a[0].rows = R4; g.a[0].cols = kD; g.a[0].ld = kD;
        g.num_segs = 1;
        g.segs[0] = {0, 0, 0, kD / kGemmBK};
        g.M = R4;
        g.w = W<__nv_bfloat16>("enc.proj.w");
        g.N = cfg_.n_feats; g.BN = cfg_.n_feats;
        g.epi.bias = W<float>("enc.proj.b");
        g.epi.row_seq = EL.row_seq_mel;
        g.epi.out_f32 = mu_rows; g.epi.ldf = cfg_.n_feats;
        if (gemm(g, st)) return -1;
    }
    launch_enc_finish(EL, mu_rows, prompt_feat, mu, cond, st);
    ++launches_;
    MD_CUDA(cudaGetLastError());
    return 0;
}

int Engine::token2mel(int B, const int32_t* n_tok, const int64_t* tokens, const int32_t* prompt_frames,
                      const float* prompt_feat, const float* embedding, int ctx, int chunk_tokens, int est_chunk_frames,
                      int n_steps, float cfg_rate, float* mel_out, cudaStream_t st, const PrefixArgs* px) {
    DeviceGuard g(device_);
    if (!finalized_) return fail("weights not finalized");
    if (!has_encoder_) return fail("encoder tensors missing: " + enc_missing_);
    if (B <= 0 || ctx < 0 || ctx > 3) return fail("token2mel: bad arguments");
    std::vector<int64_t> tok_off(B), mel_off(B), feat_off(B);
    std::vector<int32_t> lens(B);
    int64_t to = 0, mo = 0, fo = 0;
    for (int b = 0; b < B; ++b) {
        if (n_tok[b] - ctx <= 0) return fail("every utterance needs at least one token besides the look-ahead context");
        lens[b] = 4 * (n_tok[b] - ctx);
        if (prompt_frames[b] < 0 || prompt_frames[b] > lens[b]) return fail("prompt_feat longer than the decoded sequence");
        tok_off[b] = to;
        mel_off[b] = mo;
        feat_off[b] = fo;
        to += n_tok[b];
        mo += lens[b];
        fo += prompt_frames[b];
    }
    // mu | cond | spks live in engine memory between the encoder and the solve
    const size_t n_mel = (size_t)mo * cfg_.n_feats;
    ws_mu_.reserve((2 * n_mel + (size_t)B * cfg_.n_feats + 256) * 4);
    if (!ws_mu_.as<void>()) return fail("cudaMalloc(mu / cond) failed");
    float* mu = ws_mu_.as<float>();
    float* cond = mu + n_mel;
    float* spks = cond + n_mel;
    // groups of utterances so that the relative-position score buffer (8 heads x rows x 2 tmax, bf16) stays bounded
    const size_t budget = (size_t)3 << 30;
    int b0 = 0;
    while (b0 < B) {
        int b1 = b0, rows = kEncGuard, tmax = 0;
        while (b1 < B) {
            const int r2 = rows + kEncGuard + n_tok[b1], t2 = std::max(tmax, n_tok[b1] - ctx);
            const size_t bytes = (size_t)kH * 4 * round_up(r2, kGemmBM) * round_up(8 * t2 + 32, 128) * 2;
            if (b1 > b0 && bytes > budget) break;
            rows = r2;
            tmax = t2;
            ++b1;
        }
        if (encode_group(b0, b1, n_tok, tok_off.data(), mel_off.data(), prompt_frames, feat_off.data(), tokens, prompt_feat, ctx,
                         chunk_tokens, mu, cond, st))
            return -1;
        b0 = b1;
    }
    launch_spk_project(embedding, B, 192, W<float>("spk.w"), W<float>("spk.b"), cfg_.n_feats, spks, st);
    ++launches_;
    taps_["enc.mu"] = {mu, n_mel * 4};       // test hooks (mossdec_debug_read): ragged time-major [sum T_b, 80]
    taps_["enc.cond"] = {cond, n_mel * 4};
    taps_["enc.spks"] = {spks, (size_t)B * cfg_.n_feats * 4};
    if (px != nullptr && px->mode != 0) {
        // the frozen prefix must lie inside what is identical in every chunk of a stream: prompt mel frames, and prompt
        // tokens up to the encoder's 3-token look-ahead (upsample_encoder.py:81-98); whole encoder chunks
        if (chunk_tokens <= 0 || (px_frames_ % (4 * chunk_tokens)) != 0)
            return fail("token2mel: the frozen prefix must be a whole number of encoder chunks (streaming only)");
        for (int b = 0; b < B; ++b)
            if (prompt_frames[b] < px_frames_ || prompt_frames[b] / 4 < px_frames_ / 4 + 3)
                return fail("token2mel: the prompt of an utterance does not cover the frozen prefix (+ 3 look-ahead tokens)");
    }
    return flow_solve(B, lens.data(), mu, spks, cond, nullptr, 0, n_steps, cfg_rate, 1.0f, est_chunk_frames, mel_out,
                      /*time_major*/ 1, st, px);
}

}  // namespace mossdec
