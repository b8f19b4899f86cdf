// Standalone device check of the tcgen05 implicit-GEMM kernel against a host fp64 reference.
// Build: make -C moss_speech_decoder_cosy_b200/csrc test_gemm ; run on a B200.
#include "../gemm.h"

#include <cuda_bf16.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

using namespace mossdec;

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                  \
        }                                                                             \
    } while (0)

static uint32_t rng_state = 12345;
static float frand() {
    rng_state = rng_state * 1664525u + 1013904223u;
    return ((rng_state >> 8) & 0xFFFF) / 65536.f * 2.f - 1.f;
}
static float bf16r(float x) { return __bfloat162float(__float2bfloat16(x)); }

static double act_ref(double x, int act, double p, double alpha) {
    switch (act) {
        case ACT_MISH: return x * tanh(log1p(exp(x)));
        case ACT_GELU: return 0.5 * x * (1.0 + erf(x / sqrt(2.0)));
        case ACT_ELU: return x > 0 ? x : expm1(x);
        case ACT_LRELU: return x > 0 ? x : x * p;
        case ACT_SNAKE: { double s = sin(x * alpha); return x + s * s / (alpha + 1e-9); }
        case ACT_SILU: return x / (1.0 + exp(-x));
        case ACT_ABS: return fabs(x);
        default: return x;
    }
}

struct Case {
    const char* name;
    int M, N, BN;
    int a_rows[2], a_cols[2], a_ld[2];  // a_cols[1]==0 -> single A
    int num_segs;
    GemmSeg segs[kGemmMaxSegs];
    bool bias, ln1, rowseq, tbias, resid, accum, outf32, outbf16, ln2;
    bool tiled = false;   // fp32 residual / output in the lane-major tiled layout, second LayerNorm without affine (N == 256)
    int act1, act2;
    float scale;
};

static int run_case(const Case& c, int num_sms) {
    if (const char* only = getenv("GEMM_ONLY")) {   // developer aid: run the cases whose name contains this string
        if (!strstr(c.name, only)) return 0;
    }
    int total_kb = 0;
    for (int s = 0; s < c.num_segs; ++s) total_kb += c.segs[s].nkb;
    const int K = total_kb * 64;
    const int Mpad = (c.M + 127) / 128 * 128;
    const int nseq = 4;

    std::vector<float> A[2];
    std::vector<__nv_bfloat16> Ab[2];
    for (int i = 0; i < 2; ++i) {
        if (c.a_cols[i] == 0) continue;
        size_t n = (size_t)c.a_rows[i] * c.a_ld[i] + c.a_cols[i];
        A[i].resize(n);
        Ab[i].resize(n);
        for (size_t j = 0; j < n; ++j) {
            A[i][j] = bf16r(frand());
            Ab[i][j] = __float2bfloat16(A[i][j]);
        }
    }
    std::vector<float> W((size_t)c.N * K);
    std::vector<__nv_bfloat16> Wb(W.size());
    for (size_t j = 0; j < W.size(); ++j) {
        W[j] = bf16r(frand() * 0.08f);
        Wb[j] = __float2bfloat16(W[j]);
    }
    std::vector<float> bias(c.N), g1(c.N), b1(c.N), g2(c.N), b2(c.N), alpha(c.N), tb((size_t)nseq * c.N);
    for (int n = 0; n < c.N; ++n) {
        bias[n] = frand() * 0.5f;
        g1[n] = 1.f + 0.3f * frand();
        b1[n] = 0.2f * frand();
        g2[n] = 1.f + 0.3f * frand();
        b2[n] = 0.2f * frand();
        alpha[n] = 1.f + 0.5f * frand();
    }
    for (auto& x : tb) x = frand();
    std::vector<int32_t> rowseq(Mpad);
    for (int r = 0; r < Mpad; ++r) rowseq[r] = (r % 97 == 5 || r % 97 == 6) ? -1 : (r * nseq / Mpad);
    std::vector<float> resid((size_t)Mpad * c.N), outf_init((size_t)Mpad * c.N);
    for (auto& x : resid) x = frand();
    for (auto& x : outf_init) x = frand();

    // device buffers
    __nv_bfloat16 *dA[2] = {nullptr, nullptr}, *dW, *dOb;
    float *dbias, *dg1, *db1, *dg2, *db2, *dalpha, *dtb, *dresid, *dOf;
    int32_t* drs;
    for (int i = 0; i < 2; ++i)
        if (c.a_cols[i]) {
            CK(cudaMalloc(&dA[i], Ab[i].size() * 2));
            CK(cudaMemcpy(dA[i], Ab[i].data(), Ab[i].size() * 2, cudaMemcpyHostToDevice));
        }
    CK(cudaMalloc(&dW, Wb.size() * 2));
    CK(cudaMemcpy(dW, Wb.data(), Wb.size() * 2, cudaMemcpyHostToDevice));
    auto up = [&](float** d, const std::vector<float>& h) {
        CK(cudaMalloc(d, h.size() * 4));
        CK(cudaMemcpy(*d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    };
    up(&dbias, bias); up(&dg1, g1); up(&db1, b1); up(&dg2, g2); up(&db2, b2); up(&dalpha, alpha); up(&dtb, tb);
    if (c.tiled) {   // device copy of the residual in the tiled layout
        std::vector<float> rt(resid.size());
        for (int r = 0; r < Mpad; ++r)
            for (int n = 0; n < c.N; ++n) rt[tiled_f32_offset(r, n)] = resid[(size_t)r * c.N + n];
        up(&dresid, rt);
    } else {
        up(&dresid, resid);
    }
    up(&dOf, outf_init);
    CK(cudaMalloc(&drs, Mpad * 4));
    CK(cudaMemcpy(drs, rowseq.data(), Mpad * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dOb, (size_t)Mpad * c.N * 2));
    CK(cudaMemset(dOb, 0, (size_t)Mpad * c.N * 2));

    GemmProblem g;
    for (int i = 0; i < 2; ++i) {
        g.a[i].ptr = dA[i];
        g.a[i].rows = c.a_rows[i];
        g.a[i].cols = c.a_cols[i];
        g.a[i].ld = c.a_ld[i];
    }
    g.w = dW; g.M = c.M; g.N = c.N; g.BN = c.BN; g.num_segs = c.num_segs;
    for (int s = 0; s < c.num_segs; ++s) g.segs[s] = c.segs[s];
    GemmEpilogue& e = g.epi;
    if (c.bias) e.bias = dbias;
    if (c.ln1) { e.ln1_g = dg1; e.ln1_b = db1; }
    e.act1 = c.act1; e.act1_p = 0.1f; e.act1_alpha = dalpha;
    if (c.rowseq) e.row_seq = drs;
    if (c.tbias) { e.tbias = dtb; e.tbias_ld = c.N; }
    if (c.resid) { e.resid = dresid; e.ldr = c.N; }
    e.accum_f32 = c.accum; e.out_scale = c.scale;
    if (c.outf32 || c.accum) { e.out_f32 = dOf; e.ldf = c.N; }
    if (c.outbf16) { e.out_bf16 = dOb; e.ldo = c.N; }
    if (c.ln2 && !c.tiled) { e.ln2_g = dg2; e.ln2_b = db2; }
    if (c.tiled) { e.f32_tiled = 1; e.ln2_plain = c.ln2 ? 1 : 0; }
    e.act2 = c.act2; e.act2_p = 0.1f; e.act2_alpha = dalpha;

    cudaError_t err = gemm_tc_launch(g, num_sms, 0);
    if (getenv("GEMM_BENCH") && c.M > 10000) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int i = 0; i < 3; ++i) gemm_tc_launch(g, num_sms, 0);
        cudaEventRecord(e0);
        for (int i = 0; i < 10; ++i) gemm_tc_launch(g, num_sms, 0);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("[%s] %.1f us per launch\n", c.name, ms * 100.f);
    }
    if (err != cudaSuccess) { printf("[%s] launch error %s\n", c.name, cudaGetErrorString(err)); return 1; }
    err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("[%s] exec error %s\n", c.name, cudaGetErrorString(err)); exit(3); }

    std::vector<float> of((size_t)Mpad * c.N);
    std::vector<__nv_bfloat16> ob((size_t)Mpad * c.N);
    CK(cudaMemcpy(of.data(), dOf, of.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ob.data(), dOb, ob.size() * 2, cudaMemcpyDeviceToHost));

    // host reference on a subset of rows (all rows if small)
    double max_ef = 0, max_eb = 0;
    int step = c.M > 2048 ? 37 : 1;
    std::vector<double> v(c.N);
    for (int r = 0; r < c.M; r += step) {
        int seq = c.rowseq ? rowseq[r] : 0;
        bool valid = seq >= 0;
        for (int n = 0; n < c.N; ++n) {
            double acc = 0;
            int koff = 0;
            for (int s = 0; s < c.num_segs; ++s) {
                const GemmSeg& sg = c.segs[s];
                long rr = (long)r + sg.row_shift;
                if (rr >= 0 && rr < c.a_rows[sg.map]) {
                    const float* a = A[sg.map].data() + (size_t)rr * c.a_ld[sg.map];
                    const float* w = W.data() + (size_t)n * K + koff;
                    for (int k = 0; k < sg.nkb * 64; ++k) {
                        int col = sg.col0 + k;
                        if (col < c.a_cols[sg.map]) acc += (double)a[col] * w[k];
                    }
                }
                koff += sg.nkb * 64;
            }
            if (c.bias) acc += bias[n];
            v[n] = acc;
        }
        if (c.ln1) {
            double m = 0, q = 0;
            for (int n = 0; n < c.N; ++n) m += v[n];
            m /= c.N;
            for (int n = 0; n < c.N; ++n) q += (v[n] - m) * (v[n] - m);
            double rs = 1.0 / sqrt(q / c.N + 1e-5);
            for (int n = 0; n < c.N; ++n) v[n] = (v[n] - m) * rs * g1[n] + b1[n];
        }
        for (int n = 0; n < c.N; ++n) {
            double x = act_ref(v[n], c.act1, 0.1, alpha[n]);
            if (c.tbias) x += tb[(size_t)(seq < 0 ? 0 : seq) * c.N + n];
            if (c.resid) x += resid[(size_t)r * c.N + n];
            if (c.accum) x += outf_init[(size_t)r * c.N + n];
            x *= c.scale;
            if (!valid) x = 0;
            v[n] = x;
        }
        if (c.outf32 || c.accum)
            for (int n = 0; n < c.N; ++n) max_ef = fmax(max_ef, fabs(v[n] - of[c.tiled ? tiled_f32_offset(r, n) : (size_t)r * c.N + n]));
        if (c.outbf16) {
            if (c.ln2) {
                double m = 0, q = 0;
                for (int n = 0; n < c.N; ++n) m += v[n];
                m /= c.N;
                for (int n = 0; n < c.N; ++n) q += (v[n] - m) * (v[n] - m);
                double rs = 1.0 / sqrt(q / c.N + 1e-5);
                for (int n = 0; n < c.N; ++n) v[n] = valid ? (c.tiled ? (v[n] - m) * rs : (v[n] - m) * rs * g2[n] + b2[n]) : 0;
            } else if (c.act2 != ACT_NONE) {
                for (int n = 0; n < c.N; ++n) v[n] = valid ? act_ref(v[n], c.act2, 0.1, alpha[n]) : 0;
            }
            for (int n = 0; n < c.N; ++n) {
                double got = __bfloat162float(ob[(size_t)r * c.N + n]);
                double tol_scale = fmax(1.0, fabs(v[n]));
                max_eb = fmax(max_eb, fabs(v[n] - got) / tol_scale);
            }
        }
    }
    bool ok = max_ef < 2e-3 && max_eb < 1.2e-2;
    printf("[%s] M=%d N=%d K=%d  max_err_f32=%.3e  max_relerr_bf16=%.3e  %s\n", c.name, c.M, c.N, K, max_ef, max_eb,
           ok ? "PASS" : "FAIL");

    for (int i = 0; i < 2; ++i) if (dA[i]) cudaFree(dA[i]);
    cudaFree(dW); cudaFree(dOb); cudaFree(dbias); cudaFree(dg1); cudaFree(db1); cudaFree(dg2); cudaFree(db2);
    cudaFree(dalpha); cudaFree(dtb); cudaFree(dresid); cudaFree(dOf); cudaFree(drs);
    return ok ? 0 : 1;
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    printf("device %s sm_%d%d SMs=%d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
    const int sms = prop.multiProcessorCount;
    int fails = 0;

    auto base = [](const char* name, int M, int N, int BN, int K) {
        Case c{};
        c.name = name; c.M = M; c.N = N; c.BN = BN;
        c.a_rows[0] = M; c.a_cols[0] = K; c.a_ld[0] = K;
        c.num_segs = 1;
        c.segs[0] = {0, 0, 0, K / 64};
        c.bias = true; c.outbf16 = true; c.scale = 1.f;
        return c;
    };
    {   // 1. plain single k-block
        Case c = base("plain_k64", 128, 256, 256, 64); c.bias = false; c.outf32 = true;
        fails += run_case(c, sms);
    }
    {   // 2. plain K=256, ragged M
        Case c = base("plain_k256_raggedM", 300, 256, 256, 256); c.outf32 = true;
        fails += run_case(c, sms);
    }
    {   // 3. QKV-like, several n tiles
        Case c = base("qkv_n1536", 1000, 1536, 256, 256);
        fails += run_case(c, sms);
    }
    {   // 3b. the same projection at a size that takes the B-stationary schedule (6 N tiles, 157 M tiles, ragged last tile)
        Case c = base("qkv_bstat_raggedM", 20011, 1536, 256, 256);
        fails += run_case(c, sms);
    }
    {   // 4. FF1 + GELU, long persistent loop
        Case c = base("ff1_gelu_bigM", 40000, 1024, 256, 256); c.act1 = ACT_GELU;
        fails += run_case(c, sms);
    }
    {   // 5. FF2 + resid + LN2 + fp32 out
        Case c = base("ff2_resid_ln2", 5000, 256, 256, 1024); c.resid = true; c.outf32 = true; c.ln2 = true; c.rowseq = true;
        fails += run_case(c, sms);
    }
    {   // 6. causal conv k3, Cin=320, LN1+Mish+mask+tbias
        Case c = base("conv3_ln_mish", 3000, 256, 256, 320);
        c.num_segs = 3;
        c.segs[0] = {0, -2, 0, 5}; c.segs[1] = {0, -1, 0, 5}; c.segs[2] = {0, 0, 0, 5};
        c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.tbias = true;
        fails += run_case(c, sms);
    }
    {   // 7. conv k3 + LN1 + Mish + resid + LN2 (resnet block2 form)
        Case c = base("conv3_ln_mish_res_ln2", 2000, 256, 256, 256);
        c.num_segs = 3;
        c.segs[0] = {0, -2, 0, 4}; c.segs[1] = {0, -1, 0, 4}; c.segs[2] = {0, 0, 0, 4};
        c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.resid = true; c.outf32 = true; c.ln2 = true;
        fails += run_case(c, sms);
    }
    {   // 6b. same on whole tiles: LayerNorm + Mish + time bias through the register-resident epilogue (mode 3)
        Case c = base("conv3_ln_mish_lnact", 3072, 256, 256, 320);
        c.num_segs = 3;
        c.segs[0] = {0, -2, 0, 5}; c.segs[1] = {0, -1, 0, 5}; c.segs[2] = {0, 0, 0, 5};
        c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.tbias = true;
        fails += run_case(c, sms);
    }
    {   // 7b. resnet block2 on whole tiles, tiled fp32 stream, plain second LayerNorm (mode 4)
        Case c = base("conv3_ln_mish_res_ln2_lnact", 83712 / 4, 256, 256, 256);
        c.num_segs = 3;
        c.segs[0] = {0, -2, 0, 4}; c.segs[1] = {0, -1, 0, 4}; c.segs[2] = {0, 0, 0, 4};
        c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.resid = true; c.outf32 = true; c.ln2 = true; c.tiled = true;
        fails += run_case(c, sms);
    }
    if (getenv("GEMM_BENCH")) {   // config-2 sized resnet blocks (timing)
        {
            Case c = base("bench_block1_lnact", 83712, 256, 256, 256);
            c.num_segs = 3;
            c.segs[0] = {0, -2, 0, 4}; c.segs[1] = {0, -1, 0, 4}; c.segs[2] = {0, 0, 0, 4};
            c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.tbias = true;
            fails += run_case(c, sms);
        }
        {
            Case c = base("bench_block2_lnact", 83712, 256, 256, 256);
            c.num_segs = 3;
            c.segs[0] = {0, -2, 0, 4}; c.segs[1] = {0, -1, 0, 4}; c.segs[2] = {0, 0, 0, 4};
            c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.resid = true; c.outf32 = true; c.ln2 = true; c.tiled = true;
            fails += run_case(c, sms);
        }
        {
            Case c = base("bench_conv3_plain_bf16", 83712, 256, 256, 256);
            c.num_segs = 3;
            c.segs[0] = {0, -2, 0, 4}; c.segs[1] = {0, -1, 0, 4}; c.segs[2] = {0, 0, 0, 4};
            fails += run_case(c, sms);
        }
        {
            Case c = base("bench_dense_k768_lnact", 83712, 256, 256, 768);
            c.ln1 = true; c.act1 = ACT_MISH; c.rowseq = true; c.tbias = true;
            fails += run_case(c, sms);
        }
        {
            Case c = base("bench_dense_k768_bf16", 83712, 256, 256, 768);
            fails += run_case(c, sms);
        }
        {
            Case c = base("bench_qkv", 83712, 1536, 256, 256);
            fails += run_case(c, sms);
        }
        for (int m : {41856, 20928, 10496}) {   // the same projection on a half / quarter / eighth batch: output 128 / 64 / 32 MB (L2 is 126 MB)
            Case c = base(m == 41856 ? "bench_qkv_half" : (m == 20928 ? "bench_qkv_quarter" : "bench_qkv_eighth"), m, 1536, 256, 256);
            fails += run_case(c, sms);
        }
        // vocoder C = 64 stage (16 utterances x 61 440 rows): convs1 with the Snake epilogue, convs2 with the fp32 residual
        for (int k : {3, 11}) {
            Case c = base(k == 3 ? "bench_voc64_k3_snake" : "bench_voc64_k11_snake", 983040, 64, 64, 64);
            c.num_segs = k;
            for (int j = 0; j < k; ++j) c.segs[j] = {0, (j - (k - 1) / 2) * 3, 0, 1};
            c.act1 = ACT_SNAKE; c.rowseq = true;
            fails += run_case(c, sms);
        }
        {
            Case c = base("bench_voc128_k7_resid", 327680, 128, 128, 128);
            c.num_segs = 7;
            for (int j = 0; j < 7; ++j) c.segs[j] = {0, j - 3, 0, 2};
            c.rowseq = true; c.resid = true; c.outf32 = true; c.act2 = ACT_SNAKE;
            fails += run_case(c, sms);
        }
        for (int k : {3, 11}) {
            Case c = base(k == 3 ? "bench_voc64_k3_resid" : "bench_voc64_k11_resid", 983040, 64, 64, 64);
            c.num_segs = k;
            for (int j = 0; j < k; ++j) c.segs[j] = {0, j - (k - 1) / 2, 0, 1};
            c.rowseq = true; c.resid = true; c.outf32 = true;
            fails += run_case(c, sms);
        }
    }
    {   // 8. N=80 projection
        Case c = base("proj_n80", 1500, 80, 80, 256); c.outf32 = true; c.outbf16 = false; c.rowseq = true;
        fails += run_case(c, sms);
    }
    {   // 9. two A operands (channel concat), dilated symmetric taps, snake on the bf16 side, accumulate + scale
        Case c = base("dual_a_dilated_snake", 2500, 128, 128, 128);
        c.a_rows[1] = 2500; c.a_cols[1] = 128; c.a_ld[1] = 128;
        c.num_segs = 4;
        c.segs[0] = {0, -5, 0, 2}; c.segs[1] = {1, 0, 0, 2}; c.segs[2] = {0, 5, 0, 2}; c.segs[3] = {1, 3, 64, 1};
        c.accum = true; c.scale = 1.f / 3.f; c.act2 = ACT_SNAKE; c.rowseq = true;
        fails += run_case(c, sms);
    }
    {   // 10. overlapping-window A (strided conv as GEMM): ld < cols, K tail zero-filled by TMA
        Case c = base("overlap_rows_k720", 900, 256, 256, 768);
        c.a_rows[0] = 900; c.a_cols[0] = 720; c.a_ld[0] = 360;
        c.segs[0] = {0, 0, 0, 12};
        fails += run_case(c, sms);
    }
    {   // 11. small N = 64 / 16-col granularity, lrelu
        Case c = base("n64_lrelu", 4000, 64, 64, 448); c.act1 = ACT_LRELU; c.outf32 = true;
        fails += run_case(c, sms);
    }
    {   // 12. N = 32 (>=16), ELU
        Case c = base("n32_elu", 777, 32, 32, 64); c.act1 = ACT_ELU;
        fails += run_case(c, sms);
    }
    {   // 13. TMA-store epilogue with two warps per 64 columns (BN = 128), snake, masked rows, ragged M
        Case c = base("n128_snake_tma_store", 3001, 256, 128, 192); c.act1 = ACT_SNAKE; c.rowseq = true;
        fails += run_case(c, sms);
    }
    {   // 14. vocoder ResBlock second conv, C = 64: fp32 residual -> fp32 stream + Snake bf16, masked rows, ragged M (epi_vocres)
        Case c = base("vocres_n64_snake", 3001, 64, 64, 64);
        c.num_segs = 3;
        for (int j = 0; j < 3; ++j) c.segs[j] = {0, j - 1, 0, 1};
        c.rowseq = true; c.resid = true; c.outf32 = true; c.act2 = ACT_SNAKE;
        fails += run_case(c, sms);
    }
    {   // 15. same at C = 128 (two 16-column sub-tiles per warp), accumulating into the running sum with the 1/3 scale, leaky ReLU out
        Case c = base("vocres_n128_accum", 2500, 128, 128, 128);
        c.num_segs = 7;
        for (int j = 0; j < 7; ++j) c.segs[j] = {0, j - 3, 0, 2};
        c.rowseq = true; c.resid = true; c.outf32 = true; c.accum = true; c.scale = 1.f / 3.f; c.act2 = ACT_LRELU;
        fails += run_case(c, sms);
    }
    {   // 16. fp32 only (no bf16 side), no bias
        Case c = base("vocres_n64_f32only", 1000, 64, 64, 64);
        c.num_segs = 3;
        for (int j = 0; j < 3; ++j) c.segs[j] = {0, (j - 1) * 5, 0, 1};
        c.bias = false; c.outbf16 = false; c.resid = true; c.outf32 = true;
        fails += run_case(c, sms);
    }
    {   // 17. TMA-store epilogue at BN = 64 (16 columns per warp), Snake, masked rows, ragged M, no bias variant below
        Case c = base("n64_snake_tma_store", 3001, 64, 64, 64);
        c.num_segs = 3;
        for (int j = 0; j < 3; ++j) c.segs[j] = {0, (j - 1) * 3, 0, 1};
        c.act1 = ACT_SNAKE; c.rowseq = true;
        fails += run_case(c, sms);
        c.name = "n64_plain_tma_store_nobias"; c.bias = false; c.act1 = ACT_NONE; c.rowseq = false;
        fails += run_case(c, sms);
    }
    printf("%s (%d failing)\n", fails ? "SOME TESTS FAILED" : "ALL TESTS PASSED", fails);
    return fails ? 1 : 0;
}
