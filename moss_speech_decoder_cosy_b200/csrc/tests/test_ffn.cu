// Standalone device check of the fused feed-forward kernel (ffn_tc.cu) against a host fp64 reference.
// Build: make -C moss_speech_decoder_cosy_b200/csrc test_ffn ; run on a B200.
#include "../gemm.h"

#include <cuda_bf16.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
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

static uint32_t rng_state = 777;
static float frand() {
    rng_state = rng_state * 1664525u + 1013904223u;
    return ((rng_state >> 8) & 0xFFFF) / 65536.f * 2.f - 1.f;
}
static float bf16r(float x) { return __bfloat162float(__float2bfloat16(x)); }

static int run(int M, bool ln, bool store_f32, int ldo, int num_sms, int grid_cap) {
    const int C = 256, H = 1024;
    std::vector<float> x((size_t)M * C), w1((size_t)H * C), w2((size_t)C * H), b1(H), b2(C), g(C), bt(C), hres((size_t)M * C);
    std::vector<int32_t> rs(M);
    for (auto& v : x) v = bf16r(frand());
    for (auto& v : w1) v = bf16r(frand() * 0.08f);
    for (auto& v : w2) v = bf16r(frand() * 0.05f);
    for (auto& v : b1) v = frand() * 0.5f;
    for (auto& v : b2) v = frand() * 0.5f;
    for (auto& v : g) v = 1.f + 0.3f * frand();
    for (auto& v : bt) v = 0.3f * frand();
    for (auto& v : hres) v = frand() * 2.f;
    for (int r = 0; r < M; ++r) rs[r] = (r % 97 == 5 || r % 131 < 3) ? -1 : r / 300;
    auto to_bf = [](const std::vector<float>& a) {
        std::vector<__nv_bfloat16> o(a.size());
        for (size_t i = 0; i < a.size(); ++i) o[i] = __float2bfloat16(a[i]);
        return o;
    };
    auto xb = to_bf(x), w1b = to_bf(w1), w2b = to_bf(w2);
    __nv_bfloat16 *dx, *dw1, *dw2, *dout;
    float *db1, *db2, *dg, *dbt, *dh;
    int32_t* drs;
    CK(cudaMalloc(&dx, xb.size() * 2 + 65536));
    CK(cudaMalloc(&dw1, w1b.size() * 2 + 65536));
    CK(cudaMalloc(&dw2, w2b.size() * 2 + 65536));
    CK(cudaMalloc(&dout, (size_t)M * ldo * 2));
    CK(cudaMalloc(&db1, H * 4));
    CK(cudaMalloc(&db2, C * 4));
    CK(cudaMalloc(&dg, C * 4));
    CK(cudaMalloc(&dbt, C * 4));
    CK(cudaMalloc(&dh, hres.size() * 4));
    CK(cudaMalloc(&drs, M * 4));
    CK(cudaMemcpy(dx, xb.data(), xb.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw1, w1b.data(), w1b.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw2, w2b.data(), w2b.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db1, b1.data(), H * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db2, b2.data(), C * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dg, g.data(), C * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dbt, bt.data(), C * 4, cudaMemcpyHostToDevice));
    {   // the kernel keeps the residual stream in the lane-major tiled layout
        std::vector<float> ht(hres.size());
        for (int r = 0; r < M; ++r)
            for (int n = 0; n < C; ++n) ht[tiled_f32_offset(r, n)] = hres[(size_t)r * C + n];
        CK(cudaMemcpy(dh, ht.data(), ht.size() * 4, cudaMemcpyHostToDevice));
    }
    CK(cudaMemcpy(drs, rs.data(), M * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dout, 0xff, (size_t)M * ldo * 2));

    FfnProblem f;
    f.x = dx; f.M = M; f.w1 = dw1; f.b1 = db1; f.w2 = dw2; f.b2 = db2; f.hres = dh; f.row_seq = drs;
    f.ln = ln ? 1 : 0;
    f.out_bf16 = dout; f.ldo = ldo; f.store_f32 = store_f32;
    long long* dtrace = nullptr;
    CK(cudaMalloc(&dtrace, 48 * 8 * 8));
    CK(cudaMemset(dtrace, 0, 48 * 8 * 8));
    if (getenv("FFN_TRACE")) f.trace = dtrace;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    CK(cudaEventRecord(e0));
    CK(ffn_tc_launch(f, grid_cap > 0 ? grid_cap : num_sms, 0));
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);

    if (f.trace && M / 128 > 148 * 2) {
        std::vector<long long> tr(48 * 8);
        CK(cudaMemcpy(tr.data(), dtrace, tr.size() * 8, cudaMemcpyDeviceToHost));
        for (int it = 1; it < 3; ++it) {
            const long long t0 = tr[it * 48];
            printf("  tile it=%d epilogue warp: preload_done %lld |", it, tr[it * 48 + 1] - t0);
            for (int h = 0; h < 8; ++h) printf(" E1[%d] wake %lld done %lld |", h, tr[it * 48 + 2 + 2 * h] - t0, tr[it * 48 + 3 + 2 * h] - t0);
            printf(" acc2_full %lld stored %lld\n", tr[it * 48 + 18] - t0, tr[it * 48 + 19] - t0);
            printf("  tile it=%d MMA thread: x_full %lld |", it, tr[it * 48 + 32] - t0);
            for (int h = 0; h < 8; ++h) printf(" p_ready[%d] %lld |", h, tr[it * 48 + 33 + h] - t0);
            printf(" issued_all %lld | waited on weights %lld, on p_ready %lld\n", tr[it * 48 + 41] - t0, tr[it * 48 + 42], tr[it * 48 + 43]);
        }
    }
    cudaFree(dtrace);
    std::vector<float> oh(hres.size());
    std::vector<__nv_bfloat16> ob((size_t)M * ldo);
    CK(cudaMemcpy(oh.data(), dh, oh.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ob.data(), dout, ob.size() * 2, cudaMemcpyDeviceToHost));

    // host reference on a sample of rows
    double max_f = 0, max_b = 0;
    int checked = 0;
    std::vector<double> hid(H), y(C);
    for (int r = 0; r < M; r += (M > 2000 ? 37 : 1)) {
        const bool valid = rs[r] >= 0;
        for (int n = 0; n < H; ++n) {
            double a = b1[n];
            for (int k = 0; k < C; ++k) a += (double)x[(size_t)r * C + k] * w1[(size_t)n * C + k];
            const double ge = 0.5 * a * (1.0 + erf(a / sqrt(2.0)));
            hid[n] = bf16r((float)ge);
        }
        double mean = 0, var = 0;
        for (int n = 0; n < C; ++n) {
            double a = b2[n];
            for (int k = 0; k < H; ++k) a += hid[k] * w2[(size_t)n * H + k];
            a += hres[(size_t)r * C + n];
            y[n] = valid ? a : 0.0;
            mean += y[n];
        }
        mean /= C;
        for (int n = 0; n < C; ++n) var += (y[n] - mean) * (y[n] - mean);
        var /= C;
        for (int n = 0; n < C; ++n) {
            if (store_f32) max_f = fmax(max_f, fabs(oh[tiled_f32_offset(r, n)] - y[n]));
            double want = ln ? (valid ? (y[n] - mean) / sqrt(var + 1e-5) : 0.0) : y[n];
            double got = __bfloat162float(ob[(size_t)r * ldo + n]);
            max_b = fmax(max_b, fabs(got - want) / (1.0 + fabs(want)));
        }
        ++checked;
    }
    const bool ok = max_f < 2e-2 && max_b < 1.5e-2;
    const double tf = 2.0 * M * C * H * 2.0 / (ms * 1e-3) / 1e12;
    printf("ffn M=%d ln=%d f32=%d ldo=%d grid<=%d: rows checked %d  max|f32 err| %.3e  max bf16 rel err %.3e  %.3f ms (%.0f TFLOP/s incl. launch)  %s\n",
           M, (int)ln, (int)store_f32, ldo, grid_cap > 0 ? grid_cap : num_sms, checked, max_f, max_b, ms, tf, ok ? "OK" : "FAIL");
    cudaFree(dx); cudaFree(dw1); cudaFree(dw2); cudaFree(dout); cudaFree(db1); cudaFree(db2); cudaFree(dg); cudaFree(dbt);
    cudaFree(dh); cudaFree(drs);
    return ok ? 0 : 1;
}

int main() {
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int bad = 0;
    bad += run(128, true, true, 256, sms, 0);
    bad += run(128 * 5, true, true, 256, sms, 2);      // 3 / 2 tiles per CTA: phase bookkeeping across tiles
    bad += run(128 * 7, false, false, 512, sms, 3);
    bad += run(128 * 654, true, true, 256, sms, 0);    // config-2 size
    bad += run(128 * 654, true, true, 256, sms, 0);    // warm timing
    printf(bad ? "FAILED\n" : "ALL OK\n");
    return bad ? 1 : 0;
}
