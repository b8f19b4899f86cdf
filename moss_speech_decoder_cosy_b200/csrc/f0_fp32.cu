// FP32 contractions on the CUDA cores -- the dense contractions of the path that deliberately do NOT run on the tensor
// cores: the f0 predictor of the vocoder and the prompt mel front end (both feed quantities whose errors are integrated
// or logarithmically expanded downstream; both are ~1 % of the work).
//
// (1) ConvRNNF0Predictor (cosyvoice/hifigan/f0_predictor.py:31-60: 5 x {Conv1d(k3, pad 1) + ELU}, Linear(512 -> 1), abs).
// Why fp32: the NSF source integrates the predicted f0 into the sine phase over the whole utterance
// (generator.py:275-283), so a RELATIVE BIAS of the f0 predictor of 1e-5 turns into a phase error of the k-th harmonic
// of 2 pi k f0 1e-5 t.  tcgen05 accumulates its fp32 partial sums with truncation (round toward zero): measured on B200
// (profiles/r02_f0_bias.txt), the split-bf16 tensor-core version of this predictor had a mean signed relative f0 error
// of -1.9e-5 over the voiced frames of a 10 s utterance (the CPU simulation of the same split with round-to-nearest
// accumulation: +7e-7), and the source signal fell from 30 dB SNR in the first second to 10 dB in the tenth.  With
// round-to-nearest FMAs the error is unbiased and at the 1e-6 level of the reference's own fp32 convolutions.
//
// (2) matcha mel_spectrogram (third_party/Matcha-TTS/matcha/utils/audio.py:45-82) behind AudioDecoder.feat_extractor:
// reflect pad, 1920-point windowed DFT (hop 480) as an fp32 GEMM over the overlapping-window view of the padded signal,
// magnitude, Slaney mel filterbank GEMM, log(clamp(., 1e-5)).  log-mel needs the quiet bins right (-100 dB below the loud
// ones); a few hundred frames per prompt, run once per speaker.
//
//   conv_f32_kernel<kTaps, kAct> : out[m, n] = rowmask[m] * act(bias[n] + sum_{tap, c} in[m + tap - kTaps / 2, c] * Wt[tap * Cin + c, n])
//       classic register-tiled SGEMM: 128 x 128 block tile, 16-wide K slabs, 256 threads x (8 x 8) outputs, operands
//       double-buffered in shared memory, the next slab's global loads in flight during the math, packed fma.rn.f32x2.
//       Rows of different utterances are separated by zero guard rows (engine_hift.cu), so the +-1 row taps see the
//       conv's zero padding; rows outside [0, M) read as zero.  `ld_in` < Cin gives overlapping windows (STFT frames).
//   f0_classifier_kernel : f0[m] = | w . h[m, :] + b |   (one warp per row)
#include "kernels.h"
#include "ptx.cuh"

namespace mossdec {

namespace {

constexpr int kBM = 128, kBN = 128, kBK = 16, kPadA = 4;

template <int kTaps, int kAct>   // kAct: 0 = none, 1 = ELU
__global__ void __launch_bounds__(256, 2)
conv_f32_kernel(const float* __restrict__ in, int M, int64_t ld_in, int Cin, const float* __restrict__ wt, int N,
                const float* __restrict__ bias, const int32_t* __restrict__ row_seq, float* __restrict__ out) {
    __shared__ __align__(16) float sA[2][kBK][kBM + kPadA];
    __shared__ __align__(16) float sB[2][kBK][kBN];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * kBN;
    const int slabs_per_tap = Cin / kBK;
    const int nslab = kTaps * slabs_per_tap;

    // global -> register staging of one K slab: A 128 rows x 16 floats (2 float4 per thread), B 16 x 128 (2 float4)
    float4 ra[2], rb[2];
    auto load_slab = [&](int s) {
        const int tap = s / slabs_per_tap, c0 = (s % slabs_per_tap) * kBK;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int f = tid + 256 * i;
            const int r = f >> 2, kq = f & 3;
            const int row = m0 + r + tap - kTaps / 2;
            ra[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row >= 0 && row < M) ra[i] = __ldg(reinterpret_cast<const float4*>(in + (size_t)row * ld_in + c0 + kq * 4));
            const int kb = f >> 5, nq = f & 31;
            rb[i] = __ldg(reinterpret_cast<const float4*>(wt + (size_t)(s * kBK + kb) * N + n0 + nq * 4));
        }
    };
    auto store_slab = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int f = tid + 256 * i;
            const int r = f >> 2, kq = f & 3;
            sA[buf][kq * 4 + 0][r] = ra[i].x;
            sA[buf][kq * 4 + 1][r] = ra[i].y;
            sA[buf][kq * 4 + 2][r] = ra[i].z;
            sA[buf][kq * 4 + 3][r] = ra[i].w;
            const int kb = f >> 5, nq = f & 31;
            *reinterpret_cast<float4*>(&sB[buf][kb][nq * 4]) = rb[i];
        }
    };

    uint64_t acc[8][4];   // 8 rows x 8 columns as packed f32x2
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;

    load_slab(0);
    store_slab(0);
    __syncthreads();
    for (int s = 0; s < nslab; ++s) {
        const int buf = s & 1;
        if (s + 1 < nslab) load_slab(s + 1);
#pragma unroll
        for (int k = 0; k < kBK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4*>(&sA[buf][k][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&sA[buf][k][64 + ty * 4]);
            const ulonglong2 b0 = *reinterpret_cast<const ulonglong2*>(&sB[buf][k][tx * 4]);
            const ulonglong2 b1 = *reinterpret_cast<const ulonglong2*>(&sB[buf][k][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint64_t aa = f2_pack(av[i], av[i]);
                acc[i][0] = f2_fma(aa, b0.x, acc[i][0]);
                acc[i][1] = f2_fma(aa, b0.y, acc[i][1]);
                acc[i][2] = f2_fma(aa, b1.x, acc[i][2]);
                acc[i][3] = f2_fma(aa, b1.y, acc[i][3]);
            }
        }
        if (s + 1 < nslab) {
            store_slab(buf ^ 1);   // the other buffer was last read in iteration s - 1, before the barrier below
            __syncthreads();
        }
    }

    // epilogue: bias + activation (F.elu: x > 0 ? x : expm1(x)), guard rows forced to zero
    float bv[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        bv[j] = bias ? __ldg(bias + n0 + tx * 4 + j) : 0.f;
        bv[4 + j] = bias ? __ldg(bias + n0 + 64 + tx * 4 + j) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (row >= M) continue;
        const bool valid = row_seq == nullptr || __ldg(row_seq + row) >= 0;
        float v[8];
        f2_unpack(acc[i][0], v[0], v[1]);
        f2_unpack(acc[i][1], v[2], v[3]);
        f2_unpack(acc[i][2], v[4], v[5]);
        f2_unpack(acc[i][3], v[6], v[7]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float x = v[j] + bv[j];
            v[j] = valid ? (kAct == 1 ? (x > 0.f ? x : expm1f(x)) : x) : 0.f;
        }
        float* o = out + (size_t)row * N + n0 + tx * 4;
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 64) = make_float4(v[4], v[5], v[6], v[7]);
    }
}

// f0[m] = | w . h[m, :] + b | for rows that belong to a sequence, 0 on guard rows.  One warp per row, C % 128 == 0.
__global__ void f0_classifier_kernel(const float* __restrict__ h, int M, int C, const float* __restrict__ w,
                                     const float* __restrict__ b, const int32_t* __restrict__ row_seq,
                                     float* __restrict__ f0, int f0_ld) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= M) return;
    const float4* hr = reinterpret_cast<const float4*>(h + (size_t)row * C);
    const float4* wr = reinterpret_cast<const float4*>(w);
    float acc = 0.f;
    for (int i = lane; i < C / 4; i += 32) {
        const float4 a = __ldg(hr + i), c = __ldg(wr + i);
        acc = fmaf(a.x, c.x, acc);
        acc = fmaf(a.y, c.y, acc);
        acc = fmaf(a.z, c.z, acc);
        acc = fmaf(a.w, c.w, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) f0[(size_t)row * f0_ld] = (row_seq == nullptr || row_seq[row] >= 0) ? fabsf(acc + b[0]) : 0.f;
}

}  // namespace

cudaError_t launch_conv3_elu_f32(const float* in, int M, int Cin, const float* wt, int N, const float* bias,
                                 const int32_t* row_seq, float* out, cudaStream_t st) {
    if (M <= 0 || (Cin % kBK) != 0 || (N % kBN) != 0) return cudaErrorInvalidValue;
    dim3 grid((M + kBM - 1) / kBM, N / kBN);
    conv_f32_kernel<3, 1><<<grid, 256, 0, st>>>(in, M, (int64_t)Cin, Cin, wt, N, bias, row_seq, out);
    return cudaGetLastError();
}

// out [M, N] = in_view [M, K] . Wt [K, N], row m of the view starts at in + m * ld_in (ld_in < K: overlapping windows)
cudaError_t launch_gemm_f32(const float* in, int M, int64_t ld_in, int K, const float* wt, int N, float* out,
                            cudaStream_t st) {
    if (M <= 0 || (K % kBK) != 0 || (N % kBN) != 0 || (ld_in % 4) != 0) return cudaErrorInvalidValue;
    dim3 grid((M + kBM - 1) / kBM, N / kBN);
    conv_f32_kernel<1, 0><<<grid, 256, 0, st>>>(in, M, ld_in, K, wt, N, nullptr, nullptr, out);
    return cudaGetLastError();
}

cudaError_t launch_f0_classifier(const float* h, int M, int C, const float* w, const float* b, const int32_t* row_seq,
                                 float* f0, int f0_ld, cudaStream_t st) {
    if ((C % 128) != 0) return cudaErrorInvalidValue;
    f0_classifier_kernel<<<(M * 32 + 255) / 256, 256, 0, st>>>(h, M, C, w, b, row_seq, f0, f0_ld);
    return cudaGetLastError();
}

}  // namespace mossdec
