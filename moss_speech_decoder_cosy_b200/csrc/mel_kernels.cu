// Prompt mel front end (row N2): matcha mel_spectrogram (third_party/Matcha-TTS/matcha/utils/audio.py:45-82) with the
// reference's configuration n_fft = win = 1920, hop 480, 80 mels, 24 kHz, fmin 0, fmax 8000, center = False:
//   y  = reflect_pad(wav, (n_fft - hop) / 2)                                                   audio.py:58-61
//   S  = | STFT(y) |  with sqrt(re^2 + im^2 + 1e-9)                                            audio.py:63-78
//   M  = log(clamp(mel_basis @ S, 1e-5))                                                       audio.py:80-82, 23-24
// The windowed DFT is an fp32 GEMM over the overlapping-window view of y (frame t = y[480 t : 480 t + 1920], row stride
// 480) against the basis [1920, 2048] = [hann * cos | hann * -sin] (961 + 63 zero columns each); the filterbank is a
// second fp32 GEMM [T, 1024] x [1024, 128].  GEMMs: f0_fp32.cu launch_gemm_f32.  Elementwise kernels below.
#include <atomic>
#include <mutex>

#include "engine.h"
#include "gemm.h"

namespace mossdec {

namespace {

// ---- scratch of the allocation-free variant (Engine::mel_spectrogram_static, include/mossdec.h
// mossdec_mel_spectrogram_static): a file-scope device array instead of the growable cudaMalloc workspace.  Every device
// that loads libmossdec.so holds these bytes until the process exits.  Layout for Tp = kMelStaticMaxFrames rows:
//   y [Tp * 480 + 1920 + 64] | spectrum [Tp, 2048] | magnitude [Tp, 1024] | mel energies [Tp, 128]     (fp32, 1 KB aligned)
constexpr int kMelStaticMaxFrames = 512;   // == MOSSDEC_MEL_STATIC_MAX_FRAMES: 10.24 s of 24 kHz prompt
constexpr size_t mel_align1k(size_t b) { return (b + 1023) & ~size_t(1023); }
constexpr size_t kMelStaticY = mel_align1k(((size_t)kMelStaticMaxFrames * 480 + 1920 + 64) * 4);
constexpr size_t kMelStaticSpec = mel_align1k((size_t)kMelStaticMaxFrames * 2048 * 4);
constexpr size_t kMelStaticMag = mel_align1k((size_t)kMelStaticMaxFrames * 1024 * 4);
constexpr size_t kMelStaticAcc = mel_align1k((size_t)kMelStaticMaxFrames * 128 * 4);
constexpr size_t kMelStaticBytes = kMelStaticY + kMelStaticSpec + kMelStaticMag + kMelStaticAcc;   // 7 544 832 B = 7.2 MiB
static __device__ __align__(256) unsigned char g_mel_static_scratch[kMelStaticBytes];

// In host code `g_mel_static_scratch` is only a stand-in for the symbol: the device address is resolved once per device
// (with that device current) by cudaGetSymbolAddress and kept here.
uint8_t* g_mel_static_ptr[64] = {};

cudaError_t mel_static_scratch(uint8_t** out) {
    static std::atomic<uint64_t> done{0};
    static std::mutex mu;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = once_per_device(done, mu, [dev] {
        void* p = nullptr;
        const cudaError_t ge = cudaGetSymbolAddress(&p, g_mel_static_scratch);
        if (ge == cudaSuccess) g_mel_static_ptr[dev & 63] = static_cast<uint8_t*>(p);
        return ge;
    });
    if (e != cudaSuccess) return e;
    *out = g_mel_static_ptr[dev & 63];
    return *out ? cudaSuccess : cudaErrorInvalidSymbol;
}

// ypad[i] = wav[reflect(i - pad)], i in [0, L + 2 pad); zero beyond (slack for the last GEMM tile's rows)
__global__ void mel_reflect_pad_kernel(const float* __restrict__ wav, int64_t L, int pad, int64_t n_out, float* __restrict__ y) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    int64_t j = i - pad;
    if (j < 0) j = -j;
    if (j >= L) j = 2 * (L - 1) - j;
    y[i] = (i < L + 2 * pad && j >= 0 && j < L) ? wav[j] : 0.f;
}

// spec [T, 2 * ldh] = [re | im] -> mag [T, ldh]: sqrt(re^2 + im^2 + 1e-9) for k < nbins, 0 on the padding columns
__global__ void mel_magnitude_kernel(const float* __restrict__ spec, int T, int ldh, int nbins, float* __restrict__ mag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)T * ldh) return;
    const int t = (int)(i / ldh), k = (int)(i % ldh);
    const float re = spec[(size_t)t * 2 * ldh + k], im = spec[(size_t)t * 2 * ldh + ldh + k];
    mag[i] = k < nbins ? sqrtf(re * re + im * im + 1e-9f) : 0.f;
}

// acc [T, ld] (mel energies) -> out [n_mels, T] = log(max(acc, 1e-5))   (the reference's [B, 80, T] layout)
__global__ void mel_log_kernel(const float* __restrict__ acc, int T, int ld, int n_mels, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)T * n_mels) return;
    const int c = (int)(i / T), t = (int)(i % T);
    out[i] = logf(fmaxf(acc[(size_t)t * ld + c], 1e-5f));
}

}  // namespace

int Engine::mel_spectrogram(const float* wav, int64_t L, float* mel_out, cudaStream_t st) {
    DeviceGuard g(device_);
    constexpr int kFft = 1920, kHop = 480, kPad = (kFft - kHop) / 2, kBins = kFft / 2 + 1, kLdh = 1024, kMels = 80;
    const DevTensor* dft = find("mel.dft");
    const DevTensor* fb = find("mel.fb");
    if (!dft || !fb) return fail("mel front end tensors (mel.dft, mel.fb) not loaded");
    if (L <= kPad) return fail("mel_spectrogram: input shorter than the reflect padding (720 samples)");
    const int64_t T64 = (L + 2 * kPad - kFft) / kHop + 1;
    if (T64 <= 0 || T64 > (1 << 24)) return fail("mel_spectrogram: bad length");
    const int T = (int)T64;
    const int Tp = (T + 127) / 128 * 128;
    // workspace: padded signal (rows of the last tile read up to Tp * 480 + 1920), spectrum [Tp, 2048], magnitude [Tp, 1024],
    // mel energies [Tp, 128]
    const size_t n_y = (size_t)Tp * kHop + kFft + 64;
    size_t need = 0;
    auto take = [&](size_t bytes) {
        size_t o = need;
        need += (bytes + 1023) & ~size_t(1023);
        return o;
    };
    const size_t o_y = take(n_y * 4), o_spec = take((size_t)Tp * 2 * kLdh * 4), o_mag = take((size_t)Tp * kLdh * 4),
                 o_acc = take((size_t)Tp * 128 * 4);
    ws_mel_.reserve(need);
    if (!ws_mel_.as<void>()) return fail("cudaMalloc(mel workspace) failed");
    uint8_t* b = ws_mel_.as<uint8_t>();
    float *y = (float*)(b + o_y), *spec = (float*)(b + o_spec), *mag = (float*)(b + o_mag), *acc = (float*)(b + o_acc);
    mel_reflect_pad_kernel<<<(unsigned)((n_y + 255) / 256), 256, 0, st>>>(wav, L, kPad, (int64_t)n_y, y);
    cudaError_t e = launch_gemm_f32(y, T, kHop, kFft, reinterpret_cast<const float*>(dft->ptr), 2 * kLdh, spec, st);
    if (e != cudaSuccess) return fail(std::string("mel DFT gemm: ") + cudaGetErrorString(e));
    mel_magnitude_kernel<<<(unsigned)(((size_t)T * kLdh + 255) / 256), 256, 0, st>>>(spec, T, kLdh, kBins, mag);
    e = launch_gemm_f32(mag, T, kLdh, kLdh, reinterpret_cast<const float*>(fb->ptr), 128, acc, st);
    if (e != cudaSuccess) return fail(std::string("mel filterbank gemm: ") + cudaGetErrorString(e));
    mel_log_kernel<<<(unsigned)(((size_t)T * kMels + 255) / 256), 256, 0, st>>>(acc, T, 128, kMels, mel_out);
    launches_ += 5;
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("mel_spectrogram: ") + cudaGetErrorString(e));
    return T;
}

// Allocation-free twin of mel_spectrogram for callers that cannot call cudaMalloc (under stream capture, in serving loops):
// same kernels in the same order, the workspace is the file-scope array above.  Calls no cudaMalloc / cudaFree; all calls
// on a device share the array, so two calls that overlap on the GPU (different streams) race.
int Engine::mel_spectrogram_static(const float* wav, int64_t L, float* mel_out, cudaStream_t st) {
    DeviceGuard g(device_);
    constexpr int kFft = 1920, kHop = 480, kPad = (kFft - kHop) / 2, kBins = kFft / 2 + 1, kLdh = 1024, kMels = 80;
    static_assert(kMelStaticMaxFrames % 128 == 0, "scratch is sized for whole 128-row GEMM tiles");
    const DevTensor* dft = find("mel.dft");
    const DevTensor* fb = find("mel.fb");
    if (!dft || !fb) return fail("mel front end tensors (mel.dft, mel.fb) not loaded");
    if (L <= kPad) return fail("mel_spectrogram_static: input shorter than the reflect padding (720 samples)");
    const int64_t T64 = (L + 2 * kPad - kFft) / kHop + 1;
    if (T64 <= 0) return fail("mel_spectrogram_static: bad length");
    if (T64 > kMelStaticMaxFrames)
        return fail("mel_spectrogram_static: " + std::to_string(L) + " samples = " + std::to_string(T64) +
                    " frames exceed the static scratch (MOSSDEC_MEL_STATIC_MAX_FRAMES = " +
                    std::to_string(kMelStaticMaxFrames) + " frames, " + std::to_string(kMelStaticMaxFrames * kHop + kHop - 1) +
                    " samples); use mossdec_mel_spectrogram");
    const int T = (int)T64;
    const int Tp = (T + 127) / 128 * 128;
    uint8_t* b = nullptr;
    cudaError_t e = mel_static_scratch(&b);
    if (e != cudaSuccess) return fail(std::string("cudaGetSymbolAddress(mel static scratch): ") + cudaGetErrorString(e));
    float *y = (float*)b, *spec = (float*)(b + kMelStaticY), *mag = (float*)(b + kMelStaticY + kMelStaticSpec),
          *acc = (float*)(b + kMelStaticY + kMelStaticSpec + kMelStaticMag);
    const size_t n_y = (size_t)Tp * kHop + kFft + 64;
    // the array keeps the previous call's contents: y is rewritten in full (zeros past the signal) and the magnitude rows
    // of the last GEMM tile beyond T, which the filterbank GEMM loads, are cleared
    mel_reflect_pad_kernel<<<(unsigned)((n_y + 255) / 256), 256, 0, st>>>(wav, L, kPad, (int64_t)n_y, y);
    if (Tp > T) {
        e = cudaMemsetAsync(mag + (size_t)T * kLdh, 0, (size_t)(Tp - T) * kLdh * 4, st);
        if (e != cudaSuccess) return fail(std::string("mel_spectrogram_static memset: ") + cudaGetErrorString(e));
    }
    e = launch_gemm_f32(y, T, kHop, kFft, reinterpret_cast<const float*>(dft->ptr), 2 * kLdh, spec, st);
    if (e != cudaSuccess) return fail(std::string("mel DFT gemm: ") + cudaGetErrorString(e));
    mel_magnitude_kernel<<<(unsigned)(((size_t)T * kLdh + 255) / 256), 256, 0, st>>>(spec, T, kLdh, kBins, mag);
    e = launch_gemm_f32(mag, T, kLdh, kLdh, reinterpret_cast<const float*>(fb->ptr), 128, acc, st);
    if (e != cudaSuccess) return fail(std::string("mel filterbank gemm: ") + cudaGetErrorString(e));
    mel_log_kernel<<<(unsigned)(((size_t)T * kMels + 255) / 256), 256, 0, st>>>(acc, T, 128, kMels, mel_out);
    launches_ += 5;
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("mel_spectrogram_static: ") + cudaGetErrorString(e));
    return T;
}

}  // namespace mossdec
