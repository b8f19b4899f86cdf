// Bandwidth-class kernels of the HiFT vocoder (cosyvoice/hifigan/generator.py):
//   NSF harmonic source : SineGen2._f02sine / forward  generator.py:261-339, SourceModuleHnNSF2.forward :375-389
//   STFT of the source  : HiFTGenerator._stft          generator.py:504-510   (n_fft 16, hop 4, hann, center/reflect)
//   Snake + residual    : ResBlock / decode fusion     generator.py:94-101, 529-535; transformer/activation.py:73-84
//   iSTFT head          : decode tail + _istft         generator.py:545-552, 512-518
// All sequences of a group are packed on one row axis (see HiftLayout in kernels.h); frame-rate rows scale by
// 8 / 40 / 120 through the three upsampling stages, so a row index maps between stages by a multiplication.
#include "kernels.h"
#include "gemm.h"
#include "ptx.cuh"

namespace mossdec {

namespace {

__constant__ float c_hann16[16];
__constant__ float c_cos16[16];
__constant__ float c_sin16[16];

__device__ __forceinline__ float snake_f(float x, float a) {
    const float s = sinf(x * a);
    return x + (1.0f / (a + 1e-9f)) * (s * s);
}

// row -> sequence id (or -1) for a stage whose rows are `scale` x the frame rows (+ `extra` trailing rows).
__global__ void hift_row_seq_kernel(HiftLayout L, int scale, int extra, int rows, int32_t* __restrict__ out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    int lo = 0, hi = L.nseq - 1, ans = -1;
    while (lo <= hi) {  // last sequence whose first row is <= r
        const int mid = (lo + hi) >> 1;
        if ((int64_t)L.seq_start[mid] * scale <= r) {
            ans = mid;
            lo = mid + 1;
        } else {
            hi = mid - 1;
        }
    }
    if (ans >= 0) {
        const int64_t end = (int64_t)(L.seq_start[ans] + L.seq_len[ans]) * scale + extra;
        if (r >= end) ans = -1;
    }
    out[r] = ans;
}

// ragged fp32 mel [80, T_b] blocks -> bf16 rows [rows_f, 128] (columns >= 80 and guard rows are zero); `out_lo`
// receives the bf16 remainder mel - bf16(mel) for the split-bf16 f0 predictor.
__global__ void hift_pack_mel_kernel(HiftLayout L, const int32_t* __restrict__ row_seq, const float* __restrict__ mel,
                                     __nv_bfloat16* __restrict__ out, float* __restrict__ out32) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L.rows_f * 128) return;
    const int row = idx >> 7, c = idx & 127;
    const int s = row_seq[row];
    float v = 0.f;
    if (s >= 0 && c < 80) {
        const int t = row - L.seq_start[s];
        v = mel[L.mel_off[s] + (L.time_major ? (int64_t)t * 80 + c : (int64_t)c * L.seq_len[s] + t)];
    }
    out[idx] = __float2bfloat16(v);              // conv_pre operand (bf16, 128 wide)
    if (c < 80) out32[(size_t)row * 80 + c] = v;   // f0 predictor input (fp32 rows)
}

// Frame-rate phase of every harmonic: rad = (f0 * h / sr) % 1 per frame (the x480 nearest upsampling followed by
// the 1/480 linear downsampling of generator.py:275-277 returns exactly the per-frame value; the random initial
// phase only touches sample 0, which that downsampling never reads), phase = cumsum(rad) * 2 * pi, then * 480
// (generator.py:280-281).  torch's CPU cumsum accumulates float inputs in double; so does this warp scan.
// grid = nseq, block = (32, n_harm).  p480: [rows_f, 16] fp32 (columns >= n_harm unused).
__global__ void hift_phase_kernel(HiftLayout L, const float* __restrict__ f0, int f0_ld, float sample_rate,
                                  float upsample, float* __restrict__ p480) {
    const int s = blockIdx.x, h = threadIdx.y, lane = threadIdx.x;
    const int T = L.seq_len[s], start = L.seq_start[s];
    const int per = (T + 31) / 32;
    const int t0 = lane * per, t1 = min(T, t0 + per);
    const float hm = (float)(h + 1);
    double acc = 0.0;
    for (int t = t0; t < t1; ++t) {
        const float fn = f0[(size_t)(start + t) * f0_ld] * hm;
        acc += (double)fmodf(fn / sample_rate, 1.0f);
    }
    double incl = acc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const double v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    double run = incl - acc;
    const float pi_f = 3.14159265358979323846f;
    for (int t = t0; t < t1; ++t) {
        const float fn = f0[(size_t)(start + t) * f0_ld] * hm;
        run += (double)fmodf(fn / sample_rate, 1.0f);
        const float ph = __fmul_rn(__fmul_rn((float)run, 2.0f), pi_f);
        p480[(size_t)(start + t) * 16 + h] = __fmul_rn(ph, upsample);
    }
}

// Philox4x32-10 (Salmon et al. 2011) -> 4 uniform u32 per counter
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& n0, float& n1) {
    const float u = ((float)a + 0.5f) * 2.3283064365386963e-10f;  // (0, 1)
    const float v = ((float)b + 0.5f) * 2.3283064365386963e-10f;
    const float r = sqrtf(-2.0f * __logf(u));
    float sn, cs;
    __sincosf(6.283185307179586f * v, &sn, &cs);
    n0 = r * cs;
    n1 = r * sn;
}

// One thread per audio sample: linear x480 upsampling of the frame phases (F.interpolate, align_corners=False),
// sin, voiced gate, additive noise, 9 -> 1 linear + tanh (generator.py:281-283, 325-339, 383-384), then the optional
// cache_source overwrite of HiFTGenerator.inference (generator.py:579-580).
__global__ void hift_source_kernel(HiftLayout L, const float* __restrict__ f0, int f0_ld,
                                   const float* __restrict__ p480, const float* __restrict__ noise,
                                   const float* __restrict__ lin_w, const float* __restrict__ lin_b,
                                   const float* __restrict__ cache, int cache_stride,
                                   const int32_t* __restrict__ cache_len, int n_harm, float sine_amp, float noise_std,
                                   float voiced_thr, int up, uint64_t seed, float* __restrict__ src_out) {
    const int s = blockIdx.y;
    const int T = L.seq_len[s];
    const int64_t nsamp = (int64_t)T * up;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nsamp) return;
    const int64_t g = L.wav_off[s] + i;  // position inside the ragged [sum 480 T_b] signal
    if (cache != nullptr && i < cache_len[s]) {
        src_out[g] = cache[(size_t)s * cache_stride + i];
        return;
    }
    const int start = L.seq_start[s];
    const int t = (int)(i / up);
    const float f = f0[(size_t)(start + t) * f0_ld];
    const float uv = f > voiced_thr ? 1.f : 0.f;
    // area_pixel_compute_source_index: src = scale * (dst + 0.5) - 0.5, clamped at 0, scale = (float)(1 / 480)
    const float scale = (float)(1.0 / (double)up);
    float srcf = __fmaf_rn(scale, (float)i + 0.5f, -0.5f);  // torch's AVX2 kernel contracts this into one FMA
    if (srcf < 0.f) srcf = 0.f;
    const int i0 = min((int)srcf, T - 1);
    const int i1 = min(i0 + 1, T - 1);
    const float l1 = srcf - (float)i0, l0 = 1.f - l1;
    const float* pa = p480 + (size_t)(start + i0) * 16;
    const float* pb = p480 + (size_t)(start + i1) * 16;
    const float namp = uv * noise_std + (1.f - uv) * sine_amp / 3.f;
    float nz[12];
    if (noise != nullptr) {
        const float* np_ = noise + (size_t)g * n_harm;
        for (int h = 0; h < n_harm; ++h) nz[h] = np_[h];
    } else {
        const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const uint4 r = philox4x32(make_uint4((uint32_t)g, (uint32_t)((uint64_t)g >> 32), (uint32_t)q, 0x4e5346u), key);
            box_muller(r.x, r.y, nz[4 * q], nz[4 * q + 1]);
            box_muller(r.z, r.w, nz[4 * q + 2], nz[4 * q + 3]);
        }
    }
    float acc = 0.f;
    for (int h = 0; h < n_harm; ++h) {
        const float ph = __fmaf_rn(l0, pa[h], __fmul_rn(l1, pb[h]));  // same contraction as torch's CPU lerp
        const float sw = sinf(ph) * sine_amp;
        const float v = sw * uv + namp * nz[h];
        acc += v * lin_w[h];
    }
    src_out[g] = tanhf(acc + lin_b[0]);
}

// STFT-16/4 with periodic hann window, center=True (reflect pad 8): one thread per frame row of the stage-2 layout.
// out: bf16 [rows2, 24] = [real k=0..8 | imag k=0..8 | 6 zeros]; guard rows are written as zeros.
__global__ void hift_stft_kernel(HiftLayout L, const int32_t* __restrict__ row_seq2, int rows2,
                                 const float* __restrict__ src, __nv_bfloat16* __restrict__ out) {
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows2) return;
    const int s = row_seq2[row];
    uint4* o = reinterpret_cast<uint4*>(out + (size_t)row * 24);
    if (s < 0) {
        o[0] = o[1] = o[2] = make_uint4(0, 0, 0, 0);
        return;
    }
    const int m = row - L.seq_start[s] * 120;
    const int64_t n = (int64_t)L.seq_len[s] * 480;
    const float* x = src + L.wav_off[s];
    float xv[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        int64_t idx = (int64_t)4 * m + j - 8;
        if (idx < 0) idx = -idx;
        if (idx >= n) idx = 2 * (n - 1) - idx;
        xv[j] = x[idx] * c_hann16[j];
    }
    float re[9], im[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int tw = (k * j) & 15;
            a += xv[j] * c_cos16[tw];
            b -= xv[j] * c_sin16[tw];
        }
        re[k] = a;
        im[k] = b;
    }
    uint32_t w[12];
    w[0] = pack_bf16x2(re[0], re[1]); w[1] = pack_bf16x2(re[2], re[3]); w[2] = pack_bf16x2(re[4], re[5]);
    w[3] = pack_bf16x2(re[6], re[7]); w[4] = pack_bf16x2(re[8], im[0]); w[5] = pack_bf16x2(im[1], im[2]);
    w[6] = pack_bf16x2(im[3], im[4]); w[7] = pack_bf16x2(im[5], im[6]); w[8] = pack_bf16x2(im[7], im[8]);
    w[9] = w[10] = w[11] = 0u;
    o[0] = make_uint4(w[0], w[1], w[2], w[3]);
    o[1] = make_uint4(w[4], w[5], w[6], w[7]);
    o[2] = make_uint4(w[8], w[9], w[10], w[11]);
}

// x = u + si (the upsampled trunk plus the source path, generator.py:535), with the ReflectionPad1d((1, 0)) of the
// last stage folded in (row 0 of a sequence takes the trunk value of row 2, generator.py:529-530), and the Snake
// activations that open the three parallel ResBlocks (generator.py:95) emitted as bf16 GEMM operands.
// One thread per 4 channels.
__global__ void hift_fuse_snake_kernel(const int32_t* __restrict__ row_seq, const int32_t* __restrict__ seq_start,
                                       int scale, int reflect, int64_t rows, int C, const float* __restrict__ u,
                                       const float* __restrict__ si, const float* __restrict__ a0,
                                       const float* __restrict__ a1, const float* __restrict__ a2,
                                       float* __restrict__ x, __nv_bfloat16* __restrict__ o0,
                                       __nv_bfloat16* __restrict__ o1, __nv_bfloat16* __restrict__ o2) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int C4 = C >> 2;
    if (idx >= rows * C4) return;
    const int64_t row = idx / C4;
    const int c = (int)(idx % C4) * 4;
    const int s = row_seq[row];
    const size_t off = (size_t)row * C + c;
    if (s < 0) {
        *reinterpret_cast<float4*>(x + off) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<uint2*>(o0 + off) = make_uint2(0u, 0u);
        *reinterpret_cast<uint2*>(o1 + off) = make_uint2(0u, 0u);
        *reinterpret_cast<uint2*>(o2 + off) = make_uint2(0u, 0u);
        return;
    }
    size_t uoff = off;
    if (reflect && row == (int64_t)seq_start[s] * scale) uoff = off + 2 * (size_t)C;
    const float4 uv = *reinterpret_cast<const float4*>(u + uoff);
    const float4 sv = *reinterpret_cast<const float4*>(si + off);
    const float4 v = make_float4(uv.x + sv.x, uv.y + sv.y, uv.z + sv.z, uv.w + sv.w);
    *reinterpret_cast<float4*>(x + off) = v;
    const float4 al0 = *reinterpret_cast<const float4*>(a0 + c);
    const float4 al1 = *reinterpret_cast<const float4*>(a1 + c);
    const float4 al2 = *reinterpret_cast<const float4*>(a2 + c);
    *reinterpret_cast<uint2*>(o0 + off) = make_uint2(pack_bf16x2(snake_f(v.x, al0.x), snake_f(v.y, al0.y)),
                                                     pack_bf16x2(snake_f(v.z, al0.z), snake_f(v.w, al0.w)));
    *reinterpret_cast<uint2*>(o1 + off) = make_uint2(pack_bf16x2(snake_f(v.x, al1.x), snake_f(v.y, al1.y)),
                                                     pack_bf16x2(snake_f(v.z, al1.z), snake_f(v.w, al1.w)));
    *reinterpret_cast<uint2*>(o2 + off) = make_uint2(pack_bf16x2(snake_f(v.x, al2.x), snake_f(v.y, al2.y)),
                                                     pack_bf16x2(snake_f(v.z, al2.z), snake_f(v.w, al2.w)));
}

// conv_post output -> waveform: magnitude = exp(x[:9]) clipped at 100, phase = sin(x[9:]), inverse 16-point real DFT
// per frame, hann window, overlap-add with hop 4, division by the window envelope, trim 8, clamp (generator.py:547-551,
// 512-518).  A block produces kIstftTile consecutive samples of one sequence from the (tile/4 + 3) frames that touch
// them; the windowed frames are staged in shared memory.
constexpr int kIstftTile = 1024;
constexpr int kIstftFrames = kIstftTile / 4 + 3;

__global__ void __launch_bounds__(256) hift_istft_kernel(HiftLayout L, const float* __restrict__ post, int post_ld,
                                                         float limit, float* __restrict__ wav) {
    __shared__ float fr[kIstftFrames][17];
    const int s = blockIdx.y;
    const int T = L.seq_len[s];
    const int64_t nsamp = (int64_t)T * 480;
    const int64_t t0 = (int64_t)blockIdx.x * kIstftTile;
    if (t0 >= nsamp) return;
    const int nframes = T * 120 + 1;
    const int m0 = (int)(t0 / 4) - 1;
    const int64_t row0 = (int64_t)L.seq_start[s] * 120;
    for (int f = threadIdx.x; f < kIstftFrames; f += blockDim.x) {
        const int m = m0 + f;
        if (m < 0 || m >= nframes) continue;
        const float* p = post + (size_t)(row0 + m) * post_ld;
        float re[9], im[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const float mag = fminf(expf(p[k]), 100.f);
            const float ph = sinf(p[9 + k]);
            float sn, cs;
            sincosf(ph, &sn, &cs);
            re[k] = mag * cs;
            im[k] = mag * sn;
        }
#pragma unroll
        for (int n = 0; n < 16; ++n) {
            float acc = re[0] + ((n & 1) ? -re[8] : re[8]);
#pragma unroll
            for (int k = 1; k < 8; ++k) {
                const int tw = (k * n) & 15;
                acc += 2.f * (re[k] * c_cos16[tw] - im[k] * c_sin16[tw]);
            }
            fr[f][n] = acc * (1.f / 16.f) * c_hann16[n];
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < kIstftTile; j += blockDim.x) {
        const int64_t t = t0 + j;
        if (t >= nsamp) break;
        const int mlo = (int)(t / 4) - 1;
        float acc = 0.f, env = 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int m = mlo + q;
            const int n = (int)(t + 8 - 4 * (int64_t)m);
            if (m >= 0 && m < nframes && n >= 0 && n < 16) {
                acc += fr[m - m0][n];
                env += c_hann16[n] * c_hann16[n];
            }
        }
        float y = acc / env;
        y = fminf(fmaxf(y, -limit), limit);
        wav[L.wav_off[s] + t] = y;
    }
}

// speech[:n] = speech[:n] * w[:n] + prev[:n] * w[n:]   (flow_inference.py:35-43)
__global__ void fade_kernel(float* __restrict__ speech, const float* __restrict__ prev, const float* __restrict__ w,
                            int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) speech[i] = speech[i] * w[i] + prev[i] * w[n + i];
}

}  // namespace

cudaError_t hift_init_device() {
    static std::atomic<uint64_t> done{0};
    static std::mutex mu;
    return once_per_device(done, mu, [] {
        float hann[16], cs[16], sn[16];
        for (int n = 0; n < 16; ++n) {
            const double a = 2.0 * 3.14159265358979323846 * n / 16.0;
            hann[n] = (float)(0.5 - 0.5 * cos(a));  // scipy.signal.get_window("hann", 16, fftbins=True) as float32
            cs[n] = (float)cos(a);
            sn[n] = (float)sin(a);
        }
        cudaError_t e = cudaMemcpyToSymbol(c_hann16, hann, sizeof hann);
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_cos16, cs, sizeof cs);
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_sin16, sn, sizeof sn);
        return e;
    });
}

void launch_hift_row_seq(const HiftLayout& L, int scale, int extra, int rows, int32_t* out, cudaStream_t st) {
    hift_row_seq_kernel<<<(rows + 255) / 256, 256, 0, st>>>(L, scale, extra, rows, out);
}
void launch_hift_pack_mel(const HiftLayout& L, const int32_t* row_seq, const float* mel, __nv_bfloat16* out,
                          float* out32, cudaStream_t st) {
    const int n = L.rows_f * 128;
    hift_pack_mel_kernel<<<(n + 255) / 256, 256, 0, st>>>(L, row_seq, mel, out, out32);
}
void launch_hift_phase(const HiftLayout& L, const float* f0, int f0_ld, float sample_rate, float upsample, int n_harm,
                       float* p480, cudaStream_t st) {
    hift_phase_kernel<<<L.nseq, dim3(32, n_harm), 0, st>>>(L, f0, f0_ld, sample_rate, upsample, p480);
}
void launch_hift_source(const HiftLayout& L, int max_len, const float* f0, int f0_ld, const float* p480,
                        const float* noise, const float* lin_w, const float* lin_b, const float* cache,
                        int cache_stride, const int32_t* cache_len, int n_harm, float sine_amp, float noise_std,
                        float voiced_thr, int up, uint64_t seed, float* src_out, cudaStream_t st) {
    dim3 grid((unsigned)(((int64_t)max_len * up + 255) / 256), L.nseq);
    hift_source_kernel<<<grid, 256, 0, st>>>(L, f0, f0_ld, p480, noise, lin_w, lin_b, cache, cache_stride, cache_len,
                                             n_harm, sine_amp, noise_std, voiced_thr, up, seed, src_out);
}
void launch_hift_stft(const HiftLayout& L, const int32_t* row_seq2, int rows2, const float* src, __nv_bfloat16* out,
                      cudaStream_t st) {
    hift_stft_kernel<<<(rows2 + 127) / 128, 128, 0, st>>>(L, row_seq2, rows2, src, out);
}
void launch_hift_fuse_snake(const HiftLayout& L, const int32_t* row_seq, int scale, int reflect, int64_t rows, int C,
                            const float* u, const float* si, const float* a0, const float* a1, const float* a2,
                            float* x, __nv_bfloat16* o0, __nv_bfloat16* o1, __nv_bfloat16* o2, cudaStream_t st) {
    const int64_t n = rows * (C / 4);
    hift_fuse_snake_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(row_seq, L.seq_start, scale, reflect, rows, C, u,
                                                                         si, a0, a1, a2, x, o0, o1, o2);
}
void launch_hift_istft(const HiftLayout& L, int max_len, const float* post, int post_ld, float limit, float* wav,
                       cudaStream_t st) {
    dim3 grid((unsigned)(((int64_t)max_len * 480 + kIstftTile - 1) / kIstftTile), L.nseq);
    hift_istft_kernel<<<grid, 256, 0, st>>>(L, post, post_ld, limit, wav);
}
void launch_fade(float* speech, const float* prev, const float* w, int n, cudaStream_t st) {
    fade_kernel<<<(n + 255) / 256, 256, 0, st>>>(speech, prev, w, n);
}

}  // namespace mossdec
