// Persistent warp-specialised tcgen05 implicit-GEMM kernel (see gemm.h for the contract).
//
// Replaces, for the decode path, every dense contraction the reference runs through
// cuDNN/cuBLAS: CausalConv1d (cosyvoice/flow/decoder.py:36-62), the Linear layers of
// BasicTransformerBlock (matcha/models/components/transformer.py:262-314), and the HiFT
// Conv1d / ConvTranspose1d layers (cosyvoice/hifigan/generator.py:46-101,520-552).
#include "gemm.h"
#include "ptx.cuh"

#include <mutex>
#include <stdio.h>

namespace mossdec {

namespace {

constexpr int kThreads = 192;  // warp0 = TMA, warp1 = MMA (+TMEM alloc), warps 2..5 = epilogue
constexpr int kABytes = kGemmBM * kGemmBK * 2;  // 16 KB
constexpr int kBBytesMax = 256 * kGemmBK * 2;   // 32 KB
constexpr int kSmemBytes = kGemmStages * (kABytes + kBBytesMax) + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr uint32_t kTmemCols = 512;

struct KParams {
    int32_t M, N, BN;
    int32_t tiles_n, num_tiles, total_kb, num_segs;
    GemmSeg segs[kGemmMaxSegs];
    GemmEpilogue epi;
};

__device__ __forceinline__ float apply_act(float x, int act, float p, const float* alpha, int n) {
    switch (act) {
        case ACT_MISH: return mish_f(x);
        case ACT_GELU: return gelu_erf_f(x);
        case ACT_ELU: return elu_f(x);
        case ACT_LRELU: return x > 0.f ? x : x * p;
        case ACT_SNAKE: {
            float a = __ldg(alpha + n);
            float s = sinf(x * a);
            return x + s * s / (a + 1e-9f);
        }
        case ACT_SILU: return silu_f(x);
        case ACT_ABS: return fabsf(x);
        default: return x;
    }
}

__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ KParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + kGemmStages * kABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kGemmStages * (kABytes + kBBytesMax));
    uint64_t* full = bars;                     // [stages]
    uint64_t* empty = bars + kGemmStages;      // [stages]
    uint64_t* tfull = bars + 2 * kGemmStages;  // [2]
    uint64_t* tempty = tfull + 2;              // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmB);
        for (int i = 0; i < kGemmStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], 4);
        }
        mbar_fence_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const uint32_t b_bytes = (uint32_t)p.BN * kGemmBK * 2;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const int m0 = (tile / p.tiles_n) * kGemmBM;
                const int n0 = (tile % p.tiles_n) * p.BN;
                int kb = 0;
                for (int s = 0; s < p.num_segs; ++s) {
                    const GemmSeg sg = p.segs[s];
                    const CUtensorMap* tmA = sg.map ? &tmA1 : &tmA0;
                    for (int j = 0; j < sg.nkb; ++j, ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        mbar_expect_tx(&full[stage], kABytes + b_bytes);
                        tma_load_2d(sA + stage * kABytes, tmA, sg.col0 + j * kGemmBK, m0 + sg.row_shift, &full[stage]);
                        tma_load_2d(sB + stage * kBBytesMax, &tmB, kb * kGemmBK, n0, &full[stage]);
                        if (++stage == kGemmStages) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(kGemmBM, (uint32_t)p.BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t accphase = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                mbar_wait(&tempty[acc], accphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256u;
                for (int kb = 0; kb < p.total_kb; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint64_t da = umma_desc_sw128(smem_u32(sA + stage * kABytes));
                    const uint64_t db = umma_desc_sw128(smem_u32(sB + stage * kBBytesMax));
#pragma unroll
                    for (int k = 0; k < kGemmBK / 16; ++k)
                        umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
                    umma_commit(&empty[stage]);
                    if (++stage == kGemmStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&tfull[acc]);
                acc ^= 1;
                if (acc == 0) accphase ^= 1;
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue (thread == row)
        const GemmEpilogue& e = p.epi;
        const int quarter = warp & 3;  // TMEM lane quarter this warp may access
        const int nchunks = p.BN / 16;
        int acc = 0;
        uint32_t accphase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            const int m0 = (tile / p.tiles_n) * kGemmBM;
            const int n0 = (tile % p.tiles_n) * p.BN;
            const int row = m0 + quarter * 32 + lane;
            const bool inb = row < p.M;
            int seq = 0;
            if (e.row_seq != nullptr && inb) seq = __ldg(e.row_seq + row);
            const bool valid = inb && seq >= 0;
            const int ncols = min(p.BN, p.N - n0);  // columns of this tile that exist in the output

            mbar_wait(&tfull[acc], accphase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)acc * 256u;

            float mean1 = 0.f, rstd1 = 1.f;
            if (e.ln1_g != nullptr) {
                float s = 0.f, ss = 0.f;
                for (int c = 0; c < nchunks; ++c) {
                    float v[16];
                    tmem_ld16(taddr + c * 16, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        float x = v[i] + (e.bias ? __ldg(e.bias + n0 + c * 16 + i) : 0.f);
                        s += x;
                        ss += x * x;
                    }
                }
                mean1 = s / (float)p.BN;
                float var = fmaxf(ss / (float)p.BN - mean1 * mean1, 0.f);
                rstd1 = rsqrtf(var + e.ln_eps);
            }

            // value of columns [c*16, c*16+16) of this thread's row, before the bf16-side LN2/act2
            auto chunk_value = [&](int c, float* v) {
                tmem_ld16(taddr + c * 16, v);
                tmem_ld_wait();
                const int nb = n0 + c * 16;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int n = nb + i;
                    float x = v[i];
                    if (n < p.N) {
                        if (e.bias) x += __ldg(e.bias + n);
                        if (e.ln1_g) x = (x - mean1) * rstd1 * __ldg(e.ln1_g + n) + __ldg(e.ln1_b + n);
                        x = apply_act(x, e.act1, e.act1_p, e.act1_alpha, n);
                        if (e.tbias) x += __ldg(e.tbias + (size_t)(seq < 0 ? 0 : seq) * e.tbias_ld + n);
                    }
                    v[i] = x;
                }
                if (valid) {
                    if (e.resid) {
                        const float* r = e.resid + (size_t)row * e.ldr + nb;
#pragma unroll
                        for (int i = 0; i < 16; i += 4) {
                            if (nb + i < p.N) {
                                float4 t = *reinterpret_cast<const float4*>(r + i);
                                v[i] += t.x; v[i + 1] += t.y; v[i + 2] += t.z; v[i + 3] += t.w;
                            }
                        }
                    }
                    if (e.accum_f32) {
                        const float* r = e.out_f32 + (size_t)row * e.ldf + nb;
#pragma unroll
                        for (int i = 0; i < 16; i += 4) {
                            if (nb + i < p.N) {
                                float4 t = *reinterpret_cast<const float4*>(r + i);
                                v[i] += t.x; v[i + 1] += t.y; v[i + 2] += t.z; v[i + 3] += t.w;
                            }
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] *= e.out_scale;
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = 0.f;
                }
            };
            auto store_f32 = [&](int c, const float* v) {
                if (e.out_f32 != nullptr && inb) {
                    float* o = e.out_f32 + (size_t)row * e.ldf + n0 + c * 16;
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        if (c * 16 + i < ncols) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
                }
            };
            auto store_bf16 = [&](int c, const float* v) {
                if (e.out_bf16 != nullptr && inb) {
                    __nv_bfloat16* o = e.out_bf16 + (size_t)row * e.ldo + n0 + c * 16;
#pragma unroll
                    for (int i = 0; i < 16; i += 8)
                        if (c * 16 + i < ncols) {
                            uint4 t;
                            t.x = pack_bf16x2(v[i], v[i + 1]);
                            t.y = pack_bf16x2(v[i + 2], v[i + 3]);
                            t.z = pack_bf16x2(v[i + 4], v[i + 5]);
                            t.w = pack_bf16x2(v[i + 6], v[i + 7]);
                            *reinterpret_cast<uint4*>(o + i) = t;
                            if (e.out_bf16_lo != nullptr) {
                                float r[8];
#pragma unroll
                                for (int j = 0; j < 8; ++j) r[j] = v[i + j] - __bfloat162float(__float2bfloat16(v[i + j]));
                                uint4 u;
                                u.x = pack_bf16x2(r[0], r[1]);
                                u.y = pack_bf16x2(r[2], r[3]);
                                u.z = pack_bf16x2(r[4], r[5]);
                                u.w = pack_bf16x2(r[6], r[7]);
                                *reinterpret_cast<uint4*>(e.out_bf16_lo + (size_t)row * e.ldo + n0 + c * 16 + i) = u;
                            }
                        }
                }
            };

            if (e.ln2_g != nullptr) {
                // pass 1: fp32 value, statistics, fp32 store; pass 2: recompute + LN2 -> bf16
                float s = 0.f, ss = 0.f;
                for (int c = 0; c < nchunks; ++c) {
                    float v[16];
                    chunk_value(c, v);
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        s += v[i];
                        ss += v[i] * v[i];
                    }
                    store_f32(c, v);
                }
                const float mean2 = s / (float)p.BN;
                const float rstd2 = rsqrtf(fmaxf(ss / (float)p.BN - mean2 * mean2, 0.f) + e.ln_eps);
                for (int c = 0; c < nchunks; ++c) {
                    float v[16];
                    if (e.out_f32 != nullptr) {
                        // re-read what this thread stored in pass 1 (also makes resid == out_f32 in-place safe)
                        if (inb) {
                            const float* o = e.out_f32 + (size_t)row * e.ldf + n0 + c * 16;
#pragma unroll
                            for (int i = 0; i < 16; i += 4) {
                                float4 t = *reinterpret_cast<const float4*>(o + i);
                                v[i] = t.x; v[i + 1] = t.y; v[i + 2] = t.z; v[i + 3] = t.w;
                            }
                        }
                    } else {
                        chunk_value(c, v);
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int n = n0 + c * 16 + i;
                        v[i] = valid ? (v[i] - mean2) * rstd2 * __ldg(e.ln2_g + n) + __ldg(e.ln2_b + n) : 0.f;
                    }
                    store_bf16(c, v);
                }
            } else {
                for (int c = 0; c < nchunks; ++c) {
                    float v[16];
                    chunk_value(c, v);
                    store_f32(c, v);
                    if (e.act2 != ACT_NONE) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const int n = n0 + c * 16 + i;
                            if (n < p.N) v[i] = valid ? apply_act(v[i], e.act2, e.act2_p, e.act2_alpha, n) : 0.f;
                        }
                    }
                    store_bf16(c, v);
                }
            }

            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
            acc ^= 1;
            if (acc == 0) accphase ^= 1;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// 2-D bf16 tensor map: dims {cols, rows}, row stride ld elements, box {64, box_rows}, 128B swizzle.
bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)kGemmBK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr, "[mossdec] cuTensorMapEncodeTiled failed: %d (ptr=%p cols=%llu rows=%llu ld=%llu box=%u)\n",
                (int)r, ptr, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld, box_rows);
        return false;
    }
    return true;
}

}  // namespace

cudaError_t gemm_tc_launch(const GemmProblem& g, int num_sms, cudaStream_t stream) {
    if (g.M <= 0 || g.N <= 0 || g.BN < 16 || g.BN > 256 || (g.BN % 16) != 0 || g.num_segs < 1 ||
        g.num_segs > kGemmMaxSegs)
        return cudaErrorInvalidValue;
    if (g.epi.ln1_g && g.BN != g.N) return cudaErrorInvalidValue;
    if (g.epi.ln2_g && (g.BN != g.N || g.epi.accum_f32)) return cudaErrorInvalidValue;

    static std::once_flag attr_once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(attr_once, [] {
        attr_err = cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    });
    if (attr_err != cudaSuccess) return attr_err;

    KParams kp;
    kp.M = g.M;
    kp.N = g.N;
    kp.BN = g.BN;
    kp.tiles_n = (g.N + g.BN - 1) / g.BN;
    const int tiles_m = (g.M + kGemmBM - 1) / kGemmBM;
    kp.num_tiles = tiles_m * kp.tiles_n;
    kp.num_segs = g.num_segs;
    kp.total_kb = 0;
    for (int s = 0; s < g.num_segs; ++s) {
        kp.segs[s] = g.segs[s];
        kp.total_kb += g.segs[s].nkb;
    }
    kp.epi = g.epi;

    CUtensorMap tA0, tA1, tB;
    if (!encode_2d(&tA0, g.a[0].ptr, (uint64_t)g.a[0].cols, (uint64_t)g.a[0].rows, (uint64_t)g.a[0].ld, kGemmBM))
        return cudaErrorInvalidValue;
    if (g.a[1].ptr) {
        if (!encode_2d(&tA1, g.a[1].ptr, (uint64_t)g.a[1].cols, (uint64_t)g.a[1].rows, (uint64_t)g.a[1].ld, kGemmBM))
            return cudaErrorInvalidValue;
    } else {
        tA1 = tA0;
    }
    const uint64_t K = (uint64_t)kp.total_kb * kGemmBK;
    if (!encode_2d(&tB, g.w, K, (uint64_t)g.N, K, (uint32_t)g.BN)) return cudaErrorInvalidValue;

    const int grid = kp.num_tiles < num_sms ? kp.num_tiles : num_sms;
    gemm_tc_kernel<<<grid, kThreads, kSmemBytes, stream>>>(tA0, tA1, tB, kp);
    return cudaGetLastError();
}

}  // namespace mossdec
