// Self-attention of BasicTransformerBlock over packed variable-length rows
// (matcha/models/components/transformer.py:262-274 + diffusers Attention; masks: cosyvoice/flow/decoder.py:439-443,
// cosyvoice/utils/mask.py:127-158).  The additive -1e10 bias of the reference gives masked keys an exact weight of 0
// in fp32, so masking by predicate (never materialising [T, T]) is equivalent.
//
// v1: flash-style online softmax, bf16 mma.sync m16n8k16 tiles, cp.async double-buffered K/V.  64 queries x 64 keys
// per step, 4 warps (16 query rows each), head_dim 64.
#include "kernels.h"
#include "ptx.cuh"

namespace mossdec {

namespace {

constexpr int kD = 64;       // head dim
constexpr int kBQ = 64;      // queries per CTA
constexpr int kBK = 64;      // keys per step
constexpr int kTileBytes = 64 * 128;

__device__ __forceinline__ void cp_async16(uint32_t smem, const void* g, bool pred) {
    int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// byte offset of 16-byte chunk `c` of row `r` inside a [64][64] bf16 tile (XOR swizzle, conflict-free ldmatrix)
__device__ __forceinline__ uint32_t sw(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// load a [64 rows][64 dims] bf16 tile: rows row0.. of the packed matrix (row stride ld), columns col0..col0+63
__device__ __forceinline__ void load_tile(uint32_t smem, const __nv_bfloat16* base, int ld, int row0, int nvalid,
                                          int col0, int tid) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = tid + i * 128;  // 512 chunks
        const int r = idx >> 3, c = idx & 7;
        const bool ok = r < nvalid;
        const __nv_bfloat16* g = base + (size_t)(row0 + (ok ? r : 0)) * ld + col0 + c * 8;
        cp_async16(smem + sw(r, c), g, ok);
    }
}

__global__ void __launch_bounds__(128) attention_kernel(const __nv_bfloat16* __restrict__ qkv,
                                                        __nv_bfloat16* __restrict__ out,
                                                        const int32_t* __restrict__ seq_start,
                                                        const int32_t* __restrict__ seq_len, int heads, int chunk) {
    __shared__ __align__(128) uint8_t smem[kTileBytes * 5];  // Q, K0, K1, V0, V1
    const int s = blockIdx.z, h = blockIdx.y;
    const int len = seq_len[s];
    const int q0 = blockIdx.x * kBQ;
    if (q0 >= len) return;
    const int start = seq_start[s];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ld = 3 * heads * kD;
    const int ldo = heads * kD;
    const uint32_t sQ = smem_u32(smem), sK = sQ + kTileBytes, sV = sQ + 3 * kTileBytes;

    int kv_end = len;
    if (chunk > 0) {
        const int qlast = min(q0 + kBQ, len) - 1;
        kv_end = min(len, (qlast / chunk + 1) * chunk);
    }
    const int nkt = (kv_end + kBK - 1) / kBK;

    load_tile(sQ, qkv, ld, start + q0, min(kBQ, len - q0), h * kD, tid);
    load_tile(sK, qkv, ld, start, min(kBK, kv_end), heads * kD + h * kD, tid);
    load_tile(sV, qkv, ld, start, min(kBK, kv_end), 2 * heads * kD + h * kD, tid);
    cp_async_commit();

    uint32_t qf[4][4];
    float o[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const int g = lane >> 2, t4 = lane & 3;
    const int qi0 = q0 + warp * 16 + g, qi1 = qi0 + 8;  // query indices (within the sequence) of this thread's rows
    int lim0 = len, lim1 = len;
    if (chunk > 0) {
        lim0 = min(len, (qi0 / chunk + 1) * chunk);
        lim1 = min(len, (qi1 / chunk + 1) * chunk);
    }
    constexpr float kLog2e = 1.4426950408889634f;

    for (int kt = 0; kt < nkt; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < nkt) {
            const int k0n = (kt + 1) * kBK;
            load_tile(sK + (buf ^ 1) * kTileBytes, qkv, ld, start + k0n, min(kBK, kv_end - k0n), heads * kD + h * kD, tid);
            load_tile(sV + (buf ^ 1) * kTileBytes, qkv, ld, start + k0n, min(kBK, kv_end - k0n), 2 * heads * kD + h * kD, tid);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (kt == 0) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
                const int c = ks * 2 + (lane >> 4);
                ldsm_x4(sQ + sw(r, c), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
            }
        }
        const uint32_t kb = sK + buf * kTileBytes, vb = sV + buf * kTileBytes;
        // S = Q K^T  (16 x 64 per warp)
        float sc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {  // pairs of 8-key tiles
                const int mi = lane >> 3;
                const int key = np * 16 + (lane & 7) + (mi >> 1) * 8;
                const int c = ks * 2 + (mi & 1);
                uint32_t b0, b1, b2, b3;
                ldsm_x4(kb + sw(key, c), b0, b1, b2, b3);
                mma_bf16(sc[np * 2], qf[ks], b0, b1);
                mma_bf16(sc[np * 2 + 1], qf[ks], b2, b3);
            }
        }
        // mask + online softmax
        const int kbase = kt * kBK;
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int kj = kbase + nt * 8 + t4 * 2;
            if (kj >= lim0) sc[nt][0] = -INFINITY;
            if (kj + 1 >= lim0) sc[nt][1] = -INFINITY;
            if (kj >= lim1) sc[nt][2] = -INFINITY;
            if (kj + 1 >= lim1) sc[nt][3] = -INFINITY;
            mx0 = fmaxf(mx0, fmaxf(sc[nt][0], sc[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(sc[nt][2], sc[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float ms0 = mx0 == -INFINITY ? 0.f : mx0, ms1 = mx1 == -INFINITY ? 0.f : mx1;
        const float corr0 = exp2f((m0 - ms0) * kLog2e), corr1 = exp2f((m1 - ms1) * kLog2e);
        m0 = mx0;
        m1 = mx1;
        float rs0 = 0.f, rs1 = 0.f;
        uint32_t pf[4][4];  // P as A fragments: k-step j covers keys 16j..16j+15 = n-tiles 2j, 2j+1
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float p0 = exp2f((sc[nt][0] - ms0) * kLog2e), p1 = exp2f((sc[nt][1] - ms0) * kLog2e);
            const float p2 = exp2f((sc[nt][2] - ms1) * kLog2e), p3 = exp2f((sc[nt][3] - ms1) * kLog2e);
            rs0 += p0 + p1;
            rs1 += p2 + p3;
            pf[nt >> 1][(nt & 1) * 2] = pack_bf16x2(p0, p1);
            pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16x2(p2, p3);
        }
        l0 = l0 * corr0 + rs0;
        l1 = l1 * corr1 + rs1;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            o[i][0] *= corr0; o[i][1] *= corr0; o[i][2] *= corr1; o[i][3] *= corr1;
        }
        // O += P V
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {      // 16 keys per step
#pragma unroll
            for (int dp = 0; dp < 4; ++dp) {  // pairs of 8-dim tiles
                const int mi = lane >> 3;
                const int key = ks * 16 + (lane & 7) + (mi & 1) * 8;
                const int c = dp * 2 + (mi >> 1);
                uint32_t b0, b1, b2, b3;
                ldsm_x4_t(vb + sw(key, c), b0, b1, b2, b3);
                mma_bf16(o[dp * 2], pf[ks], b0, b1);
                mma_bf16(o[dp * 2 + 1], pf[ks], b2, b3);
            }
        }
        __syncthreads();
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = l0 > 0.f ? 1.f / l0 : 0.f, inv1 = l1 > 0.f ? 1.f / l1 : 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const int d = h * kD + nt * 8 + t4 * 2;
        if (qi0 < len)
            *reinterpret_cast<uint32_t*>(out + (size_t)(start + qi0) * ldo + d) = pack_bf16x2(o[nt][0] * inv0, o[nt][1] * inv0);
        if (qi1 < len)
            *reinterpret_cast<uint32_t*>(out + (size_t)(start + qi1) * ldo + d) = pack_bf16x2(o[nt][2] * inv1, o[nt][3] * inv1);
    }
}

}  // namespace

void launch_attention(const __nv_bfloat16* qkv, __nv_bfloat16* out, const int32_t* seq_start, const int32_t* seq_len,
                      int nseq, int max_len, int heads, int chunk, cudaStream_t st) {
    dim3 grid((max_len + kBQ - 1) / kBQ, heads, nseq);
    attention_kernel<<<grid, 128, 0, st>>>(qkv, out, seq_start, seq_len, heads, chunk);
}

}  // namespace mossdec
