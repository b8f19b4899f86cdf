// Implicit-GEMM core of the decode path (host-visible declarations).
//
//   out[row, n] = epilogue( sum_seg sum_c A_seg[row + shift_seg, col0_seg + c] * W[n, koff_seg + c] )
//
// Activations live channels-last ([rows, C], bf16) so a Conv1d / ConvTranspose1d tap is a
// row-shifted K-slab of the same matrix ("segment").  Rows of different utterances are packed
// back to back with zero guard rows between them, so shifted reads see the conv zero padding.
// The kernel is a persistent, warp-specialised tcgen05 kernel: TMA producer warp -> 4-stage
// smem ring -> single-thread tcgen05.mma issuer -> double-buffered fp32 accumulators in TMEM
// -> 4 epilogue warps (thread == output row, so LayerNorm over channels is thread-local).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <atomic>
#include <mutex>
#include <utility>

namespace mossdec {

enum Act : int32_t {
    ACT_NONE = 0,
    ACT_MISH = 1,
    ACT_GELU = 2,   // exact erf GELU (diffusers GELU, approximate="none")
    ACT_ELU = 3,
    ACT_LRELU = 4,  // slope in act_p
    ACT_SNAKE = 5,  // x + sin^2(alpha x) / (alpha + 1e-9), alpha per channel
    ACT_SILU = 6,
    ACT_ABS = 7,
};

constexpr int kGemmMaxSegs = 12;
constexpr int kGemmBM = 128;   // rows per tile == TMEM lanes
constexpr int kGemmBK = 64;    // bf16 elements per k-block == one 128-byte swizzle span
constexpr int kGemmStages = 4;

struct GemmSeg {
    int32_t map;        // which A tensor (0 or 1)
    int32_t row_shift;  // added to the tile's first row (conv tap offset, may be negative)
    int32_t col0;       // first channel of this slab inside A
    int32_t nkb;        // number of 64-wide k-blocks
};

struct GemmEpilogue {
    const float* bias = nullptr;       // [N]
    const float* ln1_g = nullptr;      // LayerNorm over the full row (needs N == BN), applied to acc+bias
    const float* ln1_b = nullptr;
    int32_t act1 = ACT_NONE;
    float act1_p = 0.f;
    const float* act1_alpha = nullptr; // [N] for snake
    const int32_t* row_seq = nullptr;  // [M] utterance id of each row, -1 for guard rows (output forced to 0)
    const float* tbias = nullptr;      // [num_seq, tbias_ld] added after act1 (time-embedding bias)
    int32_t tbias_ld = 0;
    const float* resid = nullptr;      // fp32 [M, ldr] added after tbias
    int32_t ldr = 0;
    int32_t f32_tiled = 0;             // resid and out_f32 use the lane-major tiled fp32 layout (tiled_f32_offset):
                                       // thread == row accesses are coalesced without staging; needs ld == 256
    int32_t accum_f32 = 0;             // v += out_f32 (before scaling)
    float out_scale = 1.f;
    float* out_f32 = nullptr;          // fp32 [M, ldf]
    int32_t ldf = 0;
    __nv_bfloat16* out_bf16 = nullptr; // bf16 [M, ldo]: LN2 / act2 applied to the fp32 value
    int32_t ldo = 0;
    __nv_bfloat16* out_bf16_lo = nullptr;  // optional bf16 [M, ldo] remainder v - bf16(v) (split-bf16 operands)
    const float* ln2_g = nullptr;
    const float* ln2_b = nullptr;
    int32_t ln2_plain = 0;             // second LayerNorm without affine (gamma / beta folded into the consumer's weights)
    int32_t act2 = ACT_NONE;
    float act2_p = 0.f;
    const float* act2_alpha = nullptr;
    float ln_eps = 1e-5f;
};

struct GemmOperandA {
    const __nv_bfloat16* ptr = nullptr;
    int64_t rows = 0;      // extent of the row dimension (reads outside are zero-filled by TMA)
    int32_t cols = 0;      // extent of the channel dimension
    int64_t ld = 0;        // row stride in elements (may be < cols for overlapping windows); ld*2 % 16 == 0
};

struct GemmProblem {
    GemmOperandA a[2];
    const __nv_bfloat16* w = nullptr;  // [N, K] K-major, K = 64 * total k-blocks
    int64_t w_ld = 0;                  // row stride of w in elements (0 = K); > K selects a column slab of a wider matrix
    int32_t M = 0;                     // output rows; buffers must be padded to a multiple of 128 rows
    int32_t N = 0;
    int32_t BN = 0;                    // n-tile (multiple of 16, <= 256)
    int32_t num_segs = 0;
    GemmSeg segs[kGemmMaxSegs];
    GemmEpilogue epi;
};

// Lane-major tiled layout of an fp32 [M, 256] workspace matrix (M % 128 == 0): the float4 holding columns 4j..4j+3 of the
// 128 rows of a tile are contiguous, so a warp whose lanes own consecutive rows reads / writes 512 contiguous bytes per
// instruction.  Used for the fp32 residual stream, which only thread == row epilogues touch.
__host__ __device__ inline size_t tiled_f32_offset(int64_t row, int col) {
    return ((size_t)((row >> 7) * 64 + (col >> 2)) * 128 + (size_t)(row & 127)) * 4 + (size_t)(col & 3);
}

// Fused transformer feed-forward (ffn_tc.cu): hres += W2 gelu(W1 x + b1) + b2 ; out = LN(hres) or bf16(hres).
struct FfnProblem {
    const __nv_bfloat16* x = nullptr;   // [M, 256] bf16 (normalised input); unused in fused mode
    // fused mode (att != nullptr): the attention out-projection + residual + LayerNorm run as a prologue of the same
    // kernel -- hres += att . Wo^T + bo, x = LayerNorm(hres) never leaves the SM (matcha transformer.py:274-297)
    const __nv_bfloat16* att = nullptr; // [M, 512] bf16 attention output
    const __nv_bfloat16* wo = nullptr;  // [256, 512] K-major
    const float* bo = nullptr;          // [256]
    int32_t M = 0;                      // multiple of 128
    const __nv_bfloat16* w1 = nullptr;  // [1024, 256] K-major
    const float* b1 = nullptr;
    const __nv_bfloat16* w2 = nullptr;  // [256, 1024] K-major
    const float* b2 = nullptr;
    float* hres = nullptr;              // [M, 256] fp32 residual stream (tiled_f32_offset layout), updated in place
    const int32_t* row_seq = nullptr;   // rows with id < 0 are forced to zero
    int32_t ln = 0;                     // 1: out = (hres - mean) * rstd (affine folded into the consumer); 0: out = bf16(hres)
    __nv_bfloat16* out_bf16 = nullptr;  // [M, ldo]
    int32_t ldo = 0;
    int32_t store_f32 = 1;              // 0: the fp32 stream is not needed afterwards
    float ln_eps = 1e-5f;
    long long* trace = nullptr;         // developer aid (tests/test_ffn.cu): clock64 stamps of CTA 0
};
cudaError_t ffn_tc_launch(const FfnProblem& p, int num_sms, cudaStream_t stream);

// Per-device one-time setup (opt-in shared-memory sizes of the kernels, constant tables): function attributes and
// __constant__ symbols live per device / context, so they are keyed by the CURRENT device ordinal under a mutex.
// The Engine constructor calls these on its device; the launchers re-check with an atomic bitmask (one relaxed load).
cudaError_t gemm_tc_init_device();
cudaError_t ffn_tc_init_device();
cudaError_t attention_ws_init_device();
cudaError_t hift_init_device();

// Enqueue on `stream`.  Returns cudaSuccess or the failing CUDA error; -1 style argument errors
// are reported as cudaErrorInvalidValue.
cudaError_t gemm_tc_launch(const GemmProblem& p, int num_sms, cudaStream_t stream);

// Runs `init` once per CUDA device (bit d of `done` = device d initialised).
template <typename F>
inline cudaError_t once_per_device(std::atomic<uint64_t>& done, std::mutex& mu, F init) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const uint64_t bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
    std::lock_guard<std::mutex> lk(mu);
    if (done.load(std::memory_order_relaxed) & bit) return cudaSuccess;
    e = init();
    if (e == cudaSuccess) done.fetch_or(bit, std::memory_order_release);
    return e;
}

// Launch with the programmatic-stream-serialization attribute (see ptx.cuh pdl_wait); off unless MOSSDEC_PDL=1 (measured: no gain on this path).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}


}  // namespace mossdec
