// Developer micro-benchmark: MUFU throughput per SM of ex2.approx.ftz.f32 vs ex2.approx.f16x2 (two MUFU.EX2.F16 in SASS)
// vs ex2.approx.ftz.bf16x2.  Prints results per clock per SM.  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned cvt_bf(unsigned x) { unsigned r; float f = __uint_as_float(x); asm volatile("cvt.rn.bf16x2.f32 %0, %1, %1;" : "=r"(r) : "f"(f)); return r; }
__device__ __forceinline__ unsigned cvt_h(unsigned x) { unsigned r; float f = __uint_as_float(x); asm volatile("cvt.rn.f16x2.f32 %0, %1, %1;" : "=r"(r) : "f"(f)); return r; }
__device__ __forceinline__ unsigned ex2f(unsigned x) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(x)); return x; }
__device__ __forceinline__ unsigned prmt(unsigned x) { asm volatile("prmt.b32 %0, %0, %0, 0x7632;" : "+r"(x)); return x; }
template <int MODE>
__global__ void k(unsigned* out, int iters) {
    unsigned a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = 0x3c003800u + threadIdx.x + i;   // small positive halves / a float near 0.0078
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(a[i]));
            if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(a[i]));
            if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(a[i]));
            if (MODE == 3) a[i] = cvt_bf(a[i]);                       // F2FP.BF16.F32.PACK_AB
            if (MODE == 4) a[i] = (i & 1) ? cvt_bf(a[i]) : ex2f(a[i]);  // one MUFU + one F2FP
            if (MODE == 5) a[i] = prmt(a[i]);                           // byte permute (truncating pack)
            if (MODE == 6) a[i] = (i & 1) ? prmt(a[i]) : ex2f(a[i]);    // one MUFU + one PRMT
            if (MODE == 7) a[i] = cvt_h(a[i]);
        }
    }
    long long t1 = clock64();
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[gridDim.x * blockDim.x] = (unsigned)(t1 - t0);
}
template <int MODE>
void run(const char* name, int results_per_instr) {
    unsigned* d;
    const int blocks = 148, threads = 1024, iters = 4096;
    cudaMalloc(&d, (blocks * threads + 1) * 4);
    k<MODE><<<blocks, threads>>>(d, 16);
    k<MODE><<<blocks, threads>>>(d, iters);
    unsigned clk;
    cudaMemcpy(&clk, d + blocks * threads, 4, cudaMemcpyDeviceToHost);
    const double instr = (double)threads * iters * 8;
    printf("%-28s %6.2f instr/clk/SM  %6.2f results/clk/SM  (%s)\n", name, instr / clk, instr * results_per_instr / clk,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}
int main() {
    run<0>("ex2.approx.ftz.f32", 1);
    run<1>("ex2.approx.f16x2", 2);
    run<2>("ex2.approx.ftz.bf16x2", 2);
    run<3>("cvt.rn.bf16x2.f32 (F2FP)", 1);
    run<4>("ex2.f32 + F2FP alternating", 1);
    run<5>("prmt.b32", 1);
    run<6>("ex2.f32 + prmt alternating", 1);
    run<7>("cvt.rn.f16x2.f32 (F2FP)", 1);
    return 0;
}
