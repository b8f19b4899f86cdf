// Developer micro-benchmark: latency and throughput of the legacy warp-level mma.sync.m16n8k16 (bf16, fp32 accumulate) and of
// ldmatrix on sm_100a.  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(float (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <int CHAINS>
__global__ void k(float* out, long long* clk, int iters) {
    float c[CHAINS][4];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
    unsigned a = 0x3f803f80u + threadIdx.x, b = 0x3f803f80u;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) mma(c[i], a, a, a, a, b, b);
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
template <int CHAINS>
void run(int warps_per_sm, const char* what) {
    float* d; long long* c; cudaMalloc(&d, 148 * 1024 * 4); cudaMalloc(&c, 8);
    const int iters = 2048;
    k<CHAINS><<<148, warps_per_sm * 32>>>(d, c, 8);
    k<CHAINS><<<148, warps_per_sm * 32>>>(d, c, iters);
    long long clk; cudaMemcpy(&clk, c, 8, cudaMemcpyDeviceToHost);
    const double n = (double)iters * CHAINS;
    printf("%-40s %2d warps/SM %d chains: %7.1f clk per mma per warp, %6.1f dense FLOP/clk/SM (%s)\n", what, warps_per_sm, CHAINS,
           clk / n, n * warps_per_sm * 4096.0 / clk, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(c);
}
int main() {
    run<1>(1, "latency (1 warp, dependent)");
    run<8>(1, "1 warp, 8 independent chains");
    run<8>(4, "4 warps, 8 chains each");
    run<8>(16, "16 warps, 8 chains each");
    run<8>(32, "32 warps, 8 chains each");
    return 0;
}
