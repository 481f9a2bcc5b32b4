// FP64 pipe microbenchmarks for B200: which inner loop shape reaches the DFMA / DMMA peak?
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// 1: 4 independent accumulators, same operands
__global__ void __launch_bounds__(256, 2) v1(double* out, int iters) {
    double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6, c[4][2] = {};
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int q = 0; q < 4; ++q) dmma(c[q][0], c[q][1], a, b);
    double s = 0; for (int q = 0; q < 4; ++q) s += c[q][0] + c[q][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 2: 4x4 tile of accumulators, 4 a + 4 b registers (no loads)
__global__ void __launch_bounds__(256, 2) v2(double* out, int iters) {
    double a[4], b[4], c[4][4][2] = {};
    for (int q = 0; q < 4; ++q) { a[q] = threadIdx.x * 1e-3 + q; b[q] = 1.0 + threadIdx.x * 1e-6 * q; }
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
            for (int y = 0; y < 4; ++y) dmma(c[x][y][0], c[x][y][1], a[x], b[y]);
    double s = 0; for (int x = 0; x < 4; ++x) for (int y = 0; y < 4; ++y) s += c[x][y][0] + c[x][y][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 3: 4x4 tile with operand fragments re-loaded from shared memory every k-step (GEMM inner loop, no global traffic)
__global__ void __launch_bounds__(256, 2) v3(double* out, int iters) {
    __shared__ double As[16 * 68], Bs[16 * 132];
    for (int i = threadIdx.x; i < 16 * 68; i += 256) As[i] = i * 1e-4;
    for (int i = threadIdx.x; i < 16 * 132; i += 256) Bs[i] = 1.0 + i * 1e-6;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lr = lane >> 2, lc = lane & 3;
    const double* Ap = As + (warp >> 2) * 32 + lr;
    const double* Bp = Bs + (warp & 3) * 32 + lr;
    double c[4][4][2] = {};
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            double a[4], b[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { a[q] = Ap[(ks * 4 + lc) * 68 + q * 8]; b[q] = Bp[(ks * 4 + lc) * 132 + q * 8]; }
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 4; ++y) dmma(c[x][y][0], c[x][y][1], a[x], b[y]);
        }
    double s = 0; for (int x = 0; x < 4; ++x) for (int y = 0; y < 4; ++y) s += c[x][y][0] + c[x][y][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 4: DFMA 8x8 register tile, operands from shared memory as 128-bit loads
template <int TR, int TC>
__global__ void __launch_bounds__(256, 2) v4(double* out, int iters) {
    __shared__ __align__(16) double As[16 * 128], Bs[16 * 128];
    for (int i = threadIdx.x; i < 16 * 128; i += 256) { As[i] = i * 1e-4; Bs[i] = 1.0 + i * 1e-6; }
    __syncthreads();
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double c[TR][TC] = {};
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            double a[TR], b[TC];
#pragma unroll
            for (int q = 0; q < TR; q += 2) { double2 v = *reinterpret_cast<const double2*>(&As[k * 128 + ty * TR + q]); a[q] = v.x; a[q + 1] = v.y; }
#pragma unroll
            for (int q = 0; q < TC; q += 2) { double2 v = *reinterpret_cast<const double2*>(&Bs[k * 128 + tx * TC + q]); b[q] = v.x; b[q + 1] = v.y; }
#pragma unroll
            for (int x = 0; x < TR; ++x)
#pragma unroll
                for (int y = 0; y < TC; ++y) c[x][y] = fma(a[x], b[y], c[x][y]);
        }
    double s = 0; for (int x = 0; x < TR; ++x) for (int y = 0; y < TC; ++y) s += c[x][y];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class F> static void run(const char* name, F launch, double flops_per_thread_iter, int iters, int blocks) {
    double* buf; cudaMalloc(&buf, sizeof(double) * blocks * 256);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0); launch(buf, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    printf("%-44s %8.2f TFLOP/s  (%.3f ms) %s\n", name, flops_per_thread_iter * iters * blocks * 256.0 / (best * 1e-3) / 1e12, best, cudaGetErrorString(cudaGetLastError()));
    cudaFree(buf);
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 2;
    run("dmma 4 acc same operands", [&](double* b, int it) { v1<<<blocks, 256>>>(b, it); }, 4 * 16.0, 1 << 14, blocks);
    run("dmma 4x4 tile, register operands", [&](double* b, int it) { v2<<<blocks, 256>>>(b, it); }, 16 * 16.0, 1 << 12, blocks);
    run("dmma 4x4 tile, smem fragments per k-step", [&](double* b, int it) { v3<<<blocks, 256>>>(b, it); }, 4 * 16 * 16.0, 1 << 10, blocks);
    run("dfma 8x8 tile, smem LDS.128", [&](double* b, int it) { v4<8, 8><<<blocks, 256>>>(b, it); }, 16 * 64 * 2.0, 1 << 10, blocks);
    run("dfma 8x4 tile, smem LDS.128", [&](double* b, int it) { v4<8, 4><<<blocks, 256>>>(b, it); }, 16 * 32 * 2.0, 1 << 10, blocks);
    run("dfma 4x4 tile, smem LDS.128", [&](double* b, int it) { v4<4, 4><<<blocks, 256>>>(b, it); }, 16 * 16 * 2.0, 1 << 10, blocks);
    return 0;
}
