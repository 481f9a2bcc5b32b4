#include <cuda_runtime.h>
__global__ void k884(double* out, int iters) {
    double a = threadIdx.x, b = threadIdx.x * 0.5, c0 = 0, c1 = 0;
    for (int i = 0; i < iters; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
    out[threadIdx.x] = c0 + c1;
}
__global__ void k16816(double* out, int iters) {
    double a[8], b[4], c[4] = {0, 0, 0, 0};
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i;
    for (int i = 0; i < 4; ++i) b[i] = threadIdx.x * 0.5 + i;
    for (int i = 0; i < iters; ++i)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    out[threadIdx.x] = c[0] + c[1] + c[2] + c[3];
}
__global__ void k1688(double* out, int iters) {
    double a[4], b[2], c[4] = {0, 0, 0, 0};
    for (int i = 0; i < 4; ++i) a[i] = threadIdx.x + i;
    for (int i = 0; i < 2; ++i) b[i] = threadIdx.x * 0.5 + i;
    for (int i = 0; i < iters; ++i)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
    out[threadIdx.x] = c[0] + c[1] + c[2] + c[3];
}
