// Dependent-issue latencies on one warp (sm_100a): what a serial chain of word visits pays per step.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lat lat.cu && ./lat
#include <cstdio>
#include <cuda_runtime.h>

#define N 512
template <int OP>
__global__ void chain(double a, double b, double *out, long long *cyc, int *idx)
{
    __shared__ double sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (double)((i * 7 + 1) & 1023);
    __syncthreads();
    double x = a + threadIdx.x;
    int j = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; i++) {
        if (OP == 0) x = fma(x, b, a);
        if (OP == 1) x = x + b;
        if (OP == 2) x = x * b;
        if (OP == 3) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x)); x = r; }
        if (OP == 4) { j = (int)sm[j & 1023]; }                       // LDS.64 + F2I chain
        if (OP == 5) x = __shfl_xor_sync(0xffffffffu, x, 1);          // 2 SHFL.32
        if (OP == 6) { int h = __double2hiint(x); x = __hiloint2double(h + 1, __double2loint(x)); }  // int op on hi word
        if (OP == 7) { j = ((volatile int *)sm)[j & 1023] ; }         // LDS.32 chain
        if (OP == 8) { x = fma(x, b, a); x = __shfl_xor_sync(0xffffffffu, x, 1); }  // DFMA + SHFL
        if (OP == 9) { x = (double)__double2hiint(x); }               // I2F.F64 chain (with the mov)
        if (OP == 11) {
            double d0, d1;
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
                         : "=d"(d0), "=d"(d1) : "d"(x), "d"(b), "d"(a), "d"(a));
            x = d0;
        }
        if (OP == 10) { float f = (float)x; asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f)); x = (double)f; }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[OP] = t1 - t0;
    out[threadIdx.x + OP * 32] = x + j;
}

// throughput: W warps per block each running 4 independent DFMA chains
__global__ void tput(double a, double b, double *out, long long *cyc)
{
    double x0 = a + threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < N; i++) { x0 = fma(x0, b, a); x1 = fma(x1, b, a); x2 = fma(x2, b, a); x3 = fma(x3, b, a); }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    out[threadIdx.x] = x0 + x1 + x2 + x3;
}

int main()
{
    double *out; long long *cyc; int *idx;
    cudaMalloc(&out, 8 * 4096); cudaMalloc(&cyc, 8 * 64); cudaMalloc(&idx, 4 * 1024);
    long long h[16];
    const char *names[] = {"DFMA", "DADD", "DMUL", "MUFU.RCP64H", "LDS.64+F2I", "SHFL.64", "hi-word IADD", "LDS.32", "DFMA+SHFL.64", "I2F.F64(hi)", "F2F+EX2+F2F", "DMMA.884 (d->a)"};
#define RUN(op) chain<op><<<1, 32>>>(1.0000001, 0.9999999, out, cyc, idx); chain<op><<<1, 32>>>(1.0000001, 0.9999999, out, cyc, idx);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(10) RUN(11)
    cudaDeviceSynchronize();
    cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    for (int i = 0; i < 12; i++) printf("%-16s %6.1f cycles per dependent step\n", names[i], (double)h[i] / N);
    for (int w = 1; w <= 16; w *= 2) {
        tput<<<1, 32 * w>>>(1.0000001, 0.9999999, out, cyc); tput<<<1, 32 * w>>>(1.0000001, 0.9999999, out, cyc);
        cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DFMA throughput, %2d warps x 4 chains: %.2f cycles per warp-DFMA per SM (%.1f lanes/clk/SM)\n", w,
               (double)h[0] / (N * 4.0 * w), 32.0 * N * 4 * w / h[0]);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
