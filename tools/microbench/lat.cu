// DEVELOPMENT TOOL: dependent-issue latencies on the device (one warp, clock64).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench/lat tools/microbench/lat.cu
#include <cstdio>
#include <cuda_runtime.h>
#define N 4096
__global__ void k(double *out, long long *cyc, double a, double b) {
    __shared__ double sm[64];
    long long t0, t1;
    double x = a + threadIdx.x * 1e-9;
    // DFMA chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = fma(x, b, a);
    t1 = clock64(); cyc[0] = t1 - t0;
    // DADD chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = x + a;
    t1 = clock64(); cyc[1] = t1 - t0;
    // MUFU.RCP64H + 1 DFMA chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x)); x = fma(r, b, a); }
    t1 = clock64(); cyc[2] = t1 - t0;
    // STS -> syncwarp -> LDS round trip (+1 DADD)
    t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N; ++i) { sm[threadIdx.x & 31] = x; __syncwarp(); x = sm[(threadIdx.x + 1) & 31] + a; __syncwarp(); }
    t1 = clock64(); cyc[3] = t1 - t0;
    // shuffle (64-bit) + DADD
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = __shfl_xor_sync(0xffffffffu, x, 1) + a;
    t1 = clock64(); cyc[4] = t1 - t0;
    // float shuffle + FMNMX
    float f = (float)x;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) f = fmaxf(f * 0.5f, __shfl_xor_sync(0xffffffffu, f, 1));
    t1 = clock64(); cyc[5] = t1 - t0;
    // LDS.64 dependent (pointer chase through smem index)
    int idx = threadIdx.x & 31;
    sm[idx] = (double)((idx + 1) & 31);
    __syncwarp();
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) idx = (int)sm[idx];
    t1 = clock64(); cyc[6] = t1 - t0;
    // FFMA chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) f = fmaf(f, 0.999f, 0.5f);
    t1 = clock64(); cyc[7] = t1 - t0;
    // DMUL + int hi-word add (exponent scaling) chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { x = x * b; x = __hiloint2double(__double2hiint(x) + 1, __double2loint(x)); x = __hiloint2double(__double2hiint(x) - 1, __double2loint(x)); }
    t1 = clock64(); cyc[8] = t1 - t0;
    out[threadIdx.x] = x + f + idx;
}
int main() {
    double *out; long long *cyc, h[9];
    cudaMalloc(&out, 256 * 8); cudaMalloc(&cyc, 9 * 8);
    for (int rep = 0; rep < 2; ++rep) { k<<<1, 32>>>(out, cyc, 0.5, 0.999); cudaDeviceSynchronize(); }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    const char *names[9] = {"DFMA dependent", "DADD dependent", "MUFU.RCP64H + DFMA", "STS+syncwarp+LDS+DADD+syncwarp",
                            "SHFL.64 + DADD", "SHFL.32 + FMNMX (+FMUL)", "LDS.64 + F2I (pointer chase)", "FFMA dependent",
                            "DMUL + 2 hi-word IADD"};
    for (int i = 0; i < 9; ++i) printf("%-36s %.1f cycles\n", names[i], (double)h[i] / N);
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
