// Developer tool: dependent-issue latencies on this GPU (cycles per op in a single-warp dependent chain).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/lat_probe.cu -o tools/lat_probe && tools/lat_probe
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(double* out, long long* cyc, double a, double b) {
    double x = a + threadIdx.x * 1e-9;
    unsigned u = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (OP == 0) x = fma(x, b, a);
            if (OP == 1) x = x * b;
            if (OP == 2) x = x + b;
            if (OP == 3) x = __shfl_xor_sync(0xffffffffu, x, 1);
            if (OP == 4) u = __reduce_max_sync(0xffffffffu, u) + 1;
            if (OP == 5) x = 1.0 / x + a;
            if (OP == 6) x = fmax(x * b, a);
            if (OP == 7) x = (x > a) ? b : x + 1.0;
        }
    }
    long long t1 = clock64();
    out[threadIdx.x] = x + u;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 256 * 8); cudaMallocManaged(&cyc, 8);
    const char* names[] = {"DFMA", "DMUL", "DADD", "SHFL.64", "REDUX.MAX.u32", "DDIV (+DADD)", "DMUL+DMNMX", "DSETP+select"};
#define RUN(OP) k<OP><<<1, 32>>>(out, cyc, 1.0000001, 0.9999999); cudaDeviceSynchronize(); k<OP><<<1, 32>>>(out, cyc, 1.0000001, 0.9999999); cudaDeviceSynchronize(); printf("%-16s %.1f cycles per dependent op\n", names[OP], cyc[0] / 4096.0);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7)
    return 0;
}
