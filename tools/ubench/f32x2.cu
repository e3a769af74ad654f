// Micro-benchmark: issue rate of scalar FMUL/FADD against the packed FMUL2/FADD2 of sm_100 (round-to-nearest, no FMA).
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float *out, float a, float b, int iters) {
    float2 x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = make_float2(threadIdx.x + i, threadIdx.x * 0.5f + i);
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) {
                x[i].x = __fadd_rn(__fmul_rn(x[i].x, a), b);
                x[i].y = __fadd_rn(__fmul_rn(x[i].y, a), b);
            } else {
                x[i] = __fadd2_rn(__fmul2_rn(x[i], aa), bb);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i].x + x[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    float *out;
    cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    for (int mode = 0; mode < 2; mode++)
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters);
            else k<1><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            const double flops = 148.0 * 8 * 256 * iters * 32.0;  // lane operations (mul + add counted separately)
            printf("mode %d (%s): %.3f ms, %.1f Glane-op/s, %.2f lane-op/clk/SM at 1.965 GHz\n", mode, mode ? "packed f32x2" : "scalar", ms,
                   flops / ms * 1e-6, flops / (ms * 1e-3) / 148 / 1.965e9);
        }
    return 0;
}
