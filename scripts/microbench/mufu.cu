// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/microbench/mufu scripts/microbench/mufu.cu
// MUFU throughput probe: ex2.approx f32 vs packed f16x2 / bf16x2 (elements per clock per SM)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(1024) probe(uint32_t* out, int iters, float seed) {
    uint32_t r[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __float_as_uint(seed * (threadIdx.x + i));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(r[i]));
            if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(r[i]));
            if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r[i]));
            if (MODE == 3) asm volatile("{.reg .b16 l, h; mov.b32 {l, h}, %0; ex2.approx.f16 l, l; mov.b32 %0, {l, h};}" : "+r"(r[i]));
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= r[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
void run(const char* name, int elems_per_instr) {
    int sms;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    uint32_t* out;
    cudaMalloc(&out, sms * 2 * 1024 * 4);
    const int iters = 4096;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    probe<MODE><<<sms * 2, 1024>>>(out, iters, 0.001f);
    cudaEventRecord(a);
    probe<MODE><<<sms * 2, 1024>>>(out, iters, 0.001f);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    const double instr = (double)sms * 2 * 1024 * iters * 8;
    const double clk = ms * 1e-3 * khz * 1e3;
    printf("%-10s %.3f ms  lane-instr/clk/SM %.2f  elements/clk/SM %.2f  (err %s)\n", name, ms, instr / clk / sms,
           instr * elems_per_instr / clk / sms, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    run<0>("f32", 1);
    run<1>("f16x2", 2);
    run<2>("bf16x2", 2);
    run<3>("f16", 1);
    return 0;
}
