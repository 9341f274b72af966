// Micro-benchmark: per-SM throughput of the instructions the attention softmax is made of (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/ubench_mufu scripts/microbench/softmax_mix.cu
// Prints warp-instructions per clock per SM for: ex2.f32, ex2.bf16x2, ex2.f16x2, cvt.bf16x2.f32, 3-input max,
// fma.f32x2, and two mixes (ex2.f32 + cvt; ex2.bf16x2 + cvt).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int kIters = 4096;
constexpr int kChains = 8;

template <int MODE>
__global__ void __launch_bounds__(512) bench(uint32_t* out, long long* clk, float seed) {
    uint32_t v[kChains];
    float f[kChains];
#pragma unroll
    for (int i = 0; i < kChains; ++i) {
        f[i] = seed * (float)(threadIdx.x + i + 1) * 1e-3f;
        v[i] = __float_as_uint(f[i]);
    }
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int i = 0; i < kChains; ++i) {
            if (MODE == 0) {
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
            } else if (MODE == 1) {
                asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(v[i]));
            } else if (MODE == 2) {
                asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(v[i]));
            } else if (MODE == 3) {
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v[i]) : "f"(f[i]), "f"(f[(i + 1) % kChains]));
                f[i] = __uint_as_float(v[i]);
            } else if (MODE == 4) {
                asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(f[(i + 1) % kChains]), "f"(seed));
            } else if (MODE == 5) {
                unsigned long long a;
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(f[i]), "f"(f[(i + 1) % kChains]));
                asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(a));
                float lo, hi;
                asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a));
                f[i] = lo + hi * 0.f;
            } else if (MODE == 6) {  // ex2.f32 x2 + one cvt (today's per-pair mix)
                float a = f[i], b = f[(i + 1) % kChains];
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b));
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v[i]) : "f"(a), "f"(b));
                f[i] = __uint_as_float(v[i]) * 1e-30f;
            } else if (MODE == 7) {  // one cvt + one ex2.bf16x2 (candidate per-pair mix)
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v[i]) : "f"(f[i]), "f"(f[(i + 1) % kChains]));
                asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(v[i]));
                f[i] = __uint_as_float(v[i] & 0x0fff0fffu);
            } else if (MODE == 8) {
                asm volatile("max.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(f[(i + 1) % kChains]));
            }
        }
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < kChains; ++i) acc ^= v[i] ^ __float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int ops_per_iter, uint32_t* out, long long* clk, int sms) {
    const int threads = 512;
    bench<MODE><<<sms, threads>>>(out, clk, 0.5f);
    cudaDeviceSynchronize();
    bench<MODE><<<sms, threads>>>(out, clk, 0.5f);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[256];
    cudaMemcpy(h, clk, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; ++i) mean += (double)h[i];
    mean /= sms;
    const double warp_instr = (double)(threads / 32) * kIters * kChains * ops_per_iter;
    printf("%-28s %s  clk/SM %.0f  warp-instr/clk/SM %.3f  (lane-ops/clk/SM %.1f)\n", name, cudaGetErrorString(e), mean,
           warp_instr / mean, warp_instr * 32 / mean);
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    uint32_t* out;
    long long* clk;
    cudaMalloc(&out, sms * 512 * 4);
    cudaMalloc(&clk, sms * 8);
    printf("%s, %d SMs\n", prop.name, sms);
    run<0>("ex2.approx.ftz.f32", 1, out, clk, sms);
    run<1>("ex2.approx.ftz.bf16x2", 1, out, clk, sms);
    run<2>("ex2.approx.f16x2", 1, out, clk, sms);
    run<3>("cvt.rn.bf16x2.f32", 1, out, clk, sms);
    run<4>("max.f32 (3 inputs)", 1, out, clk, sms);
    run<8>("max.f32 (2 inputs)", 1, out, clk, sms);
    run<5>("fma.rn.f32x2 (+movs)", 1, out, clk, sms);
    run<6>("2 ex2.f32 + cvt (per pair)", 1, out, clk, sms);
    run<7>("cvt + ex2.bf16x2 (per pair)", 1, out, clk, sms);
    return 0;
}
