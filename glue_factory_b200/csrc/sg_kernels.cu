// SuperGlue keypoint encoder, first two layers on the CUDA cores (K = 3 and K = 32 are too small for the
// tensor cores): normalise (x, y) by the image size, append the detection score, then
//   Conv1d(3 -> 32) + BatchNorm1d + ReLU -> Conv1d(32 -> 64) + BatchNorm1d + ReLU
// with the (eval-mode) batch norms folded into the weights on the host.  Output: bf16 [tokens, 64], the
// A operand of the remaining encoder layers, which run on the tcgen05 GEMM path.
// (reference: gluefactory_nonfree/superglue.py:63-102,279-288)
#include "common.cuh"
#include "lg_kernels.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) sg_kenc_head_kernel(const float* __restrict__ kpts, const float* __restrict__ scores,
                                                           const float* __restrict__ size, int n, long long total,
                                                           const float* __restrict__ w1, const float* __restrict__ b1,
                                                           const float* __restrict__ w2, const float* __restrict__ b2,
                                                           __nv_bfloat16* __restrict__ out) {
    __shared__ float sw1[32 * 3], sb1[32], sw2[64 * 32], sb2[64];
    for (int i = threadIdx.x; i < 96; i += 256) sw1[i] = w1[i];
    for (int i = threadIdx.x; i < 32; i += 256) sb1[i] = b1[i];
    for (int i = threadIdx.x; i < 2048; i += 256) sw2[i] = w2[i];
    for (int i = threadIdx.x; i < 64; i += 256) sb2[i] = b2[i];
    __syncthreads();
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / n);
    const float sw = size[2 * b], sh = size[2 * b + 1];
    const float scale = fmaxf(sw, sh) * 0.7f;
    // (kpts - size / 2) / (max(size) * 0.7)
    const float x = (kpts[2 * t] - sw / 2.f) / scale, y = (kpts[2 * t + 1] - sh / 2.f) / scale;
    const float s = scores[t];
    float h1[32];
#pragma unroll
    for (int o = 0; o < 32; ++o)
        h1[o] = fmaxf(fmaf(sw1[o * 3 + 2], s, fmaf(sw1[o * 3 + 1], y, fmaf(sw1[o * 3], x, sb1[o]))), 0.f);
    __nv_bfloat16* dst = out + t * 64;
#pragma unroll 1
    for (int o0 = 0; o0 < 64; o0 += 8) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float* wr = sw2 + (o0 + j) * 32;
            float a = sb2[o0 + j];
#pragma unroll
            for (int k = 0; k < 32; ++k) a = fmaf(wr[k], h1[k], a);
            v[j] = fmaxf(a, 0.f);
        }
        *reinterpret_cast<uint4*>(dst + o0) =
            make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    }
}

int launch_sg_kenc_head(const float* kpts, const float* scores, const float* size, int B, int n, const float* w1,
                        const float* b1, const float* w2, const float* b2, void* out, cudaStream_t stream) {
    const long long total = (long long)B * n;
    if (total == 0) return 0;
    sg_kenc_head_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(kpts, scores, size, n, total, w1, b1, w2, b2,
                                                                            reinterpret_cast<__nv_bfloat16*>(out));
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// fp32 rows -> two-term bf16 split for a near-fp32 similarity GEMM on the bf16 tensor cores:
// x = hi + lo (hi = bf16(x), lo = bf16(x - hi)); rows < split_row (image 0, the A operand) become
// [hi | hi | lo], the others (image 1, the B operand) [hi | lo | hi], so that one K = 3 * cols GEMM
// accumulates hi.hi + hi.lo + lo.hi in fp32 (the dropped lo.lo term is ~2^-16 relative).
__global__ void __launch_bounds__(256) split_hilo_kernel(const float* __restrict__ x, long long rows, int cols,
                                                         long long split_row, __nv_bfloat16* __restrict__ out) {
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;  // one thread per 4 columns
    const int per_row = cols >> 2;
    if (t >= rows * per_row) return;
    const long long row = t / per_row;
    const int c = (int)(t - row * per_row) * 4;
    const float4 v = *reinterpret_cast<const float4*>(x + row * cols + c);
    const __nv_bfloat162 h01 = __floats2bfloat162_rn(v.x, v.y), h23 = __floats2bfloat162_rn(v.z, v.w);
    const float2 f01 = __bfloat1622float2(h01), f23 = __bfloat1622float2(h23);
    const uint32_t lo0 = pack_bf16(v.x - f01.x, v.y - f01.y), lo1 = pack_bf16(v.z - f23.x, v.w - f23.y);
    const uint2 hi = make_uint2(*reinterpret_cast<const uint32_t*>(&h01), *reinterpret_cast<const uint32_t*>(&h23));
    const uint2 lo = make_uint2(lo0, lo1);
    __nv_bfloat16* o = out + row * 3 * cols + c;
    const bool a_side = row < split_row;
    *reinterpret_cast<uint2*>(o) = hi;
    *reinterpret_cast<uint2*>(o + cols) = a_side ? hi : lo;
    *reinterpret_cast<uint2*>(o + 2 * cols) = a_side ? lo : hi;
}

int launch_split_hilo(const float* x, long long rows, int cols, long long split_row, void* out, cudaStream_t stream) {
    B200_REQUIRE(cols > 0 && cols % 4 == 0, "split_hilo: cols=%d must be a positive multiple of 4", cols);
    if (rows == 0) return 0;
    split_hilo_kernel<<<(unsigned)((rows * (cols / 4) + 255) / 256), 256, 0, stream>>>(
        x, rows, cols, split_row, reinterpret_cast<__nv_bfloat16*>(out));
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_sg_split_hilo(const float* x, long long rows, long long split_row, void* out, cudaStream_t stream) {
    return launch_split_hilo(x, rows, 256, split_row, out, stream);
}

}  // namespace b200
