// Bandwidth-bound LightGlue kernels: descriptor ingest, keypoint normalisation + learnable Fourier
// encoding, row dot products (matchability / token confidence), fused dual-softmax + matchability
// log-assignment with row/column arg-max, mutual-nearest-neighbour match filter.
// (reference: gluefactory/models/matchers/lightglue.py:21-59,62-73,250-303)
#include "common.cuh"
#include "lg_kernels.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------------
// ingest: fp32 descriptors (rows, 256) -> fp32 residual stream + bf16 copy in the left half of XM
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ingest_kernel(const float* __restrict__ src, long long rows,
                                                     float* __restrict__ x32, __nv_bfloat16* __restrict__ xm,
                                                     long long ld_xm) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per 8 values
    if (i >= rows * 32) return;
    const long long row = i >> 5;
    const int col = (int)(i & 31) * 8;
    const float4 a = *reinterpret_cast<const float4*>(src + row * 256 + col);
    const float4 b = *reinterpret_cast<const float4*>(src + row * 256 + col + 4);
    *reinterpret_cast<float4*>(x32 + row * 256 + col) = a;
    *reinterpret_cast<float4*>(x32 + row * 256 + col + 4) = b;
    *reinterpret_cast<uint4*>(xm + row * ld_xm + col) =
        make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
}

int launch_ingest(const float* src, long long rows, float* x32, void* xm, long long ld_xm, cudaStream_t stream) {
    if (rows == 0) return 0;
    const long long threads = rows * 32;
    ingest_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(src, rows, x32,
                                                                         reinterpret_cast<__nv_bfloat16*>(xm), ld_xm);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// positional encoding: normalise keypoints, project with Wr (32 x 2), emit (cos, sin) per frequency.
// size == nullptr: extent-based normalisation, size = 1 + max - min over the image's keypoints.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) kpt_extent_kernel(const float* __restrict__ kpts, int n,
                                                         float* __restrict__ size_out) {
    __shared__ float smin[2][8], smax[2][8];
    const float* k = kpts + (long long)blockIdx.x * n * 2;
    float mn[2] = {INFINITY, INFINITY}, mx[2] = {-INFINITY, -INFINITY};
    for (int i = threadIdx.x; i < n; i += 256) {
#pragma unroll
        for (int d = 0; d < 2; ++d) {
            const float v = k[i * 2 + d];
            mn[d] = fminf(mn[d], v);
            mx[d] = fmaxf(mx[d], v);
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        mx[d] = warp_max(mx[d]);
        mn[d] = -warp_max(-mn[d]);
        if (lane == 0) { smin[d][warp] = mn[d]; smax[d][warp] = mx[d]; }
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        const int d = threadIdx.x;
        float a = smin[d][0], b = smax[d][0];
        for (int w = 1; w < 8; ++w) { a = fminf(a, smin[d][w]); b = fmaxf(b, smax[d][w]); }
        size_out[blockIdx.x * 2 + d] = 1.f + b - a;
    }
}

__global__ void __launch_bounds__(256) posenc_kernel(const float* __restrict__ kpts, const float* __restrict__ size,
                                                     int n, long long total, const float* __restrict__ wr,
                                                     float2* __restrict__ cs) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // (token, frequency)
    if (i >= total * 32) return;
    const long long tok = i >> 5;
    const int f = (int)(i & 31);
    const int b = (int)(tok / n);
    const float sx = size[b * 2], sy = size[b * 2 + 1];
    const float scale = fmaxf(sx, sy) / 2.f;
    const float x = (kpts[tok * 2] - sx / 2.f) / scale, y = (kpts[tok * 2 + 1] - sy / 2.f) / scale;
    const float proj = x * wr[f * 2] + y * wr[f * 2 + 1];
    float s, c;
    sincosf(proj, &s, &c);
    cs[i] = make_float2(c, s);
}

int launch_posenc(const float* kpts, const float* size, int B, int n, const float* wr, float* size_scratch,
                  void* cs, cudaStream_t stream) {
    if (B == 0 || n == 0) return 0;
    if (size == nullptr) {
        kpt_extent_kernel<<<B, 256, 0, stream>>>(kpts, n, size_scratch);
        B200_CHECK_CUDA(cudaGetLastError());
        size = size_scratch;
    }
    const long long total = (long long)B * n;
    posenc_kernel<<<(unsigned)((total * 32 + 255) / 256), 256, 0, stream>>>(kpts, size, n, total, wr,
                                                                            reinterpret_cast<float2*>(cs));
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// rowdot: z[row] = x[row, :256] . w + bias  (optionally through a sigmoid); one warp per row
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rowdot_kernel(const float* __restrict__ x, long long rows,
                                                     const float* __restrict__ w, const float* __restrict__ bias,
                                                     int apply_sigmoid, float* __restrict__ out) {
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float4* xr = reinterpret_cast<const float4*>(x + row * 256) + lane * 2;
    const float4* wv = reinterpret_cast<const float4*>(w) + lane * 2;
    const float4 a = xr[0], b = xr[1], c = __ldg(wv), d = __ldg(wv + 1);
    float s = a.x * c.x + a.y * c.y + a.z * c.z + a.w * c.w + b.x * d.x + b.y * d.y + b.z * d.z + b.w * d.w;
    s = warp_sum(s);
    if (lane == 0) {
        s += bias[0];
        out[row] = apply_sigmoid ? 1.f / (1.f + expf(-s)) : s;
    }
}

int launch_rowdot(const float* x, long long rows, const float* w, const float* bias, int apply_sigmoid, float* out,
                  cudaStream_t stream) {
    if (rows == 0) return 0;
    rowdot_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, stream>>>(x, rows, w, bias, apply_sigmoid, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
