// Match-quality metrics on the device: precision of the predicted matches against a
// ground-truth homography or relative pose, and the weighted DLT homography with its corner error.  One CTA per
// pair; everything is a reduction over at most a few thousand matches, so the kernels are latency-bound and exist
// so that the predictions never leave the GPU between the matcher and the metric.
// (reference: gluefactory/eval/utils.py:21-27,40-96,176-196; geometry/homography.py:319-347;
//  geometry/epipolar.py:7-9,32-56; kornia.geometry.homography.find_homography_dlt + normalize_points)
#include "common.cuh"
#include "io_kernels.cuh"

namespace b200 {

constexpr int kEvalThreads = 256;

template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch /*[kEvalThreads / 32]*/) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();  // scratch may still be read from a previous reduction
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    T s = 0;
    for (int w = 0; w < kEvalThreads / 32; ++w) s += scratch[w];
    return s;
}

// inverse of a 3x3 matrix by the adjugate, in double (the reference takes torch.pinverse of the fp32 matrix)
__device__ void inverse3(const double* h, double* inv) {
    const double a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], i = h[7], j = h[8];
    const double A = e * j - f * i, Bc = -(d * j - f * g), Cc = d * i - e * g;
    const double det = a * A + b * Bc + c * Cc;
    const double r = 1.0 / det;
    inv[0] = A * r; inv[1] = -(b * j - c * i) * r; inv[2] = (b * f - c * e) * r;
    inv[3] = Bc * r; inv[4] = (a * j - c * g) * r; inv[5] = -(a * f - c * d) * r;
    inv[6] = Cc * r; inv[7] = -(a * i - b * g) * r; inv[8] = (a * e - b * d) * r;
}

// from_homogeneous(to_homogeneous(p) @ T^T)
__device__ __forceinline__ void warp_point(const float* T, float x, float y, float* ox, float* oy) {
    const float u = fmaf(T[0], x, fmaf(T[1], y, T[2]));
    const float v = fmaf(T[3], x, fmaf(T[4], y, T[5]));
    const float w = fmaf(T[6], x, fmaf(T[7], y, T[8]));
    *ox = u / w;
    *oy = v / w;
}

// eval_matches_homography (eval/utils.py:76-96): out[b] = {prec@1px, prec@3px, num_matches, num_keypoints}
__global__ void __launch_bounds__(kEvalThreads) eval_homography_kernel(const float* __restrict__ kp0, const float* __restrict__ kp1,
                                                                       const long long* __restrict__ m0, int M, int N,
                                                                       const float* __restrict__ Hs, float* __restrict__ out) {
    __shared__ float T[9], Ti[9];
    __shared__ int scratch[kEvalThreads / 32];
    const int b = blockIdx.x;
    if (threadIdx.x == 0) {
        double h[9], inv[9];
        for (int i = 0; i < 9; ++i) { T[i] = Hs[b * 9 + i]; h[i] = T[i]; }
        inverse3(h, inv);
        for (int i = 0; i < 9; ++i) Ti[i] = (float)inv[i];
    }
    __syncthreads();
    int n = 0, c1 = 0, c3 = 0;
    for (int i = threadIdx.x; i < M; i += kEvalThreads) {
        const long long j = m0[(long long)b * M + i];
        if (j < 0) continue;
        const float x0 = kp0[((long long)b * M + i) * 2], y0 = kp0[((long long)b * M + i) * 2 + 1];
        const float x1 = kp1[((long long)b * N + j) * 2], y1 = kp1[((long long)b * N + j) * 2 + 1];
        float ax, ay, bx, by;
        warp_point(T, x0, y0, &ax, &ay);    // kpts0 -> image 1
        warp_point(Ti, x1, y1, &bx, &by);   // kpts1 -> image 0
        const float d01 = sqrtf((ax - x1) * (ax - x1) + (ay - y1) * (ay - y1));
        const float d10 = sqrtf((bx - x0) * (bx - x0) + (by - y0) * (by - y0));
        const float err = (d01 + d10) / 2.0f;
        ++n;
        c1 += err < 1.f;
        c3 += err < 3.f;
    }
    n = block_sum(n, scratch);
    c1 = block_sum(c1, scratch);
    c3 = block_sum(c3, scratch);
    if (threadIdx.x == 0) {
        out[b * 4 + 0] = n ? (float)c1 / (float)n : 0.f;  // .mean().nan_to_num()
        out[b * 4 + 1] = n ? (float)c3 / (float)n : 0.f;
        out[b * 4 + 2] = (float)n;
        out[b * 4 + 3] = (M + N) / 2.0f;
    }
}

// eval_matches_epipolar (eval/utils.py:40-73) for pinhole cameras:
// out[b] = {epi_prec@1e-4, epi_prec@5e-4, epi_prec@1e-3, num_matches, num_keypoints}
__global__ void __launch_bounds__(kEvalThreads) eval_epipolar_kernel(const float* __restrict__ kp0, const float* __restrict__ kp1,
                                                                     const long long* __restrict__ m0, int M, int N,
                                                                     const float* __restrict__ cam0, const float* __restrict__ cam1,
                                                                     const float* __restrict__ T01, float* __restrict__ out) {
    __shared__ float E[9], c0[4], c1[4];
    __shared__ int scratch[kEvalThreads / 32];
    const int b = blockIdx.x;
    if (threadIdx.x == 0) {
        const float* R = T01 + b * 12;
        const float tx = R[9], ty = R[10], tz = R[11];
        // E = skew(t) @ R  (epipolar.py:7-9)
        for (int c = 0; c < 3; ++c) {
            E[0 + c] = -tz * R[3 + c] + ty * R[6 + c];
            E[3 + c] = tz * R[0 + c] - tx * R[6 + c];
            E[6 + c] = -ty * R[0 + c] + tx * R[3 + c];
        }
        for (int i = 0; i < 4; ++i) { c0[i] = cam0[b * 4 + i]; c1[i] = cam1[b * 4 + i]; }
    }
    __syncthreads();
    int n = 0, a = 0, bb = 0, c = 0;
    for (int i = threadIdx.x; i < M; i += kEvalThreads) {
        const long long j = m0[(long long)b * M + i];
        if (j < 0) continue;
        // image2cam for a pinhole camera: (p - c) / f
        const float x0 = (kp0[((long long)b * M + i) * 2] - c0[2]) / c0[0], y0 = (kp0[((long long)b * M + i) * 2 + 1] - c0[3]) / c0[1];
        const float x1 = (kp1[((long long)b * N + j) * 2] - c1[2]) / c1[0], y1 = (kp1[((long long)b * N + j) * 2 + 1] - c1[3]) / c1[1];
        const float e0 = fmaf(E[0], x0, fmaf(E[1], y0, E[2]));  // E p0
        const float e1 = fmaf(E[3], x0, fmaf(E[4], y0, E[5]));
        const float e2 = fmaf(E[6], x0, fmaf(E[7], y0, E[8]));
        const float f0 = fmaf(E[0], x1, fmaf(E[3], y1, E[6]));  // E^T p1
        const float f1 = fmaf(E[1], x1, fmaf(E[4], y1, E[7]));
        const float p1Ep0 = fmaf(x1, e0, fmaf(y1, e1, e2));
        const float d0 = fmaxf(e0 * e0 + e1 * e1, 1e-6f), d1 = fmaxf(f0 * f0 + f1 * f1, 1e-6f);
        const float d = fabsf(p1Ep0) * (1.f / sqrtf(d0) + 1.f / sqrtf(d1)) / 2.f;
        ++n;
        a += d < 1e-4f;
        bb += d < 5e-4f;
        c += d < 1e-3f;
    }
    n = block_sum(n, scratch);
    a = block_sum(a, scratch);
    bb = block_sum(bb, scratch);
    c = block_sum(c, scratch);
    if (threadIdx.x == 0) {
        const float nan = __int_as_float(0x7fc00000);  // mean of an empty tensor, as torch returns it
        out[b * 5 + 0] = n ? (float)a / (float)n : nan;
        out[b * 5 + 1] = n ? (float)bb / (float)n : nan;
        out[b * 5 + 2] = n ? (float)c / (float)n : nan;
        out[b * 5 + 3] = (float)n;
        out[b * 5 + 4] = (M + N) / 2.0f;
    }
}

// ---- weighted DLT (eval_homography_dlt, eval/utils.py:176-196) ----------------------------------------------
// smallest eigenvector of a symmetric 9x9 matrix by cyclic Jacobi rotations (double)
__device__ void jacobi_smallest_eigenvector9(double* A /*[81], destroyed*/, double* vec /*[9]*/) {
    double V[81];
    for (int i = 0; i < 81; ++i) V[i] = (i / 9 == i % 9) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int p = 0; p < 9; ++p)
            for (int q = 0; q < 9; ++q) (p == q ? diag : off) += A[p * 9 + q] * A[p * 9 + q];
        if (off <= 1e-30 * diag || off == 0.0) break;
        for (int p = 0; p < 8; ++p)
            for (int q = p + 1; q < 9; ++q) {
                const double apq = A[p * 9 + q];
                if (apq == 0.0) continue;
                const double theta = (A[q * 9 + q] - A[p * 9 + p]) / (2.0 * apq);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 9; ++k) {  // A <- A J (columns p, q)
                    const double akp = A[k * 9 + p], akq = A[k * 9 + q];
                    A[k * 9 + p] = c * akp - s * akq;
                    A[k * 9 + q] = s * akp + c * akq;
                }
                for (int k = 0; k < 9; ++k) {  // A <- J^T A (rows p, q)
                    const double apk = A[p * 9 + k], aqk = A[q * 9 + k];
                    A[p * 9 + k] = c * apk - s * aqk;
                    A[q * 9 + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 9; ++k) {
                    const double vkp = V[k * 9 + p], vkq = V[k * 9 + q];
                    V[k * 9 + p] = c * vkp - s * vkq;
                    V[k * 9 + q] = s * vkp + c * vkq;
                }
            }
    }
    int best = 0;
    for (int i = 1; i < 9; ++i)
        if (A[i * 9 + i] < A[best * 9 + best]) best = i;
    for (int k = 0; k < 9; ++k) vec[k] = V[k * 9 + best];
}

__device__ __forceinline__ void mat3_mul(const double* a, const double* b, double* c) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) c[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j];
}

__global__ void __launch_bounds__(kEvalThreads) eval_dlt_kernel(const float* __restrict__ kp0, const float* __restrict__ kp1,
                                                                const long long* __restrict__ m0, const float* __restrict__ ms0,
                                                                int M, int N, const float* __restrict__ H_gt,
                                                                const float* __restrict__ image_size, float* __restrict__ H_out,
                                                                float* __restrict__ err_out) {
    __shared__ double scratch[kEvalThreads / 32];
    __shared__ double acc[45][kEvalThreads / 32];
    const int b = blockIdx.x;
    const float* k0 = kp0 + (long long)b * M * 2;
    const float* k1 = kp1 + (long long)b * N * 2;
    const long long* mm = m0 + (long long)b * M;
    // normalize_points: mean, then scale = sqrt(2) / (mean distance to the mean + eps)
    double n = 0, sx0 = 0, sy0 = 0, sx1 = 0, sy1 = 0;
    for (int i = threadIdx.x; i < M; i += kEvalThreads) {
        const long long j = mm[i];
        if (j < 0) continue;
        n += 1; sx0 += k0[i * 2]; sy0 += k0[i * 2 + 1]; sx1 += k1[j * 2]; sy1 += k1[j * 2 + 1];
    }
    n = block_sum(n, scratch);
    sx0 = block_sum(sx0, scratch); sy0 = block_sum(sy0, scratch);
    sx1 = block_sum(sx1, scratch); sy1 = block_sum(sy1, scratch);
    if (n < 4) {  // kornia asserts >= 4 points; the reference then reports the error of an all-inf matrix (NaN)
        if (threadIdx.x < 9) H_out[b * 9 + threadIdx.x] = INFINITY;
        if (threadIdx.x == 0) err_out[b] = __int_as_float(0x7fc00000);
        return;
    }
    const double mx0 = sx0 / n, my0 = sy0 / n, mx1 = sx1 / n, my1 = sy1 / n;
    double d0 = 0, d1 = 0;
    for (int i = threadIdx.x; i < M; i += kEvalThreads) {
        const long long j = mm[i];
        if (j < 0) continue;
        d0 += sqrt((k0[i * 2] - mx0) * (k0[i * 2] - mx0) + (k0[i * 2 + 1] - my0) * (k0[i * 2 + 1] - my0));
        d1 += sqrt((k1[j * 2] - mx1) * (k1[j * 2] - mx1) + (k1[j * 2 + 1] - my1) * (k1[j * 2 + 1] - my1));
    }
    d0 = block_sum(d0, scratch);
    d1 = block_sum(d1, scratch);
    const double s0 = sqrt(2.0) / (d0 / n + 1e-8), s1 = sqrt(2.0) / (d1 / n + 1e-8);
    // A^T W A, upper triangle (45 entries), rows ax = [0 0 0 -x1 -y1 -1 y2x1 y2y1 y2], ay = [x1 y1 1 0 0 0 -x2x1 -x2y1 -x2]
    double u[45];
#pragma unroll
    for (int i = 0; i < 45; ++i) u[i] = 0.0;
    for (int i = threadIdx.x; i < M; i += kEvalThreads) {
        const long long j = mm[i];
        if (j < 0) continue;
        const double w = ms0[(long long)b * M + i];
        const double x1 = (k0[i * 2] - mx0) * s0, y1 = (k0[i * 2 + 1] - my0) * s0;
        const double x2 = (k1[j * 2] - mx1) * s1, y2 = (k1[j * 2 + 1] - my1) * s1;
        const double ax[9] = {0, 0, 0, -x1, -y1, -1, y2 * x1, y2 * y1, y2};
        const double ay[9] = {x1, y1, 1, 0, 0, 0, -x2 * x1, -x2 * y1, -x2};
        int e = 0;
#pragma unroll
        for (int p = 0; p < 9; ++p)
#pragma unroll
            for (int q = p; q < 9; ++q) u[e++] += w * (ax[p] * ax[q] + ay[p] * ay[q]);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int e = 0; e < 45; ++e) {
        double v = u[e];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) acc[e][warp] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double A[81], h[9];
        int e = 0;
        for (int p = 0; p < 9; ++p)
            for (int q = p; q < 9; ++q) {
                double s = 0;
                for (int w = 0; w < kEvalThreads / 32; ++w) s += acc[e][w];
                A[p * 9 + q] = A[q * 9 + p] = s;
                ++e;
            }
        jacobi_smallest_eigenvector9(A, h);
        // H = T2^-1 (Hn T1); T = [[s, 0, -s mx], [0, s, -s my], [0, 0, 1]]
        const double T1[9] = {s0, 0, -s0 * mx0, 0, s0, -s0 * my0, 0, 0, 1};
        const double T2i[9] = {1.0 / s1, 0, mx1, 0, 1.0 / s1, my1, 0, 0, 1};
        double t[9], H[9];
        mat3_mul(h, T1, t);
        mat3_mul(T2i, t, H);
        const double norm = H[8] + 1e-8;
        float Hf[9], G[9];
        for (int i = 0; i < 9; ++i) { Hf[i] = (float)(H[i] / norm); H_out[b * 9 + i] = Hf[i]; G[i] = H_gt[b * 9 + i]; }
        // homography_corner_error (homography.py:341-347)
        const float W = image_size[b * 2], Hh = image_size[b * 2 + 1];
        const float cx[4] = {0.f, W, W, 0.f}, cy[4] = {0.f, 0.f, Hh, Hh};
        float sum = 0.f;
        for (int c = 0; c < 4; ++c) {
            float ax_, ay_, gx, gy;
            warp_point(Hf, cx[c], cy[c], &ax_, &ay_);
            warp_point(G, cx[c], cy[c], &gx, &gy);
            sum += sqrtf((ax_ - gx) * (ax_ - gx) + (ay_ - gy) * (ay_ - gy));
        }
        err_out[b] = sum / 4.f;
    }
}

int launch_eval_homography(const float* kp0, const float* kp1, const long long* m0, int B, int M, int N, const float* H,
                           float* out, cudaStream_t stream) {
    B200_REQUIRE(kp0 && kp1 && m0 && H && out, "eval_matches_homography: null argument");
    B200_REQUIRE(B >= 0 && M >= 0 && N >= 0, "eval_matches_homography: bad sizes");
    if (B == 0) return 0;
    eval_homography_kernel<<<B, kEvalThreads, 0, stream>>>(kp0, kp1, m0, M, N, H, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_eval_epipolar(const float* kp0, const float* kp1, const long long* m0, int B, int M, int N, const float* cam0,
                         const float* cam1, const float* T01, float* out, cudaStream_t stream) {
    B200_REQUIRE(kp0 && kp1 && m0 && cam0 && cam1 && T01 && out, "eval_matches_epipolar: null argument");
    B200_REQUIRE(B >= 0 && M >= 0 && N >= 0, "eval_matches_epipolar: bad sizes");
    if (B == 0) return 0;
    eval_epipolar_kernel<<<B, kEvalThreads, 0, stream>>>(kp0, kp1, m0, M, N, cam0, cam1, T01, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_eval_dlt(const float* kp0, const float* kp1, const long long* m0, const float* ms0, int B, int M, int N,
                    const float* H_gt, const float* image_size, float* H_out, float* err_out, cudaStream_t stream) {
    B200_REQUIRE(kp0 && kp1 && m0 && ms0 && H_gt && image_size && H_out && err_out, "eval_homography_dlt: null argument");
    B200_REQUIRE(B >= 0 && M >= 0 && N >= 0, "eval_homography_dlt: bad sizes");
    if (B == 0) return 0;
    eval_dlt_kernel<<<B, kEvalThreads, 0, stream>>>(kp0, kp1, m0, ms0, M, N, H_gt, image_size, H_out, err_out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
