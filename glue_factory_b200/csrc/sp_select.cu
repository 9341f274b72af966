// Bandwidth-bound SuperPoint-open kernels, part 2: threshold + per-image top-k keypoint selection
// with a total order (score descending, then linear index ascending), and descriptor sampling.
// (reference: gluefactory/models/extractors/superpoint_open.py:20-31,52-56,155-179)
#include "common.cuh"
#include "sp_kernels.cuh"

namespace b200 {

constexpr int kSelThreads = 1024;
constexpr int kSelMaxK = 16384;  // keys staged in shared memory (128 KB)

__device__ __forceinline__ uint32_t order_key(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // monotone: larger float <=> larger key
}

__device__ __forceinline__ int block_sum(int v, int* scratch /*[33]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    if (warp == 0) {
        int t = lane < (kSelThreads >> 5) ? scratch[lane] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}

__global__ void __launch_bounds__(kSelThreads) count_candidates_kernel(const float* __restrict__ map, int HW,
                                                                       float thr, int* __restrict__ counts) {
    __shared__ int scratch[33];
    const float* m = map + (long long)blockIdx.x * HW;
    int c = 0;
    for (int i = threadIdx.x; i < HW; i += kSelThreads) c += m[i] > thr;
    const int n = block_sum(c, scratch);
    if (threadIdx.x == 0) counts[blockIdx.x] = n;
}

__global__ void __launch_bounds__(kSelThreads) select_topk_kernel(const float* __restrict__ map, int H, int W,
                                                                  float thr, int k, float* __restrict__ kpts,
                                                                  float* __restrict__ kscores, int* __restrict__ kidx,
                                                                  int* __restrict__ counts,
                                                                  const int* __restrict__ only_flagged) {
    extern __shared__ unsigned long long keys[];  // [pow2 >= k]
    if (only_flagged != nullptr && only_flagged[blockIdx.x] == 0) return;  // decided by the list kernel
    __shared__ int hist[257];
    __shared__ int scratch[33];
    __shared__ int thread_cnt[kSelThreads];
    __shared__ int s_fill, s_bin, s_remaining, s_counteq;
    const int HW = H * W, b = blockIdx.x, tid = threadIdx.x;
    const float* m = map + (long long)b * HW;

    int c = 0;
    for (int i = tid; i < HW; i += kSelThreads) c += m[i] > thr;
    const int n = block_sum(c, scratch);
    const int count = n < k ? n : k;
    if (tid == 0) { counts[b] = count; s_fill = 0; }
    int P = 1;
    while (P < count) P <<= 1;
    for (int i = tid; i < P; i += kSelThreads) keys[i] = 0ull;
    __syncthreads();

    if (n <= k) {
        // every candidate is kept, in row-major order: sort by index
        for (int i = tid; i < HW; i += kSelThreads) {
            const float v = m[i];
            if (v > thr) {
                const int slot = atomicAdd(&s_fill, 1);
                keys[slot] = (static_cast<unsigned long long>(~static_cast<uint32_t>(i)) << 32) | order_key(v);
            }
        }
    } else {
        // radix select of the k-th largest score key, most significant byte first
        uint32_t prefix = 0, pmask = 0;
        int remaining = k;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            for (int i = tid; i < 257; i += kSelThreads) hist[i] = 0;
            __syncthreads();
            for (int i0 = 0; i0 < HW; i0 += kSelThreads) {
                const int i = i0 + tid;
                int bin = 256;
                if (i < HW) {
                    const float v = m[i];
                    const uint32_t key = order_key(v);
                    if (v > thr && (key & pmask) == prefix) bin = (key >> shift) & 255;
                }
                const unsigned peers = __match_any_sync(0xffffffffu, bin);
                if ((tid & 31) == __ffs(peers) - 1) atomicAdd(&hist[bin], __popc(peers));
            }
            __syncthreads();
            if (tid == 0) {
                int cum = 0, bin = 255;
                for (; bin > 0; --bin) {
                    if (cum + hist[bin] >= remaining) break;
                    cum += hist[bin];
                }
                s_bin = bin;
                s_remaining = remaining - cum;
                s_counteq = hist[bin];
            }
            __syncthreads();
            prefix |= static_cast<uint32_t>(s_bin) << shift;
            pmask |= 0xFFu << shift;
            remaining = s_remaining;
            __syncthreads();
        }
        const uint32_t T = prefix;
        const int count_eq = s_counteq;
        const bool take_all_eq = count_eq == remaining;
        for (int i = tid; i < HW; i += kSelThreads) {
            const float v = m[i];
            if (!(v > thr)) continue;
            const uint32_t key = order_key(v);
            if (key > T || (take_all_eq && key == T)) {
                const int slot = atomicAdd(&s_fill, 1);
                keys[slot] = (static_cast<unsigned long long>(key) << 32) | ~static_cast<uint32_t>(i);
            }
        }
        if (!take_all_eq) {
            // ties at the k-th score: keep the `remaining` lowest linear indices (ordered pass)
            const int chunk = (HW + kSelThreads - 1) / kSelThreads;
            const int lo = tid * chunk, hi = min(HW, lo + chunk);
            int mine = 0;
            for (int i = lo; i < hi; ++i) {
                const float v = m[i];
                mine += (v > thr) && order_key(v) == T;
            }
            thread_cnt[tid] = mine;
            __syncthreads();
            if (tid == 0) {
                int run = 0;
                for (int t = 0; t < kSelThreads; ++t) {
                    const int x = thread_cnt[t];
                    thread_cnt[t] = run;
                    run += x;
                }
            }
            __syncthreads();
            int rank = thread_cnt[tid];
            for (int i = lo; i < hi && rank < remaining; ++i) {
                const float v = m[i];
                if ((v > thr) && order_key(v) == T) {
                    const int slot = atomicAdd(&s_fill, 1);
                    keys[slot] = (static_cast<unsigned long long>(T) << 32) | ~static_cast<uint32_t>(i);
                    ++rank;
                }
            }
        }
    }
    __syncthreads();
    // bitonic sort, descending, P a power of two (padding keys are 0 = smallest)
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (P >> 1); t += kSelThreads) {
                const int lo = ((t / stride) * stride << 1) + (t % stride), hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const unsigned long long a = keys[lo], c2 = keys[hi];
                if ((a < c2) == desc) { keys[lo] = c2; keys[hi] = a; }
            }
            __syncthreads();
        }
    }
    const bool by_index = n <= k;
    for (int j = tid; j < k; j += kSelThreads) {
        const long long o = (long long)b * k + j;
        if (j < count) {
            const unsigned long long key = keys[j];
            const uint32_t idx = by_index ? ~static_cast<uint32_t>(key >> 32) : ~static_cast<uint32_t>(key);
            const int y = idx / W, x = idx - y * W;
            kpts[o * 2] = x + 0.5f;
            kpts[o * 2 + 1] = y + 0.5f;
            kscores[o] = m[idx];
            kidx[o] = static_cast<int>(idx);
        } else {
            kpts[o * 2] = 0.f;
            kpts[o * 2 + 1] = 0.f;
            kscores[o] = 0.f;
            kidx[o] = -1;
        }
    }
}

int launch_count_candidates(const float* nms_map, int B, int H, int W, float threshold, int* counts,
                            cudaStream_t stream) {
    if (B == 0) return 0;
    count_candidates_kernel<<<B, kSelThreads, 0, stream>>>(nms_map, H * W, threshold, counts);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_select_topk_flagged(const float* nms_map, int B, int H, int W, float threshold, int k, float* kpts,
                               float* kscores, int* kidx, int* counts, const int* only_flagged, cudaStream_t stream) {
    B200_REQUIRE(k >= 1 && k <= kSelMaxK, "select_topk: k=%d outside [1, %d]", k, kSelMaxK);
    if (B == 0) return 0;
    int P = 1;
    while (P < k) P <<= 1;
    const size_t smem = (size_t)P * sizeof(unsigned long long);
    static PerDeviceSize configured;
    if (smem > configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(select_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.cur() = smem;
    }
    select_topk_kernel<<<B, kSelThreads, smem, stream>>>(nms_map, H, W, threshold, k, kpts, kscores, kidx, counts,
                                                         only_flagged);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_select_topk(const float* nms_map, int B, int H, int W, float threshold, int k, float* kpts,
                       float* kscores, int* kidx, int* counts, cudaStream_t stream) {
    return launch_select_topk_flagged(nms_map, B, H, W, threshold, k, kpts, kscores, kidx, counts, nullptr, stream);
}

// =============================================================================================
// Descriptor sampling: bilinear grid_sample(align_corners=False, zero padding) of the dense NHWC
// descriptor map at the keypoints, then L2 normalisation.  One warp per keypoint, 8 channels per
// lane (C = 256), coalesced 1 KB corner reads.
// =============================================================================================
__global__ void __launch_bounds__(256) sample_descriptors_kernel(const float* __restrict__ dense, int Hc, int Wc,
                                                                 int C, const float* __restrict__ kpts, int K,
                                                                 int total, float stride_px, int mode,
                                                                 float* __restrict__ out) {
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (gw >= total) return;
    const int b = gw / K;
    // keypoints hold (x + 0.5, y + 0.5) in pixels of the full-resolution image
    const float kx = kpts[(long long)gw * 2], ky = kpts[(long long)gw * 2 + 1];
    // reference: g = (kp + 0.5) / ([w, h] * s) * 2 - 1 ; grid_sample unnormalises with ((g + 1) * size - 1) / 2
    float ix, iy;
    if (mode == 2) {
        // MagicLeap legacy sampling (gluefactory_nonfree/superpoint.py:117-133): integer keypoint k = kp - 0.5,
        // g = (k - s/2 + 0.5) / (w*s - s/2 - 0.5) * 2 - 1, grid_sample(align_corners=True)
        const float hs = stride_px * 0.5f;
        const float gx = (kx - 0.5f - hs + 0.5f) / (Wc * stride_px - hs - 0.5f) * 2.f - 1.f;
        const float gy = (ky - 0.5f - hs + 0.5f) / (Hc * stride_px - hs - 0.5f) * 2.f - 1.f;
        ix = (gx + 1.f) * 0.5f * (Wc - 1);
        iy = (gy + 1.f) * 0.5f * (Hc - 1);
    } else {
        // mode 0: SuperPoint-open, g = (k + 0.5) / (w*s) (the keypoints passed in already hold k + 0.5);
        // mode 1: nonfree "fix_sampling" (:138-149), g = k / (w*s); both align_corners=False
        const float ox = mode == 1 ? kx - 0.5f : kx, oy = mode == 1 ? ky - 0.5f : ky;
        const float gx = ox / (Wc * stride_px) * 2.f - 1.f, gy = oy / (Hc * stride_px) * 2.f - 1.f;
        ix = ((gx + 1.f) * Wc - 1.f) * 0.5f;
        iy = ((gy + 1.f) * Hc - 1.f) * 0.5f;
    }
    const float fx = floorf(ix), fy = floorf(iy);
    const int x0 = (int)fx, y0 = (int)fy;
    const float wx1 = ix - fx, wy1 = iy - fy, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
    const float* base = dense + (long long)b * Hc * Wc * C;
    float acc[8];
    float ss = 0.f;
    for (int c0 = lane * 8; c0 < C; c0 += 256) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
        for (int corner = 0; corner < 4; ++corner) {
            const int xx = x0 + (corner & 1), yy = y0 + (corner >> 1);
            const float wgt = ((corner & 1) ? wx1 : wx0) * ((corner >> 1) ? wy1 : wy0);
            if (xx < 0 || xx >= Wc || yy < 0 || yy >= Hc) continue;
            const float4* p = reinterpret_cast<const float4*>(base + ((long long)yy * Wc + xx) * C + c0);
            const float4 a = __ldg(p), d = __ldg(p + 1);
            acc[0] += a.x * wgt; acc[1] += a.y * wgt; acc[2] += a.z * wgt; acc[3] += a.w * wgt;
            acc[4] += d.x * wgt; acc[5] += d.y * wgt; acc[6] += d.z * wgt; acc[7] += d.w * wgt;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += acc[i] * acc[i];
        // C == 256: a single iteration, values stay in registers until normalised below
    }
    ss = warp_sum(ss);
    const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
    if (lane * 8 < C) {
        float4* o = reinterpret_cast<float4*>(out + (long long)gw * C + lane * 8);
        o[0] = make_float4(acc[0] * inv, acc[1] * inv, acc[2] * inv, acc[3] * inv);
        o[1] = make_float4(acc[4] * inv, acc[5] * inv, acc[6] * inv, acc[7] * inv);
    }
}

int launch_sample_descriptors(const float* dense, int B, int Hc, int Wc, int C, const float* kpts, int K,
                              int stride_px, int mode, float* out, cudaStream_t stream) {
    B200_REQUIRE(mode >= 0 && mode <= 2, "sample_descriptors: mode %d outside [0, 2]", mode);
    B200_REQUIRE(C == 256, "sample_descriptors: descriptor_dim must be 256, got %d", C);
    const int total = B * K;
    if (total == 0) return 0;
    const int blocks = (total * 32 + 255) / 256;
    sample_descriptors_kernel<<<blocks, 256, 0, stream>>>(dense, Hc, Wc, C, kpts, K, total, (float)stride_px, mode, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
