// Bandwidth-bound SuperPoint detection, fast path: register-blocked separable NMS (radius 1..4) that
// also emits a compact list of the surviving maxima, and a per-image top-k over that list.
//
// nms_list_kernel<R>: a CTA stages a 128 x 96 window of the score map (output tile + 5R halo: three
// dependent max-pools of radius R plus two mask dilations).  Float pools are separable sliding-window
// maxima held in registers (row pass: lanes along y, odd smem pitch; column pass: lanes along x);
// the max / suppression masks live as bit rows (ballot), so the two dilations cost a few word ops.
// Comparisons only => bit-exact with the reference's max_pool2d formulation
// (reference: gluefactory/models/extractors/superpoint_open.py:34-49,147-152).
//
// select_list_kernel: per image, radix select of the k-th largest composite key
// (score key << 32 | ~linear index: a total order, score descending then index ascending) over the
// candidate list, then a bitonic sort of the k winners in shared memory
// (reference: superpoint_open.py:52-56,155-179).  Images whose list cannot decide the result (list
// overflow, or a negative threshold with fewer than k positive maxima so that zero-valued pixels
// are selected) are flagged and handled by the full-map kernel in sp_select.cu.
#include "common.cuh"
#include "sp_kernels.cuh"

namespace b200 {

#ifndef NMS_EH
#define NMS_EH 96
#endif
#ifndef NMS_MINB
#define NMS_MINB 2
#endif
constexpr int kEW = 128, kEH = NMS_EH, kSeg = 16;
#ifndef NMS_THREADS
#define NMS_THREADS 768
#endif
constexpr int kNmsFastThreads = NMS_THREADS;

__device__ __forceinline__ uint32_t order_key32(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// three-input maximum (FMNMX3 on sm_100: same issue rate as the two-input form)
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// o[i] = max(v[i .. i+2R]) for i < kSeg: sliding maxima of three, combined three at a time
// (R = 3: 20 + 16 instructions per 16 outputs instead of 56 two-input ones)
template <int R>
__device__ __forceinline__ void window_max(const float (&v)[kSeg + 2 * R], float (&o)[kSeg]) {
    constexpr int n = kSeg + 2 * R;
    if constexpr (R == 1) {
#pragma unroll
        for (int i = 0; i < kSeg; ++i) o[i] = fmax3(v[i], v[i + 1], v[i + 2]);
    } else {
        float m3[n - 2];
#pragma unroll
        for (int i = 0; i < n - 2; ++i) m3[i] = fmax3(v[i], v[i + 1], v[i + 2]);
        if constexpr (R == 2) {
#pragma unroll
            for (int i = 0; i < kSeg; ++i) o[i] = fmax3(m3[i], v[i + 3], v[i + 4]);
        } else if constexpr (R == 3) {
#pragma unroll
            for (int i = 0; i < kSeg; ++i) o[i] = fmax3(m3[i], m3[i + 3], v[i + 6]);
        } else {
            static_assert(R == 4, "radius 1..4");
#pragma unroll
            for (int i = 0; i < kSeg; ++i) o[i] = fmax3(m3[i], m3[i + 3], m3[i + 6]);
        }
    }
}

template <int R>
struct NmsGeom {
    static constexpr int kPitchS = kEW + 2 * R + 1;  // odd: lanes along y hit distinct banks
    static constexpr int kPitchT = kEW + 1;
    static constexpr int kTileW = kEW - 10 * R, kTileH = kEH - 10 * R;
    static constexpr int kSFloats = kEH * kPitchS;
    static constexpr int kTFloats = (kEH + 2 * R) * kPitchT;
    static constexpr int kPitchM = kEW + 2 * R;                 // halfword masks: R zero guard columns each side
    static constexpr int kMaskHalfs = (kEH / kSeg) * kPitchM;   // per mask array (even: cleared as 32-bit words)
    static constexpr size_t kSmemBytes = (size_t)(kSFloats + kTFloats) * 4 + (size_t)2 * kMaskHalfs * 2;
};

// row pass: tmp[y][x] = max_{|dx|<=R} s(y, x+dx).  In the refinement rounds `s` already holds the suppressed scores
// (apply_supp), so the pass is the same in every round.
template <int R>
__device__ __forceinline__ void row_pass(const float* __restrict__ s, float* __restrict__ tmp) {
    using G = NmsGeom<R>;
    for (int item = threadIdx.x; item < kEH * (kEW / kSeg); item += kNmsFastThreads) {
        const int y = item % kEH, x0 = (item / kEH) * kSeg;
        float v[kSeg + 2 * R], o[kSeg];
        const float* row = s + y * G::kPitchS + x0;  // column x0-R of the image window (array is shifted by R)
#pragma unroll
        for (int j = 0; j < kSeg + 2 * R; ++j) v[j] = row[j];
        window_max<R>(v, o);
        float* dst = tmp + (y + R) * G::kPitchT + x0;
#pragma unroll
        for (int i = 0; i < kSeg; ++i) dst[i] = o[i];
    }
}

// The max / suppression masks are kept as 16-bit words ALONG y: halfword (h, x) holds rows 16 h .. 16 h + 15 of column
// x, stored [h][R guard | kEW columns | R guard].  A thread of the column pass owns exactly one halfword (no ballots,
// no shared read-modify-write between lanes), the dilation works on whole halfwords, and the suppression is applied
// by the thread that computed it.

// column pass + comparison: bit(y, x) = s == max_{|dy|<=R} tmp(y+dy, x) [and not suppressed]
// (wx0, wy0) = image coordinates of the staged window's first pixel: pixels outside the image (staged as -inf, where
// a window of nothing but padding would compare equal) are masked by position, one AND per halfword instead of a
// second comparison per pixel.
template <int R, bool ROUND>
__device__ __forceinline__ void col_pass(const float* __restrict__ s, const float* __restrict__ tmp,
                                         uint16_t* __restrict__ maskh, const uint16_t* __restrict__ supph, int wx0, int wy0,
                                         int H, int W) {
    using G = NmsGeom<R>;
    for (int item = threadIdx.x; item < kEW * (kEH / kSeg); item += kNmsFastThreads) {
        const int x = item & (kEW - 1), h = item / kEW, y0 = h * kSeg;
        const int gy = wy0 + y0, gx = wx0 + x;
        const int lo = max(0, -gy), hi = min(kSeg, H - gy);  // rows [lo, hi) of this halfword lie inside the image
        const uint32_t inside = (gx >= 0 && gx < W && hi > lo) ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
        float v[kSeg + 2 * R], o[kSeg];
#pragma unroll
        for (int j = 0; j < kSeg + 2 * R; ++j) v[j] = tmp[(y0 + j) * G::kPitchT + x];
        window_max<R>(v, o);
        uint32_t bits = 0u;
#pragma unroll
        for (int i = 0; i < kSeg; ++i) {
            const float sv = s[(y0 + i) * G::kPitchS + R + x];
            bits |= (sv == o[i]) ? (1u << i) : 0u;
        }
        bits &= inside;  // out-of-image pixels never become maxima
        uint16_t* m = maskh + h * G::kPitchM + R + x;
        if constexpr (ROUND) *m = (uint16_t)(*m | (bits & ~(uint32_t)supph[h * G::kPitchM + R + x]));
        else *m = (uint16_t)bits;
    }
}

// supp = dilate(mask, R) (max_pool of the 0/1 mask > 0), then s <- where(supp, 0, s) in place (padding stays -inf).
// The suppression mask only grows from round to round, so the second round's where(supp2, 0, s) equals
// where(supp2, 0, where(supp1, 0, s)): one float buffer serves all rounds; the original scores of the final maxima are
// re-read from global memory by the output pass.
template <int R>
__device__ __forceinline__ void dilate_apply(const uint16_t* __restrict__ maskh, uint16_t* __restrict__ supph,
                                             float* __restrict__ s) {
    using G = NmsGeom<R>;
    constexpr int kH = kEH / kSeg;
    for (int item = threadIdx.x; item < kEW * kH; item += kNmsFastThreads) {
        const int x = item & (kEW - 1), h = item / kEW;
        uint32_t up = 0u, mid = 0u, dn = 0u;  // rows above / of / below this halfword, OR over the 2R+1 columns
        const uint16_t* row = maskh + h * G::kPitchM + x;  // column x - R (guard columns are zero)
#pragma unroll
        for (int dx = 0; dx <= 2 * R; ++dx) {
            mid |= row[dx];
            if (h > 0) up |= row[dx - G::kPitchM];
            if (h < kH - 1) dn |= row[dx + G::kPitchM];
        }
        const unsigned long long c = (unsigned long long)up | ((unsigned long long)mid << 16) | ((unsigned long long)dn << 32);
        unsigned long long d = c;
#pragma unroll
        for (int k = 1; k <= R; ++k) d |= (c << k) | (c >> k);
        const uint32_t sp = (uint32_t)(d >> 16) & 0xffffu;
        supph[h * G::kPitchM + R + x] = (uint16_t)sp;
        if (sp != 0u) {
            float* px = s + (h * kSeg) * G::kPitchS + R + x;
#pragma unroll
            for (int i = 0; i < kSeg; ++i)
                if (((sp >> i) & 1u) && px[i * G::kPitchS] != -INFINITY) px[i * G::kPitchS] = 0.f;
        }
    }
}

template <int R>
__global__ void __launch_bounds__(kNmsFastThreads, NMS_MINB)
nms_list_kernel(const float* __restrict__ scores, int H, int W, int pad, const int* __restrict__ image_hw,
                float* __restrict__ out, unsigned long long* __restrict__ cand, int* __restrict__ cand_count, int cap) {
    using G = NmsGeom<R>;
    extern __shared__ uint8_t nms_fast_smem[];
    __shared__ int s_count, s_base;
    float* s = reinterpret_cast<float*>(nms_fast_smem);
    float* tmp = s + G::kSFloats;
    uint16_t* maskh = reinterpret_cast<uint16_t*>(tmp + G::kTFloats);  // [kEH / 16][kPitchM]
    uint16_t* supph = maskh + G::kMaskHalfs;
    const int b = blockIdx.z, tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int ox = blockIdx.x * G::kTileW, oy = blockIdx.y * G::kTileH;  // output tile origin
    const int x0 = ox - 5 * R, y0 = oy - 5 * R;                          // staged window origin
    const float* src = scores + (long long)b * H * W;

    // stage the window: one warp per row, lanes along x (coalesced), -inf outside the image and in the R pad columns
    static_assert(kEH % (kNmsFastThreads / 32) == 0, "whole rows per warp");
#pragma unroll
    for (int it = 0; it < kEH / (kNmsFastThreads / 32); ++it) {  // unrolled: all of a warp's row requests in flight
        const int yy = warp + it * (kNmsFastThreads / 32);
        const int y = y0 + yy;
        const bool row_ok = y >= 0 && y < H;
        const float* grow = src + (long long)y * W;
        float* srow = s + yy * G::kPitchS;
#pragma unroll
        for (int c = lane; c < G::kPitchS; c += 32) {
            const int x = x0 + c - R;
            const bool in = row_ok && c >= R && c < R + kEW && x >= 0 && x < W;
            srow[c] = in ? __ldg(grow + x) : -INFINITY;
        }
    }
    for (int i = tid; i < R * G::kPitchT; i += kNmsFastThreads) {
        tmp[i] = -INFINITY;
        tmp[(kEH + R) * G::kPitchT + i] = -INFINITY;
    }
    for (int i = tid; i < 2 * G::kMaskHalfs / 2; i += kNmsFastThreads) reinterpret_cast<uint32_t*>(maskh)[i] = 0u;
    if (tid == 0) s_count = 0;
    __syncthreads();

    row_pass<R>(s, tmp);
    __syncthreads();
    col_pass<R, false>(s, tmp, maskh, nullptr, x0, y0, H, W);
    __syncthreads();
#pragma unroll 1
    for (int round = 0; round < 2; ++round) {
        dilate_apply<R>(maskh, supph, s);
        __syncthreads();
        row_pass<R>(s, tmp);
        __syncthreads();
        col_pass<R, true>(s, tmp, maskh, supph, x0, y0, H, W);
        __syncthreads();
    }

    // bottom / right border from the per-image valid size when given (gluefactory_nonfree/superpoint.py:244-254)
    const int bh = image_hw ? image_hw[2 * b] : H, bw = image_hw ? image_hw[2 * b + 1] : W;
    float* dst = out + (long long)b * H * W;
    unsigned long long* list = cand ? cand + (long long)b * cap : nullptr;
    // candidate records of this tile are collected in shared memory (the score window and the pooling scratch are
    // free now and hold kTileW * kTileH records) and appended to the image's list with ONE global atomic per CTA
    unsigned long long* rec = reinterpret_cast<unsigned long long*>(s);
    static_assert((size_t)G::kTileW * G::kTileH * 8 <= (size_t)(G::kSFloats + G::kTFloats) * 4,
                  "candidate records must fit the window + scratch");
    // the maxima's scores come from global memory (the staged copy
    // was overwritten by the suppression): a few percent of the pixels, L2 hits
    // the tile's pixels as one linear index space (98 x 66 for R = 3: whole warps of consecutive pixels; row-wise
    // segments of 32 would leave the fourth segment of every row with 2 live lanes)
    constexpr int kTilePx = G::kTileW * G::kTileH;
    const bool has_border = pad > 0;
    for (int base = warp * 32; base < kTilePx; base += kNmsFastThreads) {
        const int i = base + lane;
        const int ty = i / G::kTileW, tx = i - ty * G::kTileW;
        const int y = oy + ty, x = ox + tx;
        const bool valid = i < kTilePx && y < H && x < W;
        float v = 0.f;
        bool keep = false;
        if (valid) {
            const int ex = tx + 5 * R, ey = ty + 5 * R;
            const int gi = y * W + x;  // < 2^31: one score map
            const bool mx = (maskh[(ey >> 4) * G::kPitchM + R + ex] >> (ey & 15)) & 1u;
            if (mx) v = __ldg(src + gi);
            if (has_border && (y < pad || x < pad || y >= bh - pad || x >= bw - pad)) v = -1.f;
            dst[gi] = v;
            keep = v > 0.f;
        }
        if (list != nullptr) {
            const uint32_t ballot = __ballot_sync(0xffffffffu, keep);
            if (ballot != 0u) {
                int slot = 0;
                if (lane == 0) slot = atomicAdd(&s_count, __popc(ballot));
                slot = __shfl_sync(0xffffffffu, slot, 0);
                if (keep)
                    rec[slot + __popc(ballot & ((1u << lane) - 1u))] =
                        (static_cast<unsigned long long>(order_key32(v)) << 32) | ~static_cast<uint32_t>(y * W + x);
            }
        }
    }
    if (list != nullptr) {
        __syncthreads();
        const int n = s_count;
        if (n > 0) {
            if (tid == 0) s_base = atomicAdd(cand_count + b, n);  // the count keeps growing past `cap`: overflow flag
            __syncthreads();
            const int base = s_base;
            for (int i = tid; i < n; i += kNmsFastThreads)
                if (base + i < cap) list[base + i] = rec[i];
        }
    }
}

template <int R>
static int launch_nms_fast(const float* scores, int B, int H, int W, int pad, const int* image_hw, float* out,
                           unsigned long long* cand, int* cand_count, int cap, cudaStream_t stream) {
    using G = NmsGeom<R>;
    auto kern = nms_list_kernel<R>;
    static PerDeviceSize configured;
    if (!configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::kSmemBytes));
        configured.cur() = 1;
    }
    dim3 grid((W + G::kTileW - 1) / G::kTileW, (H + G::kTileH - 1) / G::kTileH, B);
    kern<<<grid, kNmsFastThreads, G::kSmemBytes, stream>>>(scores, H, W, pad, image_hw, out, cand, cand_count, cap);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

bool nms_fast_supported(int radius) { return radius >= 1 && radius <= 4; }

int launch_nms_list(const float* scores, int B, int H, int W, int radius, int pad, const int* image_hw, float* out,
                    unsigned long long* cand, int* cand_count, int cap, cudaStream_t stream) {
    B200_REQUIRE(nms_fast_supported(radius), "nms (list variant): radius %d outside [1, 4]", radius);
    B200_REQUIRE((long long)H * W < (1ll << 31), "nms: image too large");
    if (B == 0) return 0;
    if (cand != nullptr) {
        B200_REQUIRE(cand_count != nullptr && cap > 0, "nms: candidate list needs a counter and a capacity");
        B200_CHECK_CUDA(cudaMemsetAsync(cand_count, 0, sizeof(int) * 2 * B, stream));
    }
    switch (radius) {
        case 1: return launch_nms_fast<1>(scores, B, H, W, pad, image_hw, out, cand, cand_count, cap, stream);
        case 2: return launch_nms_fast<2>(scores, B, H, W, pad, image_hw, out, cand, cand_count, cap, stream);
        case 3: return launch_nms_fast<3>(scores, B, H, W, pad, image_hw, out, cand, cand_count, cap, stream);
        default: return launch_nms_fast<4>(scores, B, H, W, pad, image_hw, out, cand, cand_count, cap, stream);
    }
}

// =============================================================================================
// top-k over the candidate list
// =============================================================================================
constexpr int kListThreads = 1024;
constexpr int kListMaxK = 16384;

__device__ __forceinline__ int block_sum_1024(int v, int* scratch /*[33]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    if (warp == 0) {
        int t = scratch[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}

__global__ void __launch_bounds__(kListThreads)
select_list_kernel(const float* __restrict__ map, int H, int W, float thr, int k,
                   const unsigned long long* __restrict__ cand, int* __restrict__ cand_count, int cap,
                   float* __restrict__ kpts, float* __restrict__ kscores, int* __restrict__ kidx,
                   int* __restrict__ counts) {
    extern __shared__ unsigned long long keys[];  // [pow2 >= k]
    __shared__ int hist[256];
    __shared__ int scratch[33];
    __shared__ int s_fill, s_bin, s_remaining, s_done;
    const int b = blockIdx.x, tid = threadIdx.x, B = gridDim.x;
    int* flag = cand_count + B + b;
    const int n_raw = cand_count[b];
    if (n_raw > cap) {  // list overflow: the full-map kernel decides
        if (tid == 0) *flag = 1;
        return;
    }
    const unsigned long long* list = cand + (long long)b * cap;
    const uint32_t thr_key = order_key32(thr);
    int c = 0;
    for (int i = tid; i < n_raw; i += kListThreads) c += static_cast<uint32_t>(list[i] >> 32) > thr_key;
    const int n = block_sum_1024(c, scratch);
    if (thr < 0.f && n < k) {  // zero-valued pixels would be selected: not in the list
        if (tid == 0) *flag = 1;
        return;
    }
    const int count = n < k ? n : k;
    if (tid == 0) { *flag = 0; counts[b] = count; s_fill = 0; }
    int P = 1;
    while (P < count) P <<= 1;
    for (int i = tid; i < P; i += kListThreads) keys[i] = 0ull;
    __syncthreads();
    // fewer candidates than k: the reference returns them in row-major order (no top-k call).  With a negative
    // threshold the reference's candidate set is every interior pixel, i.e. always more than k, and its
    // top-k sorts by score: n == k positive maxima then still come out score-descending.
    const bool by_index = thr >= 0.f && n <= k;
    if (by_index) {
        for (int i = tid; i < n_raw; i += kListThreads) {
            const unsigned long long e = list[i];
            if (static_cast<uint32_t>(e >> 32) > thr_key) {
                const int slot = atomicAdd(&s_fill, 1);
                keys[slot] = (e << 32) | (e >> 32);  // (~index, score key): descending sort = row-major order
            }
        }
    } else {
        // radix select of the k-th largest 64-bit composite key; keys are unique, so no tie handling
        unsigned long long prefix = 0ull, pmask = 0ull;
        int remaining = k;
        for (int pass = 0; pass < 8; ++pass) {
            const int shift = 56 - 8 * pass;
            for (int i = tid; i < 256; i += kListThreads) hist[i] = 0;
            __syncthreads();
            for (int i = tid; i < n_raw; i += kListThreads) {
                const unsigned long long e = list[i];
                if (static_cast<uint32_t>(e >> 32) > thr_key && (e & pmask) == prefix)
                    atomicAdd(&hist[(int)((e >> shift) & 255ull)], 1);
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
                s_done = hist[bin] == remaining - cum;  // the whole bin is taken: threshold = prefix
            }
            __syncthreads();
            prefix |= static_cast<unsigned long long>(s_bin) << shift;
            pmask |= 0xFFull << shift;
            remaining = s_remaining;
            const int done = s_done;
            __syncthreads();
            if (done) break;
        }
        for (int i = tid; i < n_raw; i += kListThreads) {
            const unsigned long long e = list[i];
            if (static_cast<uint32_t>(e >> 32) > thr_key && e >= prefix) {
                const int slot = atomicAdd(&s_fill, 1);
                if (slot < P) keys[slot] = e;
            }
        }
    }
    __syncthreads();
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (P >> 1); t += kListThreads) {
                const int lo = ((t / stride) * stride << 1) + (t % stride), hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const unsigned long long a = keys[lo], c2 = keys[hi];
                if ((a < c2) == desc) { keys[lo] = c2; keys[hi] = a; }
            }
            __syncthreads();
        }
    }
    const float* m = map + (long long)b * H * W;
    for (int j = tid; j < k; j += kListThreads) {
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

int launch_select_list(const float* nms_map, int B, int H, int W, float threshold, int k,
                       const unsigned long long* cand, int* cand_count, int cap, float* kpts, float* kscores,
                       int* kidx, int* counts, cudaStream_t stream) {
    B200_REQUIRE(k >= 1 && k <= kListMaxK, "select_topk: k=%d outside [1, %d]", k, kListMaxK);
    B200_REQUIRE(cand && cand_count && cap > 0, "select_topk (list variant): null candidate list");
    if (B == 0) return 0;
    int P = 1;
    while (P < k) P <<= 1;
    const size_t smem = (size_t)P * sizeof(unsigned long long);
    static PerDeviceSize configured;
    if (smem > configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(select_list_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.cur() = smem;
    }
    select_list_kernel<<<B, kListThreads, smem, stream>>>(nms_map, H, W, threshold, k, cand, cand_count, cap, kpts,
                                                          kscores, kidx, counts);
    B200_CHECK_CUDA(cudaGetLastError());
    // flagged images (cand_count[B + b] != 0) are redone from the full map
    return launch_select_topk_flagged(nms_map, B, H, W, threshold, k, kpts, kscores, kidx, counts, cand_count + B,
                                      stream);
}

}  // namespace b200
