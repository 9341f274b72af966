// Bandwidth-bound SuperPoint-open kernels: first convolution (C_in = 1), NMS + border removal,
// per-image top-k keypoint selection, bilinear descriptor sampling.
// (reference: gluefactory/models/extractors/superpoint_open.py:20-56,123-217)
#include "common.cuh"
#include "sp_kernels.cuh"

namespace b200 {

// =============================================================================================
// conv1a: gray (or RGB -> gray) image, 3x3 conv 1 -> 64, bias, ReLU, BN affine, NHWC bf16 output.
// K = 9 is useless on tensor cores, so this runs on the CUDA cores.  A thread owns 8 output channels
// (their 72 weights live in registers) and walks down a column of the tile, so a warp store covers
// 4 pixels x 128 B of contiguous NHWC output and each new pixel costs 3 shared-memory reads.
// =============================================================================================
constexpr int kC1TileW = 32, kC1TileH = 32;

__global__ void __launch_bounds__(256, 2) conv1a_kernel(const void* __restrict__ img, int u8, int C, int H, int W,
                                                        const float* __restrict__ w, const float* __restrict__ bias,
                                                        const float* __restrict__ scale, const float* __restrict__ shift,
                                                        __nv_bfloat16* __restrict__ out) {
    __shared__ float tile[kC1TileH + 2][kC1TileW + 2];
    const int tid = threadIdx.x;
    const int b = blockIdx.z, x0 = blockIdx.x * kC1TileW, y0 = blockIdx.y * kC1TileH;
    const long long img0 = (long long)b * C * H * W;
    // uint8 pixels: value / 255 correctly rounded == float32(value / 255.0) of the host conversion (image.py:121)
    auto pix = [&](long long at) {
        return u8 ? u8_unit(__ldg(static_cast<const unsigned char*>(img) + at))
                  : __ldg(static_cast<const float*>(img) + at);
    };
    for (int i = tid; i < (kC1TileH + 2) * (kC1TileW + 2); i += 256) {
        const int ty = i / (kC1TileW + 2), tx = i - ty * (kC1TileW + 2);
        const int y = y0 + ty - 1, x = x0 + tx - 1;
        float v = 0.f;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            const long long o = img0 + (long long)y * W + x;
            if (C == 3) {
                // same association order as (image * [0.299, 0.587, 0.114]).sum(1)
                v = pix(o) * 0.299f;
                v += pix(o + (long long)H * W) * 0.587f;
                v += pix(o + 2ll * H * W) * 0.114f;
            } else {
                v = pix(o);
            }
        }
        tile[ty][tx] = v;
    }
    const int cg = tid & 7, px = tid >> 3;  // channel group (8 channels), pixel column inside the tile
    float wr[8][9], bi[8], sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = cg * 8 + j;
#pragma unroll
        for (int t = 0; t < 9; ++t) wr[j][t] = __ldg(w + c * 9 + t);
        bi[j] = __ldg(bias + c);
        sc[j] = __ldg(scale + c);
        sh[j] = __ldg(shift + c);
    }
    __syncthreads();
    const int x = x0 + px;
    if (x >= W) return;
    float in[9];
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
        in[3 + dx] = tile[0][px + dx];
        in[6 + dx] = tile[1][px + dx];
    }
    const int rows = min(kC1TileH, H - y0);
    __nv_bfloat16* o = out + ((long long)(b * H + y0) * W + x) * 64 + cg * 8;
#pragma unroll 2
    for (int ty = 0; ty < rows; ++ty) {
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
            in[dx] = in[3 + dx];
            in[3 + dx] = in[6 + dx];
            in[6 + dx] = tile[ty + 2][px + dx];
        }
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float a = bi[j];
#pragma unroll
            for (int t = 0; t < 9; ++t) a = fmaf(in[t], wr[j][t], a);
            a = fmaxf(a, 0.f);
            v[j] = a * sc[j] + sh[j];
        }
        *reinterpret_cast<uint4*>(o) =
            make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
        o += (long long)W * 64;
    }
}

int launch_conv1a(const void* img, int u8, int B, int C, int H, int W, const float* w, const float* bias,
                  const float* scale, const float* shift, void* out, cudaStream_t stream) {
    B200_REQUIRE(C == 1 || C == 3, "conv1a: image must have 1 or 3 channels, got %d", C);
    if (B == 0) return 0;
    dim3 grid((W + kC1TileW - 1) / kC1TileW, (H + kC1TileH - 1) / kC1TileH, B);
    conv1a_kernel<<<grid, 256, 0, stream>>>(img, u8, C, H, W, w, bias, scale, shift,
                                            reinterpret_cast<__nv_bfloat16*>(out));
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// =============================================================================================
// NMS (plateau preserving, two refinement rounds) + border removal, one pass over the score map.
// A CTA owns a 64x64 output tile and stages it with a 5r halo in shared memory: the result at a
// pixel depends on scores up to 5r away (three dependent pools of radius r plus two mask pools).
// Comparisons only => bit-exact with the reference's max_pool2d formulation.
// =============================================================================================
constexpr int kNmsTile = 64;
constexpr int kNmsThreads = 512;

template <typename T, typename F>
__device__ __forceinline__ void pool_rows(int EH, int EW, int r, T* dst, F src) {
    // dst[y][x] = max_{|dx| <= r} src(y, x+dx), window clipped to the staged region
    for (int i = threadIdx.x; i < EH * EW; i += kNmsThreads) {
        const int y = i / EW, x = i - y * EW;
        const int lo = max(0, x - r), hi = min(EW - 1, x + r);
        T m = src(y, lo);
        for (int xx = lo + 1; xx <= hi; ++xx) {
            const T v = src(y, xx);
            m = v > m ? v : m;
        }
        dst[i] = m;
    }
}
template <typename T>
__device__ __forceinline__ T pool_col_at(const T* tmp, int EH, int EW, int r, int y, int x) {
    const int lo = max(0, y - r), hi = min(EH - 1, y + r);
    T m = tmp[lo * EW + x];
    for (int yy = lo + 1; yy <= hi; ++yy) {
        const T v = tmp[yy * EW + x];
        m = v > m ? v : m;
    }
    return m;
}

__global__ void __launch_bounds__(kNmsThreads) nms_border_kernel(const float* __restrict__ scores, int H, int W,
                                                                 int r, int pad, float* __restrict__ out) {
    extern __shared__ uint8_t nms_smem[];
    const int halo = 5 * r, E = kNmsTile + 2 * halo, EE = E * E;
    float* s = reinterpret_cast<float*>(nms_smem);
    float* tmp = s + EE;
    uint8_t* mask = reinterpret_cast<uint8_t*>(tmp + EE);
    uint8_t* supp = mask + EE;
    uint8_t* tmpb = reinterpret_cast<uint8_t*>(tmp);  // aliases tmp (used at different times)
    const int b = blockIdx.z, x0 = blockIdx.x * kNmsTile - halo, y0 = blockIdx.y * kNmsTile - halo;
    const float* src = scores + (long long)b * H * W;
    for (int i = threadIdx.x; i < EE; i += kNmsThreads) {
        const int y = y0 + i / E, x = x0 + i % E;
        s[i] = (y >= 0 && y < H && x >= 0 && x < W) ? src[(long long)y * W + x] : -INFINITY;
    }
    __syncthreads();
    // round 0: mask = (s == pool(s))
    pool_rows<float>(E, E, r, tmp, [&](int y, int x) { return s[y * E + x]; });
    __syncthreads();
    for (int i = threadIdx.x; i < EE; i += kNmsThreads) {
        const int y = i / E, x = i - y * E;
        // out-of-image pixels (-inf, the max_pool2d padding value) never become maxima
        mask[i] = s[i] != -INFINITY && s[i] == pool_col_at<float>(tmp, E, E, r, y, x);
    }
    __syncthreads();
    for (int round = 0; round < 2; ++round) {
        // supp = pool(mask) > 0
        pool_rows<uint8_t>(E, E, r, tmpb, [&](int y, int x) { return mask[y * E + x]; });
        __syncthreads();
        for (int i = threadIdx.x; i < EE; i += kNmsThreads) {
            const int y = i / E, x = i - y * E;
            supp[i] = pool_col_at<uint8_t>(tmpb, E, E, r, y, x);
        }
        __syncthreads();
        // supp_scores = supp ? 0 : s ; new maxima among them that are not suppressed join the mask
        pool_rows<float>(E, E, r, tmp, [&](int y, int x) {
            const float v = s[y * E + x];
            return (supp[y * E + x] && v != -INFINITY) ? 0.f : v;  // padding stays -inf
        });
        __syncthreads();
        for (int i = threadIdx.x; i < EE; i += kNmsThreads) {
            const int y = i / E, x = i - y * E;
            const float ss = supp[i] ? 0.f : s[i];
            const bool newmax = s[i] != -INFINITY && ss == pool_col_at<float>(tmp, E, E, r, y, x);
            mask[i] = mask[i] | (newmax && !supp[i]);
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < kNmsTile * kNmsTile; i += kNmsThreads) {
        const int ty = i / kNmsTile, tx = i - ty * kNmsTile;
        const int y = blockIdx.y * kNmsTile + ty, x = blockIdx.x * kNmsTile + tx;
        if (y >= H || x >= W) continue;
        const int e = (ty + halo) * E + tx + halo;
        float v = mask[e] ? s[e] : 0.f;
        if (pad > 0 && (y < pad || x < pad || y >= H - pad || x >= W - pad)) v = -1.f;
        out[(long long)b * H * W + (long long)y * W + x] = v;
    }
}

int launch_nms_border(const float* scores, int B, int H, int W, int radius, int pad, float* out,
                      cudaStream_t stream) {
    B200_REQUIRE(radius >= 0, "nms: radius must be >= 0");
    const int E = kNmsTile + 10 * radius;
    const size_t smem = (size_t)E * E * 10;
    B200_REQUIRE(smem <= 227 * 1024, "nms: radius %d needs %zu bytes of shared memory (max 232448)", radius, smem);
    static PerDeviceSize configured;
    if (smem > configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(nms_border_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.cur() = smem;
    }
    dim3 grid((W + kNmsTile - 1) / kNmsTile, (H + kNmsTile - 1) / kNmsTile, B);
    nms_border_kernel<<<grid, kNmsThreads, smem, stream>>>(scores, H, W, radius, pad, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
