// Bandwidth-bound kernels either side of the forward path:
//   * image pre-processing: u8 / float image -> optional separable Gaussian anti-alias blur (reflect border)
//     -> bilinear resize (half-pixel centres) -> zero square padding + padding mask
//     (reference: gluefactory/utils/image.py:12-101 ImagePreprocessor, whose resize is
//      kornia.geometry.transform.resize = gaussian_blur2d + F.interpolate(align_corners=False))
//   * export packing: keypoint rescale by 1 / scales, optional fp16 conversion and packing of a batch of
//     predictions into fixed-size per-pair records for ONE device-to-host copy per batch
//     (reference: gluefactory/utils/export_predictions.py:40-77, one .cpu() per tensor per pair)
#include "../../include/b200glue.h"
#include "common.cuh"
#include "io_kernels.cuh"

namespace b200 {

constexpr int kMaxTaps = 63;

struct PrepParams {
    const void* src;   // f32 [B,C,H,W] or u8 [B,H,W,C]
    float* tmp;        // f32 [B,C,H,W] horizontally blurred source (blur only)
    float* out;        // f32 [B,C,pad_h,pad_w]
    unsigned char* mask;  // [B,pad_h,pad_w] or null
    int B, C, H, W, out_h, out_w, pad_h, pad_w;
    int kx, ky;        // taps per direction (1 = no blur)
    float scale_y, scale_x;  // (float)H / out_h, (float)W / out_w  (F.interpolate, align_corners=False)
    float tx[kMaxTaps], ty[kMaxTaps];
};

__device__ __forceinline__ int reflect_index(int i, int n) {  // F.pad(mode="reflect"): -1 -> 1, n -> n - 2
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return i;
}

// u8 -> float exactly as numpy_image_to_torch does it (image.py:121: image / 255.0 in float64, then float32):
// u8_unit (common.cuh) reproduces that value for all 256 inputs without a division
__device__ __forceinline__ float u8_to_unit(unsigned char v) { return u8_unit(v); }

template <bool U8>
__device__ __forceinline__ float load_src(const PrepParams& p, int b, int c, int y, int x) {
    if (U8) return u8_to_unit(static_cast<const unsigned char*>(p.src)[(((long long)b * p.H + y) * p.W + x) * p.C + c]);
    return static_cast<const float*>(p.src)[(((long long)b * p.C + c) * p.H + y) * p.W + x];
}

// The three kernels below use a 3-D grid (x blocks, row, image * channel): no per-element division.

// horizontal Gaussian pass over the source: tmp[b,c,y,x] = sum_k tx[k] * src[b,c,y,reflect(x + k - r)]
template <bool U8>
__global__ void __launch_bounds__(256) prep_hblur_kernel(const __grid_constant__ PrepParams p) {
    const int x = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    const int b = blockIdx.z / p.C, c = blockIdx.z - b * p.C;
    if (x >= p.W) return;
    const int r = p.kx / 2;
    float acc = 0.f;
    for (int k = 0; k < p.kx; ++k) acc = fmaf(p.tx[k], load_src<U8>(p, b, c, y, reflect_index(x + k - r, p.W)), acc);
    p.tmp[(((long long)b * p.C + c) * p.H + y) * p.W + x] = acc;
}

// vertical Gaussian pass fused with the bilinear resize and the padding: one thread per output pixel
template <bool U8, bool BLUR>
__global__ void __launch_bounds__(256) prep_resize_kernel(const __grid_constant__ PrepParams p) {
    const int ox = blockIdx.x * 256 + threadIdx.x, oy = blockIdx.y;
    const int b = blockIdx.z / p.C, c = blockIdx.z - b * p.C;
    if (ox >= p.pad_w) return;
    const int r = p.ky / 2;
    const bool inside = oy < p.out_h && ox < p.out_w;
    if (p.mask != nullptr && c == 0) p.mask[((long long)b * p.pad_h + oy) * p.pad_w + ox] = inside ? 1 : 0;
    float v = 0.f;
    if (inside) {
        // area_pixel_compute_source_index, align_corners=False: src = scale * (dst + 0.5) - 0.5, clamped at 0
        float sy = p.scale_y * (oy + 0.5f) - 0.5f, sx = p.scale_x * (ox + 0.5f) - 0.5f;
        sy = sy < 0.f ? 0.f : sy;
        sx = sx < 0.f ? 0.f : sx;
        const int y0 = (int)sy, x0 = (int)sx;
        const int y1 = y0 + (y0 < p.H - 1 ? 1 : 0), x1 = x0 + (x0 < p.W - 1 ? 1 : 0);
        const float ly1 = sy - y0, lx1 = sx - x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
        float v00, v01, v10, v11;
        if (BLUR) {
            const float* tp = p.tmp + ((long long)b * p.C + c) * p.H * p.W;
            v00 = v01 = v10 = v11 = 0.f;
            for (int k = 0; k < p.ky; ++k) {
                const long long ra = (long long)reflect_index(y0 + k - r, p.H) * p.W;
                const long long rb = (long long)reflect_index(y1 + k - r, p.H) * p.W;
                const float w = p.ty[k];
                v00 = fmaf(w, tp[ra + x0], v00);
                v01 = fmaf(w, tp[ra + x1], v01);
                v10 = fmaf(w, tp[rb + x0], v10);
                v11 = fmaf(w, tp[rb + x1], v11);
            }
        } else {
            v00 = load_src<U8>(p, b, c, y0, x0);
            v01 = load_src<U8>(p, b, c, y0, x1);
            v10 = load_src<U8>(p, b, c, y1, x0);
            v11 = load_src<U8>(p, b, c, y1, x1);
        }
        v = ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11);
    }
    p.out[(((long long)b * p.C + c) * p.pad_h + oy) * p.pad_w + ox] = v;
}

// same-size path (no resize): conversion + padding only
template <bool U8>
__global__ void __launch_bounds__(256) prep_copy_kernel(const __grid_constant__ PrepParams p) {
    const int ox = blockIdx.x * 256 + threadIdx.x, oy = blockIdx.y;
    const int b = blockIdx.z / p.C, c = blockIdx.z - b * p.C;
    if (ox >= p.pad_w) return;
    const bool inside = oy < p.out_h && ox < p.out_w;
    if (p.mask != nullptr && c == 0) p.mask[((long long)b * p.pad_h + oy) * p.pad_w + ox] = inside ? 1 : 0;
    p.out[(((long long)b * p.C + c) * p.pad_h + oy) * p.pad_w + ox] = inside ? load_src<U8>(p, b, c, oy, ox) : 0.f;
}


int launch_prep_resize(const void* src, int src_kind, int B, int C, int H, int W, int out_h, int out_w, const float* taps_x,
                       int kx, const float* taps_y, int ky, float* tmp, float* out, int pad_h, int pad_w,
                       unsigned char* mask, int num_sms, cudaStream_t stream) {
    B200_REQUIRE(src && out, "prep_resize: null argument");
    B200_REQUIRE(src_kind == 0 || src_kind == 1, "prep_resize: src_kind must be 0 (f32 CHW) or 1 (u8 HWC)");
    B200_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && out_h >= 1 && out_w >= 1, "prep_resize: bad sizes");
    B200_REQUIRE(pad_h >= out_h && pad_w >= out_w, "prep_resize: padded size %dx%d smaller than the image %dx%d", pad_h,
                 pad_w, out_h, out_w);
    const bool blur = kx > 1 || ky > 1;
    B200_REQUIRE(kx >= 1 && ky >= 1 && kx <= kMaxTaps && ky <= kMaxTaps && (kx & 1) && (ky & 1),
                 "prep_resize: tap counts %d / %d must be odd and <= %d", kx, ky, kMaxTaps);
    B200_REQUIRE(!blur || (taps_x && taps_y && tmp), "prep_resize: the blur needs both tap arrays and the scratch image");
    B200_REQUIRE(!blur || (kx / 2 < W && ky / 2 < H), "prep_resize: reflect border needs radius < image size");
    if (B == 0) return 0;
    PrepParams p;
    memset(&p, 0, sizeof(p));
    p.src = src; p.tmp = tmp; p.out = out; p.mask = mask;
    p.B = B; p.C = C; p.H = H; p.W = W; p.out_h = out_h; p.out_w = out_w; p.pad_h = pad_h; p.pad_w = pad_w;
    p.kx = kx; p.ky = ky;
    p.scale_y = (float)H / (float)out_h;
    p.scale_x = (float)W / (float)out_w;
    if (blur) {
        memcpy(p.tx, taps_x, sizeof(float) * kx);
        memcpy(p.ty, taps_y, sizeof(float) * ky);
    }
    const bool u8 = src_kind == 1;
    B200_REQUIRE(H <= 65535 && pad_h <= 65535 && (long long)B * C <= 65535, "prep_resize: image or batch too large for the grid");
    const dim3 g_src((W + 255) / 256, H, B * C), g_out((pad_w + 255) / 256, pad_h, B * C);
    (void)num_sms;
    if (!blur && out_h == H && out_w == W) {
        if (u8) prep_copy_kernel<true><<<g_out, 256, 0, stream>>>(p);
        else prep_copy_kernel<false><<<g_out, 256, 0, stream>>>(p);
    } else if (blur) {
        if (u8) prep_hblur_kernel<true><<<g_src, 256, 0, stream>>>(p);
        else prep_hblur_kernel<false><<<g_src, 256, 0, stream>>>(p);
        prep_resize_kernel<false, true><<<g_out, 256, 0, stream>>>(p);
    } else {
        if (u8) prep_resize_kernel<true, false><<<g_out, 256, 0, stream>>>(p);
        else prep_resize_kernel<false, false><<<g_out, 256, 0, stream>>>(p);
    }
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---- export packing ------------------------------------------------------------------------------------
struct ExportParams {
    b200_export_field f[B200_EXPORT_MAX_FIELDS];
    int num_fields, B, as_half;
    unsigned char* records;
    long long record_bytes;
};

// blockIdx.y = field, blockIdx.z = pair; threads stride over the field's elements
__global__ void __launch_bounds__(256) export_pack_kernel(const __grid_constant__ ExportParams p) {
    const b200_export_field& f = p.f[blockIdx.y];
    const int b = blockIdx.z;
    unsigned char* dst = p.records + (long long)b * p.record_bytes + f.dst_offset;
    const long long n = f.elems_per_pair;
    if (f.kind == B200_EXPORT_I64) {
        const long long* s = static_cast<const long long*>(f.src) + (long long)b * n;
        long long* d = reinterpret_cast<long long*>(dst);
        for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) d[i] = s[i];
        return;
    }
    const float* s = static_cast<const float*>(f.src) + (long long)b * n;
    // keypoints: pred * (1.0 / scales) as the reference computes it (export_predictions.py:44-50): the reciprocal
    // is rounded to fp32 first, then multiplied
    float inv[2] = {1.f, 1.f};
    if (f.kind == B200_EXPORT_F32_KEYPOINTS) {
        inv[0] = __fdiv_rn(1.0f, f.scales[b * 2]);
        inv[1] = __fdiv_rn(1.0f, f.scales[b * 2 + 1]);
    }
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        float v = s[i];
        if (f.kind == B200_EXPORT_F32_KEYPOINTS) v = __fmul_rn(v, inv[i & 1]);
        if (p.as_half) reinterpret_cast<__half*>(dst)[i] = __float2half_rn(v);
        else reinterpret_cast<float*>(dst)[i] = v;
    }
}

int launch_export_pack(const b200_export_field* fields, int num_fields, int B, int as_half, void* records,
                       long long record_bytes, cudaStream_t stream) {
    B200_REQUIRE(fields && records, "export_pack: null argument");
    B200_REQUIRE(num_fields >= 1 && num_fields <= B200_EXPORT_MAX_FIELDS, "export_pack: %d fields (max %d)", num_fields,
                 B200_EXPORT_MAX_FIELDS);
    B200_REQUIRE(B >= 0 && record_bytes > 0 && record_bytes % 8 == 0, "export_pack: record size must be a multiple of 8");
    if (B == 0) return 0;
    ExportParams p;
    memset(&p, 0, sizeof(p));
    long long max_elems = 1;
    for (int i = 0; i < num_fields; ++i) {
        const b200_export_field& f = fields[i];
        B200_REQUIRE(f.src != nullptr && f.elems_per_pair >= 0 && f.dst_offset >= 0, "export_pack: field %d is malformed", i);
        B200_REQUIRE(f.kind == B200_EXPORT_F32 || f.kind == B200_EXPORT_F32_KEYPOINTS || f.kind == B200_EXPORT_I64,
                     "export_pack: field %d has unknown kind %d", i, f.kind);
        const long long elem = f.kind == B200_EXPORT_I64 ? 8 : (as_half ? 2 : 4);
        B200_REQUIRE(f.dst_offset % (f.kind == B200_EXPORT_I64 ? 8 : 4) == 0, "export_pack: field %d is misaligned", i);
        B200_REQUIRE(f.dst_offset + f.elems_per_pair * elem <= record_bytes, "export_pack: field %d overruns the record", i);
        B200_REQUIRE(f.kind != B200_EXPORT_F32_KEYPOINTS || (f.scales != nullptr && f.elems_per_pair % 2 == 0),
                     "export_pack: keypoint field %d needs scales and (x, y) pairs", i);
        p.f[i] = f;
        if (f.elems_per_pair > max_elems) max_elems = f.elems_per_pair;
    }
    p.num_fields = num_fields; p.B = B; p.as_half = as_half;
    p.records = static_cast<unsigned char*>(records);
    p.record_bytes = record_bytes;
    long long bx = (max_elems + 255) / 256;
    if (bx > 64) bx = 64;
    export_pack_kernel<<<dim3((unsigned)bx, (unsigned)num_fields, (unsigned)B), 256, 0, stream>>>(p);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
