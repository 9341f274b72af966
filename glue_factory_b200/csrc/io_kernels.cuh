#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/b200glue.h"

namespace b200 {
// image pre-processing (io_kernels.cu)
int launch_prep_resize(const void* src, int src_kind, int B, int C, int H, int W, int out_h, int out_w, const float* taps_x,
                       int kx, const float* taps_y, int ky, float* tmp, float* out, int pad_h, int pad_w,
                       unsigned char* mask, int num_sms, cudaStream_t stream);
// export packing (io_kernels.cu)
int launch_export_pack(const b200_export_field* fields, int num_fields, int B, int as_half, void* records,
                       long long record_bytes, cudaStream_t stream);
// match metrics (eval_kernels.cu)
int launch_eval_homography(const float* kp0, const float* kp1, const long long* m0, int B, int M, int N, const float* H,
                           float* out, cudaStream_t stream);
int launch_eval_epipolar(const float* kp0, const float* kp1, const long long* m0, int B, int M, int N, const float* cam0,
                         const float* cam1, const float* T01, float* out, cudaStream_t stream);
int launch_eval_dlt(const float* kp0, const float* kp1, const long long* m0, const float* ms0, int B, int M, int N,
                    const float* H_gt, const float* image_size, float* H_out, float* err_out, cudaStream_t stream);
}  // namespace b200
