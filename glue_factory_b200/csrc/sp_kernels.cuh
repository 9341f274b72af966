#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {
int launch_conv1a(const void* img, int u8, int B, int C, int H, int W, const float* w, const float* bias,
                  const float* scale, const float* shift, void* out, cudaStream_t stream);
int launch_nms_border(const float* scores, int B, int H, int W, int radius, int pad, float* out,
                      cudaStream_t stream);
int launch_select_topk(const float* nms_map, int B, int H, int W, float threshold, int k, float* kpts,
                       float* kscores, int* kidx, int* counts, cudaStream_t stream);
int launch_select_topk_flagged(const float* nms_map, int B, int H, int W, float threshold, int k, float* kpts,
                               float* kscores, int* kidx, int* counts, const int* only_flagged, cudaStream_t stream);
// fast path (sp_detect.cu): NMS radius 1..4 with a compact list of surviving maxima, top-k over the list
bool nms_fast_supported(int radius);
int launch_nms_list(const float* scores, int B, int H, int W, int radius, int pad, const int* image_hw, float* out,
                    unsigned long long* cand, int* cand_count /*[2B]*/, int cap, cudaStream_t stream);
int launch_select_list(const float* nms_map, int B, int H, int W, float threshold, int k,
                       const unsigned long long* cand, int* cand_count, int cap, float* kpts, float* kscores,
                       int* kidx, int* counts, cudaStream_t stream);
int launch_count_candidates(const float* nms_map, int B, int H, int W, float threshold, int* counts,
                            cudaStream_t stream);
int launch_sample_descriptors(const float* dense, int B, int Hc, int Wc, int C, const float* kpts, int K,
                              int stride_px, int mode, float* out, cudaStream_t stream);
}  // namespace b200
