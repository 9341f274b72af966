// Parameter block of the fused first VGG block (conv_stem.cu).
#pragma once
#include <cuda_bf16.h>

namespace b200 {

struct StemParams {
    const void* img;   // [B, C, H, W], C = 1 or 3: fp32 in [0, 1], or uint8 (u8 != 0) converted as value / 255
    int u8;
    int B, C, H, W;
    const float* w1;   // first convolution [64, 9] fp32 (device)
    // per-channel epilogue operands, carried in the kernel parameters (constant bank, no loads in the epilogues):
    // bias1 | scale1 | shift1 | bias2 | scale2 | shift2
    float par[6][64];
    int pool;
    __nv_bfloat16* out;  // NHWC bf16, [B, H(/2), W(/2), 64]
    int tiles_x, tiles_y;
};

}  // namespace b200
