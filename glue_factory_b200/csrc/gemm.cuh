// Parameter block shared by the host launcher (api) and the tcgen05 GEMM / implicit-GEMM conv kernel.
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace b200 {

enum EpiKind : int {
    EPI_BIAS_BF16 = 0,           // out16[row, col] = acc + bias (ReLU when relu != 0)
    EPI_QKV_ROTARY = 1,          // as above, rotary embedding on columns < rot_cols
    EPI_LN_GELU = 2,             // bias -> LayerNorm(512) -> GELU(erf) -> out16   (BLOCK_N == N == 512)
    EPI_RESID = 3,               // out32 = acc + bias + resid ; out16 = bf16(out32)
    EPI_F32 = 4,                 // out32[batch][row, col] = alpha * acc (+ bias)
    EPI_CONV_RELU_BN = 5,        // bias -> (ReLU) -> BN affine -> (2x2 max-pool) -> NHWC out16
    EPI_CONV_BN_L2 = 6,          // bias -> BN affine -> L2 normalise over 256 channels -> out32
    EPI_CONV_BN_SOFTMAX_D2S = 7  // bias -> BN affine -> 65-way softmax -> depth-to-space heat map
};

struct GemmParams {
    // C[batch][M, N] = A[batch][M, K] . B[batch][N, K]^T ; K multiple of 64
    int M, N, K, batch;
    int a_batch_rows, b_batch_rows;  // tensor-map row offset per batch
    // implicit-GEMM 3x3 convolution (A_MODE == 1): A is NHWC [batch, H, W, Cin], tile = 8 rows x 16 cols
    int H, W, Cin, tiles_x, tiles_y;
    // epilogue inputs (per output column unless noted)
    const float* bias;
    const float* scale;
    const float* shift;
    int relu, pool;
    float alpha;
    const float* resid;  // [M, ld32]
    const float2* cs;    // rotary (cos, sin) per row and frequency: [M, 32]
    int rot_cols;
    const float* ln_gamma;
    const float* ln_beta;
    float ln_eps;
    // outputs
    void* out16;  // bf16
    long long ld16;
    float* out32;
    long long ld32;
    long long out32_batch_stride;
    int cells_h, cells_w;  // EPI_CONV_BN_SOFTMAX_D2S: rows are (img, cell_y, cell_x)
    // optional DEVICE array [batch * m_tiles]: row tiles marked 0 are skipped by every role (adaptive LightGlue: tiles
    // that hold only pruned tokens or pairs that left the layer loop)
    const int* tile_live;
};

}  // namespace b200
