// Dense path: 3x3 convolution with C_in = 64 as an implicit GEMM on tcgen05, halo patch loaded ONCE.
//
// The generic implicit-GEMM kernel (gemm.cu, A_MODE = 1) fetches the A tile nine times, once per filter
// tap, which makes the full-resolution 64 -> 64 layers L2-bandwidth bound.  Here a CTA loads the
// (16+2) x 16-pixel halo patch of an 8 x 16 output tile with one 4-D TMA box (64 ch x 16 px x 18 rows,
// 128-byte swizzle, out-of-bounds = zero padding) and addresses the nine shifted A operands INSIDE the
// patch: UMMA row r <-> pixel (y = r / 8, x = r % 8); an 8-row core-matrix group is 8 consecutive
// pixels of one patch row (128 B apart), groups are one patch row = 16 px = 2048 B apart (SBO), and
// tap (dy, dx) only moves the descriptor start address by (dy * 16 + dx) * 128 B.  The start is then no
// longer 1024-byte aligned; measured on B200: the UMMA swizzle phase is taken from the absolute
// shared-memory address bits [7,10) (exactly what TMA used when writing), so the descriptor's
// base-offset field stays 0 (setting it to (start >> 7) & 7 gives wrong results).  The 9 x C_out x 64
// weight tiles stay resident in shared memory for the whole (persistent) CTA.
//
//   warp 0: TMA producer (weights once, then one patch per tile through a ring)
//   warp 1: TMEM allocation + single-thread tcgen05.mma issue (36 MMAs of 128 x C_out x 16 per tile)
//   warps 2-9: epilogue: thread per (pixel row, 32-column group): TMEM -> bias -> ReLU -> BN affine -> bf16 -> swizzled smem tile;
//              then thread per (output pixel, 16 B): optional 2x2 max (bf16x2) and coalesced NHWC stores.
//              The accumulator is double-buffered in TMEM so the epilogue overlaps the next tile's MMAs.
//
// (reference: gluefactory/models/extractors/superpoint_open.py:59-75,96-105: VGGBlock = conv -> ReLU -> BN,
//  MaxPool2d after the second block of the first three stages)
#include "common.cuh"
#include "gemm.cuh"

namespace b200 {

constexpr int kCpTileW = 8, kCpTileH = 16;
constexpr int kCpPatchW = 16, kCpPatchH = kCpTileH + 2;
constexpr int kCpPatchBytes = kCpPatchW * kCpPatchH * 128;  // 36 KB
// epilogue warps: the epilogue (TMEM -> bias / ReLU / BN -> bf16 -> staging -> pooled NHWC stores) of four warps took
// ~2000 of a tile's ~2400 clk (ncu: they waited for the accumulator only 16 % of the time), i.e. it, not the MMAs, set
// the pace; two warps per TMEM lane quarter split the columns.
#ifndef CP_EPI_WARPS
#define CP_EPI_WARPS 8
#endif
constexpr int kCpEpiWarps = CP_EPI_WARPS;
constexpr int kCpEpiThreads = 32 * kCpEpiWarps;
constexpr int kCpThreads = 64 + kCpEpiThreads;

template <int COUT>
struct ConvPatchCfg {
    static constexpr int kWTapBytes = COUT * 128;
    static constexpr int kWBytes = 9 * kWTapBytes;
    static constexpr int kStages = 3;
    static constexpr int kSmemBytes = kWBytes + kStages * kCpPatchBytes + 1024 + 256 + 128 * COUT * 2 + 3 * COUT * 4;
    static constexpr uint32_t kTmemCols = 2 * COUT;
    static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kCpEpiThreads) : "memory"); }
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}

template <int COUT>
__global__ void __launch_bounds__(kCpThreads, 1)
conv_patch_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                  const GemmParams p) {
    using Cfg = ConvPatchCfg<COUT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    uint8_t* sW = smem;
    uint8_t* sA = smem + Cfg::kWBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(sA + Cfg::kStages * kCpPatchBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* acc_full = empty_bar + Cfg::kStages;  // [2]
    uint64_t* acc_empty = acc_full + 2;             // [2]
    uint64_t* w_full = acc_empty + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
    uint8_t* sStage = sA + Cfg::kStages * kCpPatchBytes + 256;              // [128 px][COUT] bf16, swizzled
    float* sPar = reinterpret_cast<float*>(sStage + 128 * COUT * 2);        // bias | scale | shift

    // warp index / TMEM base through a shuffle: the compiler then treats them (and every UMMA operand derived
    // from them) as warp-uniform and issues tcgen05.mma straight from uniform registers
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const long long total_tiles = (long long)p.batch * tiles_per_img;
    // tile coordinates advance by the grid stride as a mixed-radix addition instead of divisions per tile
    struct CTile {
        int b, ty, tx;
    };
    auto decomp = [&](long long t) {
        CTile c;
        c.b = (int)(t / tiles_per_img);
        const int rem = (int)(t - (long long)c.b * tiles_per_img);
        c.ty = rem / p.tiles_x;
        c.tx = rem - c.ty * p.tiles_x;
        return c;
    };
    const CTile tile_step = decomp((long long)gridDim.x);
    auto advance = [&](CTile& c) {
        c.tx += tile_step.tx;
        c.ty += tile_step.ty;
        c.b += tile_step.b;
        if (c.tx >= p.tiles_x) { c.tx -= p.tiles_x; ++c.ty; }
        if (c.ty >= p.tiles_y) { c.ty -= p.tiles_y; ++c.b; }
    };

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_a);
        prefetch_tensormap(&tmap_w);
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(acc_full + s, 1);
            mbar_init(acc_empty + s, kCpEpiWarps);
        }
        mbar_init(w_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_wait();     // predecessor grid complete, its writes visible (prologue above overlapped its tail)
    pdl_trigger();  // successor may become resident

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(w_full, Cfg::kWBytes);
            for (int tap = 0; tap < 9; ++tap) tma_load_2d(sW + tap * Cfg::kWTapBytes, &tmap_w, w_full, tap * 64, 0);
            int stage = 0;
            uint32_t phase = 0;
            CTile tc = decomp((long long)blockIdx.x);
            for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x, advance(tc)) {
                const int b = tc.b, ty = tc.ty, tx = tc.tx;
                mbar_wait(empty_bar + stage, phase ^ 1);
                mbar_arrive_expect_tx(full_bar + stage, kCpPatchBytes);
                tma_load_4d(sA + stage * kCpPatchBytes, &tmap_a, full_bar + stage, 0, tx * kCpTileW - 1,
                            ty * kCpTileH - 1, b);
                if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // whole warp walks the tile loop (uniform control flow); one elected lane issues the MMAs / commits
#ifndef CONV_TIMING_MMA_N  // timing experiments only (results are wrong for any value but COUT): B-operand read share
#define CONV_TIMING_MMA_N COUT
#endif
        constexpr uint32_t idesc = umma_idesc(128, CONV_TIMING_MMA_N, 1u, 0u, 0u);
        mbar_wait(w_full, 0);
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        // Descriptors advance linearly: every base is 1024-aligned and shared memory ends below 256 KB, so the 14-bit
        // address field of descriptor(base + d) is that of descriptor(base) plus d >> 4, with no carry into the next
        // field.  Built from scratch per MMA (add, shift, mask, or: ~9 uniform instructions for each of the 36 MMAs of a
        // tile) the issuer warp spent 75 % of its time issuing (ncu source page), ~50 clk per 32 - 48 clk MMA.
        const uint64_t wdesc0 = umma_smem_desc_sw128(smem_u32(sW), 1024, 16);
        const uint64_t adesc0 = umma_smem_desc_sw128(smem_u32(sA), kCpPatchW * 128, 16);
        for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x) {
            mbar_wait(acc_empty + acc, acc_phase ^ 1);
            mbar_wait(full_bar + stage, phase);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * COUT;
            const uint64_t a_tile = adesc0 + (uint32_t)(stage * (kCpPatchBytes >> 4));
            if (elect_one()) {
#pragma unroll
                for (int tap = 0; tap < 9; ++tap) {
                    const int dy = tap / 3, dx = tap - dy * 3;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(d_tmem, a_tile + (uint32_t)((dy * kCpPatchW + dx) * 8 + k * 2),
                                 wdesc0 + (uint32_t)(tap * (Cfg::kWTapBytes >> 4) + k * 2), idesc, (tap | k) != 0 ? 1u : 0u);
                }
                umma_commit(empty_bar + stage);
                umma_commit(acc_full + acc);
            }
            __syncwarp();
            if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    } else {
        // ===================== epilogue (kCpEpiThreads threads) =====================
        // phase A: thread <-> pixel row: TMEM -> bias/ReLU/BN -> bf16 -> swizzled staging tile in smem
        // phase B: thread <-> (output pixel, 16-byte chunks): optional 2x2 max, coalesced NHWC stores
        const int quarter = warp & 3, grp = (warp - 2) >> 2;  // TMEM lane quarter; column group of the warp
        const int et = threadIdx.x - 64;  // 0 .. kCpEpiThreads - 1
        const int r_in_tile = quarter * 32 + lane;
        const bool pool = p.pool != 0;
        const int OH = pool ? (p.H >> 1) : p.H, OW = pool ? (p.W >> 1) : p.W;
        for (int i = et; i < COUT; i += kCpEpiThreads) {
            sPar[i] = __ldg(p.bias + i);
            sPar[COUT + i] = __ldg(p.scale + i);
            sPar[2 * COUT + i] = __ldg(p.shift + i);
        }
        epi_bar_sync();
        const float4* pb = reinterpret_cast<const float4*>(sPar);
        const float4* ps = reinterpret_cast<const float4*>(sPar + COUT);
        const float4* pt = reinterpret_cast<const float4*>(sPar + 2 * COUT);
        const bool relu = p.relu != 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        uint32_t r[32];
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out16);
        CTile tc = decomp((long long)blockIdx.x);
        for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x, advance(tc)) {
            const int b = tc.b, ty = tc.ty, tx = tc.tx;
            mbar_wait(acc_full + acc, acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + acc * COUT + ((uint32_t)(quarter * 32) << 16);
            uint8_t* my_row = sStage + r_in_tile * (COUT * 2);
#pragma unroll
            for (int c = grp; c < COUT / 32; c += kCpEpiWarps / 4) {
                tmem_ld32(taddr + c * 32, r);
                tmem_ld_wait();
                uint32_t pk[16];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 vb = pb[c * 8 + q], vs = ps[c * 8 + q], vt = pt[c * 8 + q];
                    float x0 = __uint_as_float(r[q * 4]) + vb.x, x1 = __uint_as_float(r[q * 4 + 1]) + vb.y;
                    float x2 = __uint_as_float(r[q * 4 + 2]) + vb.z, x3 = __uint_as_float(r[q * 4 + 3]) + vb.w;
                    if (relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f); }
                    pk[q * 2] = pack_bf16(fmaf(x0, vs.x, vt.x), fmaf(x1, vs.y, vt.y));
                    pk[q * 2 + 1] = pack_bf16(fmaf(x2, vs.z, vt.z), fmaf(x3, vs.w, vt.w));
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int chunk = c * 4 + q;  // 16-byte chunk inside the pixel row, swizzled by row & 7
                    *reinterpret_cast<uint4*>(my_row + ((chunk ^ (r_in_tile & 7)) << 4)) =
                        make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty + acc);  // TMEM buffer is free for the tile after next
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            epi_bar_sync();
            constexpr int kChunks = COUT / 8;  // 16-byte chunks per pixel
            if (pool) {
                // 32 pooled pixels x kChunks chunks
                for (int item = et; item < 32 * kChunks; item += kCpEpiThreads) {
                    const int pp = item / kChunks, ch = item - pp * kChunks;
                    const int oyl = pp >> 2, oxl = pp & 3;
                    const int oy = ty * (kCpTileH / 2) + oyl, ox = tx * (kCpTileW / 2) + oxl;
                    if (oy >= OH || ox >= OW) continue;
                    uint4 m;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int row = (2 * oyl + (k >> 1)) * kCpTileW + 2 * oxl + (k & 1);
                        const uint4 v = *reinterpret_cast<const uint4*>(sStage + row * (COUT * 2) + ((ch ^ (row & 7)) << 4));
                        if (k == 0) {
                            m = v;
                        } else {
                            m.x = bf16x2_max(m.x, v.x); m.y = bf16x2_max(m.y, v.y);
                            m.z = bf16x2_max(m.z, v.z); m.w = bf16x2_max(m.w, v.w);
                        }
                    }
                    *reinterpret_cast<uint4*>(out + (((long long)b * OH + oy) * OW + ox) * p.ld16 + ch * 8) = m;
                }
            } else {
                for (int item = et; item < 128 * kChunks; item += kCpEpiThreads) {
                    const int row = item / kChunks, ch = item - row * kChunks;
                    const int py = ty * kCpTileH + (row >> 3), px = tx * kCpTileW + (row & 7);
                    if (py >= OH || px >= OW) continue;
                    const uint4 v = *reinterpret_cast<const uint4*>(sStage + row * (COUT * 2) + ((ch ^ (row & 7)) << 4));
                    *reinterpret_cast<uint4*>(out + (((long long)b * OH + py) * OW + px) * p.ld16 + ch * 8) = v;
                }
            }
            epi_bar_sync();  // staging tile is rewritten by the next tile
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <int COUT>
static int launch_conv_patch_t(const CUtensorMap& ta, const CUtensorMap& tw, const GemmParams& p, int num_sms,
                               cudaStream_t stream) {
    using Cfg = ConvPatchCfg<COUT>;
    auto kern = conv_patch_kernel<COUT>;
    static PerDeviceSize configured;
    if (!configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
        configured.cur() = 1;
    }
    const long long total = (long long)p.batch * p.tiles_x * p.tiles_y;
    if (total == 0) return 0;
    const int grid = (int)(total < num_sms ? total : num_sms);
    B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kCpThreads), Cfg::kSmemBytes, stream, ta, tw, p));
    return 0;
}

int launch_conv_patch(int cout, const CUtensorMap& ta, const CUtensorMap& tw, const GemmParams& p, int num_sms,
                      cudaStream_t stream) {
    if (cout == 64) return launch_conv_patch_t<64>(ta, tw, p, num_sms, stream);
    set_last_error("conv (patch kernel): C_out=%d not instantiated", cout);
    return -1;
}

}  // namespace b200
