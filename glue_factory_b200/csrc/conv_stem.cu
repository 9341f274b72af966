// Dense path: the first VGG block of SuperPoint (conv 1 -> 64, ReLU, BN; conv 64 -> 64, ReLU, BN; 2x2 max pool)
// as ONE kernel.  The 64-channel full-resolution activation between the two convolutions (8.6 GB per step
// at 64 images of 1024 x 1024) never reaches HBM.
//
// Per 8 x 16 output tile (same halo-patch implicit GEMM as conv_patch.cu) the CTA needs the first
// convolution on the 10 x 18 halo patch.  That convolution also runs on the tensor core: K = 9 taps is
// useless as it stands, but with the fp32 image split into two bf16 terms it becomes a K = 32 GEMM of
// fp32 quality,
//     A row (patch pixel) = [hi(9) | lo(9) | hi(9) | 0(5)],   B row (channel) = [w_hi(9) | w_hi(9) | w_lo(9) | 0(5)]
//     => sum hi w_hi + lo w_hi + hi w_lo   (only lo w_lo, ~2^-18 relative, is dropped)
// so 180 pixels x 64 channels x 9 taps cost four tcgen05.mma instead of 104k FMAs.
//
//   warp 0      TMA: the nine 64 x 64 weight tiles of the second convolution, once per (persistent) CTA
//   warp 1      TMEM allocation + tcgen05.mma issue (uniform-register operands, one elected lane):
//               stem GEMM of tile i+1 (2 M-tiles x 2 K-steps), then the 36 MMAs of tile i
//   warps 2-5   epilogue of the second convolution: TMEM -> bias -> ReLU -> BN -> bf16 -> swizzled smem tile,
//               then 2x2 max (bf16x2) and coalesced NHWC stores; accumulator double-buffered in TMEM
//   warps 6-13  stem workers (two per TMEM lane quarter, 32 channels each): im2col rows of the patch (hi / lo split, 128-byte swizzled K-major rows) ->
//               [stem GEMM] -> TMEM -> bias -> ReLU -> BN -> bf16 -> the halo patch in the layout the nine
//               shifted A descriptors expect (pixel rows of 128 B, 16-pixel pitch, 128-byte swizzle);
//               patch pixels outside the image are zeros (zero padding of the SECOND convolution)
//
// (reference: gluefactory/models/extractors/superpoint_open.py:59-75,96-105, gluefactory_nonfree/superpoint.py
//  conv1a/conv1b; studied for behaviour only)
#include <type_traits>

#include "common.cuh"
#include "gemm.cuh"
#include "conv_stem.cuh"

namespace b200 {

constexpr int kStTileW = 8, kStTileH = 16;
constexpr int kStPatchPitch = 16, kStPatchH = kStTileH + 2, kStPatchW = kStTileW + 2;
constexpr int kStPatchPx = kStPatchH * kStPatchW;              // 180 pixels carry data
constexpr int kStPatchBytes = kStPatchPitch * kStPatchH * 128;  // 36 KB
constexpr int kStPatchStages = 2;
// epilogue warps of the second convolution: four.  They are busy 83 % of a tile (ncu source page: 17 % of their samples
// wait for the accumulator), the eight workers 80 %, and the schedulers issue 60 % of the cycles: eight epilogue
// warps (ST_EPI_WARPS=8, columns split in two) take issue slots from the workers and measured 3 % SLOWER
// (5.63 vs 5.47 ms per step), whereas the same change speeds up conv_patch.cu, which has no workers, by 7 %.
#ifndef ST_EPI_WARPS
#define ST_EPI_WARPS 4
#endif
constexpr int kStEpiWarps = ST_EPI_WARPS;
constexpr int kStEpiThreads = 32 * kStEpiWarps;
constexpr int kStFirstWorker = 2 + kStEpiWarps;  // must keep (warp & 3) == 2 for the first worker warp
static_assert((kStFirstWorker & 3) == 2, "worker warps start on TMEM lane quarter 2");
constexpr int kStThreads = 64 + kStEpiThreads + 256;  // producer, issuer, epilogue warps, 8 stem-worker warps
constexpr int kStWTapBytes = 64 * 128, kStWBytes = 9 * kStWTapBytes;  // second convolution: 72 KB resident
constexpr int kStABytes = 2 * 128 * 128;                               // im2col rows of the patch: 2 M-tiles
constexpr int kStW1Bytes = 64 * 128;                                   // first convolution as a [64 x 32] bf16 tile
constexpr int kStStageBytes = 128 * 64 * 2;                            // output staging tile
constexpr int kStSmemBytes = 1024 + kStWBytes + kStPatchStages * kStPatchBytes + kStABytes + kStW1Bytes + 256 +
                             kStStageBytes + 2 * 256 * 4;
constexpr int kStImgH = kStPatchH + 2, kStImgW = kStPatchW + 2;  // image pixels under the patch: 20 x 12
// [0,128): output accumulators (2 x 64)   [128,384): stem GEMM, double-buffered (2 x 2 M-tiles x 64)
constexpr uint32_t kStTmemCols = 512;
static_assert(kStSmemBytes <= 227 * 1024, "shared memory budget");


__device__ __forceinline__ void st_epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kStEpiThreads) : "memory"); }
__device__ __forceinline__ void st_work_bar() { asm volatile("bar.sync 2, 256;" ::: "memory"); }
__device__ __forceinline__ void st_build_bar() { asm volatile("bar.sync 3, 128;" ::: "memory"); }
__device__ __forceinline__ uint32_t st_bf16x2_max(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
// v -> (hi, lo) with hi = bf16(v) and lo = v - hi (still fp32; rounded to bf16 when packed)
__device__ __forceinline__ void st_split(float v, float& hi, float& lo) {
    hi = __bfloat162float(__float2bfloat16_rn(v));
    lo = v - hi;
}

template <bool U8>
__global__ void __launch_bounds__(kStThreads, 1)
conv_stem_kernel(const __grid_constant__ CUtensorMap tmap_w, const StemParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    uint8_t* sW = smem;
    uint8_t* sPatch = sW + kStWBytes;
    uint8_t* sA = sPatch + kStPatchStages * kStPatchBytes;
    uint8_t* sW1 = sA + kStABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sW1 + kStW1Bytes);
    uint64_t* w_full = bars;            // second-convolution weights landed
    // [2] im2col rows (and, the first time, the stem weight tile) written, one barrier per tile parity: with a single
    // barrier the builders could complete two phases (tiles i+1 and i+2) while the issuer was held up on acc_empty by a
    // slow epilogue, the issuer's parity wait then never returned and the builders waited for its stem GEMM - a
    // timing-dependent deadlock (seen as a hang under ncu's instrumented replay passes)
    uint64_t* a_full = bars + 13;
    uint64_t* c1_full = bars + 2;       // [2] stem GEMM retired (even / odd tiles)
    uint64_t* patch_full = bars + 4;    // [2] halo patch written
    uint64_t* patch_empty = bars + 6;   // [2] the 36 MMAs reading it retired
    uint64_t* acc_full = bars + 8;      // [2]
    uint64_t* acc_empty = bars + 10;    // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
    uint8_t* sStage = reinterpret_cast<uint8_t*>(bars) + 256;              // [128 px][64] bf16, swizzled
    float* sImg = reinterpret_cast<float*>(sStage + kStStageBytes);        // [2][20 x 12] gray pixels under the patch

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const int total_tiles = p.B * tiles_per_img;  // < 2^31, checked by the launcher
    // Tile coordinates advance by the grid stride as a mixed-radix addition: the two integer divisions of a fresh
    // decomposition (~65 dependent instructions with MUFU.RCP / F2I) sat on the per-tile critical path of the worker
    // and epilogue warps (a quarter of the workers' stall samples in the ncu source page).
    struct StTile {
        int b, ty, tx;
    };
    auto decomp = [&](int t) {
        StTile c;
        c.b = t / tiles_per_img;
        const int rem = t - c.b * tiles_per_img;
        c.ty = rem / p.tiles_x;
        c.tx = rem - c.ty * p.tiles_x;
        return c;
    };
    const StTile tile_step = decomp((int)gridDim.x);
    auto advance = [&](StTile& c) {
        c.tx += tile_step.tx;
        c.ty += tile_step.ty;
        c.b += tile_step.b;
        if (c.tx >= p.tiles_x) { c.tx -= p.tiles_x; ++c.ty; }
        if (c.ty >= p.tiles_y) { c.ty -= p.tiles_y; ++c.b; }
    };

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_w);
        mbar_init(w_full, 1);
        mbar_init(a_full, 4);
        mbar_init(a_full + 1, 4);
        for (int s = 0; s < 2; ++s) {
            mbar_init(c1_full + s, 1);
            mbar_init(patch_full + s, 8);
            mbar_init(patch_empty + s, 1);
            mbar_init(acc_full + s, 1);
            mbar_init(acc_empty + s, kStEpiWarps);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kStTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_wait();     // predecessor grid complete, its writes visible (prologue above overlapped its tail)
    pdl_trigger();  // successor may become resident

    if (warp == 0) {
        if (elect_one()) {
            mbar_arrive_expect_tx(w_full, kStWBytes);
            for (int tap = 0; tap < 9; ++tap) tma_load_2d(sW + tap * kStWTapBytes, &tmap_w, w_full, tap * 64, 0);
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const bool leader = elect_one();
#ifndef CONV_TIMING_MMA_N  // timing experiments only (results are wrong for any value but 64): B-operand read share
#define CONV_TIMING_MMA_N 64
#endif
        constexpr uint32_t idesc = umma_idesc(128, CONV_TIMING_MMA_N, 1u, 0u, 0u);
        // (descriptors are rebuilt per MMA here on purpose: with the linear form of conv_patch.cu, which issues the 36 MMAs
        // of a tile in a burst, this kernel measured 11 % SLOWER, 6.06 vs 5.45 ms per step: its workers and epilogue warps
        // pace it, and a tensor pipe that reads shared memory back to back takes bandwidth from them)
        const uint32_t w0 = smem_u32(sW), w1 = smem_u32(sW1), a_im = smem_u32(sA);
        // conv 1 -> 64 on the 180 patch pixels: 2 M-tiles x (K = 32).  Tile parity `par` selects the im2col
        // buffer (K columns [32 par, 32 par + 32) of the 128-byte rows) and the TMEM accumulator pair.
        auto issue_stem = [&](int par) {
            if (leader) {
#pragma unroll
                for (int m = 0; m < 2; ++m)
#pragma unroll
                    for (int k = 0; k < 2; ++k)
                        umma_f16(tmem_base + 128 + par * 128 + m * 64,
                                 umma_smem_desc_sw128(a_im + m * 16384 + par * 64 + k * 32, 1024, 16),
                                 umma_smem_desc_sw128(w1 + k * 32, 1024, 16), idesc, k != 0 ? 1u : 0u);
                umma_commit(c1_full + par);
            }
            __syncwarp();
        };
        if (blockIdx.x < total_tiles) {
            mbar_wait(a_full, 0);
            tc_fence_after();
            issue_stem(0);
        }
        mbar_wait(w_full, 0);
        int acc = 0;
        uint32_t acc_phase = 0;
        int i = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++i) {
            const int st = i & 1;
            if (t + gridDim.x < total_tiles) {  // stem GEMM of the next tile first: its workers are waiting
                mbar_wait(a_full + ((i + 1) & 1), (uint32_t)(((i + 1) >> 1) & 1));
                tc_fence_after();
                issue_stem((i + 1) & 1);
            }
            mbar_wait(patch_full + st, (uint32_t)((i >> 1) & 1));
            mbar_wait(acc_empty + acc, acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * 64;
            const uint32_t a0 = smem_u32(sPatch + st * kStPatchBytes);
            if (leader) {
#pragma unroll
                for (int tap = 0; tap < 9; ++tap) {
                    const int dy = tap / 3, dx = tap - dy * 3;
                    const uint32_t a_tap = a0 + (dy * kStPatchPitch + dx) * 128;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(d_tmem, umma_smem_desc_sw128(a_tap + k * 32, kStPatchPitch * 128, 16),
                                 umma_smem_desc_sw128(w0 + tap * kStWTapBytes + k * 32, 1024, 16), idesc,
                                 (tap | k) != 0 ? 1u : 0u);
                }
                umma_commit(patch_empty + st);
                umma_commit(acc_full + acc);
            }
            __syncwarp();
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    } else if (warp < kStFirstWorker) {
        // ===================== epilogue of the second convolution (kStEpiThreads threads) =====================
        // phase A: thread <-> pixel row: TMEM -> bias/ReLU/BN -> bf16 -> swizzled staging tile in smem
        // phase B: thread <-> (output pixel, 16-byte chunks): optional 2x2 max, coalesced NHWC stores
        const int quarter = warp & 3, grp = (warp - 2) >> 2;  // TMEM lane quarter; 32-column group of the warp
        const int et = threadIdx.x - 64;  // 0 .. kStEpiThreads - 1
        const int r_in_tile = quarter * 32 + lane;
        const bool pool = p.pool != 0;
        const int OH = pool ? (p.H >> 1) : p.H, OW = pool ? (p.W >> 1) : p.W;
        int acc = 0;
        uint32_t acc_phase = 0;
        uint32_t r[32];
        StTile tc = decomp((int)blockIdx.x);
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, advance(tc)) {
            const int b = tc.b, ty = tc.ty, tx = tc.tx;
            mbar_wait(acc_full + acc, acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + acc * 64 + ((uint32_t)(quarter * 32) << 16);
            uint8_t* my_row = sStage + r_in_tile * 128;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                if (kStEpiWarps == 8 && c != grp) continue;  // warp-uniform; `c` stays a compile-time constant below
                tmem_ld32(taddr + c * 32, r);
                tmem_ld_wait();
                uint32_t pk[16];
#pragma unroll
                for (int q = 0; q < 16; ++q) {  // per-channel operands straight from the constant bank
                    const int ch = c * 32 + q * 2;
                    const float x0 = fmaxf(__uint_as_float(r[q * 2]) + p.par[3][ch], 0.f);
                    const float x1 = fmaxf(__uint_as_float(r[q * 2 + 1]) + p.par[3][ch + 1], 0.f);
                    pk[q] = pack_bf16(fmaf(x0, p.par[4][ch], p.par[5][ch]), fmaf(x1, p.par[4][ch + 1], p.par[5][ch + 1]));
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
            st_epi_bar();
            if (pool) {
                // 32 pooled pixels x 8 chunks
                for (int item = et; item < 32 * 8; item += kStEpiThreads) {
                    const int pp = item >> 3, ch = item & 7;
                    const int oyl = pp >> 2, oxl = pp & 3;
                    const int oy = ty * (kStTileH / 2) + oyl, ox = tx * (kStTileW / 2) + oxl;
                    if (oy >= OH || ox >= OW) continue;
                    uint4 m;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int row = (2 * oyl + (k >> 1)) * kStTileW + 2 * oxl + (k & 1);
                        const uint4 v = *reinterpret_cast<const uint4*>(sStage + row * 128 + ((ch ^ (row & 7)) << 4));
                        if (k == 0) {
                            m = v;
                        } else {
                            m.x = st_bf16x2_max(m.x, v.x); m.y = st_bf16x2_max(m.y, v.y);
                            m.z = st_bf16x2_max(m.z, v.z); m.w = st_bf16x2_max(m.w, v.w);
                        }
                    }
                    *reinterpret_cast<uint4*>(p.out + (((long long)b * OH + oy) * OW + ox) * 64 + ch * 8) = m;
                }
            } else {
                for (int item = et; item < 128 * 8; item += kStEpiThreads) {
                    const int row = item >> 3, ch = item & 7;
                    const int py = ty * kStTileH + (row >> 3), px = tx * kStTileW + (row & 7);
                    if (py >= OH || px >= OW) continue;
                    const uint4 v = *reinterpret_cast<const uint4*>(sStage + row * 128 + ((ch ^ (row & 7)) << 4));
                    *reinterpret_cast<uint4*>(p.out + (((long long)b * OH + py) * OW + px) * 64 + ch * 8) = v;
                }
            }
            st_epi_bar();  // staging tile is rewritten by the next tile
        }
    } else {
        // ===================== stem workers (128 threads) =====================
        const int quarter = warp & 3;
        const int chalf = (warp - kStFirstWorker) >> 2;  // channels [32 chalf, 32 chalf + 32) in the TMEM -> patch pass
        const int wt = threadIdx.x - 32 * kStFirstWorker;  // 0..255: staged image pixel / patch pixel (im2col row)
        if (wt < 128) {  // first convolution as a [64 channels x K = 32] bf16 tile: [w_hi | w_hi | w_lo | 0]
            const int n = wt >> 1, h = wt & 1;
            float f[32];
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                float hi, lo;
                st_split(__ldg(p.w1 + n * 9 + k), hi, lo);
                f[k] = hi;
                f[9 + k] = hi;
                f[18 + k] = lo;
            }
#pragma unroll
            for (int k = 27; k < 32; ++k) f[k] = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if ((c >> 1) != h) continue;
                *reinterpret_cast<uint4*>(sW1 + n * 128 + ((c ^ (n & 7)) << 4)) =
                    make_uint4(pack_bf16(f[c * 8], f[c * 8 + 1]), pack_bf16(f[c * 8 + 2], f[c * 8 + 3]),
                               pack_bf16(f[c * 8 + 4], f[c * 8 + 5]), pack_bf16(f[c * 8 + 6], f[c * 8 + 7]));
            }
        }
        st_work_bar();
        // bias -> ReLU -> BN of 32 channels (operands from the constant bank), packed bf16, zeros outside the image
        auto stem_affine = [&](auto half_tag, const uint32_t (&acc)[32], bool inside, uint32_t (&pk)[16]) {
            constexpr int kCh0 = decltype(half_tag)::value * 32;
            const uint32_t mask = inside ? 0xffffffffu : 0u;  // branch-free: zero padding of the second convolution
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                const int ch = kCh0 + q * 2;
                const float x0 = fmaxf(__uint_as_float(acc[q * 2]) + p.par[0][ch], 0.f);
                const float x1 = fmaxf(__uint_as_float(acc[q * 2 + 1]) + p.par[0][ch + 1], 0.f);
                pk[q] = pack_bf16(fmaf(x0, p.par[1][ch], p.par[2][ch]), fmaf(x1, p.par[1][ch + 1], p.par[2][ch + 1])) & mask;
            }
        };
        const long long plane = (long long)p.H * p.W;
        // The TMEM lane quarters 0 / 1 hold patch pixels of both M-tiles, quarters 2 / 3 only of the first one, so
        // the warps of quarters 2 / 3 (half the TMEM -> patch work) also build the im2col rows: `bt` numbers
        // their 128 threads.
        const bool builder = quarter >= 2;
        const int bt = (((warp - kStFirstWorker) >> 2) * 2 + (quarter - 2)) * 32 + lane;
        // gray pixel e (of 20 x 12) under the patch of tile t, zero outside the image
        // gray pixel e (of 20 x 12) under the patch of tile t, zero outside the image.  Returned RAW (float bits, or
        // the byte of a single-channel uint8 image) and converted by cook() where it is consumed: the request for the
        // tile after next stays in flight for a whole tile, a conversion at the load would wait for it on the spot.
        auto px = [&](long long at) {
            return U8 ? u8_unit(__ldg(static_cast<const unsigned char*>(p.img) + at))
                      : __ldg(static_cast<const float*>(p.img) + at);
        };
        auto load_img = [&](const StTile& c, int e) -> uint32_t {
            const int b = c.b, ty = c.ty, tx = c.tx;
            const long long img0 = (long long)b * p.C * plane;
            const int iy = e / kStImgW, ix = e - iy * kStImgW;
            const int yy = ty * kStTileH - 2 + iy, xx = tx * kStTileW - 2 + ix;
            uint32_t raw = 0u;  // 0.0f and byte 0 alike
            if (e < kStImgH * kStImgW && yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) {
                const long long o = img0 + (long long)yy * p.W + xx;
                if (p.C == 3) {
                    // same association order as (image * [0.299, 0.587, 0.114]).sum(1)
                    float v = px(o) * 0.299f;
                    v += px(o + plane) * 0.587f;
                    v += px(o + 2 * plane) * 0.114f;
                    raw = __float_as_uint(v);
                } else if (U8) {
                    raw = __ldg(static_cast<const unsigned char*>(p.img) + o);
                } else {
                    raw = __float_as_uint(__ldg(static_cast<const float*>(p.img) + o));
                }
            }
            return raw;
        };
        // uint8 pixels: u8_unit(value) == float32(value / 255.0) of the host conversion (numpy_image_to_torch,
        // utils/image.py:121) for all 256 values
        auto cook = [&](uint32_t raw) { return (U8 && p.C == 1) ? u8_unit((unsigned char)raw) : __uint_as_float(raw); };
        // im2col rows of patch pixels bt, bt + 128 for the tile whose pixels are staged in img_s:
        // [hi(9) | lo(9) | hi(9) | 0], written to K columns [32 par, 32 par + 32) of the pixel's row
        // (chunks 4 par .. 4 par + 3 under the 128-byte swizzle)
        auto im2col = [&](const float* img_s, int par) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int pi = bt + j * 128;
                if (pi >= kStPatchPx) continue;
                const int pr = pi / kStPatchW, pc = pi - pr * kStPatchW;
                float f[32];
#pragma unroll
                for (int k = 27; k < 32; ++k) f[k] = 0.f;
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    float hi, lo;
                    st_split(img_s[(pr + k / 3) * kStImgW + pc + k % 3], hi, lo);
                    f[k] = hi;
                    f[9 + k] = lo;
                    f[18 + k] = hi;
                }
                uint8_t* row = sA + pi * 128;  // M-tile pi / 128, row pi % 128
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    *reinterpret_cast<uint4*>(row + (((par * 4 + c) ^ (pi & 7)) << 4)) =
                        make_uint4(pack_bf16(f[c * 8], f[c * 8 + 1]), pack_bf16(f[c * 8 + 2], f[c * 8 + 3]),
                                   pack_bf16(f[c * 8 + 4], f[c * 8 + 5]), pack_bf16(f[c * 8 + 6], f[c * 8 + 7]));
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_full + par);
        };
        uint32_t nxt0 = 0u, nxt1 = 0u;
        StTile cur = decomp((int)blockIdx.x);  // tile t of the loop below
        StTile pf = cur;                       // tile whose image pixels are requested next
        if (builder && blockIdx.x < total_tiles) {
            sImg[bt] = cook(load_img(pf, bt));
            sImg[bt + 128] = cook(load_img(pf, bt + 128));
            st_build_bar();
            im2col(sImg, 0);
            advance(pf);
            if (blockIdx.x + gridDim.x < total_tiles) {
                nxt0 = load_img(pf, bt);
                nxt1 = load_img(pf, bt + 128);
            }
        }
        uint32_t r[32];
        int i = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++i, advance(cur)) {
            const int st = i & 1;
            const int ty = cur.ty, tx = cur.tx;
            const int y0 = ty * kStTileH - 1, x0 = tx * kStTileW - 1;  // image position of patch pixel (0, 0)
            // ---- next tile first: its stem GEMM then runs while this tile's patch is written ----
            if (builder && t + gridDim.x < total_tiles) {
                sImg[(st ^ 1) * 256 + bt] = cook(nxt0);
                sImg[(st ^ 1) * 256 + bt + 128] = cook(nxt1);
                st_build_bar();  // the next tile's image pixels are staged
                im2col(sImg + (st ^ 1) * 256, st ^ 1);
                advance(pf);  // tile t + 2 grid strides
                if (t + 2 * gridDim.x < total_tiles) {  // in flight for a whole tile
                    nxt0 = load_img(pf, bt);
                    nxt1 = load_img(pf, bt + 128);
                }
            }
            mbar_wait(c1_full + st, (uint32_t)((i >> 1) & 1));  // stem GEMM of this tile retired
            tc_fence_after();
            mbar_wait(patch_empty + st, (uint32_t)(((i >> 1) & 1) ^ 1));  // the tile before last is done with the slot
            // ---- TMEM -> bias -> ReLU -> BN -> bf16 -> halo patch (zeros outside the image) ----
            uint8_t* patch = sPatch + st * kStPatchBytes;
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                if (m * 128 + quarter * 32 >= kStPatchPx) continue;  // warp-uniform: no data in these TMEM lanes
                const int pi = m * 128 + quarter * 32 + lane;
                const int pr = pi / kStPatchW, pc = pi - pr * kStPatchW;
                const int y = y0 + pr, x = x0 + pc;
                const bool inside = y >= 0 && y < p.H && x >= 0 && x < p.W;
                const uint32_t taddr = tmem_base + 128 + st * 128 + m * 64 + ((uint32_t)(quarter * 32) << 16);
                uint8_t* dst = patch + (pr * kStPatchPitch + pc) * 128;
                tmem_ld32(taddr + chalf * 32, r);
                tmem_ld_wait();
                uint32_t pk[16];
                if (chalf == 0) stem_affine(std::integral_constant<int, 0>{}, r, inside, pk);
                else stem_affine(std::integral_constant<int, 1>{}, r, inside, pk);
                if (pi < kStPatchPx) {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        *reinterpret_cast<uint4*>(dst + (((chalf * 4 + q) ^ (pc & 7)) << 4)) =
                            make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]);
                }
            }
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(patch_full + st);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<kStTmemCols>(tmem_base);
}

int launch_conv_stem(const CUtensorMap& tw, const StemParams& p, int num_sms, cudaStream_t stream) {
    static PerDeviceSize configured;
    if (!configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(conv_stem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStSmemBytes));
        B200_CHECK_CUDA(cudaFuncSetAttribute(conv_stem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStSmemBytes));
        configured.cur() = 1;
    }
    const long long total = (long long)p.B * p.tiles_x * p.tiles_y;
    if (total == 0) return 0;
    B200_REQUIRE(total < (1ll << 31), "conv_stem: too many tiles");
    const int grid = (int)(total < num_sms ? total : num_sms);
    if (p.u8) B200_CHECK_CUDA(launch_pdl(conv_stem_kernel<true>, dim3(grid), dim3(kStThreads), kStSmemBytes, stream, tw, p));
    else B200_CHECK_CUDA(launch_pdl(conv_stem_kernel<false>, dim3(grid), dim3(kStThreads), kStSmemBytes, stream, tw, p));
    return 0;
}

}  // namespace b200
