// Dense path: persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   C[M, N] = epilogue( A[M, K] . B[N, K]^T ),  bf16 operands, fp32 accumulation in TMEM.
//
// * warp 0: TMA producer (A and B tiles, 128-byte swizzle, ring of kStages smem slots)
// * warp 1: TMEM allocation + single-thread tcgen05.mma issue, tcgen05.commit to mbarriers
// * warps 2-9: epilogue, two warps per TMEM lane quarter taking alternate 32-column chunks; per-column
//   operands come from shared memory; row-major outputs are staged per warp in shared memory so global
//   loads / stores (residual, fp32 / bf16 outputs) are coalesced 128-byte row segments
// The accumulator is double buffered in TMEM (except BLOCK_N = 512) so the epilogue of tile i
// overlaps the main loop of tile i+1.  With A_MODE = 1 the A operand is gathered by 4-D TMA boxes
// from an NHWC activation tensor (implicit-GEMM 3x3 convolution; out-of-bounds box elements are
// zero-filled by TMA, which implements the zero padding).
//
// Replaces the cuDNN / cuBLAS calls behind nn.Conv2d and nn.Linear on the reference path
// (reference: gluefactory/models/extractors/superpoint_open.py:59-115,
//  gluefactory/models/matchers/lightglue.py:125-215,265-281).
#include <cstdlib>

#include "common.cuh"
#include "gemm.cuh"

namespace b200 {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int kTileH = 8, kTileW = 16;  // conv M-tile: 8 x 16 pixels
// epilogue warps: two per TMEM lane quarter (interleaved 32-column chunks); four per quarter for the
// LayerNorm + GELU epilogue, which is instruction-latency-bound per warp
constexpr int kEpiWarps = 8;
constexpr int kEpiWarpsLn = 16;
constexpr int kLnPartDefault = 128;  // measured in the step: 1.807 ms (128) vs 1.873 ms (256) per 18 launches
// the QKV + rotary epilogue is instruction-bound too (86 us per launch against 49 us of MMAs with eight warps)
#ifndef GEMM_QKV_EPI_WARPS
#define GEMM_QKV_EPI_WARPS 16
#endif
#ifndef GEMM_BIAS_EPI_WARPS
#define GEMM_BIAS_EPI_WARPS 8
#endif
#ifndef GEMM_RESID_EPI_WARPS
#define GEMM_RESID_EPI_WARPS 8
#endif
__host__ __device__ constexpr int epi_warps(int epi) {
    return epi == EPI_LN_GELU ? kEpiWarpsLn : epi == EPI_QKV_ROTARY ? GEMM_QKV_EPI_WARPS
         : epi == EPI_BIAS_BF16 ? GEMM_BIAS_EPI_WARPS : epi == EPI_RESID ? GEMM_RESID_EPI_WARPS : kEpiWarps;
}
__host__ __device__ constexpr int gemm_threads(int epi) { return 64 + 32 * epi_warps(epi); }
constexpr int kParamFloats = 3 * 768;   // per-column epilogue operands staged in smem (bias | scale/gamma | shift/beta)
// staging per epilogue warp: 32 rows x 128 B (fp32 row epilogues) or 32 rows x 64 B (bf16 rows: plain bf16 and
// LayerNorm + GELU epilogues); the LayerNorm epilogue adds the per-row statistics exchange
__host__ __device__ constexpr int stage_bytes_per_warp(int epi) { return (epi == EPI_LN_GELU || epi == EPI_BIAS_BF16) ? 2048 : 4096; }
__host__ __device__ constexpr int epi_smem_bytes(int epi) {
    return kParamFloats * 4 + epi_warps(epi) * stage_bytes_per_warp(epi) + (epi == EPI_LN_GELU ? 2 * 128 * 4 * 8 : 0);
}

template <int BLOCK_N, int EPI>
struct GemmCfg {
    static constexpr int kABytes = kBlockM * kBlockK * 2;
    static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    // the ring takes what the epilogue leaves: four 48 KB stages for the plain bf16 epilogue at 256 columns (one whole
    // K = 256 tile in flight), three for the fp32-staged ones
    static constexpr int kRingBudget = 227 * 1024 - 1024 - 256 - epi_smem_bytes(EPI);
    static constexpr int kStages = (kRingBudget / kStageBytes) > 8 ? 8 : (kRingBudget / kStageBytes);
    // LayerNorm epilogue: the 512-column row lives in TMEM as 512 / BLOCK_N accumulator stages (column parts)
    static constexpr int kAccStages = EPI == EPI_LN_GELU ? 512 / BLOCK_N : (BLOCK_N <= 256 ? 2 : 1);
    static constexpr int kAccStride = BLOCK_N <= 64 ? 64 : (BLOCK_N <= 128 ? 128 : (BLOCK_N <= 256 ? 256 : 512));
    static constexpr int kTmemCols = kAccStride * kAccStages < 32 ? 32 : kAccStride * kAccStages;
    static constexpr int kUmmaN = BLOCK_N > 256 ? 256 : BLOCK_N;
    static constexpr int kNumUmmaN = BLOCK_N / kUmmaN;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + epi_smem_bytes(EPI);
    static_assert((2 * kStages + 2 * kAccStages) * 8 + 8 <= 256, "barrier block");
    static_assert(kStageBytes % 1024 == 0, "stage must keep 1024-byte alignment for SWIZZLE_128B");
    static_assert(kTmemCols <= 512, "TMEM budget");
    static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

// GELU(erf) = x Phi(x) = max(x, 0) - a Phi(-a), a = |x|, with Phi(-a) = 2^q(a): q is the degree-6 weighted minimax
// fit of log2(erfc(a / sqrt 2) / 2) on [0, 6] (beyond 6, a is clamped: Phi(-6) = 1e-9).  Measured against float64
// over [-9, 9] in fp32 arithmetic: absolute error <= 5.1e-7, relative <= 1e-5 for x > -3 (the Abramowitz-Stegun
// 7.1.26 form used before: 5e-7 / 8e-5, with a reciprocal and an exponential per element).  ONE MUFU and 7 FMA-pipe
// instructions per element instead of ~25 for erff: this epilogue is MUFU- and issue-bound.
// The polynomial is evaluated in n = -a (odd coefficients change sign), so that the result is one FMA: n E + relu.
constexpr float kGeluQ0 = -9.999930263e-01f, kGeluQ1 = 1.151201725e+00f, kGeluQ2 = -4.587708414e-01f,
                kGeluQ3 = 5.341226608e-02f, kGeluQ4 = 8.080814034e-03f, kGeluQ5 = 7.692475338e-04f,
                kGeluQ6 = 3.309584281e-05f;
__device__ __forceinline__ float gelu_erf(float x) {
    const float n = fmaxf(-fabsf(x), -6.0f);
    float q = fmaf(n, kGeluQ6, kGeluQ5);
    q = fmaf(q, n, kGeluQ4);
    q = fmaf(q, n, kGeluQ3);
    q = fmaf(q, n, kGeluQ2);
    q = fmaf(q, n, kGeluQ1);
    q = fmaf(q, n, kGeluQ0);
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(q));
    return fmaf(n, e, fmaxf(x, 0.f));
}

// packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2 on sm_100: two fp32 operations per issue slot).  A pair lives in
// an aligned 64-bit register pair; consecutive tcgen05.ld results and float4 halves already are such pairs.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ f32x2 bc2(float a) { return pk2(a, a); }
__device__ __forceinline__ void upk2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// gelu_erf on two elements at once: 4 FMNMX + 7 packed FMA-pipe instructions + 2 MUFU per pair
__device__ __forceinline__ f32x2 gelu_erf2(f32x2 x) {
    float x0, x1;
    upk2(x, x0, x1);
    const f32x2 n = pk2(fmaxf(-fabsf(x0), -6.0f), fmaxf(-fabsf(x1), -6.0f));
    const f32x2 relu = pk2(fmaxf(x0, 0.f), fmaxf(x1, 0.f));
    f32x2 q = fma2(n, bc2(kGeluQ6), bc2(kGeluQ5));
    q = fma2(q, n, bc2(kGeluQ4));
    q = fma2(q, n, bc2(kGeluQ3));
    q = fma2(q, n, bc2(kGeluQ2));
    q = fma2(q, n, bc2(kGeluQ1));
    q = fma2(q, n, bc2(kGeluQ0));
    float q0, q1, e0, e1;
    upk2(q, q0, q1);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(q0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(q1));
    return fma2(n, pk2(e0, e1), relu);
}

__device__ __forceinline__ void store16B(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    *reinterpret_cast<uint4*>(p) = make_uint4(a, b, c, d);
}

template <int WARPS>
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(32 * WARPS) : "memory"); }

// epilogue-side view of shared memory
struct EpiSmem {
    const float* bias;    // [768]
    const float* par1;    // [768] scale / LayerNorm gamma
    const float* par2;    // [768] shift / LayerNorm beta
    uint8_t* stage;       // this warp's 4 KB staging tile
    float2* exch;         // [2][4][128] per-row partial statistics of the four column groups (two tiles in flight)
};

// ---------------------------------------------------------------------------------------------
// epilogue bodies.  `acc_taddr` addresses column 0 of this tile's accumulator for this warp's TMEM
// lanes.  Two warps share a lane quarter and take alternate 32-column chunks (`half`).
// Row-major outputs go through a per-warp staging tile so that global loads / stores are coalesced:
// phase A: thread <-> accumulator row (per-column math), phase B: 8 lanes <-> one 128-byte row segment.
// ---------------------------------------------------------------------------------------------
// explicit shared-space accesses: through the generic pointers of EpiSmem the compiler emits generic LD.E / ST.E
// (long-scoreboard latency) and, not knowing that the staging tile and the operands do not alias, serialises
// every load behind the previous store.  The loads below are issued in batches ahead of their consumers.
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

template <int BLOCK_N, int EPI>
__device__ __forceinline__ void epilogue_rows(const GemmParams& p, const EpiSmem& es, uint32_t acc_taddr, int batch,
                                              long long tile_row0, int half, int n0, int lane, uint64_t* acc_full_bar,
                                              uint32_t acc_full_phase) {
    uint32_t r[32];
    const bool has_bias = p.bias != nullptr;
    const long long row_base = (long long)batch * p.M;
    const uint32_t stage = smem_u32(es.stage);
    const uint32_t bias_s = smem_u32(es.bias);
    const int cj = lane & 7, rsub = lane >> 3;
    // residual / rotary operands of one 32-column chunk: 8 row groups x 16 B per lane.  They are requested one
    // chunk ahead (while the previous chunk is converted and stored), so their L2 / HBM latency is off the
    // per-chunk critical path; out32 may alias resid, but only within the same chunk (same columns), and
    // there the loads still precede the stores.
    float4 aux[8], aux_next[8];
    auto load_aux = [&](int c, float4 (&dst)[8]) {
        const int col0 = n0 + c * 32;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const long long row_in_batch = tile_row0 + it * 4 + rsub;
            const int col = col0 + cj * 4;
            dst[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c < BLOCK_N / 32 && col0 < p.N && row_in_batch < p.M) {
                const long long grow = row_base + row_in_batch;
                if constexpr (EPI == EPI_RESID)
                    dst[it] = *reinterpret_cast<const float4*>(p.resid + grow * p.ld32 + col);
                else if (col0 < p.rot_cols)
                    dst[it] = __ldg(reinterpret_cast<const float4*>(p.cs + grow * 32 + ((col & 63) >> 1)));
            }
        }
    };
    // first operands are requested before the accumulator is awaited.  Rotary: a warp's chunks (half, half + 2, ...)
    // all cover the same 16 frequencies of the same rows (64-column heads, 256-column tiles), so the (cos, sin)
    // slice is loaded once per tile, not per chunk.
    if constexpr (EPI == EPI_RESID || EPI == EPI_QKV_ROTARY) load_aux(half, aux);
    mbar_wait(acc_full_bar, acc_full_phase);
    tc_fence_after();
    if constexpr (EPI == EPI_BIAS_BF16) {
        // bf16 output with per-column operands only: bias (+ ReLU) and the bf16 rounding happen in the thread <-> row
        // phase (packed fp32x2 adds, bias by broadcast loads under the TMEM load), so the staging tile holds 64-byte bf16
        // rows - half the staging traffic and half the load / store instructions of the fp32 path below - and the ring
        // gets the shared memory this saves.
        const int cj4 = lane & 3, rsub8 = lane >> 2;
#pragma unroll 1
        for (int c = half; c < BLOCK_N / 32; c += epi_warps(EPI) / 4) {
            const int col0 = n0 + c * 32;
            if (col0 >= p.N) break;  // warp-uniform
            tmem_ld32(acc_taddr + c * 32, r);
            float4 b[8];
#pragma unroll
            for (int q = 0; q < 8; ++q)
                b[q] = has_bias ? lds128(bias_s + (col0 + q * 4) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                float y0, y1, y2, y3;
                upk2(add2(pk2(__uint_as_float(r[q * 4]), __uint_as_float(r[q * 4 + 1])), pk2(b[q].x, b[q].y)), y0, y1);
                upk2(add2(pk2(__uint_as_float(r[q * 4 + 2]), __uint_as_float(r[q * 4 + 3])), pk2(b[q].z, b[q].w)), y2, y3);
                if (p.relu) { y0 = fmaxf(y0, 0.f); y1 = fmaxf(y1, 0.f); y2 = fmaxf(y2, 0.f); y3 = fmaxf(y3, 0.f); }
                pk[q * 2] = pack_bf16(y0, y1);
                pk[q * 2 + 1] = pack_bf16(y2, y3);
            }
            // 64-byte rows: 16-byte chunk j of row l sits at l*64 + ((j ^ ((l >> 1) & 3)) << 4)
            const uint32_t st = stage;
            const uint32_t myrow = st + lane * 64;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                sts128(myrow + ((j ^ ((lane >> 1) & 3)) << 4), pk[j * 4], pk[j * 4 + 1], pk[j * 4 + 2], pk[j * 4 + 3]);
            __syncwarp();
            float4 v4[4];
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int rr = it * 8 + rsub8;
                v4[it] = lds128(st + rr * 64 + ((cj4 ^ ((rr >> 1) & 3)) << 4));
            }
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const long long row_in_batch = tile_row0 + it * 8 + rsub8;
                if (row_in_batch < p.M)
                    *reinterpret_cast<float4*>(reinterpret_cast<__nv_bfloat16*>(p.out16) + (row_base + row_in_batch) * p.ld16 +
                                               col0 + cj4 * 8) = v4[it];
            }
            __syncwarp();
        }
        return;
    }
#pragma unroll 1
    for (int c = half; c < BLOCK_N / 32; c += epi_warps(EPI) / 4) {
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        // phase A: thread <-> accumulator row; raw accumulators into the swizzled staging tile
        tmem_ld32(acc_taddr + c * 32, r);
        tmem_ld_wait();
        const uint32_t myrow = stage + lane * 128;
#pragma unroll
        for (int q = 0; q < 8; ++q)
            sts128(myrow + ((q ^ (lane & 7)) << 4), r[q * 4], r[q * 4 + 1], r[q * 4 + 2], r[q * 4 + 3]);
        __syncwarp();
        // phase B: 8 lanes <-> one 128-byte row segment; every staging load is in flight before the first use
        float4 v8[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            v8[it] = lds128(stage + rr * 128 + ((cj ^ (rr & 7)) << 4));
        }
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (has_bias) bv = lds128(bias_s + (col0 + cj * 4) * 4);
        if constexpr (EPI == EPI_RESID) load_aux(c + epi_warps(EPI) / 4, aux_next);  // next chunk of this warp
        const int col = col0 + cj * 4;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            float4 v = v8[it];
            if (EPI == EPI_F32) { v.x *= p.alpha; v.y *= p.alpha; v.z *= p.alpha; v.w *= p.alpha; }
            v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
            if (EPI == EPI_BIAS_BF16 && p.relu) {
                v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
            }
            const long long row_in_batch = tile_row0 + rr;
            if (row_in_batch >= p.M) continue;
            const long long grow = row_base + row_in_batch;
            if constexpr (EPI == EPI_QKV_ROTARY) {
                if (col0 < p.rot_cols) {
                    // (cos, sin) of frequencies (col & 63) / 2 and + 1
                    const float4 t = aux[it];
                    const float x0 = v.x, x1 = v.y, x2 = v.z, x3 = v.w;
                    v.x = x0 * t.x - x1 * t.y;
                    v.y = x1 * t.x + x0 * t.y;
                    v.z = x2 * t.z - x3 * t.w;
                    v.w = x3 * t.z + x2 * t.w;
                }
            }
            if constexpr (EPI == EPI_RESID) {
                const float4 t = aux[it];
                v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
                *reinterpret_cast<float4*>(p.out32 + grow * p.ld32 + col) = v;
            }
            if constexpr (EPI == EPI_F32) {
                float* o32 = p.out32 + (long long)batch * p.out32_batch_stride + row_in_batch * p.ld32 + col;
                if ((p.ld32 & 3) == 0 && col + 4 <= p.N) {
                    *reinterpret_cast<float4*>(o32) = v;
                } else {
                    if (col < p.N) o32[0] = v.x;
                    if (col + 1 < p.N) o32[1] = v.y;
                    if (col + 2 < p.N) o32[2] = v.z;
                    if (col + 3 < p.N) o32[3] = v.w;
                }
            } else {
                __nv_bfloat16* o16 = reinterpret_cast<__nv_bfloat16*>(p.out16) + grow * p.ld16 + col;
                *reinterpret_cast<uint2*>(o16) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
            }
        }
        if constexpr (EPI == EPI_RESID) {
#pragma unroll
            for (int it = 0; it < 8; ++it) aux[it] = aux_next[it];
        }
        __syncwarp();
    }
}

// Linear(512 -> 512) -> LayerNorm(512) -> GELU(erf).  The row is computed as two 256-column halves, one per
// TMEM accumulator stage, so that the MMAs of the next row tile's first half run under the second half of this
// epilogue (a single 128 x 512 fp32 accumulator would fill the TMEM and serialise MMA and epilogue).
// `grp` (0..3) = this warp's column group inside a part: 32-column chunks grp, grp + 4, ... of the part's PART_N / 32.
template <int PART_N>
__device__ __forceinline__ void ln_stats_half(const EpiSmem& es, uint32_t half_taddr, int col_base, int grp,
                                              float& sum, float& sumsq) {
    uint32_t r[32];
    const uint32_t bias_s = smem_u32(es.bias);
    f32x2 sum2 = bc2(0.f), sq2 = bc2(0.f);
#pragma unroll 1
    for (int c = grp; c < PART_N / 32; c += 4) {
        tmem_ld32(half_taddr + c * 32, r);
        float4 b[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) b[q] = lds128(bias_s + (col_base + c * 32 + q * 4) * 4);  // under the TMEM load
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const f32x2 xa = add2(pk2(__uint_as_float(r[q * 4]), __uint_as_float(r[q * 4 + 1])), pk2(b[q].x, b[q].y));
            const f32x2 xb = add2(pk2(__uint_as_float(r[q * 4 + 2]), __uint_as_float(r[q * 4 + 3])), pk2(b[q].z, b[q].w));
            sum2 = add2(sum2, add2(xa, xb));
            sq2 = fma2(xa, xa, fma2(xb, xb, sq2));
        }
    }
    float s0, s1, q0, q1;
    upk2(sum2, s0, s1);
    upk2(sq2, q0, q1);
    sum += s0 + s1;
    sumsq += q0 + q1;
}

template <int PART_N>
__device__ __forceinline__ void ln_gelu_half(const GemmParams& p, const EpiSmem& es, uint32_t half_taddr, int col_base,
                                             long long tile_row0, int grp, int lane, float rstd, float nmr) {
    uint32_t r[32];
    const uint32_t bias_s = smem_u32(es.bias), g_s = smem_u32(es.par1), be_s = smem_u32(es.par2);
    const uint32_t stage = smem_u32(es.stage);
    const f32x2 rstd2 = bc2(rstd), nmr2 = bc2(nmr);
#pragma unroll 1
    for (int c = grp; c < PART_N / 32; c += 4) {
        tmem_ld32(half_taddr + c * 32, r);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const uint32_t off = (col_base + c * 32 + q * 4) * 4;
            const float4 b = lds128(bias_s + off);
            const float4 g = lds128(g_s + off);
            const float4 be = lds128(be_s + off);
            const f32x2 xa = add2(pk2(__uint_as_float(r[q * 4]), __uint_as_float(r[q * 4 + 1])), pk2(b.x, b.y));
            const f32x2 xb = add2(pk2(__uint_as_float(r[q * 4 + 2]), __uint_as_float(r[q * 4 + 3])), pk2(b.z, b.w));
#ifndef GEMM_TIMING_LN_EPI  // timing experiments only: 1 = no GELU, 2 = neither statistics nor GELU (wrong results)
#define GEMM_TIMING_LN_EPI 0
#endif
#if GEMM_TIMING_LN_EPI == 0
            const f32x2 ya = gelu_erf2(fma2(fma2(xa, rstd2, nmr2), pk2(g.x, g.y), pk2(be.x, be.y)));
            const f32x2 yb = gelu_erf2(fma2(fma2(xb, rstd2, nmr2), pk2(g.z, g.w), pk2(be.z, be.w)));
#else
            const f32x2 ya = fma2(fma2(xa, rstd2, nmr2), pk2(g.x, g.y), pk2(be.x, be.y));
            const f32x2 yb = fma2(fma2(xb, rstd2, nmr2), pk2(g.z, g.w), pk2(be.z, be.w));
#endif
            float y0, y1, y2, y3;
            upk2(ya, y0, y1);
            upk2(yb, y2, y3);
            pk[q * 2] = pack_bf16(y0, y1);
            pk[q * 2 + 1] = pack_bf16(y2, y3);
        }
        // staging rows are 64 B (32 bf16): 16-byte chunk j of row l sits at l*64 + ((j ^ ((l >> 1) & 3)) << 4)
        const uint32_t myrow = stage + lane * 64;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            sts128(myrow + ((j ^ ((lane >> 1) & 3)) << 4), pk[j * 4], pk[j * 4 + 1], pk[j * 4 + 2], pk[j * 4 + 3]);
        __syncwarp();
        float4 v4[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int rr = it * 8 + (lane >> 2), cj = lane & 3;
            v4[it] = lds128(stage + rr * 64 + ((cj ^ ((rr >> 1) & 3)) << 4));
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int rr = it * 8 + (lane >> 2), cj = lane & 3;
            const long long grow = tile_row0 + rr;
            if (grow < p.M)
                *reinterpret_cast<float4*>(reinterpret_cast<__nv_bfloat16*>(p.out16) + grow * p.ld16 + col_base + c * 32 +
                                           cj * 8) = v4[it];
        }
        __syncwarp();
    }
}

// thread-per-row epilogues of the SuperPoint convolutions / heads (per-column operands from smem)
template <int BLOCK_N, int EPI>
__device__ __forceinline__ void epilogue_conv(const GemmParams& p, const EpiSmem& es, uint32_t acc_taddr,
                                              long long row_in_batch, bool row_valid, int half, int n0, int lane,
                                              int img, int py, int px) {
    uint32_t r[32];
    if constexpr (EPI == EPI_CONV_RELU_BN) {
        // pixel (img, py, px); with pooling the 2x2 partners are lanes ^1 (x) and ^16 (y)
        const bool store_lane = p.pool ? ((lane & 17) == 0) : true;
        const int oy = p.pool ? (py >> 1) : py, ox = p.pool ? (px >> 1) : px;
        const int OH = p.pool ? (p.H >> 1) : p.H, OW = p.pool ? (p.W >> 1) : p.W;
        const bool ok = row_valid && store_lane && oy < OH && ox < OW;
        __nv_bfloat16* o16 = reinterpret_cast<__nv_bfloat16*>(p.out16) +
                             (((long long)img * OH + oy) * OW + ox) * p.ld16 + n0;
#pragma unroll 1
        for (int c = half; c < BLOCK_N / 32; c += 2) {
            tmem_ld32(acc_taddr + c * 32, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int col = n0 + c * 32 + q * 4;
                const float4 b = *reinterpret_cast<const float4*>(es.bias + col);
                const float4 s = *reinterpret_cast<const float4*>(es.par1 + col);
                const float4 t = *reinterpret_cast<const float4*>(es.par2 + col);
                float x0 = __uint_as_float(r[q * 4]) + b.x, x1 = __uint_as_float(r[q * 4 + 1]) + b.y;
                float x2 = __uint_as_float(r[q * 4 + 2]) + b.z, x3 = __uint_as_float(r[q * 4 + 3]) + b.w;
                if (p.relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f); }
                v[q * 4] = fmaf(x0, s.x, t.x);
                v[q * 4 + 1] = fmaf(x1, s.y, t.y);
                v[q * 4 + 2] = fmaf(x2, s.z, t.z);
                v[q * 4 + 3] = fmaf(x3, s.w, t.w);
            }
            if (p.pool) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float x = v[i];
                    x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
                    x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 16));
                    v[i] = x;
                }
            }
            if (ok) {
#pragma unroll
                for (int i = 0; i < 32; i += 8)
                    store16B(o16 + c * 32 + i, pack_bf16(v[i], v[i + 1]), pack_bf16(v[i + 2], v[i + 3]),
                             pack_bf16(v[i + 4], v[i + 5]), pack_bf16(v[i + 6], v[i + 7]));
            }
        }
    } else if constexpr (EPI == EPI_CONV_BN_L2) {
        if (half != 0) return;  // one warp per lane quarter keeps the whole 256-column row
        float sumsq = 0.f;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 32; ++c) {
            tmem_ld32(acc_taddr + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int col = c * 32 + i;
                const float x = (__uint_as_float(r[i]) + es.bias[col]) * es.par1[col] + es.par2[col];
                sumsq += x * x;
            }
        }
        const float inv = 1.0f / fmaxf(sqrtf(sumsq), 1e-12f);
        float* o32 = p.out32 + row_in_batch * p.ld32;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 32; ++c) {
            tmem_ld32(acc_taddr + c * 32, r);
            tmem_ld_wait();
            if (!row_valid) continue;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                float v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = c * 32 + i + j;
                    v[j] = ((__uint_as_float(r[i + j]) + es.bias[col]) * es.par1[col] + es.par2[col]) * inv;
                }
                *reinterpret_cast<float4*>(o32 + c * 32 + i) = make_float4(v[0], v[1], v[2], v[3]);
            }
        }
    } else if constexpr (EPI == EPI_CONV_BN_SOFTMAX_D2S) {
        if (half != 0) return;
        // 65 logits per cell (columns 0..64 of an 80-wide tile)
        float v[80];
        uint32_t q[16];
        tmem_ld32(acc_taddr, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
        tmem_ld32(acc_taddr + 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) v[32 + i] = __uint_as_float(r[i]);
        tmem_ld16(acc_taddr + 64, q);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[64 + i] = __uint_as_float(q[i]);
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 65; ++i) {
            v[i] = (v[i] + es.bias[i]) * es.par1[i] + es.par2[i];
            mx = fmaxf(mx, v[i]);
        }
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < 65; ++i) {
            v[i] = expf(v[i] - mx);
            sum += v[i];
        }
        if (row_valid) {
            const int cells = p.cells_h * p.cells_w;
            const int im = (int)(row_in_batch / cells);
            const int rem = (int)(row_in_batch - (long long)im * cells);
            const int cy = rem / p.cells_w, cx = rem - cy * p.cells_w;
            const long long hw = (long long)p.cells_w * 8;
            float* o = p.out32 + ((long long)im * p.cells_h * 8 + cy * 8) * hw + cx * 8;
#pragma unroll
            for (int dy = 0; dy < 8; ++dy) {
                float* row = o + dy * hw;
                *reinterpret_cast<float4*>(row) = make_float4(v[dy * 8] / sum, v[dy * 8 + 1] / sum,
                                                              v[dy * 8 + 2] / sum, v[dy * 8 + 3] / sum);
                *reinterpret_cast<float4*>(row + 4) = make_float4(v[dy * 8 + 4] / sum, v[dy * 8 + 5] / sum,
                                                                  v[dy * 8 + 6] / sum, v[dy * 8 + 7] / sum);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------------------------
template <int BLOCK_N, int EPI, int A_MODE>
__global__ void __launch_bounds__(gemm_threads(EPI), 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const GemmParams p) {
    using Cfg = GemmCfg<BLOCK_N, EPI>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* acc_full = empty_bar + Cfg::kStages;
    uint64_t* acc_empty = acc_full + Cfg::kAccStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + Cfg::kAccStages);
    float* s_par = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + 256);
    uint8_t* s_stage = reinterpret_cast<uint8_t*>(s_par + kParamFloats);
    float2* s_exch = reinterpret_cast<float2*>(s_stage + epi_warps(EPI) * stage_bytes_per_warp(EPI));  // LayerNorm only

    // warp index / TMEM base through a shuffle: the compiler then keeps every UMMA operand in uniform registers
    // (per-lane operands cost an ELECT / R2UR loop of ~75 clk around each tcgen05.mma)
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;

    const int m_tiles = A_MODE == 1 ? p.tiles_x * p.tiles_y : (p.M + kBlockM - 1) / kBlockM;
    // LayerNorm epilogue: a tile is a full 512-column row block, produced as kHalves column halves by the same CTA
    constexpr int kHalves = EPI == EPI_LN_GELU ? 512 / BLOCK_N : 1;
    const int n_tiles = EPI == EPI_LN_GELU ? 1 : (p.N + BLOCK_N - 1) / BLOCK_N;
    const long long total_tiles = (long long)p.batch * m_tiles * n_tiles;
    const int k_blocks = p.K / kBlockK;
    // tile coordinates advance by the grid stride as a mixed-radix addition (a fresh decomposition costs 64-bit
    // divisions per tile in the producer and in every epilogue warp)
    struct GTile {
        int nt, mt, b;
    };
    auto decomp = [&](long long t) {
        GTile c;
        c.nt = (int)(t % n_tiles);
        const long long rest = t / n_tiles;
        c.mt = (int)(rest % m_tiles);
        c.b = (int)(rest / m_tiles);
        return c;
    };
    const GTile tile_step = decomp((long long)gridDim.x);
    auto advance = [&](GTile& c) {
        c.nt += tile_step.nt;
        c.mt += tile_step.mt;
        c.b += tile_step.b;
        if (c.nt >= n_tiles) { c.nt -= n_tiles; ++c.mt; }
        if (c.mt >= m_tiles) { c.mt -= m_tiles; ++c.b; }
    };

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_a);
        prefetch_tensormap(&tmap_b);
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        for (int s = 0; s < Cfg::kAccStages; ++s) {
            mbar_init(acc_full + s, 1);
            mbar_init(acc_empty + s, epi_warps(EPI));  // one arrival per epilogue warp
        }
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
        // ===================== TMA producer =====================
        // whole warp walks the loop (uniform coordinates in uniform registers); one elected lane issues
        const bool leader = elect_one();
        int stage = 0;
        uint32_t phase = 0;
        GTile tc = decomp((long long)blockIdx.x);
        for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x, advance(tc)) {
            const int nt = tc.nt, mt = tc.mt, b = tc.b;
            if (p.tile_live != nullptr && p.tile_live[b * m_tiles + mt] == 0) continue;
            for (int hh = 0; hh < kHalves; ++hh) {
            const int brow = b * p.b_batch_rows + (nt * kHalves + hh) * BLOCK_N;
            for (int kb = 0; kb < k_blocks; ++kb) {
                mbar_wait(empty_bar + stage, phase ^ 1);
                uint8_t* sa = smem + stage * Cfg::kStageBytes;
                uint8_t* sb = sa + Cfg::kABytes;
                if (leader) {
#ifndef GEMM_TIMING_LN_A_ROWS
#define GEMM_TIMING_LN_A_ROWS 128
#endif
                    mbar_arrive_expect_tx(full_bar + stage,
                                          Cfg::kStageBytes - (EPI == EPI_LN_GELU ? (128 - GEMM_TIMING_LN_A_ROWS) * 128 : 0));
                    if constexpr (A_MODE == 0) {
                        tma_load_2d(sa, &tmap_a, full_bar + stage, kb * kBlockK, b * p.a_batch_rows + mt * kBlockM);
                    } else {
                        const int cblocks = p.Cin / kBlockK;
                        const int tap = kb / cblocks, cb = kb - tap * cblocks;
                        const int ty = mt / p.tiles_x, tx = mt - ty * p.tiles_x;
                        tma_load_4d(sa, &tmap_a, full_bar + stage, cb * kBlockK, tx * kTileW + (tap % 3) - 1,
                                    ty * kTileH + (tap / 3) - 1, b);
                    }
#pragma unroll
                    for (int h = 0; h < Cfg::kNumUmmaN; ++h)
                        tma_load_2d(sb + h * Cfg::kUmmaN * kBlockK * 2, &tmap_b, full_bar + stage, kb * kBlockK,
                                    brow + h * Cfg::kUmmaN);
                }
                __syncwarp();
                if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
            }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // the whole warp walks the loop (uniform control flow); one elected lane issues MMAs and commits
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc(kBlockM, Cfg::kUmmaN, 1u, 0u, 0u);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        GTile tc = decomp((long long)blockIdx.x);
        for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x, advance(tc)) {
            if (p.tile_live != nullptr && p.tile_live[tc.b * m_tiles + tc.mt] == 0) continue;
            for (int hh = 0; hh < kHalves; ++hh) {
            mbar_wait(acc_empty + acc, acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * Cfg::kAccStride;
            for (int kb = 0; kb < k_blocks; ++kb) {
                mbar_wait(full_bar + stage, phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
                const uint32_t sb = sa + Cfg::kABytes;
                if (leader) {
#pragma unroll
                    for (int k = 0; k < kBlockK / 16; ++k) {
                        const uint64_t adesc = umma_smem_desc_sw128(sa + k * 32, 1024, 16);
#pragma unroll
                        for (int h = 0; h < Cfg::kNumUmmaN; ++h) {
                            const uint64_t bdesc =
                                umma_smem_desc_sw128(sb + h * Cfg::kUmmaN * kBlockK * 2 + k * 32, 1024, 16);
                            umma_f16(d_tmem + h * Cfg::kUmmaN, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                        }
                    }
                    umma_commit(empty_bar + stage);  // frees the smem slot once these MMAs retire
                }
                __syncwarp();
                if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
            }
            if (leader) umma_commit(acc_full + acc);
            __syncwarp();
            if (++acc == Cfg::kAccStages) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue warps (TMEM lanes 32*(warp%4) ..) =====================
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;
        // per-column operands -> smem once per CTA
        {
            const float* src1 = EPI == EPI_LN_GELU ? p.ln_gamma : p.scale;
            const float* src2 = EPI == EPI_LN_GELU ? p.ln_beta : p.shift;
            const int n = p.N < 768 ? p.N : 768;
            for (int i = et; i < n; i += 32 * epi_warps(EPI)) {
                if (p.bias != nullptr) s_par[i] = __ldg(p.bias + i);
                if (src1 != nullptr) s_par[768 + i] = __ldg(src1 + i);
                if (src2 != nullptr) s_par[1536 + i] = __ldg(src2 + i);
            }
            epi_bar_sync<epi_warps(EPI)>();
        }
        EpiSmem es;
        es.bias = s_par;
        es.par1 = s_par + 768;
        es.par2 = s_par + 1536;
        es.stage = s_stage + (warp - 2) * stage_bytes_per_warp(EPI);
        es.exch = s_exch;
        const int r_in_tile = quarter * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        GTile tc = decomp((long long)blockIdx.x);
        for (long long t = blockIdx.x; t < total_tiles; t += gridDim.x, advance(tc)) {
            const int nt = tc.nt, mt = tc.mt, b = tc.b;
            if (p.tile_live != nullptr && p.tile_live[b * m_tiles + mt] == 0) continue;
            if constexpr (EPI == EPI_LN_GELU) {
                // accumulator stage h holds columns [BLOCK_N h, BLOCK_N (h + 1)) of this row tile; every stage turns over
                // once per tile, so one phase bit serves all.  `half` is this warp's column group (0..3) here.
                // Statistics of part h run under the MMAs of part h + 1; after the last part the row is normalised
                // part by part and each part's accumulator is handed back as soon as it has been read, so the next
                // tile's MMAs start under this tile's remaining parts.  Narrow parts (128 columns) shorten the one
                // stretch nothing overlaps: last part's MMAs -> its statistics -> first part normalised.
                static_assert(Cfg::kAccStages * BLOCK_N == 512 && Cfg::kAccStride == BLOCK_N, "LayerNorm epilogue: 512 columns");
                const uint32_t lane_taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
                const long long row0 = (long long)mt * kBlockM + quarter * 32;
                float sum = 0.f, sumsq = 0.f;
#pragma unroll
                for (int h = 0; h < Cfg::kAccStages; ++h) {
                    mbar_wait(acc_full + h, acc_phase);
                    tc_fence_after();
#if !defined(GEMM_TIMING_LN_EPI) || GEMM_TIMING_LN_EPI != 2
                    ln_stats_half<BLOCK_N>(es, lane_taddr + h * BLOCK_N, h * BLOCK_N, half, sum, sumsq);
#endif
                }
                float2* exch = es.exch + (acc_phase & 1) * 512;
                exch[half * 128 + r_in_tile] = make_float2(sum, sumsq);
                epi_bar_sync<kEpiWarpsLn>();
#pragma unroll
                for (int o = 1; o < 4; ++o) {  // the row's other three column groups
                    const float2 other = exch[((half + o) & 3) * 128 + r_in_tile];
                    sum += other.x;
                    sumsq += other.y;
                }
                const float mean = sum * (1.0f / 512.f);
                const float var = fmaxf(sumsq * (1.0f / 512.f) - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.ln_eps);
                const float nmr = -mean * rstd;
#pragma unroll
                for (int h = 0; h < Cfg::kAccStages; ++h) {
                    ln_gelu_half<BLOCK_N>(p, es, lane_taddr + h * BLOCK_N, h * BLOCK_N, row0, half, lane, rstd, nmr);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_empty + h);
                }
                acc_phase ^= 1;
                continue;
            }
            const uint32_t taddr = tmem_base + acc * Cfg::kAccStride + ((uint32_t)(quarter * 32) << 16);
            if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_QKV_ROTARY || EPI == EPI_RESID || EPI == EPI_F32) {
                epilogue_rows<BLOCK_N, EPI>(p, es, taddr, b, (long long)mt * kBlockM + quarter * 32, half, nt * BLOCK_N, lane,
                                            acc_full + acc, acc_phase);
            } else {
                mbar_wait(acc_full + acc, acc_phase);
                tc_fence_after();
                long long row_in_batch = 0;
                bool row_valid;
                int py = 0, px = 0;
                if constexpr (A_MODE == 1) {
                    const int ty = mt / p.tiles_x, tx = mt - ty * p.tiles_x;
                    py = ty * kTileH + (r_in_tile >> 4);
                    px = tx * kTileW + (r_in_tile & 15);
                    row_valid = py < p.H && px < p.W;
                } else {
                    row_in_batch = (long long)mt * kBlockM + r_in_tile;
                    row_valid = row_in_batch < p.M;
                }
                epilogue_conv<BLOCK_N, EPI>(p, es, taddr, row_in_batch, row_valid, half, nt * BLOCK_N, lane, b, py, px);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty + acc);
            if (++acc == Cfg::kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// host launcher
// ---------------------------------------------------------------------------------------------
// column width of an accumulator part of the LayerNorm epilogue (also the row count of the weight TMA box):
// B200GLUE_LN_PART=128|256 for experiments
int ln_part_cols() {
    static const int v = [] {
        const char* e = getenv("B200GLUE_LN_PART");
        return (e != nullptr && atoi(e) == 256) ? 256 : (e != nullptr && atoi(e) == 128) ? 128 : kLnPartDefault;
    }();
    return v;
}

template <int BLOCK_N, int EPI, int A_MODE>
static int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int num_sms,
                      cudaStream_t stream) {
    using Cfg = GemmCfg<BLOCK_N, EPI>;
    auto kern = gemm_tcgen05_kernel<BLOCK_N, EPI, A_MODE>;
    static PerDeviceSize configured;
    if (!configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
        configured.cur() = 1;
    }
    const int m_tiles = A_MODE == 1 ? p.tiles_x * p.tiles_y : (p.M + kBlockM - 1) / kBlockM;
    const int n_tiles = EPI == EPI_LN_GELU ? 1 : (p.N + BLOCK_N - 1) / BLOCK_N;  // LayerNorm: one CTA per row tile
    const long long total = (long long)p.batch * m_tiles * n_tiles;
    if (total == 0) return 0;
    const int grid = (int)(total < num_sms ? total : num_sms);
    B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(gemm_threads(EPI)), Cfg::kSmemBytes, stream, ta, tb, p));
    return 0;
}

int launch_gemm(int block_n, int epi, int a_mode, const CUtensorMap& ta, const CUtensorMap& tb,
                const GemmParams& p, int num_sms, cudaStream_t stream) {
    B200_REQUIRE(p.K > 0 && p.K % kBlockK == 0, "gemm: K=%d must be a positive multiple of 64", p.K);
#define B200_CASE(BN, E, AM) \
    if (block_n == BN && epi == E && a_mode == AM) return launch_one<BN, E, AM>(ta, tb, p, num_sms, stream);
    B200_CASE(128, EPI_BIAS_BF16, 0)
    B200_CASE(256, EPI_BIAS_BF16, 0)
    B200_CASE(256, EPI_QKV_ROTARY, 0)
    if (block_n == 512 && epi == EPI_LN_GELU && a_mode == 0) {  // 512 columns as 128- or 256-column accumulator parts
        if (ln_part_cols() == 128) return launch_one<128, EPI_LN_GELU, 0>(ta, tb, p, num_sms, stream);
        return launch_one<256, EPI_LN_GELU, 0>(ta, tb, p, num_sms, stream);
    }
    B200_CASE(256, EPI_RESID, 0)
    B200_CASE(128, EPI_F32, 0)
    B200_CASE(256, EPI_F32, 0)
    B200_CASE(64, EPI_CONV_RELU_BN, 1)
    B200_CASE(128, EPI_CONV_RELU_BN, 1)
    B200_CASE(256, EPI_CONV_RELU_BN, 1)
    B200_CASE(256, EPI_CONV_BN_L2, 0)
    B200_CASE(80, EPI_CONV_BN_SOFTMAX_D2S, 0)
#undef B200_CASE
    set_last_error("gemm: no kernel instance for block_n=%d epi=%d a_mode=%d", block_n, epi, a_mode);
    return -1;
}

}  // namespace b200
