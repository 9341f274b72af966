// Dense path: flash-style multi-head attention on tcgen05 / TMEM for head_dim 64.
//
// One CTA per (pair of 128-query tiles, head, sequence); 1 CTA per SM (TMEM: S0 S1 O0 O1 = 384 columns).
//   warp 0      TMA producer: Q tiles once, K/V tiles through a ring of kKvStages smem slots
//   warp 1      TMEM allocation + single-thread tcgen05.mma issue
//   warps 2-9   softmax group 0 (query tile 0): two threads per query row, 64 keys of the tile each
//   warps 10-17 softmax group 1 (query tile 1)
// Per 128-key tile and group g:
//   S_g = Q_g K^T    (tcgen05.mma M=128 N=128 K=64, fp32 in TMEM)
//   the group's 256 threads pull their half S row into registers (tcgen05.ld) and release S_g at once, so
//   the next S_g is computed while this one is exponentiated; P_g -> smem as bf16 (128B-swizzled, K-major).
//   The two threads of a row (same TMEM lane quarter, warps w and w+4) agree on the row maximum through
//   smem and a 64-thread named barrier, keep partial row sums, and each rescales / writes half of O.
//   Four softmax warps per scheduler (instead of two whole-row warps) hide the FFMA -> MUFU -> FADD
//   latencies that the two-warp version left exposed.
//   O_g += P_g V     (tcgen05.mma M=128 N=64 K=128, accumulating in TMEM; V consumed token-major through
//                     an MN-major descriptor)
// Online softmax with a lazily updated reference maximum: the running maximum m only moves (and O_g /
// the row sum are rescaled, tcgen05.ld -> multiply -> tcgen05.st) when a row's new maximum exceeds it by
// more than 2^8 in the exp2 domain; otherwise P is simply taken against the stale m (values <= 256,
// harmless for bf16 / fp32) - the final O / l is mathematically unchanged.
// The two groups ping-pong: while one group runs its exponentials (the MUFU-bound part: 128x128 ex2 per
// tile against 2 x 2 MFLOP of MMA), the tensor core works on the other group's S / PV.
// Q/K/V tiles arrive by TMA from token-major [tokens, ld] bf16 matrices; a head is a 64-column slice.
// The S matrix is never materialised in HBM.  Self-attention and both directions of the bidirectional
// cross-attention use this kernel (the cross block's similarity is recomputed per direction).
//
// Replaces F.scaled_dot_product_attention / the einsum+softmax path of the reference
// (reference: gluefactory/models/matchers/lightglue.py:105-122,144-157,186-215).
#include "common.cuh"
#include "attention.cuh"

namespace b200 {

constexpr int kAttThreads = 576;
constexpr int kSoftmaxWarpsPerGroup = 8;
constexpr int kQTile = 128, kKvTile = 128, kHeadDim = 64;
constexpr int kQPerCta = 2 * kQTile;
constexpr int kTileBytes = 128 * 64 * 2;  // 16 KB: 128 rows x 64 bf16
constexpr int kKvStages = 4;
// smem: Q[2] | K[stages] | V[stages] | barriers | row exchange
constexpr int kAttXchBytes = 3 * 2 * 2 * 128 * 4;  // [2 parities + row sums][group][half][row] floats
constexpr int kAttSmemBytes = kTileBytes * (2 + 2 * kKvStages) + 1024 + 256 + kAttXchBytes;
// S0: [0,128)  S1: [128,256)  O0: [256,320)  O1: [320,384)  P0: [384,448)  P1: [448,512) (bf16 pairs)
constexpr uint32_t kAttTmemCols = 512;

#ifndef ATT_X
#define ATT_X 0
#endif
#if ATT_X & 512
__device__ long long g_att_trace[2][16][8];
#define ATT_T(role, j, slot) do { if (blockIdx.x == 1 && blockIdx.y == 1 && blockIdx.z == 3 && (j) >= 4 && (j) < 12) g_att_trace[role][j][slot] = clock64(); } while (0)
#else
#define ATT_T(role, j, slot) do {} while (0)
#endif
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__global__ void __launch_bounds__(kAttThreads, 1)
attention_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                         const __grid_constant__ CUtensorMap tmap_v, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + 2 * kTileBytes;
    uint8_t* sV = sK + kKvStages * kTileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kKvStages * kTileBytes);
    uint64_t* q_full = bars;                     // [2]
    uint64_t* k_full = q_full + 2;               // [kKvStages]
    uint64_t* v_full = k_full + kKvStages;       // [kKvStages]
    uint64_t* kv_empty = v_full + kKvStages;     // [kKvStages]
    uint64_t* s_full = kv_empty + kKvStages;     // [2]
    uint64_t* s_free = s_full + 2;               // [2]
    uint64_t* p_ready = s_free + 2;              // [2]
    uint64_t* o_full = p_ready + 2;              // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 2);
    float* xch = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

    // warp index / TMEM base / lengths go through a shuffle so that the compiler treats them, and the UMMA
    // operands derived from them, as warp-uniform: tcgen05.mma then issues straight from uniform registers
    // (a per-lane operand costs an ELECT / R2UR loop of ~75 clk around every MMA)
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int seq = blockIdx.z, head = blockIdx.y, qt = blockIdx.x;
    const AttnSeq sd = p.seqs[seq];
    const int q_len = __shfl_sync(0xffffffffu, p.len_override ? p.len_override[sd.q_len_slot] : sd.q_len, 0);
    const int kv_len = __shfl_sync(0xffffffffu, p.len_override ? p.len_override[sd.kv_len_slot] : sd.kv_len, 0);
    if (qt * kQPerCta >= q_len) return;  // uniform for the whole CTA, before any barrier / TMEM use
    const int kv_tiles = (kv_len + kKvTile - 1) / kKvTile;
    const bool g1_active = qt * kQPerCta + kQTile < q_len;  // second query tile holds at least one row

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_q);
        prefetch_tensormap(&tmap_k);
        prefetch_tensormap(&tmap_v);
        for (int g = 0; g < 2; ++g) {
            mbar_init(q_full + g, 1);
            mbar_init(s_full + g, 1);
            mbar_init(s_free + g, kSoftmaxWarpsPerGroup);
            mbar_init(p_ready + g, kSoftmaxWarpsPerGroup);
            mbar_init(o_full + g, 1);
        }
        for (int s = 0; s < kKvStages; ++s) {
            mbar_init(k_full + s, 1);
            mbar_init(v_full + s, 1);
            mbar_init(kv_empty + s, 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kAttTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int col = head * kHeadDim;
            const int qrow = sd.q_row0 + qt * kQPerCta;
            mbar_arrive_expect_tx(q_full, kTileBytes);
            tma_load_2d(sQ, &tmap_q, q_full, p.q_col0 + col, qrow);
            if (g1_active) {
                mbar_arrive_expect_tx(q_full + 1, kTileBytes);
                tma_load_2d(sQ + kTileBytes, &tmap_q, q_full + 1, p.q_col0 + col, qrow + kQTile);
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < kv_tiles; ++j) {
                const int row = sd.kv_row0 + j * kKvTile;
                mbar_wait(kv_empty + stage, phase ^ 1);
                mbar_arrive_expect_tx(k_full + stage, kTileBytes);
                tma_load_2d(sK + stage * kTileBytes, &tmap_k, k_full + stage, p.k_col0 + col, row);
                mbar_arrive_expect_tx(v_full + stage, kTileBytes);
                tma_load_2d(sV + stage * kTileBytes, &tmap_v, v_full + stage, p.v_col0 + col, row);
                if (++stage == kKvStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // the whole warp walks the loop (uniform control flow, uniform operands); one lane issues
        if (kv_tiles > 0) {
            const bool leader = elect_one();
            constexpr uint32_t idesc_s = umma_idesc(128, 128, 1u, 0u, 0u);  // Q, K both K-major
            constexpr uint32_t idesc_o = umma_idesc(128, 64, 1u, 0u, 1u);   // P K-major (TMEM), V MN-major
            const int groups = g1_active ? 2 : 1;
            auto issue_s = [&](int g, int stage) {
                const uint32_t aq = smem_u32(sQ + g * kTileBytes), bk = smem_u32(sK + stage * kTileBytes);
                if (leader) {
#pragma unroll
                    for (int k = 0; k < kHeadDim / 16; ++k)
                        umma_f16(tmem_base + g * 128, umma_smem_desc_sw128(aq + k * 32, 1024, 16),
                                 umma_smem_desc_sw128(bk + k * 32, 1024, 16), idesc_s, k != 0 ? 1u : 0u);
                    umma_commit(s_full + g);
                }
                __syncwarp();
            };
            mbar_wait(k_full, 0);
            for (int g = 0; g < groups; ++g) {
                mbar_wait(q_full + g, 0);
                tc_fence_after();
                issue_s(g, 0);
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < kv_tiles; ++j) {
                const int next_stage = stage + 1 == kKvStages ? 0 : stage + 1;
                const uint32_t next_phase = stage + 1 == kKvStages ? phase ^ 1 : phase;
                if (j + 1 < kv_tiles) {
                    mbar_wait(k_full + next_stage, next_phase);
                    for (int g = 0; g < groups; ++g) {
                        mbar_wait(s_free + g, j & 1);  // the group holds S_g(j) in registers
                        tc_fence_after();
                        if (leader) ATT_T(0, j, g * 2);
                        issue_s(g, next_stage);
                        if (leader) ATT_T(0, j, g * 2 + 1);
                    }
                }
                mbar_wait(v_full + stage, phase);
                for (int g = 0; g < groups; ++g) {
                    mbar_wait(p_ready + g, j & 1);  // P_g(j) is in TMEM (and O_g rescaled if it had to be)
                    tc_fence_after();
                    const uint32_t bv = smem_u32(sV + stage * kTileBytes);
                    if (leader) ATT_T(0, j, 4 + g * 2);
                    if (leader) {
#pragma unroll
                        for (int k = 0; k < kKvTile / 16; ++k)
                            umma_f16_ts(tmem_base + 256 + g * 64, tmem_base + 384 + g * 64 + k * 8,
                                        umma_smem_desc_sw128(bv + k * 2048, 1024, 1024), idesc_o,
                                        (j | k) != 0 ? 1u : 0u);
                        umma_commit(o_full + g);
                    }
                    __syncwarp();
                    if (leader) ATT_T(0, j, 5 + g * 2);
                }
                if (leader) umma_commit(kv_empty + stage);  // K_j / V_j are free once everything issued so far retires
                __syncwarp();
                stage = next_stage;
                phase = next_phase;
            }
        }
    } else {
        // ============ softmax + output warps: two threads per query row, 64 keys / 32 output dims each ============
        const int sw = warp - 2;
        const int g = sw >> 3;
        const int half = (sw >> 2) & 1;
        const int quarter = warp & 3;  // TMEM lane quarter this warp may touch
        const int row = quarter * 32 + lane;  // row inside the query tile == TMEM lane
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const uint32_t tmem_s = tmem_base + g * 128 + half * 64 + lane_off;
        const uint32_t tmem_o = tmem_base + 256 + g * 64 + half * 32 + lane_off;
        const uint32_t tmem_p = tmem_base + 384 + g * 64 + half * 32 + lane_off;
        const int q_idx = qt * kQPerCta + g * kQTile + row;
        const int pair_bar = 1 + g * 4 + quarter;  // named barrier shared by the row's two warps
        float* x_mine = xch + (g * 2 + half) * 128 + row;
        float* x_other = xch + (g * 2 + (half ^ 1)) * 128 + row;
        if (g == 0 || g1_active) {
            const float c = p.scale * 1.4426950408889634f;  // softmax(scale * s) via exp2
            float m_run = -INFINITY, l_run = 0.f;
            float sv[64];
            uint32_t r[32];
            for (int j = 0; j < kv_tiles; ++j) {
                const int valid = min(kKvTile, kv_len - j * kKvTile) - half * 64;
                mbar_wait(s_full + g, j & 1);
                tc_fence_after();
                if (sw == 0 && lane == 0) ATT_T(1, j, 0);
                B200_TMEM_LD32F(tmem_s, sv);
                B200_TMEM_LD32F(tmem_s + 32, sv + 32);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(s_free + g);  // S_g(j+1) may overwrite the TMEM buffer now
                if (sw == 0 && lane == 0) ATT_T(1, j, 1);
                if (valid < 64) {
#pragma unroll
                    for (int i = 0; i < 64; ++i)
                        if (i >= valid) sv[i] = -INFINITY;
                }
                float m0 = sv[0], m1 = sv[1], m2 = sv[2], m3 = sv[3];
#pragma unroll
                for (int i = 4; i < 64; i += 4) {
                    m0 = fmaxf(m0, sv[i]); m1 = fmaxf(m1, sv[i + 1]);
                    m2 = fmaxf(m2, sv[i + 2]); m3 = fmaxf(m3, sv[i + 3]);
                }
                const int par = (j & 1) * 512;
                x_mine[par] = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                const float m_tile = fmaxf(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)), x_other[par]);
                // lazy rescale: move the reference maximum only when it is off by more than 2^8
                // (both threads of a row see the same m_tile / m_run, hence take the same decision)
                if (sw == 0 && lane == 0) ATT_T(1, j, 2);
                const bool need = (m_tile - m_run) * c > 8.0f;  // always true for j == 0 (m_run = -inf)
                bool waited = false;
                if (j > 0 && __any_sync(0xffffffffu, need)) {
                    mbar_wait(o_full + g, (j - 1) & 1);  // every PV issued so far has landed in O_g
                    tc_fence_after();
                    waited = true;
                    const float m_new = need ? m_tile : m_run;
                    const float alpha = ex2_approx((m_run - m_new) * c);
                    l_run *= alpha;
                    m_run = m_new;
#pragma unroll 1
                    for (int ch = 0; ch < 4; ++ch) {  // this thread's 32 of the 64 O columns, 8 at a time
                        uint32_t t8[8];
                        tmem_ld8(tmem_o + ch * 8, t8);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 8; ++i) t8[i] = __float_as_uint(__uint_as_float(t8[i]) * alpha);
                        tmem_st8(tmem_o + ch * 8, t8);
                    }
                    tmem_st_wait();
                } else if (j == 0) {
                    m_run = m_tile;
                }
                if (j > 0 && !waited) {
                    mbar_wait(o_full + g, (j - 1) & 1);  // PV_g(j-1) has finished reading the P buffer
                    tc_fence_after();
                }
                if (sw == 0 && lane == 0) ATT_T(1, j, 3);
                const float mc = m_run * c;
                float l0 = 0.f, l1 = 0.f;
#pragma unroll
                for (int i = 0; i < 32; ++i) {  // column i of the P row = keys (2i, 2i+1) as bf16
                    const float e0 = ex2_approx(fmaf(sv[2 * i], c, -mc));
                    const float e1 = ex2_approx(fmaf(sv[2 * i + 1], c, -mc));
                    l0 += e0;
                    l1 += e1;
                    r[i] = pack_bf16(e0, e1);
                }
                tmem_st32(tmem_p, r);
                tmem_st_wait();
                l_run += l0 + l1;
                if (sw == 0 && lane == 0) ATT_T(1, j, 4);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready + g);
                if (sw == 0 && lane == 0) ATT_T(1, j, 5);
            }
            // total row sum = the two partial sums (taken against the same reference maximum)
            x_mine[1024] = l_run;
            asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
            l_run += x_other[1024];
            if (kv_tiles > 0) {
                mbar_wait(o_full + g, (kv_tiles - 1) & 1);
                tc_fence_after();
            }
            const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;  // no keys: zeros
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)(sd.q_row0 + q_idx) * p.ld_out +
                                 p.out_col0 + head * kHeadDim + half * 32;
            if (kv_tiles > 0) {
                tmem_ld32(tmem_o, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) r[i] = 0u;
            }
            if (q_idx < q_len) {
#pragma unroll
                for (int i = 0; i < 32; i += 8)
                    *reinterpret_cast<uint4*>(dst + i) = make_uint4(
                        pack_bf16(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv),
                        pack_bf16(__uint_as_float(r[i + 2]) * inv, __uint_as_float(r[i + 3]) * inv),
                        pack_bf16(__uint_as_float(r[i + 4]) * inv, __uint_as_float(r[i + 5]) * inv),
                        pack_bf16(__uint_as_float(r[i + 6]) * inv, __uint_as_float(r[i + 7]) * inv));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<kAttTmemCols>(tmem_base);
}

#if ATT_X & 512
extern "C" int b200_debug_attention_trace(long long* dst) {
    return (int)cudaMemcpyFromSymbol(dst, g_att_trace, sizeof(g_att_trace));
}
#endif

int launch_attention(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p,
                     int num_seqs, int max_q_len, int heads, cudaStream_t stream) {
    static bool configured = false;
    if (!configured) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attention_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kAttSmemBytes));
        configured = true;
    }
    if (num_seqs == 0 || max_q_len == 0) return 0;
    dim3 grid((max_q_len + kQPerCta - 1) / kQPerCta, heads, num_seqs);
    attention_tcgen05_kernel<<<grid, kAttThreads, kAttSmemBytes, stream>>>(tq, tk, tv, p);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
