// Dense path: flash-style multi-head attention on tcgen05 / TMEM for head_dim 64.
//
// One CTA per (pair of 128-query tiles, head, sequence); 1 CTA per SM (all 512 TMEM columns).
//   warp 0       TMEM allocation + TMA producer: Q tiles once, 64-key K/V tiles through a ring of smem slots
//   warps 1-16   four independent streams of four warps, one query row per thread; the first warp of a
//                stream also issues the stream's tcgen05.mma (one elected lane, uniform-register operands)
// A stream is (query tile g, key-tile parity x): it owns the even (x = 0) or odd (x = 1) 64-key tiles of
// its query tile, with its own S/P and O accumulators in TMEM and its own running (max, sum):
//   S = Q_g K_t^T      tcgen05.mma M=128 N=64 K=64, fp32 in TMEM (64 columns)
//   the stream's 128 threads pull their S row into registers (tcgen05.ld), exponentiate against the
//   running maximum and write P back over S as bf16 pairs (tcgen05.st, 32 columns)
//   O += P V_t         tcgen05.mma with the A operand read from TMEM, V consumed token-major through an
//                      MN-major descriptor, accumulating in TMEM (64 columns)
//   the next S of the stream is issued right behind PV (the tensor pipe is in order, so it may overwrite P).
// Nothing couples the streams inside the key loop (beyond the K/V ring): a stream waits on the tensor core
// for about half of its tile time, and the other three streams keep the MUFU busy meanwhile (every
// scheduler holds one warp of each stream; the streams start a quarter period apart).
// At the end the two streams of a query tile are merged like split-KV partial results:
//   O = (wA O_A + wB O_B) / (wA l_A + wB l_B),  w = 2^(c (m - max(mA, mB))).
// Online softmax with a lazily updated reference maximum: the running maximum m only moves (and O / the
// row sum are rescaled, tcgen05.ld -> multiply -> tcgen05.st) when a row's new maximum exceeds it by more
// than 2^8 in the exp2 domain; otherwise P is simply taken against the stale m (values <= 256, harmless
// for bf16 / fp32) - the final O / l is mathematically unchanged.
// Q/K/V tiles arrive by TMA from token-major [tokens, ld] bf16 matrices; a head is a 64-column slice.
// The S matrix is never materialised in HBM.  Self-attention and both directions of the bidirectional
// cross-attention use this kernel (the cross block's similarity is recomputed per direction).
//
// Replaces F.scaled_dot_product_attention / the einsum+softmax path of the reference
// (reference: gluefactory/models/matchers/lightglue.py:105-122,144-157,186-215).
#include "common.cuh"
#include "attention.cuh"

namespace b200 {

constexpr int kAttThreads = 544;
constexpr int kQTile = 128, kKvTile = 64, kHeadDim = 64;
constexpr int kQPerCta = 2 * kQTile;
constexpr int kQBytes = kQTile * kHeadDim * 2;    // 16 KB
constexpr int kKvBytes = kKvTile * kHeadDim * 2;  // 8 KB: 64 keys x 64 bf16
constexpr int kKvStages = 6;
#ifndef ATT_OFFSET_NS
#define ATT_OFFSET_NS 280
#endif
constexpr int kStreamOffsetNs = ATT_OFFSET_NS;  // ~ a quarter of the time four streams need for 64 keys each
constexpr int kCombBytes = 2 * kQTile * kHeadDim * 4;  // merge buffer: [2 query tiles][128 rows][64 fp32]
constexpr int kXchFloats = 2 * 2 * 2 * kQTile;         // [m | l][query tile][stream parity][row]
// smem: Q[2] | K[stages] V[stages] | merge buffer | barriers | (m, l) exchange
constexpr int kAttSmemBytes = 1024 + 2 * kQBytes + 2 * kKvStages * kKvBytes + kCombBytes + 256 + kXchFloats * 4;
// stream s = 2 g + x:  S/P_s: [64 s, 64 s + 64)   O_s: [256 + 64 s, 256 + 64 s + 64)
constexpr uint32_t kAttTmemCols = 512;

__device__ __forceinline__ float ex2_approx(float x) {
#ifdef ATT_NOEX2
    return x * 0.001f;
#endif
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
    uint8_t* sK = sQ + 2 * kQBytes;
    uint8_t* sV = sK + kKvStages * kKvBytes;
    float* sComb = reinterpret_cast<float*>(sV + kKvStages * kKvBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sComb) + kCombBytes);
    uint64_t* q_full = bars;                  // [2]
    uint64_t* kv_full = q_full + 2;           // [kKvStages]
    uint64_t* kv_empty = kv_full + kKvStages; // [kKvStages]
    uint64_t* s_full = kv_empty + kKvStages;  // [4]
    uint64_t* p_ready = s_full + 4;           // [4]
    uint64_t* o_done = p_ready + 4;           // [4]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 4);
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
        for (int g = 0; g < 2; ++g) mbar_init(q_full + g, 1);
        for (int s = 0; s < 4; ++s) {
            mbar_init(s_full + s, 1);
            mbar_init(p_ready + s, 4);  // one arrival per warp of the stream
            mbar_init(o_done + s, 1);
        }
        for (int s = 0; s < kKvStages; ++s) {
            mbar_init(kv_full + s, 1);
            mbar_init(kv_empty + s, g1_active ? 2 : 1);  // one commit per query tile using the K/V tile
        }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<kAttTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    if (warp == 0) {
        // ===================== TMA producer =====================
        const bool leader = elect_one();
        const int col = head * kHeadDim;
        const int qrow = sd.q_row0 + qt * kQPerCta;
        if (leader) {
            mbar_arrive_expect_tx(q_full, kQBytes);
            tma_load_2d(sQ, &tmap_q, q_full, p.q_col0 + col, qrow);
            if (g1_active) {
                mbar_arrive_expect_tx(q_full + 1, kQBytes);
                tma_load_2d(sQ + kQBytes, &tmap_q, q_full + 1, p.q_col0 + col, qrow + kQTile);
            }
        }
        __syncwarp();
        int stage = 0;
        uint32_t phase = 0;
        for (int t = 0; t < kv_tiles; ++t) {
            const int row = sd.kv_row0 + t * kKvTile;
            mbar_wait(kv_empty + stage, phase ^ 1);
            if (leader) {
                mbar_arrive_expect_tx(kv_full + stage, 2 * kKvBytes);
                tma_load_2d(sK + stage * kKvBytes, &tmap_k, kv_full + stage, p.k_col0 + col, row);
                tma_load_2d(sV + stage * kKvBytes, &tmap_v, kv_full + stage, p.v_col0 + col, row);
            }
            __syncwarp();
            if (++stage == kKvStages) { stage = 0; phase ^= 1; }
        }
    } else {
        // ============ softmax + output warps: stream (g, x), thread <-> query row ============
        const int s = (warp - 1) >> 2;
        const bool issuer = ((warp - 1) & 3) == 0;  // first warp of the stream issues its MMAs
        const int g = s >> 1, x = s & 1;
        const int quarter = warp & 3;         // TMEM lane quarter this warp may touch
        const int row = quarter * 32 + lane;  // row inside the query tile == TMEM lane
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const uint32_t tmem_s = tmem_base + s * 64 + lane_off;
        const uint32_t tmem_o = tmem_base + 256 + s * 64 + lane_off;
        const int q_idx = qt * kQPerCta + g * kQTile + row;
        const int pair_bar = 1 + g * 4 + quarter;  // named barrier shared by the row's two streams
        if (g == 0 || g1_active) {
            const float c = p.scale * 1.4426950408889634f;  // softmax(scale * s) via exp2
            float m_run = -INFINITY, l_run = 0.f;
            float sv[kKvTile];
            uint32_t r[32];
            constexpr uint32_t idesc_s = umma_idesc(128, 64, 1u, 0u, 0u);  // Q, K both K-major
            constexpr uint32_t idesc_o = umma_idesc(128, 64, 1u, 0u, 1u);  // P K-major (TMEM), V MN-major
            const bool lead_lane = elect_one();
            auto issue_s = [&](int t) {  // S_s(t) = Q_g K_t^T; whole warp calls, one lane issues
                const int stage = t % kKvStages;
                mbar_wait(kv_full + stage, (t / kKvStages) & 1);
                tc_fence_after();
                const uint32_t aq = smem_u32(sQ + g * kQBytes), bk = smem_u32(sK + stage * kKvBytes);
                if (lead_lane) {
#pragma unroll
                    for (int k = 0; k < kHeadDim / 16; ++k)
                        umma_f16(tmem_base + s * 64, umma_smem_desc_sw128(aq + k * 32, 1024, 16),
                                 umma_smem_desc_sw128(bk + k * 32, 1024, 16), idesc_s, k != 0 ? 1u : 0u);
                    umma_commit(s_full + s);
                }
                __syncwarp();
            };
            if (issuer && x < kv_tiles) {
                __nanosleep(s * kStreamOffsetNs);  // streams a quarter period apart
                mbar_wait(q_full + g, 0);
                issue_s(x);
            }
            int n = 0;  // tiles of this stream so far
            for (int t = x; t < kv_tiles; t += 2, ++n) {
                const int valid = min(kKvTile, kv_len - t * kKvTile);
                mbar_wait(s_full + s, n & 1);  // S_s(t) landed; every earlier MMA of the stream (PV too) retired
                tc_fence_after();
                B200_TMEM_LD32F(tmem_s, sv);
                B200_TMEM_LD32F(tmem_s + 32, sv + 32);
                tmem_ld_wait();
                if (valid < kKvTile) {
#pragma unroll
                    for (int i = 0; i < kKvTile; ++i)
                        if (i >= valid) sv[i] = -INFINITY;
                }
                float m0 = sv[0], m1 = sv[1], m2 = sv[2], m3 = sv[3];
#pragma unroll
                for (int i = 4; i < kKvTile; i += 4) {
                    m0 = fmaxf(m0, sv[i]); m1 = fmaxf(m1, sv[i + 1]);
                    m2 = fmaxf(m2, sv[i + 2]); m3 = fmaxf(m3, sv[i + 3]);
                }
                const float m_tile = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                // lazy rescale: move the reference maximum only when it is off by more than 2^8
                const bool need = (m_tile - m_run) * c > 8.0f;  // always true for n == 0 (m_run = -inf)
                if (n == 0) {
                    m_run = m_tile;
                } else if (__any_sync(0xffffffffu, need)) {
                    const float m_new = need ? m_tile : m_run;
                    const float alpha = ex2_approx((m_run - m_new) * c);
                    l_run *= alpha;
                    m_run = m_new;
#pragma unroll 1
                    for (int ch = 0; ch < 8; ++ch) {  // 8 columns at a time: S sits in registers meanwhile
                        uint32_t t8[8];
                        tmem_ld8(tmem_o + ch * 8, t8);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 8; ++i) t8[i] = __float_as_uint(__uint_as_float(t8[i]) * alpha);
                        tmem_st8(tmem_o + ch * 8, t8);
                    }
                }
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
                tmem_st32(tmem_s, r);  // P over the first half of S (this thread's own row)
                tmem_st_wait();
                l_run += l0 + l1;
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready + s);
                if (issuer) {
                    mbar_wait(p_ready + s, n & 1);  // the four warps' P rows are in TMEM (O rescaled if it had to be)
                    tc_fence_after();
                    const int stage = t % kKvStages;
                    const uint32_t bv = smem_u32(sV + stage * kKvBytes);
                    if (lead_lane) {
#pragma unroll
                        for (int k = 0; k < kKvTile / 16; ++k)
                            umma_f16_ts(tmem_base + 256 + s * 64, tmem_base + s * 64 + k * 8,
                                        umma_smem_desc_sw128(bv + k * 2048, 1024, 1024), idesc_o, (n | k) != 0 ? 1u : 0u);
                        umma_commit(kv_empty + stage);  // K_t / V_t free once this query tile's MMAs on them retire
                        if (t + 2 >= kv_tiles) umma_commit(o_done + s);
                    }
                    __syncwarp();
                    if (t + 2 < kv_tiles) issue_s(t + 2);  // in order behind PV: may overwrite P
                }
            }
            // ---- merge the row's two streams: O = (wA O_A + wB O_B) / (wA l_A + wB l_B) ----
            const bool have = n > 0;
            float* xm = xch + ((g * 2 + x) * kQTile + row);
            float* xo = xch + ((g * 2 + (x ^ 1)) * kQTile + row);
            xm[0] = m_run;
            xm[4 * kQTile] = l_run;
            asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
            const float m_o = xo[0], l_o = xo[4 * kQTile];
            const float m_all = fmaxf(m_run, m_o);
            const float w_me = have ? ex2_approx((m_run - m_all) * c) : 0.f;
            const float w_o = l_o > 0.f ? ex2_approx((m_o - m_all) * c) : 0.f;
            const float l_all = l_run * w_me + l_o * w_o;
            const float f = l_all > 0.f ? w_me / l_all : 0.f;  // no keys: zeros
            if (have) {
                mbar_wait(o_done + s, 0);  // the stream's last PV has landed in O_s
                tc_fence_after();
            }
            float4* comb = reinterpret_cast<float4*>(sComb) + (g * kQTile + row) * 16;
            if (x == 1) {
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    if (have) {
                        tmem_ld32(tmem_o + ch * 32, r);
                        tmem_ld_wait();
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (have)
                            v = make_float4(__uint_as_float(r[4 * i]) * f, __uint_as_float(r[4 * i + 1]) * f,
                                            __uint_as_float(r[4 * i + 2]) * f, __uint_as_float(r[4 * i + 3]) * f);
                        comb[(ch * 8 + i) ^ (row & 15)] = v;
                    }
                }
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
            } else {
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                     (long long)(sd.q_row0 + q_idx) * p.ld_out + p.out_col0 + head * kHeadDim;
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    if (have) {
                        tmem_ld32(tmem_o + ch * 32, r);
                        tmem_ld_wait();
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = 0u;
                    }
                    if (q_idx < q_len) {
#pragma unroll
                        for (int i = 0; i < 8; i += 2) {
                            const float4 a = comb[(ch * 8 + i) ^ (row & 15)], b = comb[(ch * 8 + i + 1) ^ (row & 15)];
                            *reinterpret_cast<uint4*>(dst + ch * 32 + i * 4) = make_uint4(
                                pack_bf16(__uint_as_float(r[4 * i]) * f + a.x, __uint_as_float(r[4 * i + 1]) * f + a.y),
                                pack_bf16(__uint_as_float(r[4 * i + 2]) * f + a.z, __uint_as_float(r[4 * i + 3]) * f + a.w),
                                pack_bf16(__uint_as_float(r[4 * i + 4]) * f + b.x, __uint_as_float(r[4 * i + 5]) * f + b.y),
                                pack_bf16(__uint_as_float(r[4 * i + 6]) * f + b.z, __uint_as_float(r[4 * i + 7]) * f + b.w));
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<kAttTmemCols>(tmem_base);
}

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
