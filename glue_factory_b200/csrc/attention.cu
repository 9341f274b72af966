// Dense path: flash-style multi-head attention on tcgen05 / TMEM for head_dim 64.
//
// Persistent CTAs (one per SM, all 512 TMEM columns) walk a list of work items; an item is a pair of
// 128-query tiles of one (sequence, head).
//   warp 0       TMEM allocation + TMA producer: Q tiles (double-buffered per query tile) and 64-key K/V tiles
//                through one ring of smem slots that runs ACROSS items, so the next item's Q and first K/V
//                tiles are on chip before the current item drains
//   warps 1-16   four independent streams of four warps, one query row per thread; the first warp of a
//                stream also issues the stream's tcgen05.mma (one elected lane, uniform-register operands)
// A stream is (query tile g, key-tile parity x): it owns the even (x = 0) or odd (x = 1) 64-key tiles of
// its query tile, with its own S/P and O accumulators in TMEM and its own running (max, sum):
//   S = Q_g K_t^T      tcgen05.mma M=128 N=64 K=64, fp32 in TMEM (64 columns)
//   the stream's 128 threads pull their S row into registers (tcgen05.ld), exponentiate against the
//   running maximum and write P back over S as bf16 pairs (tcgen05.st, 32 columns)
//   O += P V_t         tcgen05.mma with the A operand read from TMEM, V consumed token-major through an
//                      MN-major descriptor, accumulating in TMEM (64 columns)
//   the next S of the stream is issued right behind PV (the tensor pipe is in order, so it may overwrite P);
//   behind the item's last PV that is already the first S of the NEXT item.
// Nothing couples the streams inside the key loop (beyond the K/V ring): a stream waits on the tensor core
// for about half of its tile time, and the other three streams keep the MUFU busy meanwhile (every
// scheduler holds one warp of each stream).
// At the end of an item the two streams of a query tile are merged like split-KV partial results,
//   O = (wA O_A + wB O_B) / (wA l_A + wB l_B),  w = 2^(c (m - max(mA, mB))),
// each stream reading BOTH accumulators from TMEM for its half of the 64 output dimensions.
// Online softmax with a lazily updated reference maximum: the running maximum m only moves (and O / the
// row sum are rescaled, tcgen05.ld -> multiply -> tcgen05.st) when a row's new maximum exceeds it by more
// than 2^8 in the exp2 domain; otherwise P is simply taken against the stale m (values <= 256, harmless
// for bf16 / fp32) - the final O / l is mathematically unchanged.
// Q/K/V tiles arrive by TMA from token-major [tokens, ld] bf16 matrices; a head is a 64-column slice.
// The S matrix is never materialised in HBM.  Self-attention and both directions of the bidirectional
// cross-attention use this kernel (the cross block's similarity is recomputed per direction).
//
// Two stream layouts share the code (template parameter WIDE):
//   WIDE = false  item = 256 queries; stream = (query tile g of 2, key-tile parity x) as described above.  Twice as many,
//                 half as long items: used when the wide layout would leave SMs without work (single pairs).
//   WIDE = true   item = 512 queries; stream = query tile g of 4, every stream walks ALL key tiles of the item.  No merge,
//                 no exchange between streams at all: a stream finishes its rows alone (O / l straight from its own
//                 accumulator), K/V tiles are fetched once per 512 queries, and the per-item hand-over (the pair
//                 barriers of the merge cost the narrow layout ~14 % of an item at 2048 keys) shrinks to the wait for
//                 the stream's last PV.
//
// Replaces F.scaled_dot_product_attention / the einsum+softmax path of the reference
// (reference: gluefactory/models/matchers/lightglue.py:105-122,144-157,186-215).
#include <cstdlib>

#include "attention_impl.cuh"

namespace b200 {

template <bool WIDE>
__global__ void __launch_bounds__(kAttThreads, 1)
attention_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                         const __grid_constant__ CUtensorMap tmap_v, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    using Cfg = AttCfg<WIDE>;
    constexpr int kQTiles = Cfg::kQTiles, kQPerCta = Cfg::kQPerCta, kKvStages = Cfg::kKvStages, kKvStride = Cfg::kKvStride;
    uint8_t* sQ = smem;  // [g][buffer]
    uint8_t* sK = sQ + 2 * kQTiles * kQBytes;
    uint8_t* sV = sK + kKvStages * kKvBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kKvStages * kKvBytes);
    uint64_t* q_full = bars;                   // [g * 2 + buffer]
    uint64_t* q_empty = q_full + 2 * kQTiles;  // [g * 2 + buffer], one arrival per stream of the query tile
    uint64_t* kv_full = q_empty + 2 * kQTiles; // [kKvStages]
    uint64_t* kv_empty = kv_full + kKvStages;  // [kKvStages], one arrival per stream that reads the tile
    uint64_t* s_full = kv_empty + kKvStages;   // [4]
    uint64_t* p_ready = s_full + 4;            // [4]
    uint64_t* o_done = p_ready + 4;            // [4]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 4);
    float* xch = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);

    // warp index / TMEM base / lengths go through a shuffle so that the compiler treats them, and the UMMA
    // operands derived from them, as warp-uniform: tcgen05.mma then issues straight from uniform registers
    // (a per-lane operand costs an ELECT / R2UR loop of ~75 clk around every MMA)
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_q);
        prefetch_tensormap(&tmap_k);
        prefetch_tensormap(&tmap_v);
        for (int i = 0; i < 2 * kQTiles; ++i) {
            mbar_init(q_full + i, 1);
            mbar_init(q_empty + i, Cfg::kQUsers);
        }
        for (int i = 0; i < 4; ++i) {
            mbar_init(s_full + i, 1);
            mbar_init(p_ready + i, 4);  // one arrival per warp of the stream
            mbar_init(o_done + i, 1);
        }
        for (int s = 0; s < kKvStages; ++s) {
            mbar_init(kv_full + s, 1);
            mbar_init(kv_empty + s, Cfg::kKvUsers);
        }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<kAttTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_wait();     // predecessor grid complete, its writes visible (the prologue above overlapped its tail)
    pdl_trigger();  // successor may become resident

    if (warp == 0) {
        // ===================== TMA producer =====================
        const bool leader = elect_one();
        int kv_idx = 0;          // K/V tiles loaded so far (ring position)
        int qcnt[kQTiles];       // Q tiles loaded so far per query tile (buffer = count & 1)
#pragma unroll
        for (int g = 0; g < kQTiles; ++g) qcnt[g] = 0;
        for (int it = blockIdx.x; it < p.num_items; it += gridDim.x) {
            const AttItem a = att_item<WIDE>(p, it);
            if (!a.active || a.kv_tiles == 0) continue;
            const int col = a.head * kHeadDim;
            const int qrow = a.q_row0 + a.q_first;
#pragma unroll
            for (int g = 0; g < kQTiles; ++g) {
                if (!a.tile_active(g)) continue;
                const int slot = g * 2 + (qcnt[g] & 1);
                mbar_wait(q_empty + slot, (uint32_t)(((qcnt[g] >> 1) & 1) ^ 1));
                if (leader) {
                    mbar_arrive_expect_tx(q_full + slot, kQBytes);
                    tma_load_2d(sQ + slot * kQBytes, &tmap_q, q_full + slot, p.q_col0 + col, qrow + g * kQTile);
                }
                __syncwarp();
                ++qcnt[g];
            }
            for (int t = 0; t < a.kv_tiles; ++t, ++kv_idx) {
                const int stage = kv_idx % kKvStages;
                const int row = a.kv_row0 + t * kKvTile;
                mbar_wait(kv_empty + stage, (uint32_t)(((kv_idx / kKvStages) & 1) ^ 1));
                if (leader) {
                    mbar_arrive_expect_tx(kv_full + stage, 2 * kKvBytes);
                    tma_load_2d(sK + stage * kKvBytes, &tmap_k, kv_full + stage, p.k_col0 + col, row);
                    tma_load_2d(sV + stage * kKvBytes, &tmap_v, kv_full + stage, p.v_col0 + col, row);
                }
                __syncwarp();
            }
        }
    } else {
        // ============ stream (g, x): softmax + MMA issue + output, thread <-> query row ============
        const int s = (warp - 1) >> 2;
#if ATT_ISSUER_SPREAD
        const bool issuer = (warp & 3) == s;        // the issuers of the four streams sit on four different schedulers
#else
        const bool issuer = ((warp - 1) & 3) == 0;  // first warp of the stream issues its MMAs
#endif
        const int g = WIDE ? s : s >> 1, x = WIDE ? 0 : s & 1;
        const int quarter = warp & 3;         // TMEM lane quarter this warp may touch
        const int row = quarter * 32 + lane;  // row inside the query tile == TMEM lane
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const uint32_t tmem_s = tmem_base + s * 64 + lane_off;
        const uint32_t tmem_o = tmem_base + 256 + s * 64 + lane_off;
        const uint32_t tmem_o_other = tmem_base + 256 + (s ^ 1) * 64 + lane_off;  // narrow layout only
        const int pair_bar = 1 + g * 4 + quarter;  // named barrier shared by the row's two streams (narrow layout)
        constexpr uint32_t idesc_s = umma_idesc(128, 64, 1u, 0u, 0u);  // Q, K both K-major
        constexpr uint32_t idesc_o = umma_idesc(128, 64, 1u, 0u, 1u);  // P K-major (TMEM), V MN-major
        const bool lead_lane = elect_one();
        const float c = p.scale * 1.4426950408889634f;  // softmax(scale * s) via exp2
        // Operand descriptors: every tile base is 1024-aligned and all of shared memory lies below 256 KB, so the
        // 14-bit address field of a descriptor advances linearly: descriptor(base + d) = descriptor(base) + (d >> 4).
        // Building them from scratch (add, shift, mask, or - per MMA) put ~120 dependent uniform-datapath
        // instructions between "P is ready" and the stream's next S, i.e. on the one path the stream waits for.
        const uint64_t kdesc0 = umma_smem_desc_sw128(smem_u32(sK), 1024, 16);    // K tile, K-major
        const uint64_t vdesc0 = umma_smem_desc_sw128(smem_u32(sV), 1024, 1024);  // V tile, MN-major (token-major)
        const uint64_t qdesc0 = umma_smem_desc_sw128(smem_u32(sQ), 1024, 16);
        constexpr uint32_t kKvDescStep = kKvBytes >> 4, kQDescStep = kQBytes >> 4;
        // high word of the three descriptors (SBO 1024, version, swizzle; LBO lives in bits [16,30) of the low word and
        // is ignored for swizzled K-major tiles / equals 1024 >> 4 = 64 for the MN-major V tile, carried by the low word)
        const uint64_t desc_hi = kdesc0 & 0xffffffff00000000ull;
        // S_s = Q_g K^T for the K tile in ring slot `stage` with Q buffer `qslot`; whole warp calls, one lane issues
        auto issue_s_at = [&](int stage, uint32_t phase, int qslot) {
            mbar_wait(kv_full + stage, phase);
            tc_fence_after();
            if (lead_lane) {
                const uint64_t aq = qdesc0 + (uint32_t)qslot * kQDescStep, bk = kdesc0 + (uint32_t)stage * kKvDescStep;
#pragma unroll
                for (int k = 0; k < kHeadDim / 16; ++k)
                    umma_f16(tmem_base + s * 64, aq + k * 2, bk + k * 2, idesc_s, k != 0 ? 1u : 0u);
                umma_commit(s_full + s);
            }
            __syncwarp();
        };
        auto issue_s = [&](int idx, int qslot) { issue_s_at(idx % kKvStages, (uint32_t)((idx / kKvStages) & 1), qslot); };
        int kv_base = 0;   // ring position of the current item's first K/V tile
        int qcnt = 0;      // Q tiles of query tile g used so far (buffer = count & 1)
        int n_tot = 0;     // key tiles this stream has processed so far (phase of s_full / p_ready)
        int od_a = 0, od_b = 0;  // items in which stream (g, 0) / (g, 1) accumulated (phase of o_done)
        bool pre = false;  // the first S of the current item was issued at the end of the previous one
        float sv[kKvTile];
        uint32_t r[32];
        // the memory words of the next item are requested while this item's first tile is in flight and decoded
        // only where they are needed (pre-issue at the item's end, top of the next iteration)
        AttRaw raw = {0, 0, 0, 0}, raw_next = {0, 0, 0, 0};
        if (blockIdx.x < p.num_items) raw = att_fetch(p, blockIdx.x);
        bool next_fetched = false;
        for (int it = blockIdx.x; it < p.num_items; it += gridDim.x) {
            if (it != (int)blockIdx.x) raw = next_fetched ? raw_next : att_fetch(p, it);
            next_fetched = false;
            const AttItem a = att_decode<WIDE>(p, it, raw);
            if (!a.active) continue;
            if (!a.tile_active(g)) {
                // query tile without rows: only hand the K/V tiles of this stream back to the producer
                if (issuer) {
                    for (int t = x; t < a.kv_tiles; t += kKvStride) {
                        const int idx = kv_base + t;
                        mbar_wait(kv_full + idx % kKvStages, (uint32_t)((idx / kKvStages) & 1));
                        if (lane == 0) mbar_arrive(kv_empty + idx % kKvStages);
                        __syncwarp();
                    }
                }
                kv_base += a.kv_tiles;
                continue;
            }
            const bool has_q = a.kv_tiles > 0;
            const int qslot = g * 2 + (qcnt & 1);
            const bool have = a.kv_tiles > x;              // this stream owns at least one key tile
            const bool have_other = a.kv_tiles > (x ^ 1);  // so does the other stream of the query tile
            if (issuer && have && !pre) {
                mbar_wait(q_full + qslot, (uint32_t)((qcnt >> 1) & 1));
                issue_s(kv_base + x, qslot);
            }
            pre = false;
            // ring slot of the stream's current tile, advanced incrementally (issuer only; no division in the key loop)
            int stage = (kv_base + x) % kKvStages;
            uint32_t phase = (uint32_t)(((kv_base + x) / kKvStages) & 1);
            float m_run = -INFINITY, l_run = 0.f;
            int n = 0;  // tiles of this stream in this item
            for (int t = x; t < a.kv_tiles; t += kKvStride, ++n, ++n_tot) {
                const int valid = min(kKvTile, a.kv_len - t * kKvTile);
                mbar_wait(s_full + s, (uint32_t)(n_tot & 1));  // S_s(t) landed; every earlier MMA of the stream retired
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
                const bool need = (m_tile - m_run) * c > (float)ATT_LAZY_LOG2;  // always true for n == 0 (m_run = -inf)
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
                const f32x2 c2 = bc2(c), nmc2 = bc2(-m_run * c);
                f32x2 l2 = bc2(0.f);
#if ATT_P_STORE8
                uint32_t r8[8];
#endif
#pragma unroll
                for (int i = 0; i < 32; ++i) {  // column i of the P row = keys (2i, 2i+1) as bf16
                    constexpr int kPolyStride = ATT_POLY_PAIRS > 0 ? 32 / ATT_POLY_PAIRS : 64;
                    const bool poly = ATT_POLY_PAIRS > 0 && (i % kPolyStride) == kPolyStride - 1;
                    const f32x2 a = fma2(pk2(sv[2 * i], sv[2 * i + 1]), c2, nmc2);
                    float e0, e1;
                    if (poly) {
                        ex2_poly2(a, e0, e1);
                    } else {
                        float a0, a1;
                        upk2(a, a0, a1);
                        e0 = ex2_approx(a0);
                        e1 = ex2_approx(a1);
                    }
                    l2 = add2(l2, pk2(e0, e1));
#if ATT_P_STORE8
                    // P goes back to TMEM eight columns at a time: a 32-register store block next to the 64 S registers
                    // does not fit the 96 registers of a 544-thread CTA without spills
                    r8[i & 7] = pack_bf16(e0, e1);
                    if ((i & 7) == 7) tmem_st8(tmem_s + (i & ~7), r8);
#else
                    r[i] = pack_bf16(e0, e1);
#endif
                }
                float l0, l1;
                upk2(l2, l0, l1);
#if !ATT_P_STORE8
                tmem_st32(tmem_s, r);  // P over the first half of S (this thread's own row)
#endif
                tmem_st_wait();
                l_run += l0 + l1;
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready + s);
                if (n == 0 && it + (int)gridDim.x < p.num_items) {
                    raw_next = att_fetch(p, it + gridDim.x);  // latency hides behind the wait for the next S
                    next_fetched = true;
                }
                if (issuer) {
                    // Everything PV(t) and S(t + stride) need is prepared BEFORE the wait for P, the next K tile's
                    // arrival included (it landed long ago: the ring runs several tiles ahead), so that only the
                    // eight MMAs and their commits sit between "P is ready" and the stream's next S.
                    const bool last = t + kKvStride >= a.kv_tiles;
                    int nstage = stage + kKvStride;
                    uint32_t nphase = phase;
                    if (nstage >= kKvStages) {
                        nstage -= kKvStages;
                        nphase ^= 1u;
                    }
                    const uint64_t bv = vdesc0 + (uint32_t)stage * kKvDescStep;
                    const uint64_t bk = kdesc0 + (uint32_t)nstage * kKvDescStep;
                    const uint64_t aq = qdesc0 + (uint32_t)qslot * kQDescStep;
                    if (!last) mbar_wait(kv_full + nstage, nphase);
                    uint32_t bv_lo = (uint32_t)bv, bk_lo = (uint32_t)bk, aq_lo = (uint32_t)aq;
                    uint32_t bar_kv = smem_u32(kv_empty + stage);
                    asm volatile("" : "+r"(bv_lo), "+r"(bk_lo), "+r"(aq_lo), "+r"(bar_kv));  // computed before the wait
                    mbar_wait(p_ready + s, (uint32_t)(n_tot & 1));  // the four warps' P rows are in TMEM
                    tc_fence_after();
                    if (lead_lane) {
#pragma unroll
                        for (int k = 0; k < kKvTile / 16; ++k)
                            umma_f16_ts(tmem_base + 256 + s * 64, tmem_base + s * 64 + k * 8,
                                        (desc_hi | bv_lo) + (uint64_t)(k * 128), idesc_o, (n | k) != 0 ? 1u : 0u);
                        umma_commit_addr(bar_kv);  // K_t / V_t free once every stream's MMAs on them retire
                        if (last) {
                            umma_commit(o_done + s);
                            umma_commit(q_empty + qslot);  // this stream is done with the item's Q tile
                        } else {
#pragma unroll
                            for (int k = 0; k < kHeadDim / 16; ++k)  // in order behind PV: may overwrite P
                                umma_f16(tmem_base + s * 64, (desc_hi | aq_lo) + (uint64_t)(k * 2),
                                         (desc_hi | bk_lo) + (uint64_t)(k * 2), idesc_s, k != 0 ? 1u : 0u);
                            umma_commit(s_full + s);
                        }
                    }
                    __syncwarp();
                    stage = nstage;
                    phase = nphase;
                    if (last && it + (int)gridDim.x < p.num_items) {
                        // the next item's first S goes in right behind this item's last PV, so its start-up
                        // (Q / K arrival, first MMA) runs under the output pass below
                        const AttItem nx =
                            att_decode<WIDE>(p, it + gridDim.x, next_fetched ? raw_next : att_fetch(p, it + gridDim.x));
                        if (nx.active && nx.tile_active(g) && nx.kv_tiles > x) {
                            const int nslot = g * 2 + ((qcnt + 1) & 1);
                            mbar_wait(q_full + nslot, (uint32_t)(((qcnt + 1) >> 1) & 1));
                            issue_s(kv_base + a.kv_tiles + x, nslot);
                            pre = true;
                        }
                    }
                }
            }
            if (issuer && has_q && !have) {  // no key tile for this stream: it still releases the Q tile
                if (lane == 0) mbar_arrive(q_empty + qslot);
                __syncwarp();
            }
            if constexpr (WIDE) {
                // ---- the stream owns its rows: O / l straight from its accumulator, 32 output dimensions at a time ----
                const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;  // no keys: zeros
                if (have) {
                    mbar_wait(o_done + s, (uint32_t)(od_a & 1));  // the stream's last PV has landed
                    tc_fence_after();
                }
                const int q_idx = a.q_first + g * kQTile + row;
                __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                     (long long)(a.q_row0 + q_idx) * p.ld_out + p.out_col0 + a.head * kHeadDim;
                if (have) {
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        uint32_t rr[32];
                        tmem_ld32(tmem_o + half * 32, rr);
                        tmem_ld_wait();
                        if (q_idx < a.q_len) {
#pragma unroll
                            for (int i = 0; i < 32; i += 8)
                                *reinterpret_cast<uint4*>(dst + half * 32 + i) =
                                    make_uint4(pack_bf16(__uint_as_float(rr[i]) * inv, __uint_as_float(rr[i + 1]) * inv),
                                               pack_bf16(__uint_as_float(rr[i + 2]) * inv, __uint_as_float(rr[i + 3]) * inv),
                                               pack_bf16(__uint_as_float(rr[i + 4]) * inv, __uint_as_float(rr[i + 5]) * inv),
                                               pack_bf16(__uint_as_float(rr[i + 6]) * inv, __uint_as_float(rr[i + 7]) * inv));
                        }
                    }
                } else if (q_idx < a.q_len) {
#pragma unroll
                    for (int i = 0; i < kHeadDim; i += 8) *reinterpret_cast<uint4*>(dst + i) = make_uint4(0u, 0u, 0u, 0u);
                }
                // the accumulator reads are ordered before the stream's next PV by the fence in front of the next
                // p_ready arrival (every warp of the stream passes it before the issuer goes on)
                tc_fence_before();
            } else {
                // ---- merge the row's two streams: O = (wA O_A + wB O_B) / (wA l_A + wB l_B); this stream writes
                //      output dimensions [32 x, 32 x + 32) ----
                float* xm = xch + ((g * 2 + x) * kQTile + row);
                float* xo = xch + ((g * 2 + (x ^ 1)) * kQTile + row);
                xm[0] = m_run;
                xm[4 * kQTile] = l_run;
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                const float m_o = xo[0], l_o = xo[4 * kQTile];
                const float m_all = fmaxf(m_run, m_o);
                const float w_me = have ? ex2_approx((m_run - m_all) * c) : 0.f;
                const float w_o = have_other ? ex2_approx((m_o - m_all) * c) : 0.f;
                const float l_all = l_run * w_me + l_o * w_o;
                const float inv = l_all > 0.f ? 1.0f / l_all : 0.f;  // no keys: zeros
                const float f_me = w_me * inv, f_o = w_o * inv;
                if (have) {
                    mbar_wait(o_done + s, (uint32_t)((x == 0 ? od_a : od_b) & 1));  // the stream's last PV has landed
                    tc_fence_after();
                }
                if (have_other) {
                    mbar_wait(o_done + (s ^ 1), (uint32_t)((x == 0 ? od_b : od_a) & 1));
                    tc_fence_after();
                }
                float o[32];
                if (have) {
                    tmem_ld32(tmem_o + x * 32, r);
                    tmem_ld_wait();
    #pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]) * f_me;
                } else {
    #pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = 0.f;
                }
                if (have_other) {
                    tmem_ld32(tmem_o_other + x * 32, r);
                    tmem_ld_wait();
    #pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = fmaf(__uint_as_float(r[i]), f_o, o[i]);
                }
                const int q_idx = a.q_first + g * kQTile + row;
                if (q_idx < a.q_len) {
                    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                         (long long)(a.q_row0 + q_idx) * p.ld_out + p.out_col0 + a.head * kHeadDim + x * 32;
    #pragma unroll
                    for (int i = 0; i < 32; i += 8)
                        *reinterpret_cast<uint4*>(dst + i) =
                            make_uint4(pack_bf16(o[i], o[i + 1]), pack_bf16(o[i + 2], o[i + 3]),
                                       pack_bf16(o[i + 4], o[i + 5]), pack_bf16(o[i + 6], o[i + 7]));
                }
                tc_fence_before();
                // both streams have read both accumulators (and the exchange slots): the next item may overwrite them
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
            }
            kv_base += a.kv_tiles;
            if (has_q) ++qcnt;
            if (a.kv_tiles > 0) ++od_a;
            if (!WIDE && a.kv_tiles > 1) ++od_b;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<kAttTmemCols>(tmem_base);
}

// Layout choice: the wide layout wins ~10 % per item at full occupancy but makes half as many items; take it unless
// the narrow layout fills the SMs noticeably better (whole waves of equal items: time ~ waves x item length).
// B200GLUE_ATT_LAYOUT=wide|narrow forces one (experiments, tests).
static int att_layout_override() {  // read per call (a getenv is nanoseconds): tests switch it inside one process
    const char* e = getenv("B200GLUE_ATT_LAYOUT");
    if (!e) return -1;
    return e[0] == 'w' ? 1 : e[0] == 'n' ? 0 : -1;
}

template <bool WIDE>
static int launch_attention_as(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, AttnParams p,
                               long long items, int num_sms, cudaStream_t stream) {
    static PerDeviceSize configured;
    if (!configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attention_tcgen05_kernel<WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             AttCfg<WIDE>::kSmemBytes));
        configured.cur() = 1;
    }
    p.num_items = (int)items;
    const int grid = (int)(items < num_sms ? items : num_sms);
    B200_CHECK_CUDA(launch_pdl(attention_tcgen05_kernel<WIDE>, dim3(grid), dim3(kAttThreads), AttCfg<WIDE>::kSmemBytes, stream,
                               tq, tk, tv, p));
    return 0;
}

int launch_attention(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, AttnParams p,
                     int num_seqs, int max_q_len, int heads, int num_sms, cudaStream_t stream) {
    if (num_seqs == 0 || max_q_len == 0) return 0;
    p.heads = heads;
    const int qt_n = (max_q_len + AttCfg<false>::kQPerCta - 1) / AttCfg<false>::kQPerCta;
    const int qt_w = (max_q_len + AttCfg<true>::kQPerCta - 1) / AttCfg<true>::kQPerCta;
    const long long items_n = (long long)qt_n * heads * num_seqs, items_w = (long long)qt_w * heads * num_seqs;
    B200_REQUIRE(items_n < (1ll << 31), "attention: too many work items");
    // relative time: waves x (item length in 128-query tiles), the wide item taken at kWideGain of two narrow ones
    const double kWideGain = 0.965;  // measured: 64 x 2048 tokens 0.311 vs 0.323 ms, per-wave at 16 x 4096 likewise
    const double t_n = (double)((items_n + num_sms - 1) / num_sms) * 1.0;
    const double t_w = (double)((items_w + num_sms - 1) / num_sms) * 2.0 * kWideGain;
    bool wide = t_w <= t_n;
    // Device-side lengths = the adaptive path, whose stop / prune decisions compare against thresholds: there a batch
    // must reproduce its pairs run one by one bit for bit (tests/test_gpu_adaptive.py), so the layout, which fixes the
    // summation order of the softmax, must not depend on how many pairs share the launch.
    if (p.len_override != nullptr) wide = false;
    if (att_layout_override() >= 0) wide = att_layout_override() == 1;
    if (wide) {
        p.q_tiles = qt_w;
        return launch_attention_as<true>(tq, tk, tv, p, items_w, num_sms, stream);
    }
    p.q_tiles = qt_n;
    return launch_attention_as<false>(tq, tk, tv, p, items_n, num_sms, stream);
}

}  // namespace b200
