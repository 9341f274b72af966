// Dense path: flash attention for head_dim 64 with a DOUBLE-BUFFERED S accumulator (third layout of the attention
// kernel; the single-buffered four-stream layouts live in attention.cu).
//
// What bounds the four-stream kernel is the round trip of a stream through the tensor core: P overlays S in TMEM, so
// the stream's next S can only be issued behind P·V, and the stream's warps wait for it (25 % of all warp samples after
// every issue-path optimisation, profiles/r02_source_stalls.md).  Here a CTA runs TWO streams (one 128-query tile each,
// items of 256 queries, every stream walks all key tiles) and gives each stream THREE S/P buffers:
//   TMEM columns   stream s: S/P buffer b at [192 s + 64 b, +64)      O_s at [384 + 64 s, +64)      (all 512 used)
//   tile j of a stream (counted across work items) uses buffer j % 3.
// S_{j+3} is issued right behind P·V_j (it overwrites the buffer P_j sits in: the tensor pipe is in order), i.e. two
// tile times before the stream's warps need it.  (Two buffers = one tile time of lead were not enough: with eight
// warps a stream finishes a tile in ~500 clk, less than P·V + S + queueing behind the other stream, and 22 % of the
// warp samples were still waits for S.)  Up to three tiles of the NEXT work item are issued behind the last P·V of the
// current one, so the hand-over between items does not drain the pipeline either.
//
//   warp 0       TMEM allocation + TMA producer (Q double-buffered per query tile, K/V ring across items)
//   warps 17-18  one MMA issuer per stream (waits for the stream's P, issues P.V_t and S_{t+3}, hands K/V slots back)
//   warps 1-16   two streams of eight warps: two warps per TMEM lane quarter, a thread owns HALF a query row (32 of a
//                tile's 64 keys).  (With one thread per row, i.e. two warps
//                per scheduler, nothing hid a warp's own instruction latencies: 0.422 ms against 0.312 ms for the
//                four-stream layout.)  The two threads of a row must exponentiate against the SAME reference maximum
//                (they feed one P row of one P.V): they exchange their half-row maxima through shared memory once per
//                tile (64-thread named barrier, slots alternate with the tile parity), keep partial row sums, rescale
//                and finally normalise / store their own 32 of the 64 output dimensions.
//
// Consequences of S running ahead of P·V:
//   * `s_full` no longer implies that the stream's previous P·V retired, so the (rare) rescale of O in TMEM first waits
//     for `pv_done`, committed behind every P·V (three barriers in rotation: at most the newest P·V of a residue class
//     can be outstanding, which keeps the parity wait unambiguous);
//   * K of a ring slot is consumed three tiles before V; the slot is released by the commit behind P·V as before.
// The softmax arithmetic (lazy reference maximum, packed fp32x2 scaling and sums, polynomial exp2 on a share of the
// pairs) is that of attention.cu.
//
// (reference: gluefactory/models/matchers/lightglue.py:105-122,144-157,186-215)
#include "attention_impl.cuh"

namespace b200 {

constexpr int kDbThreads = 608;  // producer + 16 softmax warps + one issuer warp per stream
constexpr int kDbQTiles = 2, kDbQPerCta = kDbQTiles * kQTile, kDbKvStages = 8;
constexpr int kDbXchFloats = 2 * 2 * 2 * kQTile + 2 * 2 * 2 * kQTile;  // max [parity][stream][half][row] | sum [item parity][stream][half][row]
constexpr int kDbSmemBytes = 1024 + 2 * kDbQTiles * kQBytes + 2 * kDbKvStages * kKvBytes + 512 + kDbXchFloats * 4;
constexpr uint32_t kDbTmemCols = 512;
#ifndef ATT_DB_POLY_PAIRS
#define ATT_DB_POLY_PAIRS 2  // of the 16 key pairs a thread handles per tile
#endif

// one work item = (sequence, 256-query tile, head), decoded identically by every warp
struct DbItem {
    bool active;
    int q_row0, kv_row0, q_len, kv_len, kv_tiles, qt, head, q_first;
    __device__ __forceinline__ bool tile_active(int g) const { return q_first + g * kQTile < q_len; }
};

__device__ __forceinline__ DbItem db_decode(const AttnParams& p, int it, const AttRaw& r) {
    DbItem a;
    a.qt = it % p.q_tiles;
    a.head = (it / p.q_tiles) % p.heads;
    a.q_len = __shfl_sync(0xffffffffu, r.q_len, 0);
    a.kv_len = __shfl_sync(0xffffffffu, r.kv_len, 0);
    a.q_row0 = __shfl_sync(0xffffffffu, r.q_row0, 0);
    a.kv_row0 = __shfl_sync(0xffffffffu, r.kv_row0, 0);
    a.q_first = a.qt * kDbQPerCta;
    a.active = a.q_first < a.q_len;
    a.kv_tiles = (a.kv_len + kKvTile - 1) / kKvTile;
    return a;
}

__global__ void __launch_bounds__(kDbThreads, 1)
attention_db_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    uint8_t* sQ = smem;  // [g][buffer]
    uint8_t* sK = sQ + 2 * kDbQTiles * kQBytes;
    uint8_t* sV = sK + kDbKvStages * kKvBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kDbKvStages * kKvBytes);
    uint64_t* q_full = bars;                     // [g * 2 + buffer]
    uint64_t* q_empty = q_full + 4;              // [g * 2 + buffer]
    uint64_t* kv_full = q_empty + 4;             // [kDbKvStages]
    uint64_t* kv_empty = kv_full + kDbKvStages;  // [kDbKvStages], one arrival per stream
    uint64_t* s_full = kv_empty + kDbKvStages;   // [stream * 3 + S buffer]
    // [stream * 3 + S buffer], one arrival per warp.  Three in rotation like the S buffers: a warp may be up to two
    // tiles ahead of the slowest warp of its stream (S is there), and its arrival for tile t + 1 must not complete the
    // phase of tile t in the slow warp's place (with one barrier: P.V issued over an unfinished P row, then a hang).
    uint64_t* p_ready = s_full + 6;
    uint64_t* o_done = p_ready + 6;              // [stream]
    uint64_t* pv_done = o_done + 2;              // [stream * 3 + tile % 3]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 6);
    float* xch_m = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // [tile parity][stream][half][row]
    float* xch_l = xch_m + 2 * 2 * 2 * kQTile;                                        // [item parity][stream][half][row]

    // warp index / TMEM base through a shuffle: UMMA operands then stay in uniform registers
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_q);
        prefetch_tensormap(&tmap_k);
        prefetch_tensormap(&tmap_v);
        for (int i = 0; i < 4; ++i) {
            mbar_init(q_full + i, 1);
            mbar_init(q_empty + i, 1);
        }
        for (int i = 0; i < 6; ++i) {
            mbar_init(s_full + i, 1);
            mbar_init(pv_done + i, 1);
            mbar_init(p_ready + i, 8);  // one arrival per warp of the stream
        }
        for (int i = 0; i < 2; ++i) mbar_init(o_done + i, 1);
        for (int s = 0; s < kDbKvStages; ++s) {
            mbar_init(kv_full + s, 1);
            mbar_init(kv_empty + s, 2);
        }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<kDbTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    if (warp == 0) {
        // ===================== TMA producer =====================
        const bool leader = elect_one();
        int kv_idx = 0;        // K/V tiles loaded so far (ring position)
        int qcnt[2] = {0, 0};  // Q tiles loaded so far per query tile (buffer = count & 1)
        for (int it = blockIdx.x; it < p.num_items; it += gridDim.x) {
            const DbItem a = db_decode(p, it, att_fetch(p, it));
            if (!a.active || a.kv_tiles == 0) continue;
            const int col = a.head * kHeadDim;
            const int qrow = a.q_row0 + a.q_first;
#pragma unroll
            for (int g = 0; g < kDbQTiles; ++g) {
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
                const int stage = kv_idx % kDbKvStages;
                const int row = a.kv_row0 + t * kKvTile;
                mbar_wait(kv_empty + stage, (uint32_t)(((kv_idx / kDbKvStages) & 1) ^ 1));
                if (leader) {
                    mbar_arrive_expect_tx(kv_full + stage, 2 * kKvBytes);
                    tma_load_2d(sK + stage * kKvBytes, &tmap_k, kv_full + stage, p.k_col0 + col, row);
                    tma_load_2d(sV + stage * kKvBytes, &tmap_v, kv_full + stage, p.v_col0 + col, row);
                }
                __syncwarp();
            }
        }
    } else if (warp >= 17) {
        // ===================== MMA issuer of stream g = warp - 17 =====================
        // A warp of its own: as one of the softmax warps, the issuer could start its next tile only after the other
        // seven had finished the current one and the MMAs were out, so the whole stream ran at "softmax + issue" per
        // tile and the other warps idled for the difference (19 % of the samples on the wait for S).
        const int s = warp - 17, g = s;
        constexpr uint32_t idesc_s = umma_idesc(128, 64, 1u, 0u, 0u);  // Q, K both K-major
        constexpr uint32_t idesc_o = umma_idesc(128, 64, 1u, 0u, 1u);  // P K-major (TMEM), V MN-major
        const bool lead_lane = elect_one();
        // operand descriptors advance linearly (1024-aligned tiles below 256 KB), see attention.cu
        const uint64_t kdesc0 = umma_smem_desc_sw128(smem_u32(sK), 1024, 16);
        const uint64_t vdesc0 = umma_smem_desc_sw128(smem_u32(sV), 1024, 1024);
        const uint64_t qdesc0 = umma_smem_desc_sw128(smem_u32(sQ), 1024, 16);
        const uint64_t desc_hi = kdesc0 & 0xffffffff00000000ull;
        constexpr uint32_t kKvDescStep = kKvBytes >> 4, kQDescStep = kQBytes >> 4;
        int ib = 0;  // S buffer the stream's next S goes to: tiles issued so far, modulo 3
        // S for the K tile at ring position `idx` with Q buffer `qslot` into the stream's next S buffer
        auto issue_s = [&](int idx, int qslot) {
            const int stage = idx % kDbKvStages;
            mbar_wait(kv_full + stage, (uint32_t)((idx / kDbKvStages) & 1));
            tc_fence_after();
            if (lead_lane) {
                const uint64_t aq = qdesc0 + (uint32_t)qslot * kQDescStep, bk = kdesc0 + (uint32_t)stage * kKvDescStep;
#pragma unroll
                for (int k = 0; k < kHeadDim / 16; ++k)
                    umma_f16(tmem_base + s * 192 + ib * 64, aq + k * 2, bk + k * 2, idesc_s, k != 0 ? 1u : 0u);
                umma_commit(s_full + s * 3 + ib);
            }
            __syncwarp();
            ib = ib == 2 ? 0 : ib + 1;
        };
        int kv_base = 0;  // ring position of the current item's first K/V tile
        int qcnt = 0;     // Q tiles of query tile g used so far (buffer = count & 1)
        int b = 0;        // S buffer of the current tile: tiles processed so far, modulo 3
        uint32_t ph = 0;  // and its turn
        int pre = 0;      // S tiles of the current item that were issued at the end of the previous one
        for (int it = blockIdx.x; it < p.num_items; it += gridDim.x) {
            const DbItem a = db_decode(p, it, att_fetch(p, it));
            if (!a.active) continue;
            const int N = a.kv_tiles;
            if (!a.tile_active(g)) {
                // query tile without rows: only hand the K/V tiles back to the producer
                for (int t = 0; t < N; ++t) {
                    const int idx = kv_base + t;
                    mbar_wait(kv_full + idx % kDbKvStages, (uint32_t)((idx / kDbKvStages) & 1));
                    if (lane == 0) mbar_arrive(kv_empty + idx % kDbKvStages);
                    __syncwarp();
                }
                kv_base += N;
                continue;
            }
            const int qslot = g * 2 + (qcnt & 1);
            if (N > 0) {  // tiles 0..2 of the item in flight (those not issued by the previous item's tail)
                if (pre == 0) mbar_wait(q_full + qslot, (uint32_t)((qcnt >> 1) & 1));
                for (int k = pre; k < 3 && k < N; ++k) issue_s(kv_base + k, qslot);
            }
            int iss_next = 0;  // S tiles of the next item issued from this item's tail
            bool nx_known = false, nx_ok = false;
            int nx_tiles = 0;
            int stage = kv_base % kDbKvStages;  // ring slot of tile t
            for (int t = 0; t < N; ++t) {
                // Everything P.V_t and S_{t+3} need is prepared BEFORE the wait for P, the K tile's arrival included
                const bool last = t + 1 >= N;
                const bool ahead_here = t + 3 < N;  // the tile three ahead belongs to this item: the common case
                int nstage = stage + 3;
                if (nstage >= kDbKvStages) nstage -= kDbKvStages;
                const int nidx = kv_base + t + 3;
                uint32_t bv_lo = (uint32_t)(vdesc0 + (uint32_t)stage * kKvDescStep);
                uint32_t bk_lo = (uint32_t)(kdesc0 + (uint32_t)nstage * kKvDescStep);
                uint32_t aq_lo = (uint32_t)(qdesc0 + (uint32_t)qslot * kQDescStep);
                uint32_t bar_kv = smem_u32(kv_empty + stage);
                if (ahead_here) mbar_wait(kv_full + nstage, (uint32_t)((nidx / kDbKvStages) & 1));
                asm volatile("" : "+r"(bv_lo), "+r"(bk_lo), "+r"(aq_lo), "+r"(bar_kv));  // computed before the wait
                mbar_wait(p_ready + s * 3 + b, ph);  // the eight warps' P halves are in TMEM
                tc_fence_after();
                // ib == b whenever three tiles are in flight: the buffer P.V_t frees is the one S_{t+3} goes to
                if (lead_lane) {
#pragma unroll
                    for (int k = 0; k < kKvTile / 16; ++k)
                        umma_f16_ts(tmem_base + 384 + s * 64, tmem_base + s * 192 + b * 64 + k * 8,
                                    (desc_hi | bv_lo) + (uint64_t)(k * 128), idesc_o, (t | k) != 0 ? 1u : 0u);
                    umma_commit_addr(bar_kv);  // K_t / V_t free once both streams' MMAs on them retire
                    umma_commit(pv_done + s * 3 + b);
                    if (last) {
                        umma_commit(o_done + s);
                        umma_commit(q_empty + qslot);
                    }
                    if (ahead_here) {
#pragma unroll
                        for (int k = 0; k < kHeadDim / 16; ++k)
                            umma_f16(tmem_base + s * 192 + ib * 64, (desc_hi | aq_lo) + (uint64_t)(k * 2),
                                     (desc_hi | bk_lo) + (uint64_t)(k * 2), idesc_s, k != 0 ? 1u : 0u);
                        umma_commit(s_full + s * 3 + ib);
                    }
                }
                __syncwarp();
                if (ahead_here) {
                    ib = ib == 2 ? 0 : ib + 1;
                } else {
                    // tail of the item: keep three S tiles in flight with the first tiles of the NEXT item, if this
                    // stream has rows there (the immediately following item only)
                    if (!nx_known) {
                        nx_known = true;
                        if (it + (int)gridDim.x < p.num_items) {
                            const DbItem nx = db_decode(p, it + gridDim.x, att_fetch(p, it + gridDim.x));
                            nx_ok = nx.active && nx.tile_active(g) && nx.kv_tiles > 0;
                            nx_tiles = nx.kv_tiles;
                        }
                    }
                    if (nx_ok) {
                        const int nslot = g * 2 + ((qcnt + 1) & 1);
                        const int in_flight = (N - (t + 1)) + iss_next;  // S tiles issued beyond tile t
                        for (int k = in_flight; k < 3 && iss_next < nx_tiles; ++k) {
                            if (iss_next == 0) mbar_wait(q_full + nslot, (uint32_t)(((qcnt + 1) >> 1) & 1));
                            issue_s(kv_base + N + iss_next, nslot);
                            ++iss_next;
                        }
                    }
                }
                stage = stage + 1 == kDbKvStages ? 0 : stage + 1;
                if (b == 2) {
                    b = 0;
                    ph ^= 1u;
                } else {
                    ++b;
                }
            }
            kv_base += N;
            if (N > 0) ++qcnt;
            pre = iss_next;
        }
    } else {
        // ============ stream g: softmax + output, thread <-> half a query row ============
        const int s = (warp - 1) >> 3, g = s;
        const int h = ((warp - 1) >> 2) & 1;        // keys [32 h, 32 h + 32) of a tile, output dimensions [32 h, 32 h + 32)
        const int quarter = warp & 3;               // TMEM lane quarter this warp may touch
        const int row = quarter * 32 + lane;        // row inside the query tile == TMEM lane
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const uint32_t tmem_sp = tmem_base + s * 192 + lane_off;      // + 64 b: S / P buffer b (this warp's lanes)
        const uint32_t tmem_o = tmem_base + 384 + s * 64 + lane_off;
        const int pair_bar = 1 + s * 4 + quarter;  // named barrier of the two warps that share these rows
        float* my_m = xch_m + (s * 2 + h) * kQTile + row;        // + parity * 2 * 2 * kQTile
        float* other_m = xch_m + (s * 2 + (h ^ 1)) * kQTile + row;
        float* my_l = xch_l + (s * 2 + h) * kQTile + row;
        float* other_l = xch_l + (s * 2 + (h ^ 1)) * kQTile + row;
        const float c = p.scale * 1.4426950408889634f;  // softmax(scale * s) via exp2
        int n_tot = 0;                // key tiles this stream has processed so far
        int b = 0, pb = 0;            // S buffer of the current / previous tile: n_tot % 3
        uint32_t ph = 0, pph = 0;     // and its turn: (n_tot / 3) & 1
        int od = 0;                   // items in which this stream accumulated (phase of o_done)
        int items_done = 0;           // items this stream has finished (parity of the row-sum exchange slots)
        float sv[32];
        uint32_t r[16];
        AttRaw raw = {0, 0, 0, 0}, raw_next = {0, 0, 0, 0};
        if (blockIdx.x < p.num_items) raw = att_fetch(p, blockIdx.x);
        bool next_fetched = false;
        for (int it = blockIdx.x; it < p.num_items; it += gridDim.x) {
            if (it != (int)blockIdx.x) raw = next_fetched ? raw_next : att_fetch(p, it);
            next_fetched = false;
            const DbItem a = db_decode(p, it, raw);
            if (!a.active || !a.tile_active(g)) continue;
            const int N = a.kv_tiles;
            float m_run = -INFINITY, l_run = 0.f;
            for (int t = 0; t < N; ++t, ++n_tot) {
                const int valid = min(kKvTile, a.kv_len - t * kKvTile);
                mbar_wait(s_full + s * 3 + b, ph);  // S of this tile landed (issued three tiles ago)
                tc_fence_after();
                B200_TMEM_LD32F(tmem_sp + b * 64 + h * 32, sv);
                tmem_ld_wait();
                if (valid < kKvTile) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (h * 32 + i >= valid) sv[i] = -INFINITY;
                }
                float m0 = sv[0], m1 = sv[1], m2 = sv[2], m3 = sv[3];
#pragma unroll
                for (int i = 4; i < 32; i += 4) {
                    m0 = fmaxf(m0, sv[i]); m1 = fmaxf(m1, sv[i + 1]);
                    m2 = fmaxf(m2, sv[i + 2]); m3 = fmaxf(m3, sv[i + 3]);
                }
                // the row's maximum over the tile: this thread's half and its partner's (slots alternate with the tile
                // parity: a slot is rewritten two tiles later, behind the barrier of the tile in between)
                const float m_half = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                const int xo = (n_tot & 1) * (2 * 2 * kQTile);  // (tile parity, not the S buffer)
                my_m[xo] = m_half;
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                const float m_tile = fmaxf(m_half, other_m[xo]);
                // lazy rescale: move the reference maximum only when it is off by more than 2^ATT_LAZY_LOG2.  Both
                // threads of a row see the same m_tile and m_run, hence take the same decision.
                const bool need = (m_tile - m_run) * c > (float)ATT_LAZY_LOG2;
                if (t == 0) {
                    m_run = m_tile;
                } else if (__any_sync(0xffffffffu, need)) {
                    const float m_new = need ? m_tile : m_run;
                    const float alpha = ex2_approx((m_run - m_new) * c);
                    l_run *= alpha;
                    m_run = m_new;
                    // S ran ahead of P.V: the stream's previous P.V may still be accumulating into O
                    mbar_wait(pv_done + s * 3 + pb, pph);
                    tc_fence_after();
#pragma unroll 1
                    for (int ch = 0; ch < 4; ++ch) {  // this thread's 32 of the 64 output dimensions
                        uint32_t t8[8];
                        tmem_ld8(tmem_o + h * 32 + ch * 8, t8);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 8; ++i) t8[i] = __float_as_uint(__uint_as_float(t8[i]) * alpha);
                        tmem_st8(tmem_o + h * 32 + ch * 8, t8);
                    }
                }
                const f32x2 c2 = bc2(c), nmc2 = bc2(-m_run * c);
                f32x2 l2 = bc2(0.f);
#pragma unroll
                for (int i = 0; i < 16; ++i) {  // column i of this thread's P half = keys (32 h + 2i, + 1) as bf16
                    constexpr int kPolyStride = ATT_DB_POLY_PAIRS > 0 ? 16 / ATT_DB_POLY_PAIRS : 64;
                    const bool poly = ATT_DB_POLY_PAIRS > 0 && (i % kPolyStride) == kPolyStride - 1;
                    const f32x2 av = fma2(pk2(sv[2 * i], sv[2 * i + 1]), c2, nmc2);
                    float e0, e1;
                    if (poly) {
                        ex2_poly2(av, e0, e1);
                    } else {
                        float a0, a1;
                        upk2(av, a0, a1);
                        e0 = ex2_approx(a0);
                        e1 = ex2_approx(a1);
                    }
                    l2 = add2(l2, pk2(e0, e1));
                    r[i] = pack_bf16(e0, e1);
                }
                float l0, l1;
                upk2(l2, l0, l1);
                tmem_st16(tmem_sp + b * 64 + h * 16, r);  // P over this tile's S buffer: 32 keys = 16 columns per thread
                tmem_st_wait();
                l_run += l0 + l1;
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready + s * 3 + b);
                if (t == 0 && it + (int)gridDim.x < p.num_items) {
                    raw_next = att_fetch(p, it + gridDim.x);  // latency hides behind the tile
                    next_fetched = true;
                }
                pb = b;
                pph = ph;
                if (b == 2) {
                    b = 0;
                    ph ^= 1u;
                } else {
                    ++b;
                }
            }
            // ---- the stream owns its rows; the two threads of a row add their partial sums and each normalises and
            //      stores its own 32 output dimensions ----
            const bool have = N > 0;
            const int lo = (items_done & 1) * (2 * 2 * kQTile);  // slots alternate per item: rewritten two items later
            ++items_done;
            my_l[lo] = l_run;
            asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
            const float l_all = l_run + other_l[lo];
            const float inv = l_all > 0.f ? 1.0f / l_all : 0.f;  // no keys: zeros
            if (have) {
                mbar_wait(o_done + s, (uint32_t)(od & 1));  // the stream's last P.V has landed
                tc_fence_after();
            }
            const int q_idx = a.q_first + g * kQTile + row;
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                 (long long)(a.q_row0 + q_idx) * p.ld_out + p.out_col0 + a.head * kHeadDim + h * 32;
            if (have) {
                uint32_t rr[32];
                tmem_ld32(tmem_o + h * 32, rr);
                tmem_ld_wait();
                if (q_idx < a.q_len) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8)
                        *reinterpret_cast<uint4*>(dst + i) =
                            make_uint4(pack_bf16(__uint_as_float(rr[i]) * inv, __uint_as_float(rr[i + 1]) * inv),
                                       pack_bf16(__uint_as_float(rr[i + 2]) * inv, __uint_as_float(rr[i + 3]) * inv),
                                       pack_bf16(__uint_as_float(rr[i + 4]) * inv, __uint_as_float(rr[i + 5]) * inv),
                                       pack_bf16(__uint_as_float(rr[i + 6]) * inv, __uint_as_float(rr[i + 7]) * inv));
                }
            } else if (q_idx < a.q_len) {
#pragma unroll
                for (int i = 0; i < 32; i += 8) *reinterpret_cast<uint4*>(dst + i) = make_uint4(0u, 0u, 0u, 0u);
            }
            // ordered before the stream's next P.V (which overwrites O) by the fence in front of the next p_ready arrival
            tc_fence_before();
            if (have) ++od;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<kDbTmemCols>(tmem_base);
}

int launch_attention_db(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, AttnParams p, int num_seqs,
                        int max_q_len, int heads, int num_sms, cudaStream_t stream) {
    static PerDeviceSize configured;
    if (!configured.cur()) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attention_db_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDbSmemBytes));
        configured.cur() = 1;
    }
    if (num_seqs == 0 || max_q_len == 0) return 0;
    p.heads = heads;
    p.q_tiles = (max_q_len + kDbQPerCta - 1) / kDbQPerCta;
    const long long items = (long long)p.q_tiles * heads * num_seqs;
    B200_REQUIRE(items < (1ll << 31), "attention: too many work items");
    p.num_items = (int)items;
    const int grid = (int)(items < num_sms ? items : num_sms);
    attention_db_kernel<<<grid, kDbThreads, kDbSmemBytes, stream>>>(tq, tk, tv, p);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
