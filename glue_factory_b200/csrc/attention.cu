// Dense path: flash-style multi-head attention on tcgen05 / TMEM for head_dim 64.
//
// One CTA per (128-query tile, head, sequence).  Per 128-key tile:
//   S = Q K^T   (tcgen05.mma, M=128 N=128 K=64, fp32 in TMEM)
//   online softmax by 128 threads, one query row each (tcgen05.ld), P -> smem as bf16 (swizzled K-major)
//   PV = P V    (tcgen05.mma, M=128 N=64 K=128, V consumed token-major through an MN-major descriptor)
//   O = O * alpha + PV in registers.
// Q/K/V tiles arrive by TMA from token-major [tokens, ld] bf16 matrices; a head is a 64-column slice.
// The S matrix is never materialised in HBM.  Two CTAs are resident per SM so that one CTA's MMAs
// overlap the other's softmax.  Self-attention and both directions of the bidirectional
// cross-attention use this kernel (the cross block's similarity is recomputed per direction).
//
// Replaces F.scaled_dot_product_attention / the einsum+softmax path of the reference
// (reference: gluefactory/models/matchers/lightglue.py:105-122,144-157,186-215).
#include "common.cuh"
#include "attention.cuh"

namespace b200 {

constexpr int kAttThreads = 192;
constexpr int kQTile = 128, kKvTile = 128, kHeadDim = 64;
constexpr int kTileBytes = 128 * 64 * 2;  // 16 KB: 128 rows x 64 bf16
constexpr int kKvStages = 2;
// smem: Q | K[2] | V[2] | P (2 halves of 64 keys) | barriers
constexpr int kAttSmemBytes = kTileBytes * (1 + 2 * kKvStages + 2) + 1024 + 256;
constexpr uint32_t kAttTmemCols = 256;  // S: columns [0,128), PV: columns [128,192)

__global__ void __launch_bounds__(kAttThreads, 2)
attention_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                         const __grid_constant__ CUtensorMap tmap_v, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kTileBytes;
    uint8_t* sV = sK + kKvStages * kTileBytes;
    uint8_t* sP = sV + kKvStages * kTileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * kTileBytes);
    uint64_t* q_full = bars;
    uint64_t* k_full = bars + 1;               // [kKvStages]
    uint64_t* k_empty = k_full + kKvStages;    // [kKvStages]
    uint64_t* v_full = k_empty + kKvStages;    // [kKvStages]
    uint64_t* v_empty = v_full + kKvStages;    // [kKvStages]
    uint64_t* s_full = v_empty + kKvStages;
    uint64_t* p_ready = s_full + 1;
    uint64_t* o_full = p_ready + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seq = blockIdx.z, head = blockIdx.y, qt = blockIdx.x;
    const AttnSeq sd = p.seqs[seq];
    const int q_len = p.len_override ? p.len_override[sd.q_len_slot] : sd.q_len;
    const int kv_len = p.len_override ? p.len_override[sd.kv_len_slot] : sd.kv_len;
    if (qt * kQTile >= q_len) return;  // uniform for the whole CTA, before any barrier / TMEM use
    const int kv_tiles = (kv_len + kKvTile - 1) / kKvTile;

    if (warp == 0 && lane == 0) {
        prefetch_tensormap(&tmap_q);
        prefetch_tensormap(&tmap_k);
        prefetch_tensormap(&tmap_v);
        mbar_init(q_full, 1);
        for (int s = 0; s < kKvStages; ++s) {
            mbar_init(k_full + s, 1);
            mbar_init(k_empty + s, 1);
            mbar_init(v_full + s, 1);
            mbar_init(v_empty + s, 1);
        }
        mbar_init(s_full, 1);
        mbar_init(p_ready, 4);
        mbar_init(o_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kAttTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + 128;

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(q_full, kTileBytes);
            tma_load_2d(sQ, &tmap_q, q_full, p.q_col0 + head * kHeadDim, sd.q_row0 + qt * kQTile);
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < kv_tiles; ++j) {
                const int row = sd.kv_row0 + j * kKvTile;
                mbar_wait(k_empty + stage, phase ^ 1);
                mbar_arrive_expect_tx(k_full + stage, kTileBytes);
                tma_load_2d(sK + stage * kTileBytes, &tmap_k, k_full + stage, p.k_col0 + head * kHeadDim, row);
                mbar_wait(v_empty + stage, phase ^ 1);
                mbar_arrive_expect_tx(v_full + stage, kTileBytes);
                tma_load_2d(sV + stage * kTileBytes, &tmap_v, v_full + stage, p.v_col0 + head * kHeadDim, row);
                if (++stage == kKvStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc(128, 128, 1u, 0u, 0u);  // Q, K both K-major
            constexpr uint32_t idesc_o = umma_idesc(128, 64, 1u, 0u, 1u);   // P K-major, V MN-major
            mbar_wait(q_full, 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < kv_tiles; ++j) {
                // S = Q K^T ; the previous tile's S has been consumed (p_ready of tile j-1 was waited below)
                mbar_wait(k_full + stage, phase);
                tc_fence_after();
                const uint32_t aq = smem_u32(sQ), bk = smem_u32(sK + stage * kTileBytes);
#pragma unroll
                for (int k = 0; k < kHeadDim / 16; ++k)
                    umma_f16(tmem_s, umma_smem_desc_sw128(aq + k * 32, 1024, 16),
                             umma_smem_desc_sw128(bk + k * 32, 1024, 16), idesc_s, k != 0 ? 1u : 0u);
                umma_commit(k_empty + stage);
                umma_commit(s_full);
                // PV = P V once the softmax warps have written P (and finished reading the previous PV)
                mbar_wait(p_ready, j & 1);
                mbar_wait(v_full + stage, phase);
                tc_fence_after();
                const uint32_t ap = smem_u32(sP), bv = smem_u32(sV + stage * kTileBytes);
#pragma unroll
                for (int k = 0; k < kKvTile / 16; ++k)
                    umma_f16(tmem_o,
                             umma_smem_desc_sw128(ap + (k >> 2) * kTileBytes + (k & 3) * 32, 1024, 16),
                             umma_smem_desc_sw128(bv + k * 2048, 1024, 1024), idesc_o, k != 0 ? 1u : 0u);
                umma_commit(v_empty + stage);
                umma_commit(o_full);
                if (++stage == kKvStages) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ---- softmax + output warps: thread <-> query row -------------------------------------
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;  // row inside the query tile == TMEM lane
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const float c = p.scale * 1.4426950408889634f;  // softmax(scale * s) via exp2
        float m_run = -INFINITY, l_run = 0.f;
        float o[kHeadDim];
#pragma unroll
        for (int i = 0; i < kHeadDim; ++i) o[i] = 0.f;
        uint32_t r[32];
        for (int j = 0; j < kv_tiles; ++j) {
            const int valid = min(kKvTile, kv_len - j * kKvTile);
            mbar_wait(s_full, j & 1);
            tc_fence_after();
            // pass 1: row maximum
            float m_new = m_run;
#pragma unroll 1
            for (int ch = 0; ch < 4; ++ch) {
                tmem_ld32(tmem_s + lane_off + ch * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (ch * 32 + i < valid) m_new = fmaxf(m_new, __uint_as_float(r[i]));
            }
            const float alpha = exp2f((m_run - m_new) * c);
            const float mc = m_new * c;
            float l_tile = 0.f;
            // pass 2: P = exp2(s*c - m*c) -> bf16 -> swizzled smem
#pragma unroll 1
            for (int ch = 0; ch < 4; ++ch) {
                tmem_ld32(tmem_s + lane_off + ch * 32, r);
                tmem_ld_wait();
                float e[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float x = exp2f(__uint_as_float(r[i]) * c - mc);
                    e[i] = (ch * 32 + i < valid) ? x : 0.f;
                    l_tile += e[i];
                }
                uint8_t* base = sP + (ch >> 1) * kTileBytes + row * 128;
#pragma unroll
                for (int g = 0; g < 4; ++g) {  // 4 groups of 8 keys = 16 bytes each
                    const int chunk = (ch & 1) * 4 + g;
                    *reinterpret_cast<uint4*>(base + ((chunk ^ (row & 7)) << 4)) =
                        make_uint4(pack_bf16(e[g * 8], e[g * 8 + 1]), pack_bf16(e[g * 8 + 2], e[g * 8 + 3]),
                                   pack_bf16(e[g * 8 + 4], e[g * 8 + 5]), pack_bf16(e[g * 8 + 6], e[g * 8 + 7]));
                }
            }
            m_run = m_new;
            l_run = l_run * alpha + l_tile;
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_ready);
            // accumulate PV
            mbar_wait(o_full, j & 1);
            tc_fence_after();
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                tmem_ld32(tmem_o + lane_off + ch * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) o[ch * 32 + i] = o[ch * 32 + i] * alpha + __uint_as_float(r[i]);
            }
        }
        const int q_idx = qt * kQTile + row;
        if (q_idx < q_len) {
            const float inv = 1.0f / l_run;
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)(sd.q_row0 + q_idx) * p.ld_out +
                                 p.out_col0 + head * kHeadDim;
#pragma unroll
            for (int i = 0; i < kHeadDim; i += 8)
                *reinterpret_cast<uint4*>(dst + i) =
                    make_uint4(pack_bf16(o[i] * inv, o[i + 1] * inv), pack_bf16(o[i + 2] * inv, o[i + 3] * inv),
                               pack_bf16(o[i + 4] * inv, o[i + 5] * inv), pack_bf16(o[i + 6] * inv, o[i + 7] * inv));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<kAttTmemCols>(tmem_base);
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
    dim3 grid((max_q_len + kQTile - 1) / kQTile, heads, num_seqs);
    attention_tcgen05_kernel<<<grid, kAttThreads, kAttSmemBytes, stream>>>(tq, tk, tv, p);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
