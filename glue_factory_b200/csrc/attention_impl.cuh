// Pieces of the attention kernel (attention.cu) that are not the kernel itself: tile geometry, packed fp32x2
// arithmetic, the polynomial exp2, work-item decoding.
#pragma once
#include "common.cuh"
#include "attention.cuh"

namespace b200 {

constexpr int kAttThreads = 544;
constexpr int kQTile = 128, kKvTile = 64, kHeadDim = 64;
constexpr int kQBytes = kQTile * kHeadDim * 2;    // 16 KB
constexpr int kKvBytes = kKvTile * kHeadDim * 2;  // 8 KB: 64 keys x 64 bf16
constexpr int kXchFloats = 2 * 2 * 2 * kQTile;  // narrow layout: [m | l][query tile][stream parity][row]
template <bool WIDE>
struct AttCfg {
    static constexpr int kQTiles = WIDE ? 4 : 2;         // query tiles of an item
    static constexpr int kQPerCta = kQTiles * kQTile;    // queries of an item
    static constexpr int kKvStages = WIDE ? 6 : 8;       // wide: one K/V tile per round of the four streams (narrow: two)
    static constexpr int kKvStride = WIDE ? 1 : 2;       // key tiles between two tiles of a stream
    static constexpr int kKvUsers = WIDE ? 4 : 2;        // streams that consume one K/V tile
    static constexpr int kQUsers = WIDE ? 1 : 2;         // streams that consume one Q tile
    // smem: Q[query tile][2 buffers] | K[stages] V[stages] | barriers | (m, l) exchange (narrow only)
    static constexpr int kSmemBytes = 1024 + 2 * kQTiles * kQBytes + 2 * kKvStages * kKvBytes + 512 + (WIDE ? 0 : kXchFloats * 4);
};
static_assert(AttCfg<true>::kSmemBytes <= 232448 && AttCfg<false>::kSmemBytes <= 232448, "attention smem");
// stream s:  S/P_s: [64 s, 64 s + 64)   O_s: [256 + 64 s, 256 + 64 s + 64)
constexpr uint32_t kAttTmemCols = 512;

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// packed fp32x2 arithmetic (FFMA2 / FADD2 on sm_100): two fp32 operations per issue slot
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
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// 2^x on the FMA pipe for a pair of elements (Cody-Waite: round to nearest integer through the 1.5 * 2^23 trick,
// degree-3 polynomial on [-0.5, 0.5], exponent added as an integer): relative error <= 7.5e-5, far below the bf16
// rounding of P.  The exponentials bound this kernel on the 16-lane MUFU; ATT_POLY_PAIRS of the 32 key pairs of a
// tile row take this path (packed: 2 FMNMX + 6 FFMA2 / FADD2 + 2 shift-adds per pair) so that MUFU and issue
// slots are both kept busy.
#ifndef ATT_POLY_PAIRS
#define ATT_POLY_PAIRS 4
#endif
#ifndef ATT_ISSUER_SPREAD
#define ATT_ISSUER_SPREAD 0
#endif
#ifndef ATT_P_STORE8
#define ATT_P_STORE8 1
#endif
// lazy rescale: the reference maximum moves only when a row's new maximum exceeds it by more than 2^ATT_LAZY_LOG2
#ifndef ATT_LAZY_LOG2
#define ATT_LAZY_LOG2 8
#endif
__device__ __forceinline__ void ex2_poly2(f32x2 x, float& e0, float& e1) {
    float x0, x1;
    upk2(x, x0, x1);
    x = pk2(fmaxf(x0, -125.0f), fmaxf(x1, -125.0f));
    const f32x2 t = add2(x, bc2(12582912.0f));
    const f32x2 f = fma2(add2(t, bc2(-12582912.0f)), bc2(-1.0f), x);
    f32x2 q = fma2(f, bc2(0.05517173558473587f), bc2(0.2426111251115799f));
    q = fma2(q, f, bc2(0.6932609677314758f));
    q = fma2(q, f, bc2(0.9999280571937561f));
    float q0, q1, t0, t1;
    upk2(q, q0, q1);
    upk2(t, t0, t1);
    e0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
    e1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
}

// one work item, decoded identically by every warp (all fields warp-uniform)
struct AttItem {
    bool active;     // the item's first query tile holds at least one row
    int q_row0, kv_row0, q_len, kv_len, kv_tiles, qt, head;
    int q_first;     // first query of the item inside its sequence
    __device__ __forceinline__ bool tile_active(int g) const { return q_first + g * kQTile < q_len; }
};

// the four words of an item that come from memory (two dependent global loads: ~1500 clk when not hidden)
struct AttRaw {
    int q_row0, kv_row0, q_len, kv_len;
};

__device__ __forceinline__ AttRaw att_fetch(const AttnParams& p, int it) {
    const AttnSeq sd = p.seqs[it / p.q_tiles / p.heads];
    AttRaw r;
    r.q_row0 = sd.q_row0;
    r.kv_row0 = sd.kv_row0;
    r.q_len = p.len_override ? p.len_override[sd.q_len_slot] : sd.q_len;
    r.kv_len = p.len_override ? p.len_override[sd.kv_len_slot] : sd.kv_len;
    return r;
}

template <bool WIDE>
__device__ __forceinline__ AttItem att_decode(const AttnParams& p, int it, const AttRaw& r) {
    AttItem a;
    a.qt = it % p.q_tiles;
    a.head = (it / p.q_tiles) % p.heads;
    a.q_len = __shfl_sync(0xffffffffu, r.q_len, 0);
    a.kv_len = __shfl_sync(0xffffffffu, r.kv_len, 0);
    a.q_row0 = __shfl_sync(0xffffffffu, r.q_row0, 0);
    a.kv_row0 = __shfl_sync(0xffffffffu, r.kv_row0, 0);
    a.q_first = a.qt * AttCfg<WIDE>::kQPerCta;
    a.active = a.q_first < a.q_len;
    a.kv_tiles = (a.kv_len + kKvTile - 1) / kKvTile;
    return a;
}

template <bool WIDE>
__device__ __forceinline__ AttItem att_item(const AttnParams& p, int it) { return att_decode<WIDE>(p, it, att_fetch(p, it)); }


}  // namespace b200
