// C-ABI entry points (include/b200glue.h): argument checks, TMA tensor-map construction, launches.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "../../include/b200glue.h"
#include "attention.cuh"
#include "conv_stem.cuh"
#include "common.cuh"
#include "gemm.cuh"
#include "io_kernels.cuh"
#include "lg_kernels.cuh"
#include "sp_kernels.cuh"

namespace b200 {

static thread_local char g_error[512] = "";

void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int ln_part_cols();
int launch_gemm(int block_n, int epi, int a_mode, const CUtensorMap& ta, const CUtensorMap& tb,
                const GemmParams& p, int num_sms, cudaStream_t stream);
int launch_conv_patch(int cout, const CUtensorMap& ta, const CUtensorMap& tw, const GemmParams& p, int num_sms,
                      cudaStream_t stream);
int launch_conv_stem(const CUtensorMap& tw, const StemParams& p, int num_sms, cudaStream_t stream);
int launch_attention(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, AttnParams p,
                     int num_seqs, int max_q_len, int heads, int num_sms, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// driver entry point for cuTensorMapEncodeTiled (no link-time dependency on libcuda)
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

struct MapKey {
    const void* ptr;
    uint64_t d[4];
    uint64_t s[3];
    uint32_t box[4];
    uint32_t rank;
    bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
        uint64_t h = 1469598103934665603ull;
        for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
        return (size_t)h;
    }
};
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;
static std::mutex g_maps_mu;

// bf16 tensor of `rank` dims (dim 0 contiguous), 128-byte swizzle, zero fill out of bounds
static int make_map(CUtensorMap* out, const void* ptr, uint32_t rank, const uint64_t* dims,
                    const uint64_t* strides_bytes /*rank-1*/, const uint32_t* box) {
    MapKey key;
    memset(&key, 0, sizeof(key));
    key.ptr = ptr;
    key.rank = rank;
    for (uint32_t i = 0; i < rank; ++i) { key.d[i] = dims[i]; key.box[i] = box[i]; }
    for (uint32_t i = 0; i + 1 < rank; ++i) key.s[i] = strides_bytes[i];
    {
        std::lock_guard<std::mutex> lk(g_maps_mu);
        auto it = g_maps.find(key);
        if (it != g_maps.end()) { *out = it->second; return 0; }
    }
    EncodeTiledFn enc = get_encode_fn();
    B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the CUDA driver");
    B200_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand %p is not 16-byte aligned", ptr);
    for (uint32_t i = 0; i + 1 < rank; ++i)
        B200_REQUIRE(strides_bytes[i] % 16 == 0, "TMA stride %llu is not a multiple of 16 bytes",
                     (unsigned long long)strides_bytes[i]);
    cuuint64_t gd[4];
    cuuint64_t gs[3];
    cuuint32_t bx[4], es[4] = {1, 1, 1, 1};
    for (uint32_t i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; }
    for (uint32_t i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), gd, gs, bx, es,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with code %d (rank %u, dims %llu x %llu, box %u x %u)",
                 (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
    std::lock_guard<std::mutex> lk(g_maps_mu);
    if (g_maps.size() > 4096) g_maps.clear();
    g_maps.emplace(key, *out);
    return 0;
}

// [rows, cols] bf16 matrix with leading dimension ld (elements); box = 64 columns x box_rows rows
static int make_map_2d(CUtensorMap* out, const void* ptr, long long rows, long long cols, long long ld,
                       uint32_t box_rows) {
    const uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
    const uint64_t strides[1] = {(uint64_t)ld * 2};
    const uint32_t box[2] = {64, box_rows};
    return make_map(out, ptr, 2, dims, strides, box);
}

bool pdl_enabled() {
    static const bool on = [] {
        const char* v = getenv("B200_PDL");
        return v == nullptr || atoi(v) != 0;
    }();
    return on;
}

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}

static int sm_count() {  // of the current device
    static int n[kMaxDevices] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 0;
    if (n[dev] == 0 && cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n[dev] = 0;
    return n[dev];
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_abi_version(void) { return B200_ABI_VERSION; }
const char* b200_last_error(void) { return g_error; }

int b200_sm_count(int* out) {
    const int n = sm_count();
    B200_REQUIRE(n > 0, "no CUDA device available");
    *out = n;
    return 0;
}

int b200_gemm(const b200_gemm_desc* d, b200_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    B200_REQUIRE(d != nullptr, "gemm: null descriptor");
    B200_REQUIRE(d->a && d->b, "gemm: null operand");
    const int num_sms = sm_count();
    B200_REQUIRE(num_sms > 0, "gemm: no CUDA device available");
    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.M = d->M; p.N = d->N; p.K = d->K; p.batch = d->batch;
    p.a_batch_rows = d->a_batch_rows; p.b_batch_rows = d->b_batch_rows;
    p.bias = d->bias; p.scale = d->scale; p.shift = d->shift;
    p.relu = d->relu; p.pool = d->pool; p.alpha = d->alpha;
    p.resid = d->resid; p.cs = reinterpret_cast<const float2*>(d->cossin); p.rot_cols = d->rot_cols;
    p.ln_gamma = d->ln_gamma; p.ln_beta = d->ln_beta; p.ln_eps = d->ln_eps;
    p.out16 = d->out16; p.ld16 = d->ld16; p.out32 = d->out32; p.ld32 = d->ld32;
    p.out32_batch_stride = d->out32_batch_stride;
    p.cells_h = d->cells_h; p.cells_w = d->cells_w;
    p.tile_live = d->tile_live;
    B200_REQUIRE(d->batch >= 0 && d->N >= 1, "gemm: bad shape (batch=%d, N=%d)", d->batch, d->N);
    if (d->batch == 0) return 0;

    const int epi = d->epilogue;
    if (epi == B200_EPI_BIAS_BF16 || epi == B200_EPI_QKV_ROTARY || epi == B200_EPI_LN_GELU ||
        epi == B200_EPI_RESID || epi == B200_EPI_CONV_RELU_BN)
        B200_REQUIRE(d->out16 != nullptr && (d->ld16 % 8) == 0 && (reinterpret_cast<uintptr_t>(d->out16) & 15) == 0,
                     "gemm: bf16 output must be 16-byte aligned with ld %% 8 == 0");
    if (epi == B200_EPI_RESID)
        B200_REQUIRE(d->out32 && d->resid && (d->ld32 % 4) == 0 && d->N % 32 == 0, "gemm: bad residual epilogue operands");
    if (epi == B200_EPI_BIAS_BF16 || epi == B200_EPI_QKV_ROTARY)
        B200_REQUIRE(d->N % 32 == 0, "gemm: N=%d must be a multiple of 32 for bf16 epilogues", d->N);
    if (epi == B200_EPI_QKV_ROTARY)
        B200_REQUIRE(d->cossin != nullptr && d->rot_cols % 64 == 0, "gemm: rotary epilogue needs cos/sin and rot_cols %% 64 == 0");
    if (epi == B200_EPI_LN_GELU)
        B200_REQUIRE(d->N == 512 && d->block_n == 512 && d->bias && d->ln_gamma && d->ln_beta,
                     "gemm: LayerNorm epilogue needs N == block_n == 512 and bias/gamma/beta");
    if (epi == B200_EPI_F32) B200_REQUIRE(d->out32 != nullptr, "gemm: fp32 epilogue needs out32");
    B200_REQUIRE(d->N <= 768 || (d->bias == nullptr && d->scale == nullptr && d->shift == nullptr),
                 "gemm: per-column epilogue operands are staged for N <= 768 (N=%d)", d->N);
    if (epi == B200_EPI_CONV_RELU_BN || epi == B200_EPI_CONV_BN_L2 || epi == B200_EPI_CONV_BN_SOFTMAX_D2S)
        B200_REQUIRE(d->bias && d->scale && d->shift, "gemm: conv epilogues need bias, scale and shift");
    if (epi == B200_EPI_CONV_BN_L2)
        B200_REQUIRE(d->N == 256 && d->block_n == 256 && d->out32 && (d->ld32 % 4) == 0, "gemm: L2 epilogue needs N == 256");
    if (epi == B200_EPI_CONV_BN_SOFTMAX_D2S)
        B200_REQUIRE(d->N == 65 && d->block_n == 80 && d->out32 && d->cells_h > 0 && d->cells_w > 0 &&
 d->M % (d->cells_h * d->cells_w) == 0,
                     "gemm: softmax/depth-to-space epilogue needs N == 65, block_n == 80 and the cell grid");

    CUtensorMap ta, tb;
    if (d->conv && d->Cin == 64 && epi == B200_EPI_CONV_RELU_BN && d->K == 576 && d->ldb == 576 &&
        env_int("B200_CONV_PATCH", 1) != 0 &&
        (d->N == 64 || (d->N == 128 && env_int("B200_CONV_PATCH_SPLIT", 1) != 0))) {
        // C_in = 64 layers: halo patch loaded once per tile, nine taps addressed inside it (conv_patch.cu).  C_out = 128
        // (64 -> 128) runs as two passes over 64 output channels each: the patch is read twice (72 KB per tile) where the
        // generic implicit GEMM fetches its A tile nine times (144 KB), 0.76 -> see DESIGN.md 5.1
        p.H = d->H; p.W = d->W; p.Cin = 64;
        p.tiles_x = (d->W + 7) / 8; p.tiles_y = (d->H + 15) / 16;
        const uint64_t dims[4] = {64, (uint64_t)d->W, (uint64_t)d->H, (uint64_t)d->batch};
        const uint64_t strides[3] = {128, (uint64_t)d->W * 128, (uint64_t)d->H * d->W * 128};
        const uint32_t box[4] = {64, 16, 18, 1};
        if (make_map(&ta, d->a, 4, dims, strides, box) != 0) return -1;
        for (int j = 0; j < d->N / 64; ++j) {
            GemmParams pj = p;
            pj.bias = p.bias + 64 * j;
            pj.scale = p.scale + 64 * j;
            pj.shift = p.shift + 64 * j;
            pj.out16 = static_cast<__nv_bfloat16*>(p.out16) + 64 * j;  // ld16 stays the full row width
            const __nv_bfloat16* wj = static_cast<const __nv_bfloat16*>(d->b) + (size_t)64 * j * 576;
            if (make_map_2d(&tb, wj, 64, 576, 576, 64u) != 0) return -1;
            if (launch_conv_patch(64, ta, tb, pj, num_sms, stream) != 0) return -1;
        }
        return 0;
    }
    const uint32_t b_box = epi == EPI_LN_GELU ? (uint32_t)ln_part_cols() : d->block_n > 256 ? 256u : (uint32_t)d->block_n;
    if (make_map_2d(&tb, d->b, d->b_rows, d->K, d->ldb, b_box) != 0) return -1;
    int a_mode = 0;
    if (d->conv) {
        a_mode = 1;
        B200_REQUIRE(d->Cin % 64 == 0 && d->K == 9 * d->Cin, "conv: Cin=%d must be a multiple of 64 and K == 9*Cin", d->Cin);
        B200_REQUIRE(d->N % d->block_n == 0, "conv: Cout=%d must be a multiple of block_n=%d", d->N, d->block_n);
        p.H = d->H; p.W = d->W; p.Cin = d->Cin;
        p.tiles_x = (d->W + 15) / 16; p.tiles_y = (d->H + 7) / 8;
        const uint64_t dims[4] = {(uint64_t)d->Cin, (uint64_t)d->W, (uint64_t)d->H, (uint64_t)d->batch};
        const uint64_t strides[3] = {(uint64_t)d->Cin * 2, (uint64_t)d->W * d->Cin * 2,
                                     (uint64_t)d->H * d->W * d->Cin * 2};
        const uint32_t box[4] = {64, 16, 8, 1};
        if (make_map(&ta, d->a, 4, dims, strides, box) != 0) return -1;
    } else {
        B200_REQUIRE(d->M >= 1, "gemm: M=%d", d->M);
#ifndef GEMM_TIMING_LN_A_ROWS  // timing experiments only (wrong results below 128): operand-feed share of the LN GEMM
#define GEMM_TIMING_LN_A_ROWS 128
#endif
        if (make_map_2d(&ta, d->a, d->a_rows, d->K, d->lda, epi == EPI_LN_GELU ? GEMM_TIMING_LN_A_ROWS : 128) != 0) return -1;
    }
    return launch_gemm(d->block_n, epi, a_mode, ta, tb, p, num_sms, stream);
}

int b200_attention(const b200_attn_desc* d, b200_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    B200_REQUIRE(d != nullptr && d->q && d->k && d->v && d->out && d->seqs, "attention: null argument");
    B200_REQUIRE(d->heads >= 1 && d->num_seqs >= 0 && d->max_q_len >= 0, "attention: bad sizes");
    B200_REQUIRE((d->ld_out % 8) == 0 && (d->out_col0 % 8) == 0 && (reinterpret_cast<uintptr_t>(d->out) & 15) == 0,
                 "attention: output must be 16-byte aligned");
    CUtensorMap tq, tk, tv;
    if (make_map_2d(&tq, d->q, d->q_rows, d->ldq, d->ldq, 128) != 0) return -1;
    if (make_map_2d(&tk, d->k, d->k_rows, d->ldk, d->ldk, 64) != 0) return -1;  // 64-key tiles
    if (make_map_2d(&tv, d->v, d->v_rows, d->ldv, d->ldv, 64) != 0) return -1;
    AttnParams p;
    p.seqs = reinterpret_cast<const AttnSeq*>(d->seqs);
    p.len_override = d->len_override;
    p.q_col0 = d->q_col0; p.k_col0 = d->k_col0; p.v_col0 = d->v_col0;
    p.scale = d->scale;
    p.out = d->out; p.ld_out = d->ld_out; p.out_col0 = d->out_col0;
    const int num_sms = sm_count();
    B200_REQUIRE(num_sms > 0, "attention: no CUDA device available");
    return launch_attention(tq, tk, tv, p, d->num_seqs, d->max_q_len, d->heads, num_sms, stream);
}

static int conv1a_any(const void* image, int u8, int B, int C, int H, int W, const float* w, const float* bias,
                      const float* scale, const float* shift, void* out_nhwc, b200_stream_t stream) {
    return launch_conv1a(image, u8, B, C, H, W, w, bias, scale, shift, out_nhwc, static_cast<cudaStream_t>(stream));
}
int b200_sp_conv1a(const float* image, int B, int C, int H, int W, const float* w, const float* bias,
                   const float* scale, const float* shift, void* out_nhwc, b200_stream_t stream) {
    return conv1a_any(image, 0, B, C, H, W, w, bias, scale, shift, out_nhwc, stream);
}
int b200_sp_conv1a_u8(const unsigned char* image, int B, int C, int H, int W, const float* w, const float* bias,
                      const float* scale, const float* shift, void* out_nhwc, b200_stream_t stream) {
    return conv1a_any(image, 1, B, C, H, W, w, bias, scale, shift, out_nhwc, stream);
}
static int conv_stem_any(const void* image, int u8, int B, int C, int H, int W, const float* w1, const float* host_params,
                         const void* w2, int pool, void* out_nhwc, b200_stream_t stream) {
    B200_REQUIRE(image && w1 && host_params && w2 && out_nhwc, "conv_stem: null argument");
    B200_REQUIRE(C == 1 || C == 3, "conv_stem: image must have 1 or 3 channels, got %d", C);
    B200_REQUIRE(B >= 0 && H > 0 && W > 0, "conv_stem: bad sizes");
    B200_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), "conv_stem: pooling needs even H and W");
    B200_REQUIRE((reinterpret_cast<uintptr_t>(out_nhwc) & 15) == 0 && (reinterpret_cast<uintptr_t>(w2) & 15) == 0,
                 "conv_stem: output and weights must be 16-byte aligned");
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, host_params) == cudaSuccess)
        B200_REQUIRE(attr.type != cudaMemoryTypeDevice, "conv_stem: the per-channel operands must be in HOST memory");
    else
        cudaGetLastError();
    const int num_sms = sm_count();
    B200_REQUIRE(num_sms > 0, "conv_stem: no CUDA device available");
    CUtensorMap tw;
    if (make_map_2d(&tw, w2, 64, 576, 576, 64) != 0) return -1;
    StemParams p;
    p.img = image; p.u8 = u8; p.B = B; p.C = C; p.H = H; p.W = W;
    p.w1 = w1;
    memcpy(p.par, host_params, sizeof(p.par));
    p.pool = pool;
    p.out = reinterpret_cast<__nv_bfloat16*>(out_nhwc);
    p.tiles_x = (W + 7) / 8; p.tiles_y = (H + 15) / 16;
    return launch_conv_stem(tw, p, num_sms, static_cast<cudaStream_t>(stream));
}
int b200_sp_conv_stem(const float* image, int B, int C, int H, int W, const float* w1, const float* host_params,
                      const void* w2, int pool, void* out_nhwc, b200_stream_t stream) {
    return conv_stem_any(image, 0, B, C, H, W, w1, host_params, w2, pool, out_nhwc, stream);
}
int b200_sp_conv_stem_u8(const unsigned char* image, int B, int C, int H, int W, const float* w1,
                         const float* host_params, const void* w2, int pool, void* out_nhwc, b200_stream_t stream) {
    return conv_stem_any(image, 1, B, C, H, W, w1, host_params, w2, pool, out_nhwc, stream);
}
int b200_sp_nms_border(const float* scores, int B, int H, int W, int nms_radius, int remove_borders, float* out,
                       b200_stream_t stream) {
    return launch_nms_border(scores, B, H, W, nms_radius, remove_borders, out, static_cast<cudaStream_t>(stream));
}
int b200_sp_nms_border_list(const float* scores, int B, int H, int W, int nms_radius, int remove_borders,
                            const int* image_hw, float* out, unsigned long long* candidates, int* candidate_state,
                            int capacity, b200_stream_t stream) {
    return launch_nms_list(scores, B, H, W, nms_radius, remove_borders, image_hw, out, candidates, candidate_state,
                           capacity, static_cast<cudaStream_t>(stream));
}
int b200_sp_select_topk_list(const float* nms_map, int B, int H, int W, float threshold, int k,
                             const unsigned long long* candidates, int* candidate_state, int capacity,
                             float* keypoints, float* scores, int* indices, int* counts, b200_stream_t stream) {
    return launch_select_list(nms_map, B, H, W, threshold, k, candidates, candidate_state, capacity, keypoints, scores,
                              indices, counts, static_cast<cudaStream_t>(stream));
}
int b200_sp_count_candidates(const float* nms_map, int B, int H, int W, float threshold, int* counts,
                             b200_stream_t stream) {
    return launch_count_candidates(nms_map, B, H, W, threshold, counts, static_cast<cudaStream_t>(stream));
}
int b200_sp_select_topk(const float* nms_map, int B, int H, int W, float threshold, int k, float* keypoints,
                        float* scores, int* indices, int* counts, b200_stream_t stream) {
    return launch_select_topk(nms_map, B, H, W, threshold, k, keypoints, scores, indices, counts,
                              static_cast<cudaStream_t>(stream));
}
int b200_sp_sample_descriptors(const float* dense, int B, int Hc, int Wc, int C, const float* keypoints, int K,
                               int stride, float* out, b200_stream_t stream) {
    return launch_sample_descriptors(dense, B, Hc, Wc, C, keypoints, K, stride, 0, out, static_cast<cudaStream_t>(stream));
}
int b200_sp_sample_descriptors_mode(const float* dense, int B, int Hc, int Wc, int C, const float* keypoints, int K,
                                    int stride, int mode, float* out, b200_stream_t stream) {
    return launch_sample_descriptors(dense, B, Hc, Wc, C, keypoints, K, stride, mode, out,
                                     static_cast<cudaStream_t>(stream));
}
int b200_lg_ingest(const float* desc, long long rows, float* x32, void* xm, long long ld_xm, b200_stream_t stream) {
    return launch_ingest(desc, rows, x32, xm, ld_xm, static_cast<cudaStream_t>(stream));
}
int b200_lg_posenc(const float* keypoints, const float* image_size, int B, int n, const float* wr,
                   float* size_scratch, void* cossin, b200_stream_t stream) {
    return launch_posenc(keypoints, image_size, B, n, wr, size_scratch, cossin, static_cast<cudaStream_t>(stream));
}
int b200_lg_rowdot(const float* x32, long long rows, const float* w, const float* bias, int apply_sigmoid, float* out,
                   b200_stream_t stream) {
    return launch_rowdot(x32, rows, w, bias, apply_sigmoid, out, static_cast<cudaStream_t>(stream));
}
size_t b200_lg_assignment_workspace(int B, int M, int N) { return assignment_workspace_bytes(B, M, N); }
int b200_lg_assignment(const float* sim, const float* z0, const float* z1, int B, int M, int N,
                       float filter_threshold, void* workspace, float* scores, long long* matches0,
                       long long* matches1, float* mscores0, float* mscores1, b200_stream_t stream) {
    return launch_assignment(sim, z0, z1, B, M, N, filter_threshold, workspace, scores, matches0, matches1, mscores0,
                             mscores1, static_cast<cudaStream_t>(stream));
}
size_t b200_lg_assignment_static_capacity(void) { return assignment_static_capacity(); }
int b200_lg_assignment_static(const float* sim, const float* z0, const float* z1, int B, int M, int N,
                              float filter_threshold, float* scores, long long* matches0, long long* matches1,
                              float* mscores0, float* mscores1, b200_stream_t stream) {
    B200_REQUIRE(sim && z0 && z1 && scores && matches0 && matches1 && mscores0 && mscores1,
                 "assignment_static: null argument");
    return launch_assignment_static(sim, z0, z1, B, M, N, filter_threshold, scores, matches0, matches1, mscores0,
                                    mscores1, static_cast<cudaStream_t>(stream));
}

int b200_sg_keypoint_encoder_head(const float* keypoints, const float* scores, const float* image_size, int B, int n,
                                  const float* w1, const float* b1, const float* w2, const float* b2, void* out,
                                  b200_stream_t stream) {
    B200_REQUIRE(keypoints && scores && image_size && w1 && b1 && w2 && b2 && out, "keypoint encoder: null argument");
    return launch_sg_kenc_head(keypoints, scores, image_size, B, n, w1, b1, w2, b2, out, static_cast<cudaStream_t>(stream));
}
int b200_sg_split_hilo(const float* x, long long rows, long long split_row, void* out, b200_stream_t stream) {
    B200_REQUIRE(x && out, "split_hilo: null argument");
    return launch_sg_split_hilo(x, rows, split_row, out, static_cast<cudaStream_t>(stream));
}
int b200_split_hilo(const float* x, long long rows, int cols, long long split_row, void* out, b200_stream_t stream) {
    B200_REQUIRE(x && out, "split_hilo: null argument");
    return launch_split_hilo(x, rows, cols, split_row, out, static_cast<cudaStream_t>(stream));
}
size_t b200_nn_match_workspace(int B, int M, int N) { return nn_match_workspace_bytes(B, M, N); }
int b200_nn_match(const float* sim, int B, int M, int N, float ratio_thresh_sq, float distance_thresh_sq, int mutual_check,
                  void* workspace, float* log_assignment, long long* matches0, long long* matches1, float* mscores0,
                  float* mscores1, b200_stream_t stream) {
    B200_REQUIRE(sim && workspace && log_assignment && matches0 && matches1 && mscores0 && mscores1,
                 "nn_match: null argument");
    return launch_nn_match(sim, B, M, N, ratio_thresh_sq, distance_thresh_sq, mutual_check, workspace, log_assignment,
                           matches0, matches1, mscores0, mscores1, static_cast<cudaStream_t>(stream));
}
size_t b200_sg_sinkhorn_workspace(int B, int M, int N) { return sinkhorn_workspace_bytes(B, M, N); }
int b200_sg_sinkhorn_assignment(const float* cost, const float* bin_score, int B, int M, int N, int iters,
                                float filter_threshold, void* workspace, float* scores, long long* matches0,
                                long long* matches1, float* mscores0, float* mscores1, b200_stream_t stream) {
    B200_REQUIRE(cost && bin_score && workspace && scores && matches0 && matches1 && mscores0 && mscores1,
                 "sinkhorn: null argument");
    return launch_sinkhorn_assignment(cost, bin_score, B, M, N, iters, filter_threshold, workspace, scores, matches0,
                                      matches1, mscores0, mscores1, static_cast<cudaStream_t>(stream));
}

int b200_lg_count_below(const float* conf, int n, float threshold, int* out, b200_stream_t stream) {
    return launch_count_below(conf, n, threshold, out, static_cast<cudaStream_t>(stream));
}
int b200_lg_prune(const float* match, const float* conf, int n0, int n1, float match_thr, float conf_thr,
                  int* pos_scratch, int* counts, const float* x_src, float* x_dst, const void* xm_src, void* xm_dst,
                  long long ld_xm, const void* cs_src, void* cs_dst, const int* ind_src, int* ind_dst,
                  long long* prune0, long long* prune1, b200_stream_t stream) {
    B200_REQUIRE(n0 >= 0 && n1 >= 0 && match && pos_scratch && counts, "prune: bad arguments");
    return launch_prune(match, conf, n0, n1, match_thr, conf_thr, pos_scratch, counts, x_src, x_dst, xm_src, xm_dst,
                        ld_xm, cs_src, cs_dst, ind_src, ind_dst, prune0, prune1, static_cast<cudaStream_t>(stream));
}
int b200_lg_scatter_matches(const long long* m0, const long long* m1, const float* ms0, const float* ms1,
                            const int* ind, int n0, int n1, long long* out_m0, long long* out_m1, float* out_ms0,
                            float* out_ms1, b200_stream_t stream) {
    return launch_scatter_matches(m0, m1, ms0, ms1, ind, n0, n1, out_m0, out_m1, out_ms0, out_ms1,
                                  static_cast<cudaStream_t>(stream));
}

int b200_lg_adaptive_decide(const float* conf, int B, int M, int N, float threshold, float depth_conf, int layer, int* state,
                            b200_stream_t stream) {
    return launch_adaptive_decide(conf, B, M, N, threshold, depth_conf, layer, state, static_cast<cudaStream_t>(stream));
}
int b200_lg_adaptive_prune(const float* match, const float* conf, int B, int M, int N, float match_thr, float conf_thr,
                           int* state, int* pos_scratch, int* new_counts, const float* x_src, float* x_dst,
                           const void* xm_src, void* xm_dst, long long ld_xm, const void* cs_src, void* cs_dst,
                           const int* ind_src, int* ind_dst, long long* prune0, long long* prune1, b200_stream_t stream) {
    return launch_adaptive_prune(match, conf, B, M, N, match_thr, conf_thr, state, pos_scratch, new_counts, x_src, x_dst,
                                 xm_src, xm_dst, ld_xm, cs_src, cs_dst, ind_src, ind_dst, prune0, prune1,
                                 static_cast<cudaStream_t>(stream));
}
int b200_lg_adaptive_commit(int* state, const int* new_counts, int B, int M, int N, int* lens, int* tile_live,
                            int* hist_next, b200_stream_t stream) {
    return launch_adaptive_commit(state, new_counts, B, M, N, lens, tile_live, hist_next, static_cast<cudaStream_t>(stream));
}
int b200_lg_adaptive_snapshot(int* state, int B, int M, int N, int final_pass, const float* md, float* md_out,
                              const float* z, float* z_out, const float* x, float* x_out, const int* ind, int* ind_out,
                              int* final_len, b200_stream_t stream) {
    return launch_adaptive_snapshot(state, B, M, N, final_pass, md, md_out, z, z_out, x, x_out, ind, ind_out, final_len,
                                    static_cast<cudaStream_t>(stream));
}
int b200_lg_assignment_live(const float* sim, const float* z0, const float* z1, int B, int M, int N, const int* live,
                            float filter_threshold, void* workspace, float* scores, long long* matches0,
                            long long* matches1, float* mscores0, float* mscores1, b200_stream_t stream) {
    B200_REQUIRE(sim && z0 && z1 && workspace && scores && matches0 && matches1 && mscores0 && mscores1,
                 "assignment_live: null argument");
    return launch_assignment_live(sim, z0, z1, B, M, N, live, filter_threshold, workspace, scores, matches0, matches1,
                                  mscores0, mscores1, static_cast<cudaStream_t>(stream));
}
int b200_lg_adaptive_scatter(const long long* m0, const long long* m1, const float* ms0, const float* ms1, const int* ind,
                             const int* final_len, int B, int M, int N, long long* out_m0, long long* out_m1,
                             float* out_ms0, float* out_ms1, b200_stream_t stream) {
    return launch_adaptive_scatter(m0, m1, ms0, ms1, ind, final_len, B, M, N, out_m0, out_m1, out_ms0, out_ms1,
                                   static_cast<cudaStream_t>(stream));
}

int b200_prep_resize(const void* src, int src_kind, int B, int C, int H, int W, int out_h, int out_w, const float* taps_x,
                     int kx, const float* taps_y, int ky, float* tmp, float* out, int pad_h, int pad_w,
                     unsigned char* mask, b200_stream_t stream) {
    const int num_sms = sm_count();
    B200_REQUIRE(num_sms > 0, "prep_resize: no CUDA device available");
    return launch_prep_resize(src, src_kind, B, C, H, W, out_h, out_w, taps_x, kx, taps_y, ky, tmp, out, pad_h, pad_w, mask,
                              num_sms, static_cast<cudaStream_t>(stream));
}
int b200_export_pack(const b200_export_field* fields, int num_fields, int B, int as_half, void* records,
                     long long record_bytes, b200_stream_t stream) {
    return launch_export_pack(fields, num_fields, B, as_half, records, record_bytes, static_cast<cudaStream_t>(stream));
}
int b200_eval_matches_homography(const float* keypoints0, const float* keypoints1, const long long* matches0, int B, int M,
                                 int N, const float* H_0to1, float* out, b200_stream_t stream) {
    return launch_eval_homography(keypoints0, keypoints1, matches0, B, M, N, H_0to1, out, static_cast<cudaStream_t>(stream));
}
int b200_eval_matches_epipolar(const float* keypoints0, const float* keypoints1, const long long* matches0, int B, int M,
                               int N, const float* cam0, const float* cam1, const float* T_0to1, float* out,
                               b200_stream_t stream) {
    return launch_eval_epipolar(keypoints0, keypoints1, matches0, B, M, N, cam0, cam1, T_0to1, out,
                                static_cast<cudaStream_t>(stream));
}
int b200_eval_homography_dlt(const float* keypoints0, const float* keypoints1, const long long* matches0,
                             const float* mscores0, int B, int M, int N, const float* H_gt, const float* image_size,
                             float* H_out, float* err_out, b200_stream_t stream) {
    return launch_eval_dlt(keypoints0, keypoints1, matches0, mscores0, B, M, N, H_gt, image_size, H_out, err_out,
                           static_cast<cudaStream_t>(stream));
}

}  // extern "C"
