/* b200glue -- C ABI of the B200 (sm_100a) kernels behind Glue Factory's SuperPoint-open + LightGlue
 * forward path.  Plain pointers and sizes only; no torch / pybind types.
 *
 * Conventions:
 *  - every entry returns 0 on success, a negative code on failure; b200_last_error() gives the
 *    thread-local message;
 *  - all pointers are DEVICE pointers unless stated; outputs and workspaces are allocated by the caller;
 *  - entries never allocate device memory, never synchronise, and enqueue on the given stream only.
 *
 * Each entry names the reference code it replaces (paths relative to the reference repository).
 */
#ifndef B200GLUE_H_
#define B200GLUE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 9

typedef void* b200_stream_t; /* cudaStream_t */

int b200_abi_version(void);
const char* b200_last_error(void);
/* number of SMs of the current device (persistent kernels size their grid with it) */
int b200_sm_count(int* out);

/* ---- epilogue kinds of the dense tcgen05 GEMM ------------------------------------------------ */
enum {
    B200_EPI_BIAS_BF16 = 0,      /* nn.Linear: out16 = A.W^T + bias                                   */
    B200_EPI_QKV_ROTARY = 1,     /* SelfBlock Wqkv + rotary on q,k      (lightglue.py:36-43,144-154)  */
    B200_EPI_LN_GELU = 2,        /* ffn[0..2]: Linear -> LayerNorm -> GELU (lightglue.py:138-143)     */
    B200_EPI_RESID = 3,          /* ffn[3] + residual add               (lightglue.py:157,213-214)    */
    B200_EPI_F32 = 4,            /* MatchAssignment similarity einsum   (lightglue.py:276)            */
    B200_EPI_CONV_RELU_BN = 5,   /* VGGBlock conv -> ReLU -> BN (+ MaxPool2d) (superpoint_open.py:59-105) */
    B200_EPI_CONV_BN_L2 = 6,     /* descriptor head 1x1 + BN + F.normalize (superpoint_open.py:129-131) */
    B200_EPI_CONV_BN_SOFTMAX_D2S = 7 /* detector head 1x1 + BN + softmax + depth-to-space (:134-140)  */
};

/* C[batch][M,N] = epilogue(A[batch][M,K] . B[batch][N,K]^T), bf16 operands (row-major, K contiguous). */
typedef struct {
    /* operands */
    const void* a;      /* bf16, rows x K, leading dimension lda (elements)                             */
    long long a_rows;   /* total rows addressable through `a` (all batches)                            */
    long long lda;
    const void* b;      /* bf16, rows x K, leading dimension ldb                                        */
    long long b_rows;
    long long ldb;
    int M, N, K, batch;
    int a_batch_rows, b_batch_rows; /* row offset between consecutive batches in a / b                */
    int block_n;        /* N tile: 64, 80, 128, 256 or 512 (must match an instantiated kernel)         */
    int epilogue;       /* B200_EPI_*                                                                   */
    /* implicit-GEMM 3x3 convolution: set conv=1, `a` is NHWC [batch,H,W,Cin], K = 9*Cin               */
    int conv, H, W, Cin;
    /* epilogue operands (device pointers, fp32 unless noted; unused ones NULL)                        */
    const float* bias;
    const float* scale;
    const float* shift;
    int relu, pool;
    float alpha;
    const float* resid;     /* [M, ld32]                                                               */
    const void* cossin;     /* float2 [M, 32] (cos, sin) per row and rotary frequency                  */
    int rot_cols;
    const float* ln_gamma;
    const float* ln_beta;
    float ln_eps;
    void* out16;            /* bf16 output                                                             */
    long long ld16;
    float* out32;           /* fp32 output                                                             */
    long long ld32;
    long long out32_batch_stride;
    int cells_h, cells_w;   /* D2S epilogue: rows are (image, cell_y, cell_x)                          */
    const int* tile_live;   /* optional DEVICE array [batch * ceil(M/128)]: 128-row tiles marked 0 are skipped
                               (adaptive LightGlue: pruned tokens, pairs that left the layer loop); NULL = all */
} b200_gemm_desc;

/* nn.Linear / nn.Conv2d / einsum on the path -> tcgen05 GEMM (dense path). */
int b200_gemm(const b200_gemm_desc* d, b200_stream_t stream);

/* One attention problem (see b200_attention). */
typedef struct {
    int q_row0, q_len, kv_row0, kv_len;
    int q_len_slot, kv_len_slot; /* indices into len_override when it is non-NULL                      */
} b200_attn_seq;

/* softmax(scale * Q K^T) V for head_dim 64, heads laid out as 64-column slices of token-major bf16
 * matrices.  Replaces Attention.forward / the CrossBlock einsum path (lightglue.py:105-122,186-215). */
typedef struct {
    const void* q; long long q_rows, ldq; int q_col0;
    const void* k; long long k_rows, ldk; int k_col0;
    const void* v; long long v_rows, ldv; int v_col0;
    const b200_attn_seq* seqs; /* DEVICE array [num_seqs]                                              */
    int num_seqs, max_q_len, heads;
    const int* len_override;   /* optional DEVICE array of live sequence lengths (adaptive pruning)     */
    float scale;
    void* out; long long ld_out; int out_col0; /* bf16 [tokens, ld_out]                                */
} b200_attn_desc;

int b200_attention(const b200_attn_desc* d, b200_stream_t stream);

/* ---- SuperPoint-open bandwidth kernels (superpoint_open.py) ---------------------------------- */
/* first VGG block, C_in = 1 (or RGB -> gray first, :125-127): conv3x3 + bias + ReLU + BN, NHWC bf16 out */
int b200_sp_conv1a(const float* image, int B, int C, int H, int W, const float* w /*[64,9]*/,
                   const float* bias, const float* scale, const float* shift, void* out_nhwc,
                   b200_stream_t stream);
/* the whole first VGG block in one kernel (:59-75,96-105 of superpoint_open.py; conv1a / conv1b of
 * gluefactory_nonfree/superpoint.py:166-167,276-277): conv3x3 1 -> 64 (+ReLU+BN) and conv3x3 64 -> 64 (+ReLU+BN),
 * optional 2x2 max pool; the intermediate activation stays on chip.  w2 is the second convolution packed
 * as bf16 [64, 9*64] (tap-major, then input channel), as b200_gemm's conv mode takes it.  host_params is
 * a HOST array float[6][64] = bias1 | scale1 | shift1 | bias2 | scale2 | shift2: the per-channel operands travel
 * in the kernel parameters (constant bank), so the epilogues issue no loads for them. */
int b200_sp_conv_stem(const float* image, int B, int C, int H, int W, const float* w1 /*[64,9] device*/,
                      const float* host_params /*[6][64] host*/, const void* w2, int pool, void* out_nhwc,
                      b200_stream_t stream);
/* the two entries above on uint8 images [B,C,H,W] (pixels as read from disk): the conversion value / 255 of
 * numpy_image_to_torch (gluefactory/utils/image.py:112-121) is done in the load, bit-identical to the host conversion,
 * so a loader ships a quarter of the bytes */
int b200_sp_conv1a_u8(const unsigned char* image, int B, int C, int H, int W, const float* w, const float* bias,
                      const float* scale, const float* shift, void* out_nhwc, b200_stream_t stream);
int b200_sp_conv_stem_u8(const unsigned char* image, int B, int C, int H, int W, const float* w1,
                         const float* host_params, const void* w2, int pool, void* out_nhwc, b200_stream_t stream);
/* batched_nms (:34-49) + border removal (:147-152); out has the shape of scores */
int b200_sp_nms_border(const float* scores, int B, int H, int W, int nms_radius, int remove_borders,
                       float* out, b200_stream_t stream);
/* Fast path of the two entries above and below for nms_radius 1..4: the NMS pass also appends every
 * surviving positive maximum to a per-image list candidates[B][capacity] (64-bit records: monotone score
 * key << 32 | ~linear index), counted in candidate_state[0..B) (candidate_state has 2*B ints and is
 * zeroed by the call); b200_sp_select_topk_list then selects from the list, and re-does from the full map
 * only the images whose list cannot decide the result (overflow, or zero-score pixels among the top k). */
/* image_hw: optional DEVICE array [B][2] = (h, w) of the valid image region: the bottom / right border is
 * then taken from it per image (gluefactory_nonfree/superpoint.py:244-254); NULL = the full map. */
int b200_sp_nms_border_list(const float* scores, int B, int H, int W, int nms_radius, int remove_borders,
                            const int* image_hw, float* out, unsigned long long* candidates, int* candidate_state,
                            int capacity, b200_stream_t stream);
int b200_sp_select_topk_list(const float* nms_map, int B, int H, int W, float threshold, int k,
                             const unsigned long long* candidates, int* candidate_state, int capacity,
                             float* keypoints, float* scores, int* indices, int* counts, b200_stream_t stream);
/* number of pixels above the detection threshold per image (:155-160) */
int b200_sp_count_candidates(const float* nms_map, int B, int H, int W, float threshold, int* counts,
                             b200_stream_t stream);
/* threshold + select_top_k_keypoints (:52-56,155-179): keypoints (x+0.5, y+0.5), scores, linear
 * indices, per-image count = min(candidates, k).  Sorted by (score desc, index asc) when more than k
 * candidates exist, row-major order otherwise. */
int b200_sp_select_topk(const float* nms_map, int B, int H, int W, float threshold, int k,
                        float* keypoints /*[B,k,2]*/, float* scores /*[B,k]*/, int* indices /*[B,k]*/,
                        int* counts /*[B]*/, b200_stream_t stream);
/* sample_descriptors (:20-31): dense is NHWC fp32 [B,Hc,Wc,256]; out [B,K,256] */
int b200_sp_sample_descriptors(const float* dense, int B, int Hc, int Wc, int C, const float* keypoints,
                               int K, int stride, float* out, b200_stream_t stream);

/* same with the grid convention selectable: mode 0 = SuperPoint-open (above), 1 = gluefactory_nonfree
 * sample_descriptors_fix_sampling (superpoint.py:138-149), 2 = its legacy sampling, align_corners=True
 * (superpoint.py:117-133).  keypoints always hold (x + 0.5, y + 0.5). */
int b200_sp_sample_descriptors_mode(const float* dense, int B, int Hc, int Wc, int C, const float* keypoints,
                                    int K, int stride, int mode, float* out, b200_stream_t stream);

/* ---- LightGlue bandwidth kernels (lightglue.py) ---------------------------------------------- */
/* fp32 descriptors -> fp32 residual stream x32 [rows,256] and bf16 copy xm[:, :256] (ld_xm elements) */
int b200_lg_ingest(const float* desc, long long rows, float* x32, void* xm, long long ld_xm,
                   b200_stream_t stream);
/* normalize_keypoints (:21-33) + LearnableFourierPositionalEncoding (:46-59).  image_size [B,2] (w,h)
 * or NULL (extent based; size_scratch [B,2] is then used).  out: float2 [B*n, 32] = (cos, sin). */
int b200_lg_posenc(const float* keypoints, const float* image_size, int B, int n, const float* wr /*[32,2]*/,
                   float* size_scratch, void* cossin, b200_stream_t stream);
/* matchability / token-confidence logits (:62-73,283-284): out[row] = x[row,:256].w + bias, optional sigmoid */
int b200_lg_rowdot(const float* x32, long long rows, const float* w, const float* bias, int apply_sigmoid,
                   float* out, b200_stream_t stream);
/* sigmoid_log_double_softmax (:250-262) + filter_matches (:287-303) */
size_t b200_lg_assignment_workspace(int B, int M, int N);
int b200_lg_assignment(const float* sim /*[B,M,N]*/, const float* z0 /*[B,M]*/, const float* z1 /*[B,N]*/,
                       int B, int M, int N, float filter_threshold, void* workspace,
                       float* scores /*[B,M+1,N+1]*/, long long* matches0, long long* matches1,
                       float* mscores0, float* mscores1, b200_stream_t stream);

/* Allocation-free variant of b200_lg_assignment.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results as b200_lg_assignment, but the scratch lives in a
 * file-scope device array of B200_LG_ASSIGNMENT_STATIC_MAX_BYTES inside the library (held by every GPU that loads
 * it, until exit), so the call performs no cudaMalloc / cudaFree and takes no workspace.  Problems with
 * b200_lg_assignment_workspace(B, M, N) > B200_LG_ASSIGNMENT_STATIC_MAX_BYTES are rejected (-1, message in
 * b200_last_error()); 4 MiB covers e.g. B = 4 at 2048 x 2048 or B = 1 at 4096 x 4096 keypoints.  Every call on a
 * device shares that one array: calls that overlap on the GPU (different streams) race; serialise them.
 * b200_lg_assignment_static_capacity() returns the limit in bytes. */
#define B200_LG_ASSIGNMENT_STATIC_MAX_BYTES (4u << 20)
size_t b200_lg_assignment_static_capacity(void);
int b200_lg_assignment_static(const float* sim /*[B,M,N]*/, const float* z0 /*[B,M]*/, const float* z1 /*[B,N]*/,
                              int B, int M, int N, float filter_threshold, float* scores /*[B,M+1,N+1]*/,
                              long long* matches0, long long* matches1, float* mscores0, float* mscores1,
                              b200_stream_t stream);

/* ---- nearest-neighbour matcher (gluefactory/models/matchers/nearest_neighbor_matcher.py) ------ */
/* fp32 [rows,cols] -> bf16 [rows,3*cols] two-term split feeding a K = 3*cols similarity GEMM with near-fp32
 * accuracy (the einsum at :52): rows < split_row: [hi|hi|lo] (A side), the others: [hi|lo|hi] (B side). */
int b200_split_hilo(const float* x, long long rows, int cols, long long split_row, void* out, b200_stream_t stream);
/* find_nn (:14-25) in both directions, mutual_check (:28-36), matching scores and log_assignment (:59-64) from the
 * similarity sim [B,M,N]: nearest neighbour by largest similarity (lowest index on ties), squared distance
 * 2 (1 - sim); the ratio test (d0 <= ratio_thresh_sq * d1) and the distance test (d0 <= distance_thresh_sq) are
 * disabled when the threshold is <= 0.  log_assignment [B,M+1,N+1] = log_softmax over rows + over columns, last
 * row / column zero; matches int64, -1 = unmatched; mscores 1.0 for a match, else 0.0. */
size_t b200_nn_match_workspace(int B, int M, int N);
int b200_nn_match(const float* sim, int B, int M, int N, float ratio_thresh_sq, float distance_thresh_sq,
                  int mutual_check, void* workspace, float* log_assignment, long long* matches0, long long* matches1,
                  float* mscores0, float* mscores1, b200_stream_t stream);


/* ---- SuperGlue configuration (gluefactory_nonfree/superglue.py) ------------------------------- */
/* normalize_keypoints (:75-86) + the first two KeypointEncoder layers (:89-102) with folded BatchNorm1d:
 * relu(W2 . relu(W1 . (x, y, score) + b1) + b2) -> bf16 [B*n, 64].  image_size [B,2] = (w, h). */
int b200_sg_keypoint_encoder_head(const float* keypoints, const float* scores, const float* image_size, int B, int n,
                                  const float* w1 /*[32,3]*/, const float* b1, const float* w2 /*[64,32]*/,
                                  const float* b2, void* out, b200_stream_t stream);
/* fp32 [rows,256] -> bf16 [rows,768] two-term split feeding a K = 768 similarity GEMM with near-fp32 accuracy:
 * rows < split_row: [hi|hi|lo] (A side), the others: [hi|lo|hi] (B side).  Used for the score einsum (:284-285). */
int b200_sg_split_hilo(const float* x, long long rows, long long split_row, void* out, b200_stream_t stream);
/* log_optimal_transport (:173-201) on cost [B,M,N] with the dustbin score *bin_score (device scalar),
 * `iters` Sinkhorn iterations in one persistent cooperative kernel, then the mutual-NN filter (:288-298).
 * scores [B,M+1,N+1]; matches int64, -1 = unmatched.  Precondition: |cost| < ~60 (fp32 exp range). */
size_t b200_sg_sinkhorn_workspace(int B, int M, int N);
int b200_sg_sinkhorn_assignment(const float* cost, const float* bin_score, int B, int M, int N, int iters,
                                float filter_threshold, void* workspace, float* scores, long long* matches0,
                                long long* matches1, float* mscores0, float* mscores1, b200_stream_t stream);

/* ---- adaptive depth / width (batch size 1; lightglue.py:470-511,527-552) ---------------------- */
/* check_if_stop: *out = number of entries of conf[0..n) below `threshold` (out is zeroed first) */
int b200_lg_count_below(const float* conf, int n, float threshold, int* out, b200_stream_t stream);
/* get_pruning_mask + index_select compaction: keep = match > match_thr OR (conf != NULL AND conf <= conf_thr).
 * Tokens are rows [0,n0) (image 0) and [n0,n0+n1) (image 1) of x_src (fp32 [.,256]), xm_src (bf16, ld_xm),
 * cs_src (float2 [.,32]) and ind_src (original indices).  Kept rows are packed, order preserved, image 1
 * right behind image 0, into the *_dst buffers; counts[0..1] = kept per image; prune{0,1}[orig] += 1. */
int b200_lg_prune(const float* match, const float* conf, int n0, int n1, float match_thr, float conf_thr,
                  int* pos_scratch /*[n0+n1]*/, int* counts /*[2]*/, const float* x_src, float* x_dst,
                  const void* xm_src, void* xm_dst, long long ld_xm, const void* cs_src, void* cs_dst,
                  const int* ind_src, int* ind_dst, long long* prune0, long long* prune1, b200_stream_t stream);
/* scatter matches of the pruned problem back to original keypoint indices (lightglue.py:499-509);
 * the outputs must be pre-filled with -1 / 0 */
int b200_lg_scatter_matches(const long long* m0, const long long* m1, const float* ms0, const float* ms1,
                            const int* ind /*[n0+n1]*/, int n0, int n1, long long* out_m0, long long* out_m1,
                            float* out_ms0, float* out_ms1, b200_stream_t stream);

/* ---- adaptive depth / width for a BATCH of pairs, no host synchronisation (lightglue.py:470-511,527-552) -------
 * Token rows: image 0 of pair b = [b M, b M + M), image 1 = [B M + b N, ...); the live tokens of a segment are its
 * first m_b / n_b rows.  state: int32 [5 B] = m | n | active | exit_layer | just_exited (caller initialises
 * m = M, n = N, active = 1, exit_layer = n_layers - 1, just_exited = 0). */
/* check_if_stop (:541-552) per active pair on conf = sigmoid(token confidence) of every row: a pair whose ratio
 * 1 - below / (M + N) exceeds depth_conf gets active = 0, exit_layer = layer, just_exited = 1 */
int b200_lg_adaptive_decide(const float* conf, int B, int M, int N, float threshold, float depth_conf, int layer,
                            int* state, b200_stream_t stream);
/* get_pruning_mask (:532-539) + ordered compaction inside every active pair's segments (src -> dst buffers, as
 * b200_lg_prune), prune counters [B,M] / [B,N] += 1 for kept keypoints; new lengths go to new_counts [2 B]
 * (committed by b200_lg_adaptive_commit).  conf may be NULL (no depth confidence). */
int b200_lg_adaptive_prune(const float* match, const float* conf, int B, int M, int N, float match_thr, float conf_thr,
                           int* state, int* pos_scratch /*[B (M+N)]*/, int* new_counts, const float* x_src, float* x_dst,
                           const void* xm_src, void* xm_dst, long long ld_xm, const void* cs_src, void* cs_dst,
                           const int* ind_src, int* ind_dst, long long* prune0, long long* prune1, b200_stream_t stream);
/* commit new_counts (NULL: keep the lengths) and refresh lens [2 B] (live lengths, 0 for inactive pairs: the
 * len_override of b200_attention), tile_live [ceil(B (M+N) / 128)] (b200_gemm_desc.tile_live) and, if not NULL,
 * hist_next [2 B] (lengths entering the next layer, for the executed-FLOPs report) */
int b200_lg_adaptive_commit(int* state, const int* new_counts, int B, int M, int N, int* lens, int* tile_live,
                            int* hist_next, b200_stream_t stream);
/* keep what the final assignment needs of the pairs that just left the loop (final_pass = 0; clears just_exited) or
 * that are still active after the last layer (final_pass = 1): projected descriptors md, matchability logits z,
 * residual stream x, index map ind -> *_out (same row layout), live lengths -> final_len [2 B] */
int b200_lg_adaptive_snapshot(int* state, int B, int M, int N, int final_pass, const float* md, float* md_out,
                              const float* z, float* z_out, const float* x, float* x_out, const int* ind, int* ind_out,
                              int* final_len, b200_stream_t stream);
/* b200_lg_assignment on padded [B,M,N] problems with DEVICE live sizes live[b] x live[B+b]; the dustbin column / row
 * are written at indices live[B+b] / live[b], so scores[b, :m+1, :n+1] is the reference result; matches beyond
 * the live block are -1 / 0.  The caller zero-fills `scores`. */
int b200_lg_assignment_live(const float* sim, const float* z0, const float* z1, int B, int M, int N, const int* live,
                            float filter_threshold, void* workspace, float* scores, long long* matches0,
                            long long* matches1, float* mscores0, float* mscores1, b200_stream_t stream);
/* b200_lg_scatter_matches for every pair at once (index map and lengths as snapshot left them) */
int b200_lg_adaptive_scatter(const long long* m0, const long long* m1, const float* ms0, const float* ms1,
                             const int* ind, const int* final_len, int B, int M, int N, long long* out_m0,
                             long long* out_m1, float* out_ms0, float* out_ms1, b200_stream_t stream);

/* ---- pre-processing before the path (gluefactory/utils/image.py:12-101, datasets/image_pairs.py:59-90) -------- */
/* ImagePreprocessor.__call__: resize (kornia.geometry.transform.resize = optional separable Gaussian anti-alias blur
 * with reflect border, then F.interpolate bilinear, align_corners=False), zero padding to pad_h x pad_w (square_pad,
 * :55-66) and the optional padding mask.  src_kind 0: float32 [B,C,H,W] in [0,1]; 1: uint8 [B,H,W,C] as cv2 reads
 * it, converted like numpy_image_to_torch (:112-121, value / 255.0).  taps_x / taps_y are HOST arrays with the kx / ky
 * (odd, <= 63) normalised Gaussian taps; kx = ky = 1 means no blur (taps may be NULL).  tmp: float32 [B,C,H,W]
 * scratch for the horizontal pass (blur only).  out: float32 [B,C,pad_h,pad_w]; mask: uint8 [B,pad_h,pad_w] or NULL. */
int b200_prep_resize(const void* src, int src_kind, int B, int C, int H, int W, int out_h, int out_w,
                     const float* taps_x, int kx, const float* taps_y, int ky, float* tmp, float* out, int pad_h,
                     int pad_w, unsigned char* mask, b200_stream_t stream);

/* ---- export step after the path (gluefactory/utils/export_predictions.py:40-77) ------------------------------ */
/* Packs a batch of predictions into fixed-size per-pair records on the device, so that ONE device-to-host copy per
 * batch replaces the reference's `.cpu()` per tensor per pair.  A field copies elems_per_pair elements of pair b from
 * src + b * elems_per_pair to records + b * record_bytes + dst_offset.  Kinds: F32 copy; F32_KEYPOINTS = (x, y)
 * pairs multiplied by 1 / scales[b] (the renormalisation at :44-50; scales is a DEVICE array [B,2]); I64 copy.
 * as_half converts the float fields to fp16 (:66-70).  `fields` is a HOST array. */
enum { B200_EXPORT_F32 = 0, B200_EXPORT_F32_KEYPOINTS = 1, B200_EXPORT_I64 = 2 };
#define B200_EXPORT_MAX_FIELDS 16
typedef struct {
    const void* src;
    const float* scales;
    long long elems_per_pair;
    long long dst_offset; /* bytes inside a record; multiple of 4 (8 for I64) */
    int kind;
    int reserved;
} b200_export_field;
int b200_export_pack(const b200_export_field* fields, int num_fields, int B, int as_half, void* records,
                     long long record_bytes, b200_stream_t stream);

/* ---- match metrics on the device (gluefactory/eval/utils.py) ------------------------------------------------- */
/* eval_matches_homography (:76-96, sym_homography_error geometry/homography.py:319-328): per pair
 * out[b] = {prec@1px, prec@3px, num_matches, num_keypoints}.  H_0to1: float32 [B,9] row-major. */
int b200_eval_matches_homography(const float* keypoints0 /*[B,M,2]*/, const float* keypoints1 /*[B,N,2]*/,
                                 const long long* matches0 /*[B,M]*/, int B, int M, int N, const float* H_0to1,
                                 float* out /*[B,4]*/, b200_stream_t stream);
/* eval_matches_epipolar (:40-73, generalized_epi_dist essential=True, geometry/epipolar.py:32-56,75-85) for pinhole
 * cameras: cam = (fx, fy, cx, cy) per pair, T_0to1 = R (row-major 9) | t (3).
 * out[b] = {epi_prec@1e-4, epi_prec@5e-4, epi_prec@1e-3, num_matches, num_keypoints} (NaN precisions without matches) */
int b200_eval_matches_epipolar(const float* keypoints0, const float* keypoints1, const long long* matches0, int B,
                               int M, int N, const float* cam0 /*[B,4]*/, const float* cam1 /*[B,4]*/,
                               const float* T_0to1 /*[B,12]*/, float* out /*[B,5]*/, b200_stream_t stream);
/* eval_homography_dlt (:176-196): kornia's find_homography_dlt on the matched points weighted by the matching
 * scores (point normalisation, A^T W A, eigenvector of the smallest eigenvalue), then homography_corner_error against
 * H_gt (geometry/homography.py:341-347).  H_out [B,9], err_out [B]; fewer than 4 matches: inf matrix, NaN error. */
int b200_eval_homography_dlt(const float* keypoints0, const float* keypoints1, const long long* matches0,
                             const float* mscores0 /*[B,M]*/, int B, int M, int N, const float* H_gt /*[B,9]*/,
                             const float* image_size /*[B,2] (w,h)*/, float* H_out, float* err_out,
                             b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GLUE_H_ */
