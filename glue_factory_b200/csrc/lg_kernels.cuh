#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace b200 {
int launch_ingest(const float* src, long long rows, float* x32, void* xm, long long ld_xm, cudaStream_t stream);
int launch_posenc(const float* kpts, const float* size, int B, int n, const float* wr, float* size_scratch,
                  void* cs, cudaStream_t stream);
int launch_rowdot(const float* x, long long rows, const float* w, const float* bias, int apply_sigmoid, float* out,
                  cudaStream_t stream);
size_t assignment_workspace_bytes(int B, int M, int N);
int launch_assignment(const float* sim, const float* z0, const float* z1, int B, int M, int N, float threshold,
                      void* workspace, float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                      cudaStream_t stream);
int launch_assignment_live(const float* sim, const float* z0, const float* z1, int B, int M, int N, const int* live,
                           float threshold, void* workspace, float* scores, long long* m0, long long* m1, float* ms0,
                           float* ms1, cudaStream_t stream);
int launch_assignment_finalize(const float* sim, float simw, int B, int M, int N, float threshold, void* workspace,
                               float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                               cudaStream_t stream, const int* live = nullptr);
// adaptive depth / width for a batch of pairs, state on the device (lg_adaptive_batched.cu)
int launch_adaptive_decide(const float* conf, int B, int M, int N, float thr, float depth_conf, int layer, int* st,
                           cudaStream_t stream);
int launch_adaptive_prune(const float* match, const float* conf, int B, int M, int N, float match_thr, float conf_thr,
                          int* st, int* pos, int* newc, const float* x_src, float* x_dst, const void* xm_src, void* xm_dst,
                          long long ld_xm, const void* cs_src, void* cs_dst, const int* ind_src, int* ind_dst,
                          long long* prune0, long long* prune1, cudaStream_t stream);
int launch_adaptive_commit(int* st, const int* newc, int B, int M, int N, int* lens, int* tile_live, int* hist_next,
                           cudaStream_t stream);
int launch_adaptive_snapshot(int* st, int B, int M, int N, int final_pass, const float* md, float* md_out, const float* z,
                             float* z_out, const float* x, float* x_out, const int* ind, int* ind_out, int* final_len,
                             cudaStream_t stream);
int launch_adaptive_scatter(const long long* m0, const long long* m1, const float* ms0, const float* ms1, const int* ind,
                            const int* final_len, int B, int M, int N, long long* out_m0, long long* out_m1, float* out_ms0,
                            float* out_ms1, cudaStream_t stream);
size_t assignment_static_capacity();
int launch_assignment_static(const float* sim, const float* z0, const float* z1, int B, int M, int N, float threshold,
                             float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                             cudaStream_t stream);
size_t nn_match_workspace_bytes(int B, int M, int N);
int launch_nn_match(const float* sim, int B, int M, int N, float ratio_sq, float dist_sq, int mutual, void* workspace,
                    float* scores, long long* m0, long long* m1, float* ms0, float* ms1, cudaStream_t stream);
void assignment_terms(void* workspace, int B, int M, int N, float** rowterm, float** colterm);
// SuperGlue: log-domain Sinkhorn + mutual-NN filter (sg_sinkhorn.cu)
size_t sinkhorn_workspace_bytes(int B, int M, int N);
int launch_sinkhorn_assignment(const float* cost, const float* bin_score, int B, int M, int N, int iters,
                               float threshold, void* workspace, float* scores, long long* m0, long long* m1,
                               float* ms0, float* ms1, cudaStream_t stream);
int launch_sg_kenc_head(const float* kpts, const float* scores, const float* size, int B, int n, const float* w1,
                        const float* b1, const float* w2, const float* b2, void* out, cudaStream_t stream);
int launch_split_hilo(const float* x, long long rows, int cols, long long split_row, void* out, cudaStream_t stream);
int launch_sg_split_hilo(const float* x, long long rows, long long split_row, void* out, cudaStream_t stream);
int launch_count_below(const float* conf, int n, float thr, int* out, cudaStream_t stream);
int launch_prune(const float* match, const float* conf, int n0, int n1, float match_thr, float conf_thr, int* pos,
                 int* counts, const float* x_src, float* x_dst, const void* xm_src, void* xm_dst, long long ld_xm,
                 const void* cs_src, void* cs_dst, const int* ind_src, int* ind_dst, long long* prune0,
                 long long* prune1, cudaStream_t stream);
int launch_scatter_matches(const long long* m0, const long long* m1, const float* ms0, const float* ms1, const int* ind,
                           int n0, int n1, long long* out_m0, long long* out_m1, float* out_ms0, float* out_ms1,
                           cudaStream_t stream);
}  // namespace b200
