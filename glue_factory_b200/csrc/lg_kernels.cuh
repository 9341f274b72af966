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
}  // namespace b200
