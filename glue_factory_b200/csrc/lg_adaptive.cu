// Bandwidth-bound LightGlue kernels, part 3: adaptive depth (token confidence count) and adaptive
// width (point pruning by ordered stream compaction), plus the scatter of matches back to the
// original keypoint indices.  Batch size 1, as the reference asserts.
// (reference: gluefactory/models/matchers/lightglue.py:470-511,532-552)
#include "common.cuh"
#include "lg_kernels.cuh"

namespace b200 {

// number of tokens whose confidence is below the layer threshold (check_if_stop, :541-552)
__global__ void __launch_bounds__(256) count_below_kernel(const float* __restrict__ conf, int n, float thr,
                                                          int* __restrict__ out) {
    int c = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) c += conf[i] < thr;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// get_pruning_mask (:532-539) + ordered compaction map.  One CTA per image (blockIdx.x = image).
// keep = matchability > match_thr  OR  (conf given AND conf <= conf_thr).  pos[t] = new row or -1.
__global__ void __launch_bounds__(1024) prune_scan_kernel(const float* __restrict__ match, const float* __restrict__ conf,
                                                          int n0, int n1, float match_thr, float conf_thr,
                                                          int* __restrict__ pos, int* __restrict__ counts) {
    __shared__ int warp_tot[32];
    __shared__ int running;
    const int img = blockIdx.x, n = img == 0 ? n0 : n1, off = img == 0 ? 0 : n0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int t = base + threadIdx.x;
        bool keep = false;
        if (t < n) {
            keep = match[off + t] > match_thr;
            if (conf != nullptr) keep = keep || (conf[off + t] <= conf_thr);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_tot[warp] = __popc(bal);
        __syncthreads();
        int before = running;
        for (int w = 0; w < warp; ++w) before += warp_tot[w];
        if (t < n) pos[off + t] = keep ? before + __popc(bal & ((1u << lane) - 1)) : -1;
        __syncthreads();
        if (threadIdx.x == 0) {
            int s = running;
            for (int w = 0; w < 32; ++w) s += warp_tot[w];
            running = s;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) counts[img] = running;
}

// move the kept tokens (one warp per token): fp32 residual row, bf16 copy (left half of XM), rotary
// table row and original-index entry; image 1 is packed right behind image 0's kept rows.
__global__ void __launch_bounds__(256) prune_gather_kernel(const int* __restrict__ pos, const int* __restrict__ counts,
                                                           int n0, int n1, const float* __restrict__ x_src,
                                                           float* __restrict__ x_dst, const __nv_bfloat16* __restrict__ xm_src,
                                                           __nv_bfloat16* __restrict__ xm_dst, long long ld_xm,
                                                           const float2* __restrict__ cs_src, float2* __restrict__ cs_dst,
                                                           const int* __restrict__ ind_src, int* __restrict__ ind_dst,
                                                           long long* __restrict__ prune0, long long* __restrict__ prune1) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (t >= n0 + n1) return;
    const int p = pos[t];
    if (p < 0) return;
    const bool second = t >= n0;
    const long long d = second ? counts[0] + p : p;
    const float4* xs = reinterpret_cast<const float4*>(x_src + (long long)t * 256) + lane * 2;
    float4* xd = reinterpret_cast<float4*>(x_dst + d * 256) + lane * 2;
    xd[0] = xs[0];
    xd[1] = xs[1];
    *(reinterpret_cast<uint4*>(xm_dst + d * ld_xm) + lane) = *(reinterpret_cast<const uint4*>(xm_src + (long long)t * ld_xm) + lane);
    cs_dst[d * 32 + lane] = cs_src[(long long)t * 32 + lane];
    if (lane == 0) {
        const int orig = ind_src[t];
        ind_dst[d] = orig;
        if (second) prune1[orig] += 1; else prune0[orig] += 1;
    }
}

// scatter matches of the pruned problem back to original indices (:499-509)
__global__ void __launch_bounds__(256) scatter_matches_kernel(const long long* __restrict__ m0, const long long* __restrict__ m1,
                                                              const float* __restrict__ ms0, const float* __restrict__ ms1,
                                                              const int* __restrict__ ind, int n0, int n1,
                                                              long long* __restrict__ out_m0, long long* __restrict__ out_m1,
                                                              float* __restrict__ out_ms0, float* __restrict__ out_ms1) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int* ind0 = ind;
    const int* ind1 = ind + n0;
    if (t < n0) {
        const long long j = m0[t];
        out_m0[ind0[t]] = j < 0 ? -1 : ind1[j];
        out_ms0[ind0[t]] = ms0[t];
    } else if (t < n0 + n1) {
        const int u = t - n0;
        const long long i = m1[u];
        out_m1[ind1[u]] = i < 0 ? -1 : ind0[i];
        out_ms1[ind1[u]] = ms1[u];
    }
}

int launch_count_below(const float* conf, int n, float thr, int* out, cudaStream_t stream) {
    B200_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(int), stream));
    if (n > 0) count_below_kernel<<<(n + 255) / 256 < 64 ? (n + 255) / 256 : 64, 256, 0, stream>>>(conf, n, thr, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_prune(const float* match, const float* conf, int n0, int n1, float match_thr, float conf_thr, int* pos,
                 int* counts, const float* x_src, float* x_dst, const void* xm_src, void* xm_dst, long long ld_xm,
                 const void* cs_src, void* cs_dst, const int* ind_src, int* ind_dst, long long* prune0,
                 long long* prune1, cudaStream_t stream) {
    prune_scan_kernel<<<2, 1024, 0, stream>>>(match, conf, n0, n1, match_thr, conf_thr, pos, counts);
    const int total = n0 + n1;
    if (total > 0)
        prune_gather_kernel<<<(total * 32 + 255) / 256, 256, 0, stream>>>(
            pos, counts, n0, n1, x_src, x_dst, reinterpret_cast<const __nv_bfloat16*>(xm_src),
            reinterpret_cast<__nv_bfloat16*>(xm_dst), ld_xm, reinterpret_cast<const float2*>(cs_src),
            reinterpret_cast<float2*>(cs_dst), ind_src, ind_dst, prune0, prune1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_scatter_matches(const long long* m0, const long long* m1, const float* ms0, const float* ms1, const int* ind,
                           int n0, int n1, long long* out_m0, long long* out_m1, float* out_ms0, float* out_ms1,
                           cudaStream_t stream) {
    const int total = n0 + n1;
    if (total > 0)
        scatter_matches_kernel<<<(total + 255) / 256, 256, 0, stream>>>(m0, m1, ms0, ms1, ind, n0, n1, out_m0, out_m1,
                                                                         out_ms0, out_ms1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
