// Adaptive depth (early exit) and width (point pruning) for a BATCH of pairs without host synchronisation.
// (reference: gluefactory/models/matchers/lightglue.py:470-511,527-552, which asserts batch size 1 and reads the stop
//  decision and the kept indices on the host every layer; here every pair carries its state on the device)
//
// Token layout (fixed for the whole pass): image 0 of pair b owns rows [b M, b M + M), image 1 rows
// [B M + b N, B M + b N + N); the LIVE tokens of a segment are its first m_b / n_b rows (pruning compacts inside the
// segment, order preserved).  State (int32, B entries each, contiguous): m | n | active | exit_layer | just_exited.
//   lens[2 B]      live lengths fed to b200_attention through len_override (0 for pairs that left the loop)
//   tile_live[T/128] 1 when a 128-row GEMM tile holds a live token (b200_gemm skips the others)
//   hist[L][2 B]   live lengths entering every layer (for the executed-FLOPs report)
#include "common.cuh"
#include "lg_kernels.cuh"

namespace b200 {

struct AdaptiveState {
    int* m;
    int* n;
    int* active;
    int* exit_layer;
    int* just;
};

__host__ __device__ inline AdaptiveState state_of(int* st, int B) {
    AdaptiveState s;
    s.m = st; s.n = st + B; s.active = st + 2 * B; s.exit_layer = st + 3 * B; s.just = st + 4 * B;
    return s;
}

// check_if_stop (:541-552) per pair: stop when 1 - (#confidences below the threshold) / (M + N) > depth_confidence.
// conf holds sigmoid(token confidence) for every row; only the live rows of an active pair count.
__global__ void __launch_bounds__(256) adaptive_decide_kernel(const float* __restrict__ conf, int B, int M, int N, float thr,
                                                              float depth_conf, int layer, int* __restrict__ st) {
    __shared__ int part[8];
    const int b = blockIdx.x;
    const AdaptiveState s = state_of(st, B);
    if (!s.active[b]) return;
    const int m = s.m[b], n = s.n[b];
    const float* c0 = conf + (long long)b * M;
    const float* c1 = conf + (long long)B * M + (long long)b * N;
    int c = 0;
    for (int i = threadIdx.x; i < m; i += 256) c += c0[i] < thr;
    for (int i = threadIdx.x; i < n; i += 256) c += c1[i] < thr;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int below = 0;
        for (int w = 0; w < 8; ++w) below += part[w];
        // fp32 arithmetic of the reference: 1.0 - sum.float() / num_points, compared with the fp32 threshold
        const float ratio = 1.0f - __fdiv_rn((float)below, (float)(M + N));
        if (ratio > depth_conf) {
            s.active[b] = 0;
            s.exit_layer[b] = layer;
            s.just[b] = 1;
        }
    }
}

// get_pruning_mask (:532-539) + ordered compaction map; grid (2 images, B pairs), 1024 threads.
// keep = matchability > match_thr OR (conf given AND conf <= conf_thr); pos[row] = new position inside the segment, or -1
__global__ void __launch_bounds__(1024) adaptive_scan_kernel(const float* __restrict__ match, const float* __restrict__ conf,
                                                             int B, int M, int N, float match_thr, float conf_thr,
                                                             const int* __restrict__ st, int* __restrict__ pos,
                                                             int* __restrict__ newc /*[2B]*/) {
    __shared__ int warp_tot[32];
    __shared__ int running;
    const int img = blockIdx.x, b = blockIdx.y;
    const AdaptiveState s = state_of(const_cast<int*>(st), B);
    const int n = img == 0 ? s.m[b] : s.n[b];
    if (!s.active[b]) {
        if (threadIdx.x == 0) newc[img * B + b] = n;
        return;
    }
    const long long off = img == 0 ? (long long)b * M : (long long)B * M + (long long)b * N;
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
            int tot = running;
            for (int w = 0; w < 32; ++w) tot += warp_tot[w];
            running = tot;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) newc[img * B + b] = running;
}

// move the kept tokens of the active pairs (one warp per row): fp32 residual row, bf16 copy (left half of XM), rotary
// table row, original-index entry; prune counters of the kept keypoints += 1 (:485,493)
__global__ void __launch_bounds__(256) adaptive_gather_kernel(const int* __restrict__ pos, const int* __restrict__ st, int B,
                                                              int M, int N, const float* __restrict__ x_src,
                                                              float* __restrict__ x_dst, const __nv_bfloat16* __restrict__ xm_src,
                                                              __nv_bfloat16* __restrict__ xm_dst, long long ld_xm,
                                                              const float2* __restrict__ cs_src, float2* __restrict__ cs_dst,
                                                              const int* __restrict__ ind_src, int* __restrict__ ind_dst,
                                                              long long* __restrict__ prune0, long long* __restrict__ prune1) {
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long T = (long long)B * (M + N);
    if (row >= T) return;
    const bool second = row >= (long long)B * M;
    const int b = second ? (int)((row - (long long)B * M) / N) : (int)(row / M);
    const long long seg = second ? (long long)B * M + (long long)b * N : (long long)b * M;
    const AdaptiveState s = state_of(const_cast<int*>(st), B);
    if (!s.active[b]) return;
    const int t = (int)(row - seg);
    if (t >= (second ? s.n[b] : s.m[b])) return;
    const int p = pos[row];
    if (p < 0) return;
    const long long d = seg + p;
    const float4* xs = reinterpret_cast<const float4*>(x_src + row * 256) + lane * 2;
    float4* xd = reinterpret_cast<float4*>(x_dst + d * 256) + lane * 2;
    xd[0] = xs[0];
    xd[1] = xs[1];
    *(reinterpret_cast<uint4*>(xm_dst + d * ld_xm) + lane) = *(reinterpret_cast<const uint4*>(xm_src + row * ld_xm) + lane);
    cs_dst[d * 32 + lane] = cs_src[row * 32 + lane];
    if (lane == 0) {
        const int orig = ind_src[row];
        ind_dst[d] = orig;
        if (second) prune1[(long long)b * N + orig] += 1; else prune0[(long long)b * M + orig] += 1;
    }
}

// the rows of pairs that did NOT prune this layer (they left the loop) stay where they are: nothing to do, their
// final data was snapshot at exit.  Commit the new lengths and refresh lens / tile_live / hist for the next layer.
__global__ void __launch_bounds__(256) adaptive_commit_kernel(int* __restrict__ st, const int* __restrict__ newc, int B, int M,
                                                              int N, int* __restrict__ lens, int* __restrict__ tile_live,
                                                              int num_tiles, int* __restrict__ hist_next) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const AdaptiveState s = state_of(st, B);
    __shared__ int dummy;
    (void)dummy;
    // phase 1 (first B threads of block 0 would race with the tile pass of other blocks): every thread derives the
    // committed lengths itself from (st, newc), the last block-0 threads write them back
    auto len_of = [&](int img, int b) {
        const int cur = img == 0 ? s.m[b] : s.n[b];
        return newc != nullptr && s.active[b] ? newc[img * B + b] : cur;
    };
    if (t < num_tiles) {
        // live when any row of [128 t, 128 t + 128) is a live row of an active pair
        const long long r0 = (long long)t * 128;
        const long long T = (long long)B * (M + N);
        int live = 0;
        for (long long r = r0; r < r0 + 128 && r < T && !live;) {
            const bool second = r >= (long long)B * M;
            const int b = second ? (int)((r - (long long)B * M) / N) : (int)(r / M);
            const long long seg = second ? (long long)B * M + (long long)b * N : (long long)b * M;
            const int len = s.active[b] ? len_of(second ? 1 : 0, b) : 0;
            if (r - seg < len) live = 1;
            r = seg + (second ? N : M);  // next segment
        }
        tile_live[t] = live;
    }
    if (t < 2 * B) {
        const int img = t / B, b = t % B;
        const int len = s.active[b] ? len_of(img, b) : 0;
        lens[t] = len;
        if (hist_next != nullptr) hist_next[t] = len;
    }
}

// second half of the commit: write the new lengths into the state (separate launch: the kernel above reads the old
// state from every block)
__global__ void __launch_bounds__(256) adaptive_store_kernel(int* __restrict__ st, const int* __restrict__ newc, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const AdaptiveState s = state_of(st, B);
    if (s.active[b]) {
        s.m[b] = newc[b];
        s.n[b] = newc[B + b];
    }
}

// snapshot the data the final assignment needs for the pairs that leave the loop at this layer (just_exited), or,
// with `final` set, for the pairs still active after the last layer: projected descriptors, matchability logits,
// residual stream (ref_descriptors), index map, live lengths.  One warp per row.
__global__ void __launch_bounds__(256) adaptive_snapshot_kernel(int* __restrict__ st, int B, int M, int N, int final_pass,
                                                                const float* __restrict__ md, float* __restrict__ md_out,
                                                                const float* __restrict__ z, float* __restrict__ z_out,
                                                                const float* __restrict__ x, float* __restrict__ x_out,
                                                                const int* __restrict__ ind, int* __restrict__ ind_out,
                                                                int* __restrict__ final_len /*[2B]*/) {
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long T = (long long)B * (M + N);
    if (row >= T) return;
    const bool second = row >= (long long)B * M;
    const int b = second ? (int)((row - (long long)B * M) / N) : (int)(row / M);
    const long long seg = second ? (long long)B * M + (long long)b * N : (long long)b * M;
    const AdaptiveState s = state_of(st, B);
    const bool take = final_pass ? s.active[b] != 0 : s.just[b] != 0;
    if (!take) return;
    const int len = second ? s.n[b] : s.m[b];
    const int t = (int)(row - seg);
    if (t == 0 && lane == 0) final_len[(second ? B : 0) + b] = len;
    if (t >= len) return;
    const float4* a = reinterpret_cast<const float4*>(md + row * 256) + lane * 2;
    float4* ao = reinterpret_cast<float4*>(md_out + row * 256) + lane * 2;
    ao[0] = a[0];
    ao[1] = a[1];
    const float4* c = reinterpret_cast<const float4*>(x + row * 256) + lane * 2;
    float4* co = reinterpret_cast<float4*>(x_out + row * 256) + lane * 2;
    co[0] = c[0];
    co[1] = c[1];
    if (lane == 0) {
        z_out[row] = z[row];
        ind_out[row] = ind[row];
    }
}

__global__ void __launch_bounds__(256) adaptive_clear_just_kernel(int* __restrict__ st, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) state_of(st, B).just[b] = 0;
}

// scatter the matches of the compacted problems back to original keypoint indices (:499-509), all pairs at once;
// outputs pre-filled with -1 / 0.  m0 / ms0 are [B, M] in compacted index space (rows >= live length ignored).
__global__ void __launch_bounds__(256) adaptive_scatter_kernel(const long long* __restrict__ m0, const long long* __restrict__ m1,
                                                               const float* __restrict__ ms0, const float* __restrict__ ms1,
                                                               const int* __restrict__ ind, const int* __restrict__ final_len,
                                                               int B, int M, int N, long long* __restrict__ out_m0,
                                                               long long* __restrict__ out_m1, float* __restrict__ out_ms0,
                                                               float* __restrict__ out_ms1) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int* ind0 = ind + (long long)b * M;
    const int* ind1 = ind + (long long)B * M + (long long)b * N;
    const int m = final_len[b], n = final_len[B + b];
    if (t < m) {
        const long long j = m0[(long long)b * M + t];
        out_m0[(long long)b * M + ind0[t]] = j < 0 ? -1 : ind1[j];
        out_ms0[(long long)b * M + ind0[t]] = ms0[(long long)b * M + t];
    } else if (t >= M && t - M < n) {
        const int u = t - M;
        const long long i = m1[(long long)b * N + u];
        out_m1[(long long)b * N + ind1[u]] = i < 0 ? -1 : ind0[i];
        out_ms1[(long long)b * N + ind1[u]] = ms1[(long long)b * N + u];
    }
}

// ---- host launchers ------------------------------------------------------------------------------------
int launch_adaptive_decide(const float* conf, int B, int M, int N, float thr, float depth_conf, int layer, int* st,
                           cudaStream_t stream) {
    B200_REQUIRE(conf && st && B >= 0 && M >= 1 && N >= 1, "adaptive_decide: bad arguments");
    if (B == 0) return 0;
    adaptive_decide_kernel<<<B, 256, 0, stream>>>(conf, B, M, N, thr, depth_conf, layer, st);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_adaptive_prune(const float* match, const float* conf, int B, int M, int N, float match_thr, float conf_thr,
                          int* st, int* pos, int* newc, const float* x_src, float* x_dst, const void* xm_src, void* xm_dst,
                          long long ld_xm, const void* cs_src, void* cs_dst, const int* ind_src, int* ind_dst,
                          long long* prune0, long long* prune1, cudaStream_t stream) {
    B200_REQUIRE(match && st && pos && newc && x_src && x_dst && xm_src && xm_dst && cs_src && cs_dst && ind_src && ind_dst &&
                     prune0 && prune1,
                 "adaptive_prune: null argument");
    B200_REQUIRE(B >= 0 && M >= 1 && N >= 1, "adaptive_prune: bad sizes");
    if (B == 0) return 0;
    adaptive_scan_kernel<<<dim3(2, B), 1024, 0, stream>>>(match, conf, B, M, N, match_thr, conf_thr, st, pos, newc);
    const long long T = (long long)B * (M + N);
    adaptive_gather_kernel<<<(unsigned)((T * 32 + 255) / 256), 256, 0, stream>>>(
        pos, st, B, M, N, x_src, x_dst, reinterpret_cast<const __nv_bfloat16*>(xm_src),
        reinterpret_cast<__nv_bfloat16*>(xm_dst), ld_xm, reinterpret_cast<const float2*>(cs_src),
        reinterpret_cast<float2*>(cs_dst), ind_src, ind_dst, prune0, prune1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_adaptive_commit(int* st, const int* newc, int B, int M, int N, int* lens, int* tile_live, int* hist_next,
                           cudaStream_t stream) {
    B200_REQUIRE(st && lens && tile_live, "adaptive_commit: null argument");
    if (B == 0) return 0;
    const long long T = (long long)B * (M + N);
    const int num_tiles = (int)((T + 127) / 128);
    const int work = num_tiles > 2 * B ? num_tiles : 2 * B;
    adaptive_commit_kernel<<<(work + 255) / 256, 256, 0, stream>>>(st, newc, B, M, N, lens, tile_live, num_tiles, hist_next);
    if (newc != nullptr) adaptive_store_kernel<<<(B + 255) / 256, 256, 0, stream>>>(st, newc, B);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_adaptive_snapshot(int* st, int B, int M, int N, int final_pass, const float* md, float* md_out, const float* z,
                             float* z_out, const float* x, float* x_out, const int* ind, int* ind_out, int* final_len,
                             cudaStream_t stream) {
    B200_REQUIRE(st && md && md_out && z && z_out && x && x_out && ind && ind_out && final_len, "adaptive_snapshot: null argument");
    if (B == 0) return 0;
    const long long T = (long long)B * (M + N);
    adaptive_snapshot_kernel<<<(unsigned)((T * 32 + 255) / 256), 256, 0, stream>>>(st, B, M, N, final_pass, md, md_out, z,
                                                                                  z_out, x, x_out, ind, ind_out, final_len);
    if (!final_pass) adaptive_clear_just_kernel<<<(B + 255) / 256, 256, 0, stream>>>(st, B);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int launch_adaptive_scatter(const long long* m0, const long long* m1, const float* ms0, const float* ms1, const int* ind,
                            const int* final_len, int B, int M, int N, long long* out_m0, long long* out_m1, float* out_ms0,
                            float* out_ms1, cudaStream_t stream) {
    B200_REQUIRE(m0 && m1 && ms0 && ms1 && ind && final_len && out_m0 && out_m1 && out_ms0 && out_ms1,
                 "adaptive_scatter: null argument");
    if (B == 0) return 0;
    adaptive_scatter_kernel<<<dim3((M + N + 255) / 256, B), 256, 0, stream>>>(m0, m1, ms0, ms1, ind, final_len, B, M, N, out_m0,
                                                                              out_m1, out_ms0, out_ms1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200
