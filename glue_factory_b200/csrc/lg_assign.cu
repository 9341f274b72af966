// Bandwidth-bound LightGlue kernels, part 2: log-assignment from the similarity matrix
//   scores[:M,:N] = log_softmax_rows(sim) + log_softmax_cols(sim) + logsigmoid(z0) + logsigmoid(z1)^T
//   scores[:M, N] = logsigmoid(-z0),  scores[M, :N] = logsigmoid(-z1)
// plus row / column arg-max of the inner block and the mutual-nearest-neighbour filter.
// The similarity matrix is read twice (statistics pass, finalise pass) and the scores written once;
// both passes use 64 x 256 tiles with one thread per column, so every global access is coalesced and
// column statistics stay in registers while row statistics go through warp shuffles.
// (reference: gluefactory/models/matchers/lightglue.py:250-262,265-281,287-303)
#include "common.cuh"
#include "lg_kernels.cuh"

namespace b200 {

constexpr int kATileM = 64, kATileN = 256;

__device__ __forceinline__ float log_sigmoid(float z) {
    return fminf(z, 0.f) - log1pf(expf(-fabsf(z)));
}

struct AssignWs {
    float2* row_part;  // [B, M, tiles_n]   (max, sumexp) then (best value, best column)
    float2* col_part;  // [B, tiles_m, N]
    float* rowterm;    // [B, M]  logsigmoid(z0) - row LSE
    float* colterm;    // [B, N]
    float* max0;       // [B, M]
    int* arg0;         // [B, M]
    int* arg1;         // [B, N]
};

__host__ __device__ inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

static AssignWs carve(void* ws, int B, int M, int N) {
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    uint8_t* p = reinterpret_cast<uint8_t*>(ws);
    AssignWs w;
    w.row_part = reinterpret_cast<float2*>(p); p += align256((size_t)B * M * tn * sizeof(float2));
    w.col_part = reinterpret_cast<float2*>(p); p += align256((size_t)B * tm * N * sizeof(float2));
    w.rowterm = reinterpret_cast<float*>(p); p += align256((size_t)B * M * 4);
    w.colterm = reinterpret_cast<float*>(p); p += align256((size_t)B * N * 4);
    w.max0 = reinterpret_cast<float*>(p); p += align256((size_t)B * M * 4);
    w.arg0 = reinterpret_cast<int*>(p); p += align256((size_t)B * M * 4);
    w.arg1 = reinterpret_cast<int*>(p);
    return w;
}

size_t assignment_workspace_bytes(int B, int M, int N) {
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    return align256((size_t)B * M * tn * 8) + align256((size_t)B * tm * N * 8) + 3 * align256((size_t)B * M * 4) +
           2 * align256((size_t)B * N * 4);
}

// pass 1: per-tile (max, sum exp) for rows and columns
__global__ void __launch_bounds__(256) assign_stats_kernel(const float* __restrict__ sim, int M, int N, AssignWs w) {
    __shared__ float2 srow[kATileM][8];
    const int b = blockIdx.z, tm = blockIdx.y, tn = blockIdx.x;
    const int tiles_n = gridDim.x, tiles_m = gridDim.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = tn * kATileN + threadIdx.x;
    const bool col_ok = j < N;
    const float* base = sim + (long long)b * M * N;
    float cm = -INFINITY, cs = 0.f;
    for (int r = 0; r < kATileM; ++r) {
        const int i = tm * kATileM + r;
        const float v = (col_ok && i < M) ? base[(long long)i * N + j] : -INFINITY;
        const float wm = warp_max(v);
        const float e = (v == -INFINITY) ? 0.f : expf(v - wm);
        const float ws = warp_sum(e);
        if (lane == 0) srow[r][warp] = make_float2(wm, ws);
        if (v != -INFINITY) {
            if (v > cm) {
                cs = cs * expf(cm - v) + 1.f;  // exp(-inf) = 0 on the first element
                cm = v;
            } else {
                cs += expf(v - cm);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < kATileM) {
        const int i = tm * kATileM + threadIdx.x;
        if (i < M) {
            float m = -INFINITY;
#pragma unroll
            for (int q = 0; q < 8; ++q) m = fmaxf(m, srow[threadIdx.x][q].x);
            float s = 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float2 t = srow[threadIdx.x][q];
                if (t.x != -INFINITY) s += t.y * expf(t.x - m);
            }
            w.row_part[((long long)b * M + i) * tiles_n + tn] = make_float2(m, s);
        }
    }
    if (col_ok) w.col_part[((long long)b * tiles_m + tm) * N + j] = make_float2(cm, cs);
}

// combine the partial statistics, fold in the matchability terms, write the dustbin row / column
__global__ void __launch_bounds__(256) assign_combine_kernel(int M, int N, int tiles_m, int tiles_n,
                                                             const float* __restrict__ z0, const float* __restrict__ z1,
                                                             AssignWs w, float* __restrict__ scores) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    float* sc = scores + (long long)b * (M + 1) * (N + 1);
    if (t < M) {
        const float2* p = w.row_part + ((long long)b * M + t) * tiles_n;
        float m = -INFINITY;
        for (int q = 0; q < tiles_n; ++q) m = fmaxf(m, p[q].x);
        float s = 0.f;
        for (int q = 0; q < tiles_n; ++q) s += p[q].y * expf(p[q].x - m);
        const float z = z0[(long long)b * M + t];
        w.rowterm[(long long)b * M + t] = log_sigmoid(z) - (m + logf(s));
        sc[(long long)t * (N + 1) + N] = log_sigmoid(-z);
    } else if (t < M + N) {
        const int j = t - M;
        float m = -INFINITY;
        for (int q = 0; q < tiles_m; ++q) m = fmaxf(m, w.col_part[((long long)b * tiles_m + q) * N + j].x);
        float s = 0.f;
        for (int q = 0; q < tiles_m; ++q) {
            const float2 v = w.col_part[((long long)b * tiles_m + q) * N + j];
            s += v.y * expf(v.x - m);
        }
        const float z = z1[(long long)b * N + j];
        w.colterm[(long long)b * N + j] = log_sigmoid(z) - (m + logf(s));
        sc[(long long)M * (N + 1) + j] = log_sigmoid(-z);
    } else if (t == M + N) {
        sc[(long long)M * (N + 1) + N] = 0.f;
    }
}

// pass 2: write the scores, track per-tile row / column maxima (lowest index wins ties)
__global__ void __launch_bounds__(256) assign_finalize_kernel(const float* __restrict__ sim, float simw, int M, int N,
                                                              AssignWs w, float* __restrict__ scores) {
    __shared__ float2 srow[kATileM][8];
    const int b = blockIdx.z, tm = blockIdx.y, tn = blockIdx.x;
    const int tiles_n = gridDim.x, tiles_m = gridDim.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = tn * kATileN + threadIdx.x;
    const bool col_ok = j < N;
    const float* base = sim + (long long)b * M * N;
    float* sc = scores + (long long)b * (M + 1) * (N + 1);
    const float ct = col_ok ? w.colterm[(long long)b * N + j] : 0.f;
    float cbest = -INFINITY;
    int cidx = 0;
    for (int r = 0; r < kATileM; ++r) {
        const int i = tm * kATileM + r;
        float v = -INFINITY;
        if (col_ok && i < M) {
            v = simw * base[(long long)i * N + j] + w.rowterm[(long long)b * M + i] + ct;
            sc[(long long)i * (N + 1) + j] = v;
            if (v > cbest) { cbest = v; cidx = i; }
        }
        float bv = v;
        int bi = j;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) srow[r][warp] = make_float2(bv, __int_as_float(bi));
    }
    __syncthreads();
    if (threadIdx.x < kATileM) {
        const int i = tm * kATileM + threadIdx.x;
        if (i < M) {
            float2 best = srow[threadIdx.x][0];
#pragma unroll
            for (int q = 1; q < 8; ++q) {
                const float2 t = srow[threadIdx.x][q];
                if (t.x > best.x) best = t;  // warps are in ascending column order: strict > keeps the first
            }
            w.row_part[((long long)b * M + i) * tiles_n + tn] = best;
        }
    }
    if (col_ok) w.col_part[((long long)b * tiles_m + tm) * N + j] = make_float2(cbest, __int_as_float(cidx));
}

__global__ void __launch_bounds__(256) assign_argmax_kernel(int M, int N, int tiles_m, int tiles_n, AssignWs w) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < M) {
        const float2* p = w.row_part + ((long long)b * M + t) * tiles_n;
        float2 best = p[0];
        for (int q = 1; q < tiles_n; ++q)
            if (p[q].x > best.x) best = p[q];
        w.max0[(long long)b * M + t] = best.x;
        w.arg0[(long long)b * M + t] = __float_as_int(best.y);
    } else if (t < M + N) {
        const int j = t - M;
        float2 best = w.col_part[((long long)b * tiles_m) * N + j];
        for (int q = 1; q < tiles_m; ++q) {
            const float2 v = w.col_part[((long long)b * tiles_m + q) * N + j];
            if (v.x > best.x) best = v;
        }
        w.arg1[(long long)b * N + j] = __float_as_int(best.y);
    }
}

// mutual nearest neighbours + threshold (lightglue.py:287-303); int64 matches, -1 = unmatched
__global__ void __launch_bounds__(256) filter_matches_kernel(int M, int N, float th, AssignWs w,
                                                             long long* __restrict__ m0, long long* __restrict__ m1,
                                                             float* __restrict__ ms0, float* __restrict__ ms1) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int* a0 = w.arg0 + (long long)b * M;
    const int* a1 = w.arg1 + (long long)b * N;
    const float* mx = w.max0 + (long long)b * M;
    if (t < M) {
        const int j = a0[t];
        const bool mutual = a1[j] == t;
        const float s = mutual ? expf(mx[t]) : 0.f;
        ms0[(long long)b * M + t] = s;
        m0[(long long)b * M + t] = (mutual && s > th) ? j : -1;
    } else if (t < M + N) {
        const int j = t - M;
        const int i = a1[j];
        const bool mutual = a0[i] == j;  // then keypoint i is mutual as well and its score is exp(max0[i])
        const float s = mutual ? expf(mx[i]) : 0.f;
        ms1[(long long)b * N + j] = s;
        m1[(long long)b * N + j] = (mutual && s > th) ? i : -1;
    }
}

int launch_assignment(const float* sim, const float* z0, const float* z1, int B, int M, int N, float threshold,
                      void* workspace, float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                      cudaStream_t stream) {
    B200_REQUIRE(M >= 1 && N >= 1, "assignment: needs at least one keypoint per image (M=%d, N=%d)", M, N);
    if (B == 0) return 0;
    const AssignWs w = carve(workspace, B, M, N);
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    dim3 tiles(tn, tm, B), lin((M + N + 1 + 255) / 256, B);
    assign_stats_kernel<<<tiles, 256, 0, stream>>>(sim, M, N, w);
    assign_combine_kernel<<<lin, 256, 0, stream>>>(M, N, tm, tn, z0, z1, w, scores);
    B200_CHECK_CUDA(cudaGetLastError());
    return launch_assignment_finalize(sim, 2.0f, B, M, N, threshold, workspace, scores, m0, m1, ms0, ms1, stream);
}

// scores[:M,:N] = simw * sim + rowterm + colterm (terms in the workspace), then arg-max + mutual-NN filter
int launch_assignment_finalize(const float* sim, float simw, int B, int M, int N, float threshold, void* workspace,
                               float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                               cudaStream_t stream) {
    const AssignWs w = carve(workspace, B, M, N);
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    dim3 tiles(tn, tm, B), lin((M + N + 1 + 255) / 256, B);
    assign_finalize_kernel<<<tiles, 256, 0, stream>>>(sim, simw, M, N, w, scores);
    assign_argmax_kernel<<<lin, 256, 0, stream>>>(M, N, tm, tn, w);
    filter_matches_kernel<<<lin, 256, 0, stream>>>(M, N, threshold, w, m0, m1, ms0, ms1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

void assignment_terms(void* workspace, int B, int M, int N, float** rowterm, float** colterm) {
    const AssignWs w = carve(workspace, B, M, N);
    *rowterm = w.rowterm;
    *colterm = w.colterm;
}

}  // namespace b200
