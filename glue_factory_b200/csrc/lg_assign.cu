// Bandwidth-bound LightGlue kernels, part 2: log-assignment from the similarity matrix
//   scores[:M,:N] = log_softmax_rows(sim) + log_softmax_cols(sim) + logsigmoid(z0) + logsigmoid(z1)^T
//   scores[:M, N] = logsigmoid(-z0),  scores[M, :N] = logsigmoid(-z1)
// plus row / column arg-max of the inner block and the mutual-nearest-neighbour filter.
// The similarity matrix is read twice (statistics pass, finalise pass) and the scores written once;
// both passes use 64 x 256 tiles with one thread per column, so every global access is coalesced; column
// statistics stay in registers, row statistics are taken after a transposition through shared memory.
// (reference: gluefactory/models/matchers/lightglue.py:250-262,265-281,287-303)
#include <mutex>

#include "../../include/b200glue.h"
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
    float2* row_top;   // nearest-neighbour matcher only: [B, M, tiles_n] (arg-max, second largest)
    float2* col_top;   //                                 [B, tiles_m, N]
};

__host__ __device__ inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

static AssignWs carve(void* ws, int B, int M, int N, bool nn = false) {
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    uint8_t* p = reinterpret_cast<uint8_t*>(ws);
    AssignWs w;
    w.row_part = reinterpret_cast<float2*>(p); p += align256((size_t)B * M * tn * sizeof(float2));
    w.col_part = reinterpret_cast<float2*>(p); p += align256((size_t)B * tm * N * sizeof(float2));
    w.rowterm = reinterpret_cast<float*>(p); p += align256((size_t)B * M * 4);
    w.colterm = reinterpret_cast<float*>(p); p += align256((size_t)B * N * 4);
    w.max0 = reinterpret_cast<float*>(p); p += align256((size_t)B * M * 4);
    w.arg0 = reinterpret_cast<int*>(p); p += align256((size_t)B * M * 4);
    w.arg1 = reinterpret_cast<int*>(p); p += align256((size_t)B * N * 4);
    w.row_top = w.col_top = nullptr;
    if (nn) {
        w.row_top = reinterpret_cast<float2*>(p); p += align256((size_t)B * M * tn * sizeof(float2));
        w.col_top = reinterpret_cast<float2*>(p);
    }
    return w;
}

size_t assignment_workspace_bytes(int B, int M, int N) {
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    return align256((size_t)B * M * tn * 8) + align256((size_t)B * tm * N * 8) + 3 * align256((size_t)B * M * 4) +
           2 * align256((size_t)B * N * 4);
}

// ---- tile machinery shared by both passes ----------------------------------------------------------
// A CTA owns a 64 x 256 tile, processed as two 32-row halves.  Column side: thread t holds column t's 32
// values in registers (32 independent coalesced loads in flight).  Row side: the half tile is transposed
// through shared memory; thread t then owns row t / 8 and the 32 columns (t % 8) + 8 k, so a row statistic
// costs 32 register operations + a 3-step shuffle merge instead of a 5-step warp reduction per element.
// Row stride 264 floats (= 8 mod 32 banks): both the column-wise stores and the strided row reads are
// bank-conflict free.
constexpr int kAHalf = 32, kAStride = 264;

__device__ __forceinline__ float finite_or_zero(float m) { return m == -INFINITY ? 0.f : m; }

struct Top2 {  // running (max, arg-max, second largest); scanning in ascending index order keeps the lowest index
    float v1 = -INFINITY, v2 = -INFINITY;
    int i1 = 0;
    __device__ __forceinline__ void push(float v, int idx) {
        v2 = fmaxf(v2, fminf(v1, v));
        i1 = v > v1 ? idx : i1;
        v1 = fmaxf(v1, v);
    }
    __device__ __forceinline__ void merge(float ov1, int oi1, float ov2) {
        v2 = fmaxf(fmaxf(v2, ov2), fminf(v1, ov1));
        if (ov1 > v1 || (ov1 == v1 && oi1 < i1)) i1 = oi1;
        v1 = fmaxf(v1, ov1);
    }
};

// pass 1: per-tile (max, sum exp) for rows and columns; TOP2 adds (arg-max, second largest) for the
// nearest-neighbour matcher
template <bool TOP2>
__global__ void __launch_bounds__(256) assign_stats_kernel(const float* __restrict__ sim, int M, int N, AssignWs w,
                                                           const int* __restrict__ live) {
    __shared__ float tile[kAHalf * kAStride];
    const int b = blockIdx.z, tm = blockIdx.y, tn = blockIdx.x;
    const int tiles_n = gridDim.x, tiles_m = gridDim.y;
    // `live` (adaptive pruning): only the first Mb rows / Nb columns of the padded [M, N] block exist
    const int Mb = live ? live[b] : M, Nb = live ? live[gridDim.z + b] : N;
    if (tm * kATileM >= Mb || tn * kATileN >= Nb) return;
    const int j = tn * kATileN + threadIdx.x;
    const bool col_ok = j < Nb;
    const int rrow = threadIdx.x >> 3, seg = threadIdx.x & 7;
    const float* base = sim + (long long)b * M * N;
    float cm = -INFINITY, cs = 0.f;
    Top2 ctop;
    for (int h = 0; h < kATileM / kAHalf; ++h) {
        const int i0 = tm * kATileM + h * kAHalf;
        if (i0 >= Mb) break;
        float v[kAHalf];
#pragma unroll
        for (int r = 0; r < kAHalf; ++r) v[r] = (col_ok && i0 + r < Mb) ? __ldg(base + (long long)(i0 + r) * N + j) : -INFINITY;
        // column side
        float m = cm;
#pragma unroll
        for (int r = 0; r < kAHalf; ++r) {
            m = fmaxf(m, v[r]);
            if (TOP2) ctop.push(v[r], i0 + r);
        }
        const float ms = finite_or_zero(m);
        float s = 0.f;
#pragma unroll
        for (int r = 0; r < kAHalf; ++r) s += __expf(v[r] - ms);
        cs = cs * __expf(cm - ms) + s;
        cm = m;
        // row side
#pragma unroll
        for (int r = 0; r < kAHalf; ++r) tile[r * kAStride + threadIdx.x] = v[r];
        __syncthreads();
        float rm = -INFINITY;
        Top2 rtop;
#pragma unroll
        for (int k = 0; k < kAHalf; ++k) {
            v[k] = tile[rrow * kAStride + seg + 8 * k];
            rm = fmaxf(rm, v[k]);
            if (TOP2) rtop.push(v[k], tn * kATileN + seg + 8 * k);
        }
        const float rms = finite_or_zero(rm);
        float rs = 0.f;
#pragma unroll
        for (int k = 0; k < kAHalf; ++k) rs += __expf(v[k] - rms);
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const float om = __shfl_xor_sync(0xffffffffu, rm, o), os = __shfl_xor_sync(0xffffffffu, rs, o);
            const float nm = fmaxf(rm, om), nms = finite_or_zero(nm);
            rs = rs * __expf(rm - nms) + os * __expf(om - nms);
            rm = nm;
            if (TOP2) {
                const float ov1 = __shfl_xor_sync(0xffffffffu, rtop.v1, o), ov2 = __shfl_xor_sync(0xffffffffu, rtop.v2, o);
                const int oi1 = __shfl_xor_sync(0xffffffffu, rtop.i1, o);
                rtop.merge(ov1, oi1, ov2);
            }
        }
        if (seg == 0 && i0 + rrow < Mb) {
            const long long at = ((long long)b * M + i0 + rrow) * tiles_n + tn;
            w.row_part[at] = make_float2(rm, rs);
            if (TOP2) w.row_top[at] = make_float2(__int_as_float(rtop.i1), rtop.v2);
        }
        __syncthreads();
    }
    if (col_ok) {
        const long long at = ((long long)b * tiles_m + tm) * N + j;
        w.col_part[at] = make_float2(cm, cs);
        if (TOP2) w.col_top[at] = make_float2(__int_as_float(ctop.i1), ctop.v2);
    }
}

// combine the partial statistics, fold in the matchability terms, write the dustbin row / column
__global__ void __launch_bounds__(256) assign_combine_kernel(int M, int N, int tiles_m, int tiles_n,
                                                             const float* __restrict__ z0, const float* __restrict__ z1,
                                                             AssignWs w, float* __restrict__ scores,
                                                             const int* __restrict__ live) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    float* sc = scores + (long long)b * (M + 1) * (N + 1);
    // with `live` the dustbin column / row sit right behind the live block (indices Nb / Mb of the padded matrix),
    // so that scores[b, :Mb+1, :Nb+1] is the reference's (Mb+1) x (Nb+1) result
    const int Mb = live ? live[b] : M, Nb = live ? live[gridDim.y + b] : N;
    const int tn_live = (Nb + kATileN - 1) / kATileN, tm_live = (Mb + kATileM - 1) / kATileM;
    if (t < Mb) {
        const float2* p = w.row_part + ((long long)b * M + t) * tiles_n;
        float m = -INFINITY;
        for (int q = 0; q < tn_live; ++q) m = fmaxf(m, p[q].x);
        float s = 0.f;
        for (int q = 0; q < tn_live; ++q) s += p[q].y * expf(p[q].x - m);
        const float z = z0[(long long)b * M + t];
        w.rowterm[(long long)b * M + t] = log_sigmoid(z) - (m + logf(s));
        sc[(long long)t * (N + 1) + Nb] = log_sigmoid(-z);
    } else if (t >= M && t < M + Nb) {
        const int j = t - M;
        float m = -INFINITY;
        for (int q = 0; q < tm_live; ++q) m = fmaxf(m, w.col_part[((long long)b * tiles_m + q) * N + j].x);
        float s = 0.f;
        for (int q = 0; q < tm_live; ++q) {
            const float2 v = w.col_part[((long long)b * tiles_m + q) * N + j];
            s += v.y * expf(v.x - m);
        }
        const float z = z1[(long long)b * N + j];
        w.colterm[(long long)b * N + j] = log_sigmoid(z) - (m + logf(s));
        sc[(long long)Mb * (N + 1) + j] = log_sigmoid(-z);
    } else if (t == M + N) {
        sc[(long long)Mb * (N + 1) + Nb] = 0.f;
    }
}

// pass 2: write the scores; ARGMAX also tracks per-tile row / column maxima (lowest index wins ties)
template <bool ARGMAX>
__global__ void __launch_bounds__(256, ARGMAX ? 3 : 2) assign_finalize_kernel(const float* __restrict__ sim, float simw, int M, int N,
                                                              AssignWs w, float* __restrict__ scores,
                                                              const int* __restrict__ live) {
    __shared__ float tile[ARGMAX ? kAHalf * kAStride : 1];
    __shared__ float srt[kATileM];
    const int b = blockIdx.z, tm = blockIdx.y, tn = blockIdx.x;
    const int tiles_n = gridDim.x, tiles_m = gridDim.y;
    const int Mb = live ? live[b] : M, Nb = live ? live[gridDim.z + b] : N;
    if (tm * kATileM >= Mb || tn * kATileN >= Nb) return;
    const int j = tn * kATileN + threadIdx.x;
    const bool col_ok = j < Nb;
    const int rrow = threadIdx.x >> 3, seg = threadIdx.x & 7;
    const float* base = sim + (long long)b * M * N;
    float* sc = scores + (long long)b * (M + 1) * (N + 1);
    const float ct = col_ok ? w.colterm[(long long)b * N + j] : 0.f;
    if (threadIdx.x < kATileM) {
        const int i = tm * kATileM + threadIdx.x;
        srt[threadIdx.x] = i < Mb ? w.rowterm[(long long)b * M + i] : 0.f;
    }
    __syncthreads();
    float cbest = -INFINITY;
    int cidx = 0;
    for (int h = 0; h < kATileM / kAHalf; ++h) {
        const int i0 = tm * kATileM + h * kAHalf;
        if (i0 >= Mb) break;
        float v[kAHalf];
#pragma unroll
        for (int r = 0; r < kAHalf; ++r) v[r] = (col_ok && i0 + r < Mb) ? __ldg(base + (long long)(i0 + r) * N + j) : -INFINITY;
#pragma unroll
        for (int r = 0; r < kAHalf; ++r) {
            v[r] = fmaf(simw, v[r], srt[h * kAHalf + r] + ct);  // -inf stays -inf outside the matrix
            if (col_ok && i0 + r < Mb) sc[(long long)(i0 + r) * (N + 1) + j] = v[r];
            if (ARGMAX) {
                cidx = v[r] > cbest ? i0 + r : cidx;
                cbest = fmaxf(cbest, v[r]);
            }
        }
        if (ARGMAX) {
#pragma unroll
            for (int r = 0; r < kAHalf; ++r) tile[r * kAStride + threadIdx.x] = v[r];
            __syncthreads();
            float bv = -INFINITY;
            int bi = tn * kATileN + seg;
#pragma unroll
            for (int k = 0; k < kAHalf; ++k) {
                const float u = tile[rrow * kAStride + seg + 8 * k];
                bi = u > bv ? tn * kATileN + seg + 8 * k : bi;
                bv = fmaxf(bv, u);
            }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (seg == 0 && i0 + rrow < Mb)
                w.row_part[((long long)b * M + i0 + rrow) * tiles_n + tn] = make_float2(bv, __int_as_float(bi));
            __syncthreads();
        }
    }
    if (ARGMAX && col_ok) w.col_part[((long long)b * tiles_m + tm) * N + j] = make_float2(cbest, __int_as_float(cidx));
}

__global__ void __launch_bounds__(256) assign_argmax_kernel(int M, int N, int tiles_m, int tiles_n, AssignWs w,
                                                            const int* __restrict__ live) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int Mb = live ? live[b] : M, Nb = live ? live[gridDim.y + b] : N;
    const int tn_live = (Nb + kATileN - 1) / kATileN, tm_live = (Mb + kATileM - 1) / kATileM;
    if (t < Mb) {
        const float2* p = w.row_part + ((long long)b * M + t) * tiles_n;
        float2 best = p[0];
        for (int q = 1; q < tn_live; ++q)
            if (p[q].x > best.x) best = p[q];
        w.max0[(long long)b * M + t] = best.x;
        w.arg0[(long long)b * M + t] = __float_as_int(best.y);
    } else if (t >= M && t < M + Nb) {
        const int j = t - M;
        float2 best = w.col_part[((long long)b * tiles_m) * N + j];
        for (int q = 1; q < tm_live; ++q) {
            const float2 v = w.col_part[((long long)b * tiles_m + q) * N + j];
            if (v.x > best.x) best = v;
        }
        w.arg1[(long long)b * N + j] = __float_as_int(best.y);
    }
}

// mutual nearest neighbours + threshold (lightglue.py:287-303); int64 matches, -1 = unmatched
__global__ void __launch_bounds__(256) filter_matches_kernel(int M, int N, float th, AssignWs w,
                                                             long long* __restrict__ m0, long long* __restrict__ m1,
                                                             float* __restrict__ ms0, float* __restrict__ ms1,
                                                             const int* __restrict__ live) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int* a0 = w.arg0 + (long long)b * M;
    const int* a1 = w.arg1 + (long long)b * N;
    const float* mx = w.max0 + (long long)b * M;
    const int Mb = live ? live[b] : M, Nb = live ? live[gridDim.y + b] : N;
    if ((t < M && t >= Mb) || (t >= M && t < M + N && t - M >= Nb)) {  // padding behind the live block
        if (t < M) { m0[(long long)b * M + t] = -1; ms0[(long long)b * M + t] = 0.f; }
        else { m1[(long long)b * N + t - M] = -1; ms1[(long long)b * N + t - M] = 0.f; }
        return;
    }
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
    return launch_assignment_live(sim, z0, z1, B, M, N, nullptr, threshold, workspace, scores, m0, m1, ms0, ms1, stream);
}

// same on padded [B, M, N] problems whose live sizes (Mb, Nb) are DEVICE values live[b], live[B + b] (adaptive
// pruning without host synchronisation); live == nullptr: every pair is M x N
int launch_assignment_live(const float* sim, const float* z0, const float* z1, int B, int M, int N, const int* live,
                           float threshold, void* workspace, float* scores, long long* m0, long long* m1, float* ms0,
                           float* ms1, cudaStream_t stream) {
    B200_REQUIRE(M >= 1 && N >= 1, "assignment: needs at least one keypoint per image (M=%d, N=%d)", M, N);
    if (B == 0) return 0;
    const AssignWs w = carve(workspace, B, M, N);
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    dim3 tiles(tn, tm, B), lin((M + N + 1 + 255) / 256, B);
    assign_stats_kernel<false><<<tiles, 256, 0, stream>>>(sim, M, N, w, live);
    assign_combine_kernel<<<lin, 256, 0, stream>>>(M, N, tm, tn, z0, z1, w, scores, live);
    B200_CHECK_CUDA(cudaGetLastError());
    return launch_assignment_finalize(sim, 2.0f, B, M, N, threshold, workspace, scores, m0, m1, ms0, ms1, stream, live);
}

// scores[:M,:N] = simw * sim + rowterm + colterm (terms in the workspace), then arg-max + mutual-NN filter
int launch_assignment_finalize(const float* sim, float simw, int B, int M, int N, float threshold, void* workspace,
                               float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                               cudaStream_t stream, const int* live) {
    const AssignWs w = carve(workspace, B, M, N);
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    dim3 tiles(tn, tm, B), lin((M + N + 1 + 255) / 256, B);
    assign_finalize_kernel<true><<<tiles, 256, 0, stream>>>(sim, simw, M, N, w, scores, live);
    assign_argmax_kernel<<<lin, 256, 0, stream>>>(M, N, tm, tn, w, live);
    filter_matches_kernel<<<lin, 256, 0, stream>>>(M, N, threshold, w, m0, m1, ms0, ms1, live);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---- allocation-free variant of launch_assignment ---------------------------------------------------
// Callers that cannot provide a workspace (CUDA graph capture, serving loops that forbid cudaMalloc) get the
// scratch from a file-scope device array instead.  Every GPU that loads the library holds these bytes until
// exit.  In HOST code `g_assign_scratch` is only a stand-in for the symbol, not a device address: the address
// is looked up once per device with cudaGetSymbolAddress (device current) and carved like a caller workspace.
// All calls share the one array and it keeps the previous call's contents; the assignment kernels write every
// partial / term / arg-max slot before reading it, so nothing has to start at zero.  Calls that overlap on
// the GPU (two streams, or two host threads on different streams) race on it.
constexpr size_t kStaticScratchBytes = B200_LG_ASSIGNMENT_STATIC_MAX_BYTES;
static __device__ __align__(256) unsigned char g_assign_scratch[kStaticScratchBytes];

static void* g_assign_scratch_ptr[kMaxDevices];  // per device, filled on first use
static std::mutex g_assign_scratch_mu;

static int static_scratch(void** out) {
    int dev = -1;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "assignment_static: device index %d is out of range", dev);
    std::lock_guard<std::mutex> lk(g_assign_scratch_mu);
    if (g_assign_scratch_ptr[dev] == nullptr) {
        void* ptr = nullptr;
        B200_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_assign_scratch));
        B200_REQUIRE(ptr != nullptr, "assignment_static: cudaGetSymbolAddress returned a null address");
        g_assign_scratch_ptr[dev] = ptr;
    }
    *out = g_assign_scratch_ptr[dev];
    return 0;
}

size_t assignment_static_capacity() { return kStaticScratchBytes; }

int launch_assignment_static(const float* sim, const float* z0, const float* z1, int B, int M, int N, float threshold,
                             float* scores, long long* m0, long long* m1, float* ms0, float* ms1,
                             cudaStream_t stream) {
    B200_REQUIRE(B >= 0 && M >= 1 && N >= 1, "assignment_static: bad sizes (B=%d, M=%d, N=%d)", B, M, N);
    const size_t need = assignment_workspace_bytes(B, M, N);
    B200_REQUIRE(need <= kStaticScratchBytes,
                 "assignment_static: B=%d, M=%d, N=%d needs %zu bytes of scratch, the built-in array holds %zu; "
                 "use b200_lg_assignment with a caller workspace",
                 B, M, N, need, kStaticScratchBytes);
    void* ws = nullptr;
    if (static_scratch(&ws) != 0) return -2;
    return launch_assignment(sim, z0, z1, B, M, N, threshold, ws, scores, m0, m1, ms0, ms1, stream);
}

// ---- nearest-neighbour matcher (reference: gluefactory/models/matchers/nearest_neighbor_matcher.py:14-72) ----
size_t nn_match_workspace_bytes(int B, int M, int N) {
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    return assignment_workspace_bytes(B, M, N) + align256((size_t)B * M * tn * 8) + align256((size_t)B * tm * N * 8);
}

// find_nn (:14-25) on the merged statistics: nearest neighbour kept when its squared distance 2 (1 - sim) passes
// the ratio test against the second nearest and the absolute threshold (each disabled when <= 0)
__device__ __forceinline__ int nn_decide(float best, int idx, float second, float ratio_sq, float dist_sq) {
    const float d0 = 2.f * (1.f - best), d1 = 2.f * (1.f - second);
    bool ok = true;
    if (ratio_sq > 0.f) ok = ok && d0 <= ratio_sq * d1;
    if (dist_sq > 0.f) ok = ok && d0 <= dist_sq;
    return ok ? idx : -1;
}

// merge the per-tile statistics: log-sum-exp terms of both softmaxes, one-way nearest neighbours, zero
// dustbin row / column of the log-assignment
__global__ void __launch_bounds__(256) nn_combine_kernel(int M, int N, int tiles_m, int tiles_n, float ratio_sq,
                                                         float dist_sq, AssignWs w, float* __restrict__ scores) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    float* sc = scores + (long long)b * (M + 1) * (N + 1);
    if (t < M) {
        const long long at = ((long long)b * M + t) * tiles_n;
        float m = -INFINITY;
        for (int q = 0; q < tiles_n; ++q) m = fmaxf(m, w.row_part[at + q].x);
        float s = 0.f;
        Top2 top;
        for (int q = 0; q < tiles_n; ++q) {
            const float2 p = w.row_part[at + q], u = w.row_top[at + q];
            s += p.y * expf(p.x - m);
            top.merge(p.x, __float_as_int(u.x), u.y);
        }
        w.rowterm[(long long)b * M + t] = -(m + logf(s));
        w.arg0[(long long)b * M + t] = nn_decide(top.v1, top.i1, top.v2, ratio_sq, dist_sq);
        sc[(long long)t * (N + 1) + N] = 0.f;
    } else if (t < M + N) {
        const int j = t - M;
        float m = -INFINITY;
        for (int q = 0; q < tiles_m; ++q) m = fmaxf(m, w.col_part[((long long)b * tiles_m + q) * N + j].x);
        float s = 0.f;
        Top2 top;
        for (int q = 0; q < tiles_m; ++q) {
            const float2 p = w.col_part[((long long)b * tiles_m + q) * N + j], u = w.col_top[((long long)b * tiles_m + q) * N + j];
            s += p.y * expf(p.x - m);
            top.merge(p.x, __float_as_int(u.x), u.y);
        }
        w.colterm[(long long)b * N + j] = -(m + logf(s));
        w.arg1[(long long)b * N + j] = nn_decide(top.v1, top.i1, top.v2, ratio_sq, dist_sq);
        sc[(long long)M * (N + 1) + j] = 0.f;
    } else if (t == M + N) {
        sc[(long long)M * (N + 1) + N] = 0.f;
    }
}

// mutual_check (:28-36) and the 0 / 1 matching scores (:63-64); int64 matches, -1 = unmatched
__global__ void __launch_bounds__(256) nn_mutual_kernel(int M, int N, int mutual, AssignWs w, long long* __restrict__ m0,
                                                        long long* __restrict__ m1, float* __restrict__ ms0,
                                                        float* __restrict__ ms1) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int* a0 = w.arg0 + (long long)b * M;
    const int* a1 = w.arg1 + (long long)b * N;
    if (t < M) {
        int j = a0[t];
        if (mutual && j >= 0 && a1[j] != t) j = -1;
        m0[(long long)b * M + t] = j;
        ms0[(long long)b * M + t] = j >= 0 ? 1.f : 0.f;
    } else if (t < M + N) {
        const int c = t - M;
        int i = a1[c];
        if (mutual && i >= 0 && a0[i] != c) i = -1;
        m1[(long long)b * N + c] = i;
        ms1[(long long)b * N + c] = i >= 0 ? 1.f : 0.f;
    }
}

int launch_nn_match(const float* sim, int B, int M, int N, float ratio_sq, float dist_sq, int mutual, void* workspace,
                    float* scores, long long* m0, long long* m1, float* ms0, float* ms1, cudaStream_t stream) {
    B200_REQUIRE(M >= 1 && N >= 1, "nn_match: needs at least one keypoint per image (M=%d, N=%d)", M, N);
    B200_REQUIRE(!(ratio_sq > 0.f) || (M >= 2 && N >= 2), "nn_match: the ratio test needs two candidates (M=%d, N=%d)", M, N);
    if (B == 0) return 0;
    const AssignWs w = carve(workspace, B, M, N, true);
    const int tn = (N + kATileN - 1) / kATileN, tm = (M + kATileM - 1) / kATileM;
    dim3 tiles(tn, tm, B), lin((M + N + 1 + 255) / 256, B);
    assign_stats_kernel<true><<<tiles, 256, 0, stream>>>(sim, M, N, w, nullptr);
    nn_combine_kernel<<<lin, 256, 0, stream>>>(M, N, tm, tn, ratio_sq, dist_sq, w, scores);
    assign_finalize_kernel<false><<<tiles, 256, 0, stream>>>(sim, 2.0f, M, N, w, scores, nullptr);
    nn_mutual_kernel<<<lin, 256, 0, stream>>>(M, N, mutual, w, m0, m1, ms0, ms1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

void assignment_terms(void* workspace, int B, int M, int N, float** rowterm, float** colterm) {
    const AssignWs w = carve(workspace, B, M, N);
    *rowterm = w.rowterm;
    *colterm = w.colterm;
}

}  // namespace b200
