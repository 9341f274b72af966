// Log-domain Sinkhorn for the SuperGlue configuration: one persistent cooperative kernel for all
// iterations and all pairs of the batch.
//
//   couplings Z (M+1) x (N+1): cost in the inner block, the learned bin score in the last row / column
//   iterate  u_i = log_mu_i - LSE_j(Z_ij + v_j) ;  v_j = log_nu_j - LSE_i(Z_ij + u_i)
//   result   Z_ij + u_i + v_j - norm,  norm = -log(M + N)
// (reference: gluefactory_nonfree/superglue.py:173-201)
//
// The reference makes 2 full passes over Z per iteration.  Here Z is cut into blocks of R rows; a CTA
// owns a contiguous range of blocks, stages a block in shared memory ONCE per iteration (cp.async,
// double buffered; blocks stay resident when a CTA owns at most two) and uses it twice:
//   row pass   : exact max-stabilised LSE over the block's rows with v of the previous iteration -> u
//   column pass: with the fresh u, partial column sums  sum_i exp(Z_ij + u_i - c_j),  c_j = log_nu_j -
//                v_j being last iteration's column LSE (a stabiliser that is exact in real arithmetic:
//                v_new = v_old - log(total column sum); Sinkhorn updates are non-expansive, so the
//                exponent stays far from overflow for |cost| < ~60).
// Per-CTA partial column sums go to a small global buffer; after ONE grid barrier per iteration every
// CTA re-derives v for its pairs by summing the partials in a fixed order (deterministic, no atomics).
// Algorithmic traffic: iters x (M+1)(N+1) x 4 B (half of the reference's two passes) and
// 2 x iters x (M+1)(N+1) exponentials.
#include "common.cuh"
#include "lg_kernels.cuh"

namespace b200 {

constexpr int kSkThreads = 512;
constexpr int kSkMaxCols = 16;  // columns per thread: N + 1 <= 8192
constexpr int kSkRowF4 = 9;    // float4 per lane in the row pass: a warp's column part holds <= 1152 columns

struct SinkhornParams {
    const float* cost;   // [B, M, N]
    const float* bin;    // device scalar (bin_score)
    int B, M, N, iters, R, nblk, pitch, kmax;
    float* U;            // [B, M+1]      row potentials of the current iteration
    float* V;            // [2, B, N+1]   column potentials (double buffered by iteration parity)
    float* P;            // [2, B, kmax, N+1] partial column sums
    unsigned int* barrier;
    float* rowterm;      // [B, M]  u_i - norm          (consumed by the finalize / arg-max kernels)
    float* colterm;      // [B, N]  v_j
    float* scores;       // [B, M+1, N+1]: this kernel writes the dustbin row and column
};

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ int sk_owner(long long g, long long total, int G) {
    // CTA c owns global blocks [c * total / G, (c + 1) * total / G)
    int c = (int)((g * G) / total);
    while (c + 1 < G && ((long long)(c + 1) * total) / G <= g) ++c;
    while (c > 0 && ((long long)c * total) / G > g) --c;
    return c;
}

__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        unsigned int seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
        } while (seen < target);
        __threadfence();
    }
    __syncthreads();
}

constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;

__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Stage block gb (R rows of the couplings) into dst.  Fast path (N % 4 == 0): one thread issues one bulk
// copy per cost row, completion on an mbarrier; otherwise 4-byte cp.async by all threads.  The dustbin
// column / row are plain stores.  The caller separates this from earlier reads of dst by a __syncthreads.
template <int R>
__device__ __forceinline__ void sk_load_block(const SinkhornParams& p, long long gb, float* dst, uint64_t* bar,
                                              float alpha, bool bulk) {
    const int Mp = p.M + 1, Np = p.N + 1;
    const int pr = (int)(gb / p.nblk), rb = (int)(gb - (long long)pr * p.nblk);
    const int i0 = rb * R;
    const int cost_rows = max(0, min(R, p.M - i0));  // rows of this block that come from the cost matrix
    const float* src = p.cost + ((long long)pr * p.M + i0) * p.N;
    if (bulk) {
        if (threadIdx.x == 0) {
            mbar_arrive_expect_tx(bar, (uint32_t)cost_rows * (uint32_t)p.N * 4u);
            for (int r = 0; r < cost_rows; ++r) bulk_load(dst + r * p.pitch, src + (long long)r * p.N, (uint32_t)p.N * 4u, bar);
        }
    } else {
        for (int r = 0; r < cost_rows; ++r)
            for (int q = threadIdx.x; q < p.N; q += kSkThreads) cp_async4(dst + r * p.pitch + q, src + (long long)r * p.N + q);
        cp_async_commit();
    }
    if (threadIdx.x < cost_rows) dst[threadIdx.x * p.pitch + p.N] = alpha;
    if (i0 + R > p.M && i0 <= p.M) {  // the dustbin row lives in this block
        float* row = dst + (p.M - i0) * p.pitch;
        for (int j = threadIdx.x; j < Np; j += kSkThreads) row[j] = alpha;
    }
    (void)Mp;
}
template <int R>
__global__ void __launch_bounds__(kSkThreads, 1) sinkhorn_kernel(const SinkhornParams p) {
    extern __shared__ __align__(16) float sk_smem[];
    const int Mp = p.M + 1, Np = p.N + 1, pitch = p.pitch;
    float* sv2 = sk_smem + 2 * R * pitch;   // [pitch] v_j * log2(e) of the current pair
    float* soff = sv2 + pitch;              // [pitch] (v_j - log_nu_j) * log2(e): minus the column stabiliser
    float* cu = soff + pitch;               // [16] u_i * log2(e) of the current block
    float2* sred = reinterpret_cast<float2*>(cu + 16);                // [16] per-warp (max, sum), log2 domain
    uint64_t* full = reinterpret_cast<uint64_t*>(sred + 16);          // [2] bulk-copy completion per buffer
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = gridDim.x, c = blockIdx.x;
    const long long total = (long long)p.B * p.nblk;
    const long long g0 = (c * total) / G, g1 = ((c + 1) * total) / G;
    const int nown = (int)(g1 - g0);
    const bool resident = nown <= 2;
    const bool bulk = (p.N & 3) == 0 && (reinterpret_cast<uintptr_t>(p.cost) & 15) == 0;
    const float alpha = __ldg(p.bin);
    const float norm = -logf((float)(p.M + p.N));
    const float log_mu_bin = logf((float)p.N) + norm, log_nu_bin = logf((float)p.M) + norm;
    const long long v_stride = (long long)p.B * Np, p_stride = (long long)p.B * p.kmax * Np;
    const int ncolk = (Np + kSkThreads - 1) / kSkThreads;

    if (tid == 0) {
        mbar_init(full, 1);
        mbar_init(full + 1, 1);
        fence_barrier_init();
    }
    __syncthreads();
    uint32_t phase0 = 0, phase1 = 0;  // parity of the next completion of each buffer
    if (nown > 0) sk_load_block<R>(p, g0, sk_smem, full, alpha, bulk);
    if (resident && nown == 2) sk_load_block<R>(p, g0 + 1, sk_smem + R * pitch, full + 1, alpha, bulk);
    int cur = 0;
    float acc[kSkMaxCols];

    for (int t = 1; t <= p.iters + 1; ++t) {
        const bool last_round = t == p.iters + 1;  // only derives the final v and writes the outputs
        int cur_pair = -1, kidx = 0;
        for (long long gb = g0; gb < g1; ++gb) {
            const int pr = (int)(gb / p.nblk), rb = (int)(gb - (long long)pr * p.nblk);
            const bool designated = gb == (long long)pr * p.nblk;  // this CTA owns the pair's first block
            if (pr != cur_pair) {
                if (cur_pair >= 0 && !last_round) {
                    float* dst = p.P + (t & 1) * p_stride + ((long long)cur_pair * p.kmax + kidx) * Np;
#pragma unroll
                    for (int k = 0; k < kSkMaxCols; ++k) {
                        const int j = tid + k * kSkThreads;
                        if (j < Np) dst[j] = acc[k];
                    }
                }
                cur_pair = pr;
                const int first = sk_owner((long long)pr * p.nblk, total, G);
                const int kcount = sk_owner((long long)pr * p.nblk + p.nblk - 1, total, G) - first + 1;
                kidx = c - first;
                __syncthreads();  // everyone is done with the previous pair's sv2 / soff
                for (int j = tid; j < Np; j += kSkThreads) {
                    float v = 0.f;
                    if (t >= 2) {
                        const float vprev = t >= 3 ? p.V[((t - 2) & 1) * v_stride + (long long)pr * Np + j] : 0.f;
                        const float* part = p.P + ((t - 1) & 1) * p_stride + (long long)pr * p.kmax * Np + j;
                        float s = 0.f;
                        for (int k = 0; k < kcount; ++k) s += part[(long long)k * Np];
                        v = vprev - logf(s);
                        if (designated) p.V[((t - 1) & 1) * v_stride + (long long)pr * Np + j] = v;
                    }
                    sv2[j] = v * kLog2e;
                    soff[j] = (v - (j < p.N ? norm : log_nu_bin)) * kLog2e;
                }
#pragma unroll
                for (int k = 0; k < kSkMaxCols; ++k) acc[k] = 0.f;
                __syncthreads();
                if (last_round && designated) {
                    // column potentials and the dustbin row of the result
                    const float uM = p.U[(long long)pr * Mp + p.M];
                    float* srow = p.scores + ((long long)pr * Mp + p.M) * Np;
                    for (int j = tid; j < Np; j += kSkThreads) {
                        const float v = sv2[j] * kLn2;
                        if (j < p.N) p.colterm[(long long)pr * p.N + j] = v;
                        srow[j] = alpha + uM + v - norm;
                    }
                }
            }
            if (last_round) {
                // row potentials and the dustbin column for the rows of this block
                for (int r = tid; r < R; r += kSkThreads) {
                    const int i = rb * R + r;
                    if (i < p.M) {
                        const float u = p.U[(long long)pr * Mp + i];
                        p.rowterm[(long long)pr * p.M + i] = u - norm;
                        p.scores[((long long)pr * Mp + i) * Np + p.N] = alpha + u + sv2[p.N] * kLn2 - norm;
                    }
                }
                continue;
            }
            const int sel = resident ? (int)(gb - g0) : cur;
            const float* z = sk_smem + sel * R * pitch;
            if (!resident || t == 1) {
                if (bulk) {
                    mbar_wait(full + sel, sel ? phase1 : phase0);
                    if (sel) phase1 ^= 1; else phase0 ^= 1;
                } else {
                    cp_async_wait_all();
                }
            }
            __syncthreads();  // block data visible; the other buffer is no longer read by anyone
            if (!resident) {
                const bool more = gb + 1 < g1 || t < p.iters;
                if (more)
                    sk_load_block<R>(p, gb + 1 < g1 ? gb + 1 : g0, sk_smem + (cur ^ 1) * R * pitch, full + (cur ^ 1), alpha, bulk);
            }
            // ---- row pass: warp <-> (row, column part); the lane's slice of (Z + v) log2(e) sits in registers ----
            {
                constexpr int parts = 16 / R;
                const int r = warp % R, part = warp / R;
                const int i = rb * R + r;
                const int chunk = (((Np + parts - 1) / parts) + 3) & ~3;
                const int j0 = part * chunk, j1 = min(Np, j0 + chunk);
                float m = -INFINITY, s = 0.f;
                if (i < Mp && j0 < j1) {
                    const float* zr = z + r * pitch + j0;
                    const float* vr = sv2 + j0;
                    const int len = j1 - j0, nf4 = (len + 3) >> 2;
                    const bool ragged = (len & 3) != 0;  // only the last part of a row
                    float vals[4 * kSkRowF4];
#pragma unroll
                    for (int it = 0; it < kSkRowF4; ++it) {
                        const int q = it * 32 + lane;
                        float4 a = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
                        if (q < nf4) {
                            const float4 zz = *reinterpret_cast<const float4*>(zr + 4 * q);
                            const float4 vv = *reinterpret_cast<const float4*>(vr + 4 * q);
                            a = make_float4(fmaf(zz.x, kLog2e, vv.x), fmaf(zz.y, kLog2e, vv.y), fmaf(zz.z, kLog2e, vv.z),
                                            fmaf(zz.w, kLog2e, vv.w));
                            if (ragged && q == nf4 - 1) {
                                if (4 * q + 1 >= len) a.y = -INFINITY;
                                if (4 * q + 2 >= len) a.z = -INFINITY;
                                if (4 * q + 3 >= len) a.w = -INFINITY;
                            }
                        }
                        vals[4 * it] = a.x; vals[4 * it + 1] = a.y; vals[4 * it + 2] = a.z; vals[4 * it + 3] = a.w;
                    }
                    float m0 = vals[0], m1 = vals[1], m2 = vals[2], m3 = vals[3];
#pragma unroll
                    for (int q = 1; q < kSkRowF4; ++q) {
                        m0 = fmaxf(m0, vals[4 * q]); m1 = fmaxf(m1, vals[4 * q + 1]);
                        m2 = fmaxf(m2, vals[4 * q + 2]); m3 = fmaxf(m3, vals[4 * q + 3]);
                    }
                    m = warp_max(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)));
                    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
                    for (int q = 0; q < kSkRowF4; ++q) {
                        s0 += ex2f(vals[4 * q] - m); s1 += ex2f(vals[4 * q + 1] - m);
                        s2 += ex2f(vals[4 * q + 2] - m); s3 += ex2f(vals[4 * q + 3] - m);
                    }
                    s = warp_sum((s0 + s1) + (s2 + s3));
                }
                if (lane == 0) sred[warp] = make_float2(m, s);
            }
            __syncthreads();
            if (tid < R) {
                constexpr int parts = 16 / R;
                const int i = rb * R + tid;
                float mm = -INFINITY;
#pragma unroll
                for (int q = 0; q < parts; ++q) mm = fmaxf(mm, sred[q * R + tid].x);
                float ss = 0.f;
#pragma unroll
                for (int q = 0; q < parts; ++q) {
                    const float2 t2 = sred[q * R + tid];
                    if (t2.x > -INFINITY) ss += t2.y * ex2f(t2.x - mm);
                }
                float u = 0.f;
                if (i < Mp) {
                    u = (i < p.M ? norm : log_mu_bin) - (mm + log2f(ss)) * kLn2;
                    p.U[(long long)pr * Mp + i] = u;
                }
                cu[tid] = u * kLog2e;
            }
            __syncthreads();
            // ---- column pass: thread <-> columns tid, tid + 512, ...: partial sums of exp(Z_ij + u_i - c_j) ----
            {
                const int rcount = min(R, Mp - rb * R);
                float ur[R];
#pragma unroll
                for (int r = 0; r < R; ++r) ur[r] = cu[r];
#pragma unroll
                for (int k = 0; k < kSkMaxCols; ++k) {
                    if (k < ncolk) {  // uniform
                        const int j = tid + k * kSkThreads;
                        const bool ok = j < Np;
                        const int jj = ok ? j : 0;
                        const float off = soff[jj];
                        const float* zc = z + jj;
                        float a0 = 0.f, a1 = 0.f;
                        if (rcount == R) {
#pragma unroll
                            for (int r = 0; r < R; r += 2) {
                                a0 += ex2f(fmaf(zc[r * pitch], kLog2e, ur[r] + off));
                                a1 += ex2f(fmaf(zc[(r + 1) * pitch], kLog2e, ur[r + 1] + off));
                            }
                        } else {
#pragma unroll
                            for (int r = 0; r < R; ++r)
                                if (r < rcount) a0 += ex2f(fmaf(zc[r * pitch], kLog2e, ur[r] + off));
                        }
                        if (ok) acc[k] += a0 + a1;
                    }
                }
            }
            cur ^= 1;
        }
        if (cur_pair >= 0 && !last_round) {
            float* dst = p.P + (t & 1) * p_stride + ((long long)cur_pair * p.kmax + kidx) * Np;
#pragma unroll
            for (int k = 0; k < kSkMaxCols; ++k) {
                const int j = tid + k * kSkThreads;
                if (j < Np) dst[j] = acc[k];
            }
        }
        if (!last_round) grid_barrier(p.barrier, (unsigned int)G * (unsigned int)t);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int sk_owner_host(long long g, long long total, int G) {
    int c = (int)((g * G) / total);
    while (c + 1 < G && ((long long)(c + 1) * total) / G <= g) ++c;
    while (c > 0 && ((long long)c * total) / G > g) --c;
    return c;
}

struct SkPlan {
    int R, nblk, pitch, grid, kmax;
    size_t smem;
};

static int sk_plan(int B, int M, int N, SkPlan* out) {
    int dev = 0, sms = 0;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    B200_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    SkPlan pl;
    pl.pitch = (N + 1 + 3) & ~3;
    pl.R = 16;
    while (pl.R > 2 && ((size_t)(2 * pl.R + 2) * pl.pitch * 4 + 512 > 200 * 1024 ||
                        (N + 1 + 16 / pl.R - 1) / (16 / pl.R) > 128 * kSkRowF4 - 4))
        pl.R >>= 1;
    pl.smem = (size_t)(2 * pl.R + 2) * pl.pitch * 4 + 512;
    B200_REQUIRE(pl.smem <= 227 * 1024, "sinkhorn: N=%d needs %zu bytes of shared memory", N, pl.smem);
    pl.nblk = (M + 1 + pl.R - 1) / pl.R;
    const long long total = (long long)B * pl.nblk;
    pl.grid = (int)(total < sms ? total : sms);
    pl.kmax = 1;
    for (int b = 0; b < B; ++b) {
        const int k = sk_owner_host((long long)b * pl.nblk + pl.nblk - 1, total, pl.grid) -
                      sk_owner_host((long long)b * pl.nblk, total, pl.grid) + 1;
        if (k > pl.kmax) pl.kmax = k;
    }
    *out = pl;
    return 0;
}

static size_t sk_align(size_t x) { return (x + 255) & ~size_t(255); }

size_t sinkhorn_workspace_bytes(int B, int M, int N) {
    SkPlan pl;
    if (B <= 0 || M < 1 || N < 1 || sk_plan(B, M, N, &pl) != 0) return 0;
    return assignment_workspace_bytes(B, M, N) + sk_align(256) + sk_align((size_t)B * (M + 1) * 4) +
           sk_align((size_t)2 * B * (N + 1) * 4) + sk_align((size_t)2 * B * pl.kmax * (N + 1) * 4);
}

int launch_sinkhorn_assignment(const float* cost, const float* bin_score, int B, int M, int N, int iters,
                               float threshold, void* workspace, float* scores, long long* m0, long long* m1,
                               float* ms0, float* ms1, cudaStream_t stream) {
    B200_REQUIRE(M >= 1 && N >= 1 && iters >= 0, "sinkhorn: bad sizes (M=%d, N=%d, iters=%d)", M, N, iters);
    B200_REQUIRE(N + 1 <= kSkMaxCols * kSkThreads, "sinkhorn: N=%d exceeds %d columns", N, kSkMaxCols * kSkThreads - 1);
    if (B == 0) return 0;
    SkPlan pl;
    if (sk_plan(B, M, N, &pl) != 0) return -1;
    uint8_t* w = reinterpret_cast<uint8_t*>(workspace) + assignment_workspace_bytes(B, M, N);
    SinkhornParams p;
    p.cost = cost; p.bin = bin_score;
    p.B = B; p.M = M; p.N = N; p.iters = iters;
    p.R = pl.R; p.nblk = pl.nblk; p.pitch = pl.pitch; p.kmax = pl.kmax;
    p.barrier = reinterpret_cast<unsigned int*>(w); w += sk_align(256);
    p.U = reinterpret_cast<float*>(w); w += sk_align((size_t)B * (M + 1) * 4);
    p.V = reinterpret_cast<float*>(w); w += sk_align((size_t)2 * B * (N + 1) * 4);
    p.P = reinterpret_cast<float*>(w);
    p.scores = scores;
    assignment_terms(workspace, B, M, N, &p.rowterm, &p.colterm);
    // barrier counter and U (u = 0 when iters == 0) start from zero
    B200_CHECK_CUDA(cudaMemsetAsync(p.barrier, 0, sk_align(256) + sk_align((size_t)B * (M + 1) * 4), stream));
    void* kern = nullptr;
    switch (pl.R) {
        case 16: kern = reinterpret_cast<void*>(sinkhorn_kernel<16>); break;
        case 8: kern = reinterpret_cast<void*>(sinkhorn_kernel<8>); break;
        case 4: kern = reinterpret_cast<void*>(sinkhorn_kernel<4>); break;
        default: kern = reinterpret_cast<void*>(sinkhorn_kernel<2>); break;
    }
    B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
    int per_sm = 0;
    B200_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessorWithFlags(&per_sm, kern, kSkThreads, pl.smem, 0));
    B200_REQUIRE(per_sm >= 1, "sinkhorn: kernel does not fit on an SM");
    void* args[] = {&p};
    B200_CHECK_CUDA(cudaLaunchCooperativeKernel(kern, dim3(pl.grid), dim3(kSkThreads), args, pl.smem, stream));
    // inner block Z_ij + u_i + v_j - norm, row / column arg-max, mutual-NN filter (shared with LightGlue)
    return launch_assignment_finalize(cost, 1.0f, B, M, N, threshold, workspace, scores, m0, m1, ms0, ms1, stream);
}

}  // namespace b200
