// K6 (ranks / centred-rank weights), K7 (fitness-weighted reduction over noise-table rows),
// K8 (fused L2 + SGD/Adam update).  Reference: algorithm/nic_nes/nic_nes_master.py:123-136,170-221
// and algorithm/nic_nes/optimizers.py:15-22,44-47,78-83.
#include "common.cuh"
#include <cstdlib>
#include <cstring>
#include <cmath>

namespace nes {

// ---- K6 -----------------------------------------------------------------------------------------
// rank[i] = #{j : x[j] < x[i]} + #{j < i : x[j] == x[i]}   (== stable argsort ranks; the reference's
// default argsort leaves tie order unspecified).  O(n^2) compares, n = 2F <= ~40k.
__global__ void __launch_bounds__(256) k_ranks(const double* __restrict__ x, int64_t n, int64_t* __restrict__ ranks) {
    __shared__ double tile[1024];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const double xi = i < n ? x[i] : 0.0;
    int64_t cnt = 0;
    for (int64_t base = 0; base < n; base += 1024) {
        const int m = (int)((n - base) < 1024 ? (n - base) : 1024);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) tile[t] = x[base + t];
        __syncthreads();
        if (i < n) {
#pragma unroll 8
            for (int t = 0; t < m; ++t) {
                const double xj = tile[t];
                cnt += (xj < xi) || (xj == xi && (base + t) < i);
            }
        }
    }
    if (i < n) ranks[i] = cnt;
}
// centred = rank/(2F-1) - 0.5 (float64, same operation order as nic_nes_master.py:191-193);
// w_i = centred[i,0] - centred[i,1] (:179), cast to float32 where the reference does (:219)
__global__ void k_rank_weights(const int64_t* __restrict__ ranks, int64_t F, double* __restrict__ centered, float* __restrict__ w) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= F) return;
    const double denom = (double)(2 * F - 1);
    const double a = __dsub_rn(__ddiv_rn((double)ranks[2 * i], denom), 0.5);
    const double b = __dsub_rn(__ddiv_rn((double)ranks[2 * i + 1], denom), 0.5);
    if (centered) { centered[2 * i] = a; centered[2 * i + 1] = b; }
    if (w) w[i] = (float)__dsub_rn(a, b);
}

int launch_ranks(Ctx* c, const double* d_x, int64_t n, int64_t* d_ranks) {
    if (n == 0) return 0;
    k_ranks<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_x, n, d_ranks);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}
int launch_rank_weights(Ctx* c, const double* d_fit, int64_t F, double* d_centered, float* d_w) {
    if (F == 0) return 0;
    if (launch_ranks(c, d_fit, 2 * F, c->d_ranks)) return 1;
    k_rank_weights<<<(unsigned)((F + 255) / 256), 256, 0, c->stream>>>(c->d_ranks, F, d_centered, d_w);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

// ---- K6 without any allocation ------------------------------------------------------------------
// nes_rank_weights_static (capi.cu): the same two launches as launch_rank_weights with the int64 rank scratch in a file-scope device
// array instead of the context's growable buffer, for callers that cannot reach cudaMalloc (CUDA graph capture, serving loops).  Every
// GPU that loads the library holds the array (kStaticRankCap * 8 B = 512 KB) until exit.  Every call shares it: k_ranks overwrites
// entries [0, 2F) before k_rank_weights reads them, nothing in it has to start at zero, and two calls that overlap on the GPU (different
// streams, or two contexts on one device) race on it.
static __device__ __align__(256) long long g_static_ranks[kStaticRankCap];

// host side: `g_static_ranks` here is a host stand-in for the symbol, not a device address -- the address comes from
// cudaGetSymbolAddress, once per device, with that device current
static int static_ranks_ptr(int device, int64_t** out) {
    static int64_t* ptr[64] = {nullptr};
    NES_CHECK(device >= 0 && device < 64, "device index out of range");
    if (!ptr[device]) {
        int cur = -1;
        NES_CUDA(cudaGetDevice(&cur));
        if (cur != device) NES_CUDA(cudaSetDevice(device));
        void* a = nullptr;
        NES_CUDA(cudaGetSymbolAddress(&a, g_static_ranks));
        NES_CHECK(a != nullptr, "no device address for the static rank scratch");
        ptr[device] = reinterpret_cast<int64_t*>(a);
        if (cur != device) NES_CUDA(cudaSetDevice(cur));
    }
    *out = ptr[device];
    return 0;
}
int launch_rank_weights_static(Ctx* c, const double* d_fit, int64_t F, double* d_centered, float* d_w) {
    NES_CHECK(2 * F <= kStaticRankCap, "nes_rank_weights_static: F exceeds NES_STATIC_RANK_MAX_F (32768 antithetic pairs)");
    if (F == 0) return 0;
    int64_t* d_ranks = nullptr;
    if (static_ranks_ptr(c->device, &d_ranks)) return 1;
    k_ranks<<<(unsigned)((2 * F + 255) / 256), 256, 0, c->stream>>>(d_fit, 2 * F, d_ranks);
    k_rank_weights<<<(unsigned)((F + 255) / 256), 256, 0, c->stream>>>(d_ranks, F, d_centered, d_w);
    c->launches += 2;
    NES_CUDA(cudaGetLastError());
    return 0;
}

// ---- K7 -----------------------------------------------------------------------------------------
// g[p] = f32(sigma) * (sum_i w_i * eps_i[p]) / (2F), eps_i = table[off_i + pstart_k + (p - start_k)].
// HBM-bound: reads F * D * 2 bytes (fp16 table) once, writes 4*D.  One thread owns 8 consecutive
// elements (one 128-bit load per row); the row loop is unrolled so 16 independent loads are in flight per thread (measured on B200,
// tools/k7_sweep.py: 8 -> 16 loads +7 %, 32 no further gain; the block size does not matter).
struct GradPlan { int n; int64_t vec_prefix[kMaxTensors + 1]; };

__device__ __forceinline__ uint4 ld_stream_128(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void fma8(float (&acc)[8], const uint4& v, float w) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        float2 f = __half22float2(h[e]);
        acc[2 * e]     = fmaf(w, f.x, acc[2 * e]);
        acc[2 * e + 1] = fmaf(w, f.y, acc[2 * e + 1]);
    }
}

constexpr int kGradChunk = 512;
// Rows in table order.  A noise row is D_pad elements long and the F offsets lie L / F apart on average: at F = 5000 on the 2^32-element
// table every table element belongs to ~16 rows (F * D_pad / L).  Summed in the order they were sampled, those 16 reads are hours apart
// and every one comes from HBM.  Sorted by offset, the rows that share a table line are neighbours in the loop: thread column p at row i
// reads table[off_i + p], and column p - (off_{i+1} - off_i) reads the same line one row later -- an L2 hit as long as both columns
// are resident at the same time.  k_sort_rows puts (offset, weight) in ascending offset order (ties: by index, so the order is a
// function of the offsets alone and identical on every rank); O(F^2) compares like k_ranks.
__global__ void __launch_bounds__(256) k_sort_rows(const int64_t* __restrict__ off, const float* __restrict__ w, int64_t F,
                                                   int64_t* __restrict__ off_sorted, float* __restrict__ w_sorted) {
    __shared__ int64_t tile[1024];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t oi = i < F ? off[i] : 0;
    int64_t cnt = 0;
    for (int64_t base = 0; base < F; base += 1024) {
        const int m = (int)((F - base) < 1024 ? (F - base) : 1024);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) tile[t] = off[base + t];
        __syncthreads();
        if (i < F) {
#pragma unroll 8
            for (int t = 0; t < m; ++t) {
                const int64_t oj = tile[t];
                cnt += (oj < oi) || (oj == oi && (base + t) < i);
            }
        }
    }
    if (i < F) { off_sorted[cnt] = oi; w_sorted[cnt] = w[i]; }
}

// One thread owns NC vectors of 8 consecutive elements, 256 vectors apart (a CTA covers NC * 2048 consecutive elements of every row), and
// keeps U rows of all of them in flight.  NC > 1 widens the band of columns that is resident on the chip at one time (148 SMs x resident
// threads x NC x 8 elements), which is what turns the overlap of neighbouring sorted rows into L2 hits.
template <int NC, int U>
__global__ void __launch_bounds__(256, (NC * U >= 16 && NC > 1) ? 1 : 2) k_gradient(const __half* __restrict__ table, const int64_t* __restrict__ off,
                                                  const float* __restrict__ w, int64_t F, GradPlan plan,
                                                  const Layout* __restrict__ lay, float sigma, float* __restrict__ g, int64_t v_lo, int64_t v_hi, int span) {
    __shared__ int64_t s_off[kGradChunk];
    __shared__ float s_w[kGradChunk];
    // CTA b owns vectors [v_lo + b * span, v_lo + (b + 1) * span) below v_hi, span <= 256 * NC (the host balances span so that every
    // launch is one full wave of co-resident CTAs, see launch_gradient)
    const int64_t v0 = v_lo + (int64_t)blockIdx.x * span + threadIdx.x;
    uint32_t pos[NC]; bool act[NC];            // only the row positions (< D_pad) stay live through the loop; the tensor lookup is redone at the end
    float acc[NC][8];
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const int64_t v = v0 + c * 256;
        act[c] = (c * 256 + (int)threadIdx.x) < span && v < v_hi;
        int k = 0;
        if (act[c]) { while (v >= plan.vec_prefix[k + 1]) ++k; }
        pos[c] = (uint32_t)(lay->t[k].pstart + (act[c] ? (v - plan.vec_prefix[k]) * 8 : 0));
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[c][e] = 0.f;
    }
    for (int64_t c0 = 0; c0 < F; c0 += kGradChunk) {
        const int m = (int)((F - c0) < kGradChunk ? (F - c0) : kGradChunk);
        __syncthreads();
        for (int i = threadIdx.x; i < m; i += blockDim.x) { s_off[i] = off[c0 + i]; s_w[i] = w[c0 + i]; }
        __syncthreads();
        int i = 0;
        for (; i + U <= m; i += U) {
            uint4 buf[U][NC];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t o = s_off[i + u];
#pragma unroll
                for (int c = 0; c < NC; ++c) if (act[c]) buf[u][c] = ld_stream_128(table + (o + pos[c]));
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const float wi = s_w[i + u];
#pragma unroll
                for (int c = 0; c < NC; ++c) if (act[c]) fma8(acc[c], buf[u][c], wi);
            }
        }
        for (; i < m; ++i) {
            const int64_t o = s_off[i]; const float wi = s_w[i];
#pragma unroll
            for (int c = 0; c < NC; ++c) if (act[c]) { const uint4 b = ld_stream_128(table + (o + pos[c])); fma8(acc[c], b, wi); }
        }
    }
    const float denom = (float)(2 * F);
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        if (!act[c]) continue;
        const int64_t v = v0 + c * 256;
        int k = 0;
        while (v >= plan.vec_prefix[k + 1]) ++k;
        const TensorDesc t = lay->t[k];
        const int64_t j0 = (v - plan.vec_prefix[k]) * 8;
#pragma unroll
        for (int e = 0; e < 8; ++e)
            if (j0 + e < t.count) g[t.start + j0 + e] = __fdiv_rn(__fmul_rn(sigma, acc[c][e]), denom);
    }
}

// with a safe-mutation scale the reference sums w_i * (sigma eps_i (.) s); the scale is common to all rows, so it is applied
// once to the sum (agrees to float32 round-off, within the 1e-5 the update is held to)
__global__ void __launch_bounds__(256) k_scale_vec(float* __restrict__ g, const float* __restrict__ s, int64_t D, int div) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D; i += (int64_t)gridDim.x * blockDim.x) g[i] = apply_scale(g[i], s[i], div);
}
static int sm_count_of(Ctx* c) {
    static int n[64] = {0};
    if (c->device < 0 || c->device >= 64) return 148;
    if (!n[c->device]) { int v = 0; if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, c->device) != cudaSuccess || v < 1) { cudaGetLastError(); v = 148; } n[c->device] = v; }
    return n[c->device];
}
int launch_gradient(Ctx* c, const int64_t* d_off, const float* d_w, int64_t F, float sigma, float* d_g, int shard, int n_shards) {
    GradPlan plan; plan.n = c->lay.n; plan.vec_prefix[0] = 0;
    for (int k = 0; k < c->lay.n; ++k) plan.vec_prefix[k + 1] = plan.vec_prefix[k] + (c->lay.t[k].count + 7) / 8;
    const int64_t total = plan.vec_prefix[plan.n];
    if (F == 0) { NES_CUDA(cudaMemsetAsync(d_g, 0, sizeof(float) * c->lay.D, c->stream)); return 0; }
    // multi-GPU: the parameter axis is cut into n_shards runs of vectors; every element is summed over all F rows by one rank in the
    // single-GPU order (table order, whatever the launch shape), the other ranks hold zeros, so the all-reduce that follows reproduces the
    // single-GPU vector bit for bit
    NES_CHECK(n_shards >= 1 && shard >= 0 && shard < n_shards, "bad gradient shard");
    // NES_K7 = "<vectors per thread>x<rows in flight>" and NES_K7_SORT=0 are experiment switches (tools/k7_sweep.py).  Measured on B200
    // (F = 5000, D = 13.4 M, 2^32-element table, one box): rows in the order they were sampled 18.4 ms = 7.3 TB/s whatever the shape (ncu:
    // 134.7 GB from DRAM for 134.1 GB algorithmic); rows sorted by offset: 1x16 17.9, 2x4 15.1, 4x2 13.0, 8x1 12.4 ms = 10.8 TB/s on the
    // algorithmic bytes (1.35 of the 8 TB/s nominal HBM rate: most reads are L2 hits; the same kernel with all rows packed into 256 MB,
    // i.e. its L2 -> SM ceiling, takes 11.7 ms).  More rows in flight per thread at the same width cost registers, hence residency: 4x4 /
    // 8x2 fall back to 19 ms.  Holding the rows in flight in shared memory instead (cp.async into a private ring per thread, 8 vectors x
    // 3 - 6 rows, 128 - 192 KB in flight per SM) was tried and lost: 14.3 - 16.6 ms, same 11.5 ms ceiling.  The shape follows
    // from the band width below (8x1 whenever a band of 2 CTAs per SM needs more than 1024 vectors per CTA).
    int env_nc = 0, env_u = 0, ksort = 1;            // read on every call: the tests walk through the shapes in one process
    {
        const char* u = getenv("NES_K7");
        if (u && strlen(u) >= 3 && u[1] == 'x') { env_nc = u[0] - '0'; env_u = atoi(u + 2); }
        const char* so = getenv("NES_K7_SORT");
        if (so && atoi(so) == 0) ksort = 0;
    }
    // this rank's slice of the parameter axis, in vectors of 8 elements
    const int64_t per = (total + n_shards - 1) / n_shards;
    const int64_t s_lo = std::min(total, (int64_t)shard * per), s_hi = std::min(total, s_lo + per);
    if (n_shards > 1) NES_CUDA(cudaMemsetAsync(d_g, 0, sizeof(float) * c->lay.D, c->stream));
    if (c->grad_timed) NES_CUDA(cudaEventRecord(c->ev_a, c->stream));
    if (ksort && s_hi > s_lo) {
        k_sort_rows<<<(unsigned)((F + 255) / 256), 256, 0, c->stream>>>(d_off, d_w, F, c->d_off_sorted, c->d_w_sorted);
        c->launches++;
        d_off = c->d_off_sorted; d_w = c->d_w_sorted;
    }
    // One launch per BAND of columns, every band exactly one wave of co-resident CTAs (2 per SM) with the vectors spread evenly over
    // them.  A single launch of ceil(vectors / 2048) CTAs leaves a ragged last wave (D = 5 M: 305 CTAs on 296 slots took two full
    // waves) and, over many waves, CTAs that start whenever a slot frees up: the resident CTAs then sit at unrelated rows and stop
    // sharing table lines (D = 50 M: 6.5 TB/s).  Bands keep every resident CTA on the same rows.
    // Band width: a table line is first touched by the band's highest column and last by its lowest, (band width / offset spacing) rows
    // later; everything the band reads in between (that many rows x the band's bytes) has to stay in L2, so the width that still shares
    // grows with the square root of the spacing.  Measured at D = 13.4 M on the 2^32 table: F = 5000 (spacing 0.86 M elements) 4.85 M
    // columns = 8 vectors per thread is best (12.3 ms; 2.4 M columns 13.0); F = 20000 (spacing 0.21 M) 2.4 M columns (41 ms; 4.85 M: 60 ms).
    const int slots = 2 * sm_count_of(c);
    const int64_t n_vec = s_hi - s_lo;
    const double spacing = (double)std::max<int64_t>(1, c->table_len - c->lay.D_pad) / (double)F;
    const int64_t span_max = std::min<int64_t>(2048, std::max<int64_t>(256, (int64_t)(5200.0 * sqrt(spacing) / (8.0 * slots))));
    const int64_t waves = std::max<int64_t>(1, (n_vec + (int64_t)slots * span_max - 1) / ((int64_t)slots * span_max));
    int span = (int)((n_vec + waves * slots - 1) / (waves * slots));                 // vectors per CTA, <= 2048
    if (span < 1) span = 1;
    int knc = env_nc, ku = env_u;
    if (!knc) {                             // as many vectors per thread as the band needs, and rows in flight so that a thread keeps 8 - 16 loads going
        static const int u_of[9] = {0, 16, 4, 4, 2, 2, 2, 1, 1};
        knc = (span + 255) / 256;
        ku = u_of[knc];
    }
    if (span > 256 * knc) span = 256 * knc;                                           // a narrower shape was forced: more bands
#define NES_K7_LAUNCH(NC, U) k_gradient<NC, U><<<grid, 256, 0, c->stream>>>(c->table, d_off, d_w, F, plan, c->d_lay, sigma, d_g, v_lo, s_hi, span)
    for (int64_t v_lo = s_lo; v_lo < s_hi; v_lo += (int64_t)slots * span) {
        const int64_t left = s_hi - v_lo;
        const unsigned grid = (unsigned)std::min<int64_t>(slots, (left + span - 1) / span);
        if (knc == 1 && ku == 16) NES_K7_LAUNCH(1, 16);
        else if (knc == 1 && ku == 8) NES_K7_LAUNCH(1, 8);
        else if (knc == 2 && ku == 8) NES_K7_LAUNCH(2, 8);
        else if (knc == 2 && ku == 4) NES_K7_LAUNCH(2, 4);
        else if (knc == 4 && ku == 2) NES_K7_LAUNCH(4, 2);
        else if (knc == 4 && ku == 4) NES_K7_LAUNCH(4, 4);
        else if (knc == 8 && ku == 2) NES_K7_LAUNCH(8, 2);
        else if (knc == 8 && ku == 1) NES_K7_LAUNCH(8, 1);
        else if (knc == 3 && ku == 4) NES_K7_LAUNCH(3, 4);
        else if (knc == 5 && ku == 2) NES_K7_LAUNCH(5, 2);
        else if (knc == 6 && ku == 2) NES_K7_LAUNCH(6, 2);
        else if (knc == 7 && ku == 1) NES_K7_LAUNCH(7, 1);
        else { NES_CHECK(false, "NES_K7: unsupported shape (1x16 1x8 2x8 2x4 3x4 4x4 4x2 5x2 6x2 7x1 8x2 8x1)"); }
        c->launches++;
    }
#undef NES_K7_LAUNCH
    if (c->scale32) { k_scale_vec<<<148 * 8, 256, 0, c->stream>>>(d_g, c->scale32, c->lay.D, c->scale_div); c->launches++; }
    if (c->grad_timed) NES_CUDA(cudaEventRecord(c->ev_b, c->stream));
    NES_CUDA(cudaGetLastError());
    return 0;
}

// ---- K8 -----------------------------------------------------------------------------------------
// One pass: G = -g + l2*theta ; Adam / SGD-momentum step in float64 with the reference's operation
// order (no FMA contraction) ; theta64' = theta64 + step ; theta32' = f32(theta64') ; partial sums of
// |step|^2 and |theta|^2 for Optimizer.update's ratio (optimizers.py:19).
// GMODE says where G comes from and which dtype it has IN THE REFERENCE, because numpy keeps
// (1-beta)*G and G*G in float32 when G is a float32 array (python-scalar * float32 array):
//   0 fused, theta has float64 provenance: G = -(double)g + l2*theta               (float64; iterations >= 2)
//   1 fused, theta has float32 provenance: G = f32(-g + f32(l2)*f32(theta))        (float32; first iteration)
//   2 direct float32 G (Optimizer.update(globalg) with a float32 array)
//   3 direct float64 G
template <int KIND, int GMODE>
__global__ void __launch_bounds__(256) k_opt_update(const float* __restrict__ g, const double* __restrict__ G64,
                                                    double* __restrict__ t64, float* __restrict__ t32,
                                                    double* __restrict__ m, double* __restrict__ v, int64_t D, double l2,
                                                    double a_or_lr, double b1, double b2, double eps,
                                                    double* __restrict__ partials) {
    double s_step = 0.0, s_theta = 0.0;
    const float l2f = (float)l2;
    const double omb1 = 1.0 - b1, omb2 = 1.0 - b2;
    const float omb1f = (float)omb1, omb2f = (float)omb2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D; i += (int64_t)gridDim.x * blockDim.x) {
        const double th = t64[i];
        double G, gm, gv;       // G, (1-b1)*G, (1-b2)*G*G  with the reference's dtypes
        if (GMODE == 1 || GMODE == 2) {
            const float Gf = (GMODE == 1) ? __fadd_rn(-g[i], __fmul_rn(l2f, (float)th)) : g[i];
            G = (double)Gf;
            gm = (double)__fmul_rn(omb1f, Gf);
            gv = (double)__fmul_rn(omb2f, __fmul_rn(Gf, Gf));
        } else {
            G = (GMODE == 0) ? __dadd_rn(-(double)g[i], __dmul_rn(l2, th)) : G64[i];
            gm = __dmul_rn(omb1, G);
            gv = __dmul_rn(omb2, __dmul_rn(G, G));
        }
        double step;
        if (KIND == NES_OPT_ADAM) {
            const double mi = __dadd_rn(__dmul_rn(b1, m[i]), gm);
            const double vi = __dadd_rn(__dmul_rn(b2, v[i]), gv);
            m[i] = mi; v[i] = vi;
            step = __ddiv_rn(__dmul_rn(-a_or_lr, mi), __dadd_rn(__dsqrt_rn(vi), eps));
        } else {
            const double vi = __dadd_rn(__dmul_rn(b1, v[i]), gm);
            v[i] = vi;
            step = __dmul_rn(-a_or_lr, vi);
        }
        const double nt = __dadd_rn(th, step);
        t64[i] = nt;
        t32[i] = (float)nt;
        s_step = fma(step, step, s_step);
        s_theta = fma(th, th, s_theta);
    }
    __shared__ double sh[2][8];
    for (int o = 16; o > 0; o >>= 1) {
        s_step += __shfl_xor_sync(0xffffffffu, s_step, o);
        s_theta += __shfl_xor_sync(0xffffffffu, s_theta, o);
    }
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sh[0][wid] = s_step; sh[1][wid] = s_theta; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0;
        for (int q = 0; q < 8; ++q) { a += sh[0][q]; b += sh[1][q]; }
        partials[2 * blockIdx.x] = a; partials[2 * blockIdx.x + 1] = b;
    }
}
// one block: the per-block partial sums of |step|^2 and |theta|^2 reduced in a fixed order (thread t sums partials t, t + 256, ...; then a
// shuffle tree), so the ratio is reproducible run to run
__global__ void __launch_bounds__(256) k_norm_finalize(const double* __restrict__ partials, int nblocks, double* __restrict__ norms) {
    __shared__ double sh[2][8];
    double a = 0, b = 0;
    for (int q = threadIdx.x; q < nblocks; q += 256) { a += partials[2 * q]; b += partials[2 * q + 1]; }
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = a; sh[1][threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = 0; b = 0;
        for (int q = 0; q < 8; ++q) { a += sh[0][q]; b += sh[1][q]; }
        norms[0] = sqrt(a); norms[1] = sqrt(b);
    }
}

template <int KIND>
static void opt_dispatch(Ctx* c, int gmode, const float* d_g, const double* d_G64, double l2, double a_or_lr) {
    const int grid = c->norm_blocks;
#define NES_OPT_LAUNCH(GM) k_opt_update<KIND, GM><<<grid, 256, 0, c->stream>>>(d_g, d_G64, c->theta64, c->theta32, c->opt_m, c->opt_v, \
        c->lay.D, l2, a_or_lr, c->opt_b1, c->opt_b2, c->opt_eps, c->d_norm_partials)
    switch (gmode) { case 0: NES_OPT_LAUNCH(0); break; case 1: NES_OPT_LAUNCH(1); break; case 2: NES_OPT_LAUNCH(2); break; default: NES_OPT_LAUNCH(3); }
#undef NES_OPT_LAUNCH
}

// direct: 0 = fused from the gradient estimate g (nic_nes_master.py:123-133), 1 = d_g IS globalg (float32), 2 = d_G64 is globalg (float64)
int launch_opt_update(Ctx* c, const float* d_g, const double* d_G64, int direct, double l2, double* ratio_out) {
    NES_CHECK(c->opt_kind == NES_OPT_ADAM || c->opt_kind == NES_OPT_SGD, "nes_opt_init has not been called");
    if (refresh_theta64(c, 0)) return 1;
    c->opt_t += 1;
    const int gmode = direct == 0 ? (c->theta_is_f32[0] ? 1 : 0) : (direct == 1 ? 2 : 3);
    if (c->opt_kind == NES_OPT_ADAM) {
        const double t = (double)c->opt_t;
        const double a = c->opt_stepsize * sqrt(1.0 - pow(c->opt_b2, t)) / (1.0 - pow(c->opt_b1, t));
        opt_dispatch<NES_OPT_ADAM>(c, gmode, d_g, d_G64, l2, a);
    } else {
        opt_dispatch<NES_OPT_SGD>(c, gmode, d_g, d_G64, l2, c->opt_stepsize);
    }
    c->launches++;
    NES_CUDA(cudaGetLastError());
    k_norm_finalize<<<1, 256, 0, c->stream>>>(c->d_norm_partials, c->norm_blocks, c->d_norms);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    c->theta_is_f32[0] = 0;
    if (launch_theta_shadow(c, 0)) return 1;
    if (ratio_out) {
        double h[2];
        NES_CUDA(cudaMemcpyAsync(h, c->d_norms, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
        NES_CUDA(cudaStreamSynchronize(c->stream));
        *ratio_out = h[0] / h[1];
    }
    return 0;
}

}  // namespace nes
