// K1: shared noise table, offset / parent sampling, theta copies, member materialisation.
//
// The reference has no noise table: every worker draws torch.empty_like(theta).normal_(0, sigma)
// (algorithm/nets.py:92-102) and ships the float32 vector through Redis
// (nic_nes_worker.py:156-161).  Here eps lives once in HBM as an fp16 table addressed by an int64
// offset; the specification is the header of oracle/noise.py and this file is its CUDA statement.
#include "common.cuh"

namespace nes {

__global__ void k_fill_f32(float* __restrict__ p, int64_t n, float v);

// ---- K1: table fill ---------------------------------------------------------------------------
// one thread = one Philox block = 4 table elements (one 8-byte store).  Box-Muller in float64 so
// the fp16 results equal the numpy restatement bit for bit.
__global__ void __launch_bounds__(256) k_fill_table(__half* __restrict__ table, int64_t n_blocks, uint32_t k0, uint32_t k1) {
    const double two_m32 = 2.3283064365386963e-10;           // 2^-32
    const double two_pi  = 6.283185307179586;                // float64(2*pi), as (2.0*math.pi) in the spec
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_blocks; b += (int64_t)gridDim.x * blockDim.x) {
        uint32_t r[4];
        philox4x32_10((uint32_t)b, (uint32_t)((uint64_t)b >> 32), 0u, 0u, k0, k1, r);
        __half h[4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            double u1 = __dmul_rn(__dadd_rn((double)r[2 * j], 0.5), two_m32);
            double u2 = __dmul_rn(__dadd_rn((double)r[2 * j + 1], 0.5), two_m32);
            double rad = sqrt(__dmul_rn(-2.0, log(u1)));
            double ang = __dmul_rn(two_pi, u2);
            double s, c;
            sincos(ang, &s, &c);
            h[2 * j]     = __float2half_rn(__double2float_rn(__dmul_rn(rad, c)));
            h[2 * j + 1] = __float2half_rn(__double2float_rn(__dmul_rn(rad, s)));
        }
        uint2 packed;
        packed.x = (uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16);
        packed.y = (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16);
        *reinterpret_cast<uint2*>(table + 4 * b) = packed;
    }
}

int launch_fill_table(Ctx* c) {
    int64_t n_blocks = c->table_len / 4;
    int grid = 148 * 16;
    k_fill_table<<<grid, 256, 0, c->stream>>>(c->table, n_blocks, (uint32_t)c->seed, (uint32_t)(c->seed >> 32));
    c->launches++;
    NES_CUDA(cudaGetLastError());
    if (c->is_fc) return launch_bound_norms(c, -1);
    return 0;
}

// ---- offsets / parents ------------------------------------------------------------------------
__global__ void k_sample_offsets(int64_t* __restrict__ out, int64_t F, uint32_t it_lo, uint32_t it_hi,
                                 uint32_t k0, uint32_t k1, uint64_t n_slots, int64_t G) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= F) return;
    uint32_t r[4];
    philox4x32_10((uint32_t)i, it_lo, it_hi, kOffsetDomain, k0, k1, r);
    uint64_t slot = ((uint64_t)r[0] * n_slots) >> 32;
    out[i] = (int64_t)slot * G;
}
__global__ void k_sample_parents(int32_t* __restrict__ out, int64_t F, uint32_t it_lo, uint32_t it_hi,
                                 uint32_t k0, uint32_t k1, uint32_t n_parents) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= F) return;
    uint32_t r[4];
    philox4x32_10((uint32_t)i, it_lo, it_hi, kParentDomain, k0, k1, r);
    out[i] = (int32_t)(((uint64_t)r[0] * n_parents) >> 32);
}

// tournament selection (nic_es_worker.py:150-158: rs.choice(len(parents), min(len(parents), tournament_size), replace=False).min();
// the parents are sorted best first, so the winner is the lowest index drawn).  A uniformly random t-subset by Floyd's algorithm:
// for j = n-t .. n-1: r = draw in [0, j]; take r unless it is already in the set, else take j.  Draw d of offspring i is word d % 4 of
// philox(counter = (i, lo32(it), hi32(it) + ((d / 4) << 16), "TOUR"), key(seed)), mapped by (r * (j + 1)) >> 32.  Spec: oracle/noise.py.
constexpr int kTournMaxParents = 4096;
__global__ void k_sample_parents_tournament(int32_t* __restrict__ out, int64_t F, uint32_t it_lo, uint32_t it_hi,
                                            uint32_t k0, uint32_t k1, uint32_t n, uint32_t t) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= F) return;
    uint64_t mask[kTournMaxParents / 64];
    for (uint32_t w = 0; w < (n + 63) / 64; ++w) mask[w] = 0ull;
    uint32_t best = n, r4[4] = {0, 0, 0, 0};
    for (uint32_t d = 0; d < t; ++d) {
        if ((d & 3) == 0) philox4x32_10((uint32_t)i, it_lo, it_hi + ((d >> 2) << 16), kTournDomain, k0, k1, r4);
        const uint32_t j = n - t + d;
        uint32_t pick = (uint32_t)(((uint64_t)r4[d & 3] * (uint64_t)(j + 1)) >> 32);
        if ((mask[pick >> 6] >> (pick & 63)) & 1ull) pick = j;
        mask[pick >> 6] |= 1ull << (pick & 63);
        best = pick < best ? pick : best;
    }
    out[i] = (int32_t)best;
}

int launch_sample_parents_tournament(Ctx* c, uint64_t iteration, int64_t F, int32_t n_parents, int32_t tournament_size, int32_t* d_out) {
    if (F == 0) return 0;
    NES_CHECK(n_parents >= 1 && n_parents <= kTournMaxParents, "tournament selection handles 1..4096 parents");
    NES_CHECK(tournament_size >= 1, "tournament_size must be >= 1");
    NES_CHECK((iteration >> 48) == 0, "iteration must be below 2^48");
    const int32_t t = tournament_size < n_parents ? tournament_size : n_parents;
    k_sample_parents_tournament<<<(unsigned)((F + 127) / 128), 128, 0, c->stream>>>(d_out, F, (uint32_t)iteration, (uint32_t)(iteration >> 32),
                                                                                  (uint32_t)c->seed, (uint32_t)(c->seed >> 32), (uint32_t)n_parents, (uint32_t)t);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

int launch_sample_offsets(Ctx* c, uint64_t iteration, int64_t F, int64_t* d_out) {
    if (F == 0) return 0;
    int64_t n_slots = (c->table_len - c->lay.D_pad) / c->lay.G + 1;
    NES_CHECK(n_slots >= 1 && n_slots < (1ll << 32), "noise table too small (or too large) for this layout");
    k_sample_offsets<<<(unsigned)((F + 255) / 256), 256, 0, c->stream>>>(d_out, F, (uint32_t)iteration, (uint32_t)(iteration >> 32),
                                                                       (uint32_t)c->seed, (uint32_t)(c->seed >> 32), (uint64_t)n_slots, c->lay.G);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}
int launch_sample_parents(Ctx* c, uint64_t iteration, int64_t F, int32_t n_parents, int32_t* d_out) {
    if (F == 0) return 0;
    NES_CHECK(n_parents >= 1, "n_parents must be >= 1");
    k_sample_parents<<<(unsigned)((F + 255) / 256), 256, 0, c->stream>>>(d_out, F, (uint32_t)iteration, (uint32_t)(iteration >> 32),
                                                                       (uint32_t)c->seed, (uint32_t)(c->seed >> 32), (uint32_t)n_parents);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

// ---- theta copies -----------------------------------------------------------------------------
// split x*scale into fp16 hi + lo (hi = rn(x*scale), lo = rn(x*scale - hi))
__device__ __forceinline__ void split_half(float x, float scale, __half& hi, __half& lo) {
    float xs = x * scale;
    hi = __float2half_rn(xs);
    lo = __float2half_rn(xs - __half2float(hi));
}

// theta32 [D] (flat) -> theta_hi / theta_lo [D_pad] (padded layout, zero in the padding)
__global__ void __launch_bounds__(256) k_theta_shadow(const float* __restrict__ th, __half* __restrict__ hi, __half* __restrict__ lo,
                                                      const Layout* __restrict__ lay) {
    const int k = blockIdx.y;
    const TensorDesc t = lay->t[k];
    const int64_t pend = (k + 1 < lay->n) ? lay->t[k + 1].pstart : lay->D_pad;
    const int64_t span = pend - t.pstart;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < span; j += (int64_t)gridDim.x * blockDim.x) {
        __half a = __float2half_rn(0.f), b = a;
        if (j < t.count) split_half(th[t.start + j], kThetaScale, a, b);
        hi[t.pstart + j] = a;
        lo[t.pstart + j] = b;
    }
}
int launch_theta_shadow(Ctx* c, int parent) {
    dim3 grid(148 * 2, c->lay.n);
    k_theta_shadow<<<grid, 256, 0, c->stream>>>(c->theta32 + (int64_t)parent * c->lay.D,
                                                c->theta_hi + (int64_t)parent * c->lay.D_pad,
                                                c->theta_lo + (int64_t)parent * c->lay.D_pad, c->d_lay);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    if (c->mut_kind == NES_MUT_PROPORTIONAL && launch_prop_scale(c, parent)) return 1;
    if (c->is_fc) return launch_bound_norms(c, parent);
    return 0;
}

template <typename T>
__global__ void __launch_bounds__(256) k_set_theta(const T* __restrict__ src, double* __restrict__ t64, float* __restrict__ t32, int64_t D) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D; i += (int64_t)gridDim.x * blockDim.x) {
        T v = src[i];
        t64[i] = (double)v;
        t32[i] = (float)v;
    }
}
int launch_set_theta(Ctx* c, int parent, const void* d_src, bool is_f64) {
    const int64_t D = c->lay.D;
    int grid = (int)((D + 255) / 256 < 148 * 8 ? (D + 255) / 256 : 148 * 8);
    if (grid < 1) grid = 1;
    if (is_f64) k_set_theta<double><<<grid, 256, 0, c->stream>>>((const double*)d_src, c->theta64 + parent * D, c->theta32 + parent * D, D);
    else        k_set_theta<float><<<grid, 256, 0, c->stream>>>((const float*)d_src, c->theta64 + parent * D, c->theta32 + parent * D, D);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    c->theta_is_f32[parent] = is_f64 ? 0 : 1;
    c->theta64_stale[parent] = 0;
    return launch_theta_shadow(c, parent);
}

// ---- member materialisation: PolicyNet.evolve (algorithm/nets.py:101-119) ----------------------
// noise = f32(sigma) * f32(eps) ; member = theta + noise  (sign<0: theta - noise, nic_nes_worker.py:151)
__global__ void __launch_bounds__(256) k_evolve(const float* __restrict__ th, const __half* __restrict__ row, const Layout* __restrict__ lay,
                                                float sign, float sigma, float* __restrict__ member, float* __restrict__ noise,
                                                double* __restrict__ dst64, float* __restrict__ dst32, const float* __restrict__ scale, int scale_div) {
    const TensorDesc t = lay->t[blockIdx.y];
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < t.count; j += (int64_t)gridDim.x * blockDim.x) {
        float nz = mut_noise(sigma, __half2float(row[t.pstart + j]), scale, t.start + j, scale_div);
        float th0 = th[t.start + j];
        float m = sign > 0.f ? __fadd_rn(th0, nz) : __fsub_rn(th0, nz);
        if (noise)  noise[t.start + j] = nz;
        if (member) member[t.start + j] = m;
        if (dst32) { dst32[t.start + j] = m; dst64[t.start + j] = (double)m; }
    }
}
int launch_evolve(Ctx* c, int parent, int64_t offset, float sign, float sigma, float* d_member, float* d_noise, int dst_parent) {
    const int64_t D = c->lay.D;
    NES_CHECK(offset >= 0 && offset + c->lay.D_pad <= c->table_len && offset % c->lay.G == 0, "offset outside the noise table or not slot aligned");
    NES_CHECK(dst_parent != parent, "dst_parent must differ from the source parent");
    NES_CHECK(!(c->mut_kind == NES_MUT_SENSITIVITY && sigma != 0.f && c->scale_stale[parent]) || d_noise != nullptr || d_member != nullptr,
              "stale sensitivity: call nes_set_sensitivity for this parent first");
    dim3 grid(148 * 2, c->lay.n);
    k_evolve<<<grid, 256, 0, c->stream>>>(c->theta32 + parent * D, c->table + offset, c->d_lay, sign, sigma, d_member, d_noise,
                                          dst_parent >= 0 ? c->theta64 + dst_parent * D : nullptr,
                                          dst_parent >= 0 ? c->theta32 + dst_parent * D : nullptr,
                                          c->scale32 ? c->scale32 + parent * D : nullptr, c->scale_div);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    if (dst_parent >= 0) {
        c->theta_is_f32[dst_parent] = 1; c->theta64_stale[dst_parent] = 0;
        if (c->mut_kind == NES_MUT_SENSITIVITY) {          // the new parent's sensitivity is not its slot's old one (see evolve_batch_post)
            k_fill_f32<<<148 * 2, 256, 0, c->stream>>>(c->scale32 + (int64_t)dst_parent * D, D, nanf(""));
            c->launches++;
            c->scale_stale[dst_parent] = 1;
        }
        return launch_theta_shadow(c, dst_parent);
    }
    return 0;
}

// ---- NIC-ES survivor materialisation, all survivors in one launch: item i becomes parent dst[i] = theta32[src[i]] + sigma_i * eps(off_i)
// (nic_es_worker.py:188 without the .pth file); writes the float32 copy and the padded split-fp16 shadow in one pass (the float64 master
// copy is marked stale and rebuilt only if somebody asks for it)
__global__ void __launch_bounds__(256) k_widen(const float* __restrict__ src, double* __restrict__ dst, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[i] = (double)src[i];
}
int refresh_theta64(Ctx* c, int parent) {
    if (!c->theta64_stale[parent]) return 0;
    const int64_t D = c->lay.D;
    k_widen<<<148 * 8, 256, 0, c->stream>>>(c->theta32 + (int64_t)parent * D, c->theta64 + (int64_t)parent * D, D);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    c->theta64_stale[parent] = 0;
    return 0;
}

__global__ void __launch_bounds__(256) k_evolve_batch(float* theta32, __half* __restrict__ th_hi,
                                                      __half* __restrict__ th_lo, const __half* __restrict__ table, const Layout* __restrict__ lay,
                                                      const int64_t* __restrict__ offs, const float* __restrict__ sigmas, const int64_t* __restrict__ srcdst,
                                                      const float* __restrict__ scale, int scale_div) {
    const int k = blockIdx.y;
    const int item = blockIdx.z;
    const TensorDesc t = lay->t[k];
    const int64_t D = lay->D, D_pad = lay->D_pad;
    const int64_t pend = (k + 1 < lay->n) ? lay->t[k + 1].pstart : D_pad;
    const int64_t span = pend - t.pstart;
    const int src = (int)(srcdst[item] & 0xffffffffll), dst = (int)(srcdst[item] >> 32);
    const float sigma = sigmas[item];
    const float* th = theta32 + (int64_t)src * D + t.start;
    const __half* row = table + offs[item] + t.pstart;
    const float* sc = scale ? scale + (int64_t)src * D : nullptr;
    float* d32 = theta32 + (int64_t)dst * D + t.start;           // dst slots are disjoint from src slots (checked by the host)
    __half* dhi = th_hi + (int64_t)dst * D_pad + t.pstart;
    __half* dlo = th_lo + (int64_t)dst * D_pad + t.pstart;
    // 8 elements per thread through 16-byte accesses where the tensor starts on a 16-byte boundary in every array
    const bool vec = ((reinterpret_cast<uintptr_t>(th) | reinterpret_cast<uintptr_t>(d32) |
                       reinterpret_cast<uintptr_t>(row) | reinterpret_cast<uintptr_t>(dhi)) & 15) == 0;
    const int64_t n8 = vec ? t.count / 8 : 0;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    constexpr int U = 4;                                          // vectors in flight per thread: all loads of a batch before the first store
    for (int64_t v0 = tid; v0 < n8; v0 += U * nth) {
        uint4 ev[U]; float4 ta[U], tb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = (v0 + u * nth) * 8;
            if (v0 + u * nth < n8) {
                ev[u] = __ldg(reinterpret_cast<const uint4*>(row + j));
                ta[u] = __ldg(reinterpret_cast<const float4*>(th + j)); tb[u] = __ldg(reinterpret_cast<const float4*>(th + j) + 1);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = (v0 + u * nth) * 8;
            if (v0 + u * nth >= n8) break;
            const float tv[8] = {ta[u].x, ta[u].y, ta[u].z, ta[u].w, tb[u].x, tb[u].y, tb[u].z, tb[u].w};
            const __half* eh = reinterpret_cast<const __half*>(&ev[u]);
            __align__(16) float m[8];
            __align__(16) __half hh[8];
            __align__(16) __half hl[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                m[e] = __fadd_rn(tv[e], mut_noise(sigma, __half2float(eh[e]), sc, t.start + j + e, scale_div));
                split_half(m[e], kThetaScale, hh[e], hl[e]);
            }
            __stcs(reinterpret_cast<float4*>(d32 + j), *reinterpret_cast<float4*>(&m[0]));
            __stcs(reinterpret_cast<float4*>(d32 + j) + 1, *reinterpret_cast<float4*>(&m[4]));
            __stcs(reinterpret_cast<uint4*>(dhi + j), *reinterpret_cast<uint4*>(hh));
            __stcs(reinterpret_cast<uint4*>(dlo + j), *reinterpret_cast<uint4*>(hl));
        }
    }
    for (int64_t j = n8 * 8 + tid; j < span; j += nth) {          // ragged tail, unaligned tensors, zero padding of the shadow
        __half a = __float2half_rn(0.f), b = a;
        if (j < t.count) {
            const float m = __fadd_rn(th[j], mut_noise(sigma, __half2float(row[j]), sc, t.start + j, scale_div));
            d32[j] = m;
            split_half(m, kThetaScale, a, b);
        }
        dhi[j] = a;
        dlo[j] = b;
    }
}
// bookkeeping after k_evolve_batch wrote `n` destination slots (host view of the slots): master-copy flags, safe-mutation scale of the
// new parents, HYBRID bound norms.  Under NES_MUT_SENSITIVITY the sensitivity of a NEW parent is not the one its slot held before
// (the reference recomputes it per (task, parent), nic_es_worker.py:172-173): the slot's divisors are poisoned with NaN and flagged
// stale until nes_set_sensitivity supplies them, so a rollout from such a slot cannot silently use a foreign sensitivity.
__global__ void __launch_bounds__(256) k_fill_f32(float* __restrict__ p, int64_t n, float v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
static int evolve_batch_post(Ctx* c, int n, const int32_t* dst) {
    for (int i = 0; i < n; ++i) {
        c->theta_is_f32[dst[i]] = 1;
        c->theta64_stale[dst[i]] = 1;          // the float64 master copy is rebuilt on demand (refresh_theta64): NIC-ES never reads it
        if (c->mut_kind == NES_MUT_PROPORTIONAL && launch_prop_scale(c, dst[i])) return 1;
        if (c->mut_kind == NES_MUT_SENSITIVITY) {
            k_fill_f32<<<148 * 2, 256, 0, c->stream>>>(c->scale32 + (int64_t)dst[i] * c->lay.D, c->lay.D, nanf(""));
            c->launches++;
            c->scale_stale[dst[i]] = 1;
        }
    }
    NES_CUDA(cudaGetLastError());
    if (c->is_fc) return launch_bound_norms_batch(c, dst, n);
    return 0;
}

int launch_evolve_batch(Ctx* c, int n, const int32_t* src, const int64_t* offsets, const float* sigmas, const int32_t* dst) {
    NES_CHECK(n >= 0 && n <= 65535, "too many items");
    if (n == 0) return 0;
    std::vector<uint8_t> is_src(c->n_parents, 0), is_dst(c->n_parents, 0);
    std::vector<int64_t> packed(n);
    for (int i = 0; i < n; ++i) {
        NES_CHECK(src[i] >= 0 && src[i] < c->n_parents && dst[i] >= 0 && dst[i] < c->n_parents, "parent index out of range");
        NES_CHECK(offsets[i] >= 0 && offsets[i] + c->lay.D_pad <= c->table_len && offsets[i] % c->lay.G == 0, "offset outside the noise table or not slot aligned");
        NES_CHECK(!is_dst[dst[i]], "a destination slot appears twice");
        NES_CHECK(!(c->mut_kind == NES_MUT_SENSITIVITY && sigmas[i] != 0.f && c->scale_stale[src[i]]), "stale sensitivity: call nes_set_sensitivity for this parent first");
        is_src[src[i]] = 1; is_dst[dst[i]] = 1;
        packed[i] = (int64_t)src[i] | ((int64_t)dst[i] << 32);
    }
    for (int p = 0; p < c->n_parents; ++p) NES_CHECK(!(is_src[p] && is_dst[p]), "destination slots must be disjoint from source slots");
    if (ensure_small(c, n)) return 1;
    NES_CUDA(cudaMemcpyAsync(c->d_offsets, offsets, sizeof(int64_t) * n, cudaMemcpyDefault, c->stream));
    NES_CUDA(cudaMemcpyAsync(c->d_weights, sigmas, sizeof(float) * n, cudaMemcpyDefault, c->stream));
    NES_CUDA(cudaMemcpyAsync(c->d_ranks, packed.data(), sizeof(int64_t) * n, cudaMemcpyDefault, c->stream));
    NES_CUDA(cudaStreamSynchronize(c->stream));            // `packed` is a stack-lifetime host buffer
    dim3 grid(148, c->lay.n, n);
    k_evolve_batch<<<grid, 256, 0, c->stream>>>(c->theta32, c->theta_hi, c->theta_lo, c->table, c->d_lay, c->d_offsets, c->d_weights, c->d_ranks, c->scale32, c->scale_div);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return evolve_batch_post(c, n, dst);
}

// ---- NIC-ES truncation selection + survivor materialisation without a host round trip (ESMaster.selection nic_es_master.py:155-167,
// ESIteration.record_parents nic_es/iteration.py:83-93, nic_es_worker.py:188).  k_es_select ranks the n fitnesses in the order of
// Python's stable sort with reverse=True (equal fitnesses keep their submission order) and turns the n_keep best into evolve items
//   dst_base + r  <-  theta[parents[i_r]] + sigma * eps(offsets[i_r])          (i_r = offspring at rank r)
// which one k_evolve_batch launch then writes; order_out[r] = i_r.  NaN fitnesses rank last.
__global__ void __launch_bounds__(256) k_es_select(const double* __restrict__ fit, const int64_t* __restrict__ offs, const int32_t* __restrict__ pars,
                                                   int64_t n, int n_keep, float sigma, int dst_base, int n_slots,
                                                   int64_t* __restrict__ item_off, float* __restrict__ item_sigma, int64_t* __restrict__ item_srcdst,
                                                   int64_t* __restrict__ order_out) {
    __shared__ double tile[1024];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto key = [](double v) { return v != v ? -INFINITY : v; };
    const double xi = i < n ? key(fit[i]) : 0.0;
    int64_t cnt = 0;
    for (int64_t base = 0; base < n; base += 1024) {
        const int m = (int)((n - base) < 1024 ? (n - base) : 1024);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) tile[t] = key(fit[base + t]);
        __syncthreads();
        if (i < n) {
#pragma unroll 8
            for (int t = 0; t < m; ++t) {
                const double xj = tile[t];
                cnt += (xj > xi) || (xj == xi && (base + t) < i);
            }
        }
    }
    if (i < n && cnt < n_keep) {
        int src = pars ? pars[i] : 0;
        // never read through a wild index, never read a slot this launch writes (the host cannot see device arrays): such a member
        // falls back to the first slot outside the destination range
        if (src < 0 || src >= n_slots || (src >= dst_base && src < dst_base + n_keep)) src = dst_base > 0 ? 0 : dst_base + n_keep;
        item_off[cnt] = offs[i];
        item_sigma[cnt] = sigma;
        item_srcdst[cnt] = (int64_t)src | ((int64_t)(dst_base + cnt) << 32);
        order_out[cnt] = i;
    }
}

int launch_es_select_evolve(Ctx* c, const double* d_fit, const int64_t* d_off, const int32_t* d_par, int64_t n, int n_keep, float sigma,
                            int dst_base, int64_t* d_order) {
    NES_CHECK(n >= 1 && n_keep >= 0 && n_keep <= n && n_keep <= 65535, "bad selection size");
    NES_CHECK(dst_base >= 0 && dst_base + n_keep <= c->n_parents && n_keep < c->n_parents, "destination slots out of range");
    if (n_keep == 0) return 0;
    if (ensure_small(c, n_keep)) return 1;
    k_es_select<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_fit, d_off, d_par, n, n_keep, sigma, dst_base, c->n_parents,
                                                                  c->d_offsets, c->d_weights, c->d_ranks, d_order);
    dim3 grid(148, c->lay.n, n_keep);
    k_evolve_batch<<<grid, 256, 0, c->stream>>>(c->theta32, c->theta_hi, c->theta_lo, c->table, c->d_lay, c->d_offsets, c->d_weights, c->d_ranks, c->scale32, c->scale_div);
    c->launches += 2;
    NES_CUDA(cudaGetLastError());
    std::vector<int32_t> dst(n_keep);
    for (int i = 0; i < n_keep; ++i) dst[i] = dst_base + i;
    return evolve_batch_post(c, n_keep, dst.data());
}

// one parent slot copied onto another in stream order (an elite entering the podium, tools/podium.py:43-62, is a file copy there)
int copy_parent(Ctx* c, int src, int dst) {
    NES_CHECK(src >= 0 && src < c->n_parents && dst >= 0 && dst < c->n_parents, "parent index out of range");
    if (src == dst) return 0;
    const int64_t D = c->lay.D, Dp = c->lay.D_pad;
    NES_CUDA(cudaMemcpyAsync(c->theta32 + (int64_t)dst * D, c->theta32 + (int64_t)src * D, sizeof(float) * D, cudaMemcpyDeviceToDevice, c->stream));
    NES_CUDA(cudaMemcpyAsync(c->theta_hi + (int64_t)dst * Dp, c->theta_hi + (int64_t)src * Dp, sizeof(__half) * Dp, cudaMemcpyDeviceToDevice, c->stream));
    NES_CUDA(cudaMemcpyAsync(c->theta_lo + (int64_t)dst * Dp, c->theta_lo + (int64_t)src * Dp, sizeof(__half) * Dp, cudaMemcpyDeviceToDevice, c->stream));
    if (!c->theta64_stale[src]) NES_CUDA(cudaMemcpyAsync(c->theta64 + (int64_t)dst * D, c->theta64 + (int64_t)src * D, sizeof(double) * D, cudaMemcpyDeviceToDevice, c->stream));
    c->theta64_stale[dst] = c->theta64_stale[src];
    c->theta_is_f32[dst] = c->theta_is_f32[src];
    NES_CUDA(cudaMemcpyAsync(c->d_theta_rowmax + dst, c->d_theta_rowmax + src, sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
    if (c->scale32) {
        NES_CUDA(cudaMemcpyAsync(c->scale32 + (int64_t)dst * D, c->scale32 + (int64_t)src * D, sizeof(float) * D, cudaMemcpyDeviceToDevice, c->stream));
        NES_CUDA(cudaMemcpyAsync(c->d_scale_max + dst, c->d_scale_max + src, sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
        c->scale_stale[dst] = c->scale_stale[src];
    }
    return 0;
}


// ---- SM-PROPORTIONAL scale (algorithm/nets.py:107-111): params = theta.clone(); params[params == 0] = mean(|theta|); noise *= |params|
// The mean is accumulated in float64 in a fixed order and rounded to float32 (torch's float32 pairwise mean agrees to ~1 ulp;
// it only enters through parameters that are exactly zero, i.e. untrained biases).
__global__ void __launch_bounds__(256) k_abs_partial(const float* __restrict__ th, int64_t D, double* __restrict__ partials) {
    __shared__ double sh[8];
    double a = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < D; i += (int64_t)gridDim.x * 256) a += (double)fabsf(th[i]);
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) { double t = 0; for (int q = 0; q < 8; ++q) t += sh[q]; partials[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(256) k_prop_scale(const float* __restrict__ th, int64_t D, const double* __restrict__ partials, int n_partials,
                                                    float* __restrict__ scale, float* __restrict__ scale_max) {
    double tot = 0.0;
    for (int q = 0; q < n_partials; ++q) tot += partials[q];
    const float mean = (float)(tot / (double)D);
    float mx = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < D; i += (int64_t)gridDim.x * 256) {
        const float v = th[i];
        const float s = v == 0.0f ? fabsf(mean) : fabsf(v);
        scale[i] = s; mx = fmaxf(mx, s);
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int*>(scale_max), __float_as_int(mx));     // mx >= 0
}
int launch_prop_scale(Ctx* c, int parent) {
    const int64_t D = c->lay.D;
    NES_CHECK(c->scale32 && c->d_scale_max && c->d_abs_partials, "proportional mutation buffers missing");
    const float* th = c->theta32 + (int64_t)parent * D;
    k_abs_partial<<<kAbsBlocks, 256, 0, c->stream>>>(th, D, c->d_abs_partials);
    NES_CUDA(cudaMemsetAsync(c->d_scale_max + parent, 0, sizeof(float), c->stream));
    k_prop_scale<<<kAbsBlocks, 256, 0, c->stream>>>(th, D, c->d_abs_partials, kAbsBlocks, c->scale32 + (int64_t)parent * D, c->d_scale_max + parent);
    c->launches += 2;
    NES_CUDA(cudaGetLastError());
    return 0;
}
int set_mutation(Ctx* c, int kind) {
    NES_CHECK(kind == NES_MUT_DEFAULT || kind == NES_MUT_PROPORTIONAL || kind == NES_MUT_SENSITIVITY || kind == NES_MUT_SCALE, "unknown mutation kind");
    if (kind == c->mut_kind) return 0;
    NES_CUDA(cudaStreamSynchronize(c->stream));
    if (c->scale32) {
        cudaFree(c->scale32); cudaFree(c->d_scale_max); cudaFree(c->d_abs_partials);
        c->scale32 = nullptr; c->d_scale_max = nullptr; c->d_abs_partials = nullptr;
    }
    c->mut_kind = kind;
    c->scale_div = kind == NES_MUT_SENSITIVITY ? 1 : 0;
    for (int p = 0; p < c->n_parents; ++p) c->scale_stale[p] = 0;
    if (kind == NES_MUT_DEFAULT) return 0;
    NES_CUDA(cudaMalloc((void**)&c->scale32, sizeof(float) * (size_t)c->n_parents * c->lay.D));
    NES_CUDA(cudaMalloc((void**)&c->d_scale_max, sizeof(float) * c->n_parents));
    NES_CUDA(cudaMalloc((void**)&c->d_abs_partials, sizeof(double) * kAbsBlocks));
    if (kind == NES_MUT_PROPORTIONAL) {
        for (int p = 0; p < c->n_parents; ++p) if (launch_prop_scale(c, p)) return 1;
    } else {                                   // sensitivities / multipliers of 1 until nes_set_sensitivity supplies them
        k_fill_f32<<<148 * 8, 256, 0, c->stream>>>(c->scale32, (int64_t)c->n_parents * c->lay.D, 1.0f);
        k_fill_f32<<<1, 256, 0, c->stream>>>(c->d_scale_max, c->n_parents, 1.0f);
        c->launches += 2;
        NES_CUDA(cudaGetLastError());
    }
    return 0;
}

// sensitivities of one parent (parent < 0: of every parent): noise /= sens.  scale_max = max 1/sens bounds the multiplier for the
// HYBRID error margin.
__global__ void __launch_bounds__(256) k_inv_max(const float* __restrict__ s, int64_t D, float* __restrict__ out, int div) {
    float mx = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D; i += (int64_t)gridDim.x * blockDim.x) mx = fmaxf(mx, fabsf(div ? __fdiv_rn(1.0f, s[i]) : s[i]));
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int*>(out), __float_as_int(mx));
}
int set_sensitivity(Ctx* c, int parent, const float* sens) {
    NES_CHECK((c->mut_kind == NES_MUT_SENSITIVITY || c->mut_kind == NES_MUT_SCALE) && c->scale32, "call nes_set_mutation(NES_MUT_SENSITIVITY or NES_MUT_SCALE) first");
    NES_CHECK(sens != nullptr && parent < c->n_parents, "bad sensitivity argument");
    const int64_t D = c->lay.D;
    for (int p = (parent < 0 ? 0 : parent); p < (parent < 0 ? c->n_parents : parent + 1); ++p) {
        NES_CUDA(cudaMemcpyAsync(c->scale32 + (int64_t)p * D, sens, sizeof(float) * D, cudaMemcpyDefault, c->stream));
        NES_CUDA(cudaMemsetAsync(c->d_scale_max + p, 0, sizeof(float), c->stream));
        k_inv_max<<<kAbsBlocks, 256, 0, c->stream>>>(c->scale32 + (int64_t)p * D, D, c->d_scale_max + p, c->scale_div);
        c->launches++;
        c->scale_stale[p] = 0;
    }
    NES_CUDA(cudaGetLastError());
    NES_CUDA(cudaStreamSynchronize(c->stream));          // sens may be a pageable host buffer
    return 0;
}

}  // namespace nes
