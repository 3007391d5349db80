// HYBRID vocabulary projection: helpers around the fp16-operand logit GEMM (k_member_dense<DEN_LOGITC>, dense.cuh).
//
// The reference picks the next token as argmax of float32 log-softmax(logit(h)) (captioning/nets.py:202-208).  Only the
// arg-max and its log-prob are needed, so the 64 % of the FLOPs spent in the vocabulary projection do not need
// split-fp16 accuracy everywhere: single-fp16 operands give every logit with a RIGOROUS error bound
//     |l~ - l| <= (2 * 2^-11 + 2^-22) |x|_2 |w_j|_2  +  fp32 accumulation,      w_j = theta_j +- sigma eps_j
// (x and the materialised w_j are each rounded to 11 significant bits, Cauchy-Schwarz), with |w_j|_2 <= |theta_j|_2 +
// sigma |eps_j|_2, |theta_j|_2 <= max row norm of logit.weight (k_rowmax_theta, per parent) and |eps_j|_2 <= max row norm
// of the [L/R, R] view of the table (k_rowmax_eps).  Every column within margin = 2 * bound of the running max is appended
// to the row's candidate list (a superset of the columns within the margin of the final max, since the running max only
// rises) and k_refine re-evaluates the survivors in fp32 from the float32 theta and the exact fp16 eps -- the same member
// weight theta +- f32(sigma)*eps the reference forms (algorithm/nets.py:113).  Rows whose list overflowed flag their
// unit, which is then recomputed by the split-fp16 kernel (masked launch), so the result never depends on the bound
// being tight, only on it being an upper bound.
#pragma once
#include "dense.cuh"

namespace nes {

__device__ __forceinline__ float warp_sum(float v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
__device__ __forceinline__ float warp_max(float v) { for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o)); return v; }
// Loads written as asm statements at the head of a phase, so that the front end keeps them together and ahead of the arithmetic.  ncu's
// source view of the tail before this: the second half of h, the candidate entries, the second half of the candidate's weight row and of
// the embedding row were each loaded after the first use of the previous load -- eight dependent memory trips per row, 72 % of the issue
// slots stalled on them (long scoreboard 9.5 cycles per issue).  With these, the halves of a row travel together: 5.7 cycles per issue,
// 318 -> 251 us per launch under ncu, 6.8 -> 6.3 ms per generation.  ptxas still sinks the list heads / entries below the conversion of
// h; forcing the order with ld.volatile / ld.relaxed.gpu costs more than the trip it saves (8.4 / 7.1 ms), so plain loads it is.
__device__ __forceinline__ uint4 ldp_128(const void* p) {
    uint4 r; asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r;
}
__device__ __forceinline__ uint2 ldp_64(const void* p) { uint2 r; asm volatile("ld.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p)); return r; }
__device__ __forceinline__ uint32_t ldp_32(const void* p) { uint32_t r; asm volatile("ld.global.u32 %0, [%1];" : "=r"(r) : "l"(p)); return r; }
__device__ __forceinline__ uint16_t ldp_16(const void* p) { uint16_t r; asm volatile("ld.global.u16 %0, [%1];" : "=h"(r) : "l"(p)); return r; }
__device__ __forceinline__ void atomic_max_pos(float* addr, float v) { atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v)); }   // v >= 0

// max over the rows of logit.weight [V1, R] (float32 theta of one parent) of the row 2-norm; *out zeroed by the caller
__global__ void __launch_bounds__(256) k_rowmax_theta(const float* __restrict__ w, int64_t n_rows, int R, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    float best = 0.f;
    for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < n_rows; r += (int64_t)gridDim.x * 8) {
        float s = 0.f;
        for (int k = lane; k < R; k += 32) { const float v = w[r * R + k]; s = fmaf(v, v, s); }
        best = fmaxf(best, warp_sum(s));
    }
    if (lane == 0) atomic_max_pos(out, sqrtf(best) * 1.0001f);
}
__global__ void k_zero_slots(const int64_t* __restrict__ srcdst, int n, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[(int)(srcdst[i] >> 32)] = 0.f;
}
// the same for several parents at once: blockIdx.y = item, parents[item] = parent slot
__global__ void __launch_bounds__(256) k_rowmax_theta_batch(const float* __restrict__ theta32, int64_t D, int64_t w_start, int64_t n_rows, int R,
                                                            const int64_t* __restrict__ srcdst, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int par = (int)(srcdst[blockIdx.y] >> 32);
    const float* w = theta32 + (int64_t)par * D + w_start;
    float best = 0.f;
    for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < n_rows; r += (int64_t)gridDim.x * 8) {
        float s = 0.f;
        for (int k = lane; k < R; k += 32) { const float v = w[r * R + k]; s = fmaf(v, v, s); }
        best = fmaxf(best, warp_sum(s));
    }
    if (lane == 0) atomic_max_pos(out + par, sqrtf(best) * 1.0001f);
}
// max over the rows of the [L/R, R] view of the fp16 noise table of the row 2-norm
__global__ void __launch_bounds__(256) k_rowmax_eps(const __half* __restrict__ table, int64_t n_rows, int R, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    float best = 0.f;
    for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < n_rows; r += (int64_t)gridDim.x * 8) {
        float s = 0.f;
        for (int k = lane * 8; k < R; k += 256) {
            const uint4 v = *reinterpret_cast<const uint4*>(table + r * R + k);
            const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
            for (int e = 0; e < 4; ++e) { const float2 f = __half22float2(h[e]); s = fmaf(f.x, f.x, s); s = fmaf(f.y, f.y, s); }
        }
        best = fmaxf(best, warp_sum(s));
    }
    if (lane == 0) atomic_max_pos(out, sqrtf(best) * 1.0001f);
}

struct MarginParams {
    const __half *h_hi, *h_lo; int R; int rows, rows_pad; int64_t n_members;
    const int32_t* parent; const float* theta_rowmax; const float* eps_rowmax; const float* scale_max; float sigma; float* margin;
};
// one warp per (member, row): margin = 2 * error bound of the fp16-operand logits of that row
__global__ void __launch_bounds__(128) k_row_margin(const MarginParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (wid >= p.n_members * p.rows) return;
    const int64_t member = wid / p.rows;
    const int64_t grow = member * p.rows_pad + (wid % p.rows);
    float s = 0.f;
    for (int k = lane * 8; k < p.R; k += 256) {
        const uint4 a = *reinterpret_cast<const uint4*>(p.h_hi + grow * p.R + k), b = *reinterpret_cast<const uint4*>(p.h_lo + grow * p.R + k);
        const __half2 *ha = reinterpret_cast<const __half2*>(&a), *hb = reinterpret_cast<const __half2*>(&b);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 fa = __half22float2(ha[e]), fb = __half22float2(hb[e]);
            const float x0 = (fa.x + fb.x) * (1.0f / kActScale), x1 = (fa.y + fb.y) * (1.0f / kActScale);
            s = fmaf(x0, x0, s); s = fmaf(x1, x1, s);
        }
    }
    s = warp_sum(s);
    if (lane == 0) {
        const int par = p.parent ? p.parent[member] : 0;
        const float smax = p.scale_max ? p.scale_max[par] : 1.f;            // |eps_j (.) s|_2 <= max(s) |eps_j|_2
        const float wsum = 2.f * (p.theta_rowmax[par] + fabsf(p.sigma) * p.eps_rowmax[0] * smax);
        // 2^-11 relative rounding of both operands, +10 % for the fp32 accumulation of K <= 4096 terms, x2 (two logits compared)
        // ... plus an absolute floor for the fp16 subnormal range (|32 h_k| or |256 w_jk| below 6e-5 round with an absolute, not a
        // relative, error of 2^-25: at most sqrt(R) 2^-30 |w_j|_2 per logit, far below 1e-7)
        p.margin[grow] = 2.f * 1.1f * 4.8828125e-4f * sqrtf(s) * 1.0001f * (wsum + 1e-4f) + 1e-7f;
    }
}

struct RefineParams {
    const CandHead* heads; const CandEntry* entries; const float* margin;
    const __half *h_hi, *h_lo; int R, V1; int rows, rows_pad, m_tiles; int64_t n_members;
    const int64_t* off; const float* sign; const int32_t* parent;
    const float* theta32; const __half* table; int64_t D; const float* scale; int scale_div;
    int64_t w_start, w_pstart, b_start, b_pstart;     // logit.weight / logit.bias: flat and padded starts
    float sigma;
    int32_t* tok; float* lp;                          // [member*rows_pad + row]: arg-max column and its log-prob
    uint8_t* unit_flag;                               // [n_members*m_tiles]: 1 -> recompute this unit with the split-fp16 kernel
    unsigned long long* stats;                        // nullable: [0] rows refined, [1] candidates evaluated, [2] rows overflowed
    int force_overflow;                               // NES_DEBUG bit 6 (64): every row takes the fallback (tests)
    int n_split;                                      // lists of a row: n_split * kCandLists (DenseParams::n_split; 1 for full-size launches)
};
// one warp per (member, row): filter the two half-row candidate lists by the final threshold, re-evaluate every survivor
// in fp32 with the member weight theta +- f32(sigma)*eps (algorithm/nets.py:113), pick the reference's argmax of float32
// log-probs (lowest index among equal rounded log-probs, captioning/nets.py:202-208)
// -> false if the row's lists overflowed (its unit is flagged for the fallback), else idx / best = arg-max column and its log-prob
// SCALED = p.scale is set (safe-mutation scale): a template flag, so that the plain kernels carry none of the per-element scale code
template <bool SCALED>
__device__ __forceinline__ bool refine_row(const RefineParams& p, int64_t member, int r, int64_t grow, int lane, int wq, int (*s_idx)[32],
                                           int& idx_out, float& best_out) {
    // ---- phase 1: every load whose address does not depend on another load (the row's hidden state, margin, list heads and -- whether or
    // not they are all in use, kCandCap >= 32 -- the first 32 entries of each list; the member's sign / offset / parent), pinned
    const bool one = p.n_split == 1;
    uint4 hraw[2][2];                                  // h_hi / h_lo, two chunks of 8 (R <= 512); a lane past R reads the row start and ignores it
#pragma unroll
    for (int c2 = 0; c2 < 2; ++c2) {
        const int k = c2 * 256 + lane * 8;
        const int ks = k < p.R ? k : 0;
        hraw[c2][0] = ldp_128(p.h_hi + grow * p.R + ks);
        hraw[c2][1] = ldp_128(p.h_lo + grow * p.R + ks);
    }
    CandHead hd[kCandLists];
    CandEntry pre[kCandLists];
#pragma unroll
    for (int hf = 0; hf < kCandLists; ++hf) {
        const int64_t l = one ? grow * kCandLists + hf : grow * (int64_t)p.n_split * kCandLists;      // (split lists: read below; any valid address here)
        const uint4 hv = ldp_128(p.heads + l);
        hd[hf].mx = __uint_as_float(hv.x); hd[hf].sum = __uint_as_float(hv.y); hd[hf].n = (int32_t)hv.z; hd[hf].pad = 0;
        const uint2 ev = ldp_64(p.entries + l * (int64_t)kCandCap + lane);
        pre[hf].v = __uint_as_float(ev.x); pre[hf].idx = (int32_t)ev.y;
    }
    const float mg = __uint_as_float(ldp_32(p.margin + grow));
    const float sgn_m = __uint_as_float(ldp_32(p.sign + member));
    const uint2 off2 = ldp_64(p.off + member);
    const int par = p.parent ? (int)ldp_32(p.parent + member) : 0;
    const int64_t off = (int64_t)(((uint64_t)off2.y << 32) | off2.x);
    const float ssig = sgn_m > 0.f ? p.sigma : -p.sigma;               // theta - fl(sigma eps) == theta + fl(-sigma eps), bit for bit
    // this lane's slice of the row's hidden state, rebuilt from the split halves once
    float xs[2][8];
#pragma unroll
    for (int c2 = 0; c2 < 2; ++c2) {
        const bool in = c2 * 256 + lane * 8 < p.R;
        const __half2 *ha = reinterpret_cast<const __half2*>(&hraw[c2][0]), *hb = reinterpret_cast<const __half2*>(&hraw[c2][1]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 fa = __half22float2(ha[e]), fb = __half22float2(hb[e]);
            xs[c2][2 * e] = in ? (fa.x + fb.x) * (1.0f / kActScale) : 0.f; xs[c2][2 * e + 1] = in ? (fa.y + fb.y) * (1.0f / kActScale) : 0.f;
        }
    }
    float M = -INFINITY, S = 0.f;
    bool overflow = p.force_overflow != 0;
    int count = 0;
    if (one) {
#pragma unroll
        for (int hf = 0; hf < kCandLists; ++hf) {
            M = fmaxf(M, hd[hf].mx);
            overflow = overflow || hd[hf].n > kCandCap;
        }
        const float thr = M - mg;
        if (!overflow) {
#pragma unroll
            for (int hf = 0; hf < kCandLists; ++hf) {
                const int n = hd[hf].n;
                const CandEntry* ent = p.entries + (grow * kCandLists + hf) * (int64_t)kCandCap;
                for (int e0 = 0; e0 < n; e0 += 32) {
                    const int e = e0 + lane;
                    CandEntry ce = pre[hf];
                    if (e0 > 0) { ce.v = -INFINITY; ce.idx = 0; if (e < n) ce = ent[e]; }
                    const bool ok = e < n && ce.v >= thr;
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    const int pos = count + __popc(bal & ((1u << lane) - 1u));
                    if (ok && pos < 32) s_idx[wq][pos] = ce.idx;
                    count += __popc(bal);
                }
            }
        }
#pragma unroll
        for (int hf = 0; hf < kCandLists; ++hf) S += hd[hf].sum * expf(hd[hf].mx - M);
    } else {
        // few-unit launches: the row's vocabulary sweep was cut into n_split ranges with kCandLists lists each (dense.cuh)
        const int n_lists = p.n_split * kCandLists;
        const CandHead* hd = p.heads + grow * n_lists;
        for (int l = lane; l < n_lists; l += 32) { const CandHead h = hd[l]; M = fmaxf(M, h.mx); overflow = overflow || h.n > kCandCap; }
        M = warp_max(M);
        overflow = __any_sync(0xffffffffu, overflow);
        for (int l = lane; l < n_lists; l += 32) { const CandHead h = hd[l]; if (h.mx > -INFINITY) S += h.sum * expf(h.mx - M); }
        S = warp_sum(S);
        const float thr = M - mg;
        if (!overflow) {
            // one lane per list (32 lists at a time), entry e of all of them in one step: the lists are short (1 - 3 entries), so this is
            // a handful of dependent loads instead of one per list.  A list whose own maximum is below the threshold holds nothing.
            for (int l0 = 0; l0 < n_lists; l0 += 32) {
                const int l = l0 + lane;
                int n = 0;
                if (l < n_lists) { const CandHead h = hd[l]; n = h.mx >= thr ? h.n : 0; }
                const CandEntry* ent = p.entries + (grow * n_lists + (l < n_lists ? l : 0)) * (int64_t)kCandCap;
                const int nmax = __reduce_max_sync(0xffffffffu, n);
                for (int e = 0; e < nmax; ++e) {
                    CandEntry ce; ce.v = -INFINITY; ce.idx = 0;
                    if (e < n) ce = ent[e];
                    const bool ok = e < n && ce.v >= thr;
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    const int pos = count + __popc(bal & ((1u << lane) - 1u));
                    if (ok && pos < 32) s_idx[wq][pos] = ce.idx;
                    count += __popc(bal);
                }
            }
        }
    }
    overflow = overflow || count > 32;
    if (overflow) {                                            // more candidates than the lists / one warp hold
        if (lane == 0) { p.unit_flag[member * p.m_tiles + r / 128] = 1; if (p.stats) atomicAdd(&p.stats[2], 1ull); }
        return false;
    }
    __syncwarp();
    const int my_idx = lane < count ? s_idx[wq][lane] : -1;
    const float* wbase = p.theta32 + (int64_t)par * p.D + p.w_start;
    const bool aligned = (reinterpret_cast<uintptr_t>(wbase) & 15) == 0;     // R is a multiple of 64: every row start shares it
    float my_val = -INFINITY;
    for (int c = 0; c < count; ++c) {
        const int j = __shfl_sync(0xffffffffu, my_idx, c);
        const float* tw = wbase + (int64_t)j * p.R;
        const __half* ew = p.table + off + p.w_pstart + (int64_t)j * p.R;
        const float* sw = SCALED ? p.scale + (int64_t)par * p.D + p.w_start + (int64_t)j * p.R : nullptr;
        float acc = 0.f;
        // ---- phase 2, per candidate: both chunks of its weight row (fp16 eps, float32 theta) and its bias pair, pinned, then the arithmetic
        uint4 evs[2]; float tvs[2][8];
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
            const int k = c2 * 256 + lane * 8;
            const int ks = k < p.R ? k : 0;
            evs[c2] = ldp_128(ew + ks);
            if (aligned) {
                const uint4 t0 = ldp_128(tw + ks), t1 = ldp_128(tw + ks + 4);
                tvs[c2][0] = __uint_as_float(t0.x); tvs[c2][1] = __uint_as_float(t0.y); tvs[c2][2] = __uint_as_float(t0.z); tvs[c2][3] = __uint_as_float(t0.w);
                tvs[c2][4] = __uint_as_float(t1.x); tvs[c2][5] = __uint_as_float(t1.y); tvs[c2][6] = __uint_as_float(t1.z); tvs[c2][7] = __uint_as_float(t1.w);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) tvs[c2][e] = __uint_as_float(ldp_32(tw + ks + e));
            }
        }
        const float tb = __uint_as_float(ldp_32(p.theta32 + (int64_t)par * p.D + p.b_start + j));
        const __half eb = __ushort_as_half(ldp_16(p.table + off + p.b_pstart + j));
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
            const int k = c2 * 256 + lane * 8;
            if (k < p.R) {
                const __half2* he = reinterpret_cast<const __half2*>(&evs[c2]);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 ef = __half22float2(he[e]);
                    float n0 = __fmul_rn(ssig, ef.x), n1 = __fmul_rn(ssig, ef.y);
                    if (SCALED) { n0 = apply_scale(n0, sw[k + 2 * e], p.scale_div); n1 = apply_scale(n1, sw[k + 2 * e + 1], p.scale_div); }
                    acc = fmaf(xs[c2][2 * e], __fadd_rn(tvs[c2][2 * e], n0), acc);
                    acc = fmaf(xs[c2][2 * e + 1], __fadd_rn(tvs[c2][2 * e + 1], n1), acc);
                }
            }
        }
        acc = warp_sum(acc);
        const float nzb = mut_noise(ssig, __half2float(eb), SCALED ? p.scale : nullptr, (int64_t)par * p.D + p.b_start + j, p.scale_div);
        const float v = acc + __fadd_rn(tb, nzb);
        if (lane == c) my_val = v;
    }
    const float vmax = warp_max(my_val);
    // log_softmax(x) = (x - max) - log(sum exp(x - max)); the sum comes from the approximate logits (relative to M)
    const float lse = __fadd_rn(__fsub_rn(M, vmax), logf(S));
    float best = -INFINITY; int idx = 0x7fffffff;
    if (my_idx >= 0) { best = __fsub_rn(__fsub_rn(my_val, vmax), lse); idx = my_idx; }
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
    }
    if (p.stats && lane == 0) { atomicAdd(&p.stats[0], 1ull); atomicAdd(&p.stats[1], (unsigned long long)count); }
    idx_out = idx; best_out = best;
    return true;
}
__global__ void __launch_bounds__(128) k_refine(const RefineParams p) {
    __shared__ int s_idx[4][32];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t wid = (int64_t)blockIdx.x * 4 + wq;
    if (wid >= p.n_members * p.rows) return;
    const int64_t member = wid / p.rows;
    const int r = (int)(wid % p.rows);
    const int64_t grow = member * p.rows_pad + r;
    int idx; float best;
    const bool ok = p.scale ? refine_row<true>(p, member, r, grow, lane, wq, s_idx, idx, best) : refine_row<false>(p, member, r, grow, lane, wq, s_idx, idx, best);
    if (ok && lane == 0) { p.tok[grow] = idx; p.lp[grow] = best; }
}

// Fallback of the HYBRID projection when member weights carry a safe-mutation scale (the factored split-fp16 kernel cannot
// scale eps tiles on their way through TMA): every logit of every row of a flagged unit in plain fp32 on the CUDA cores, then
// the reference's log_softmax / argmax (captioning/nets.py:202-208).  One block per row; rare (candidate list overflow).
__global__ void __launch_bounds__(256) k_exact_rows(const RefineParams p) {
    extern __shared__ float s_logit[];                    // [V1]
    __shared__ float s_red[8]; __shared__ int s_idx[8];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t wid = blockIdx.x;
    const int64_t member = wid / p.rows;
    const int r = (int)(wid % p.rows);
    if (!p.unit_flag[member * p.m_tiles + r / 128]) return;
    const int64_t grow = member * p.rows_pad + r;
    const float sgn = p.sign[member];
    const int64_t off = p.off[member];
    const int par = p.parent ? p.parent[member] : 0;
    for (int j = wq; j < p.V1; j += 8) {
        const float* tw = p.theta32 + (int64_t)par * p.D + p.w_start + (int64_t)j * p.R;
        const __half* ew = p.table + off + p.w_pstart + (int64_t)j * p.R;
        float acc = 0.f;
        for (int k = lane; k < p.R; k += 32) {
            const float x = (__half2float(p.h_hi[grow * p.R + k]) + __half2float(p.h_lo[grow * p.R + k])) * (1.0f / kActScale);
            const float nz = mut_noise(p.sigma, __half2float(ew[k]), p.scale, (int64_t)par * p.D + p.w_start + (int64_t)j * p.R + k, p.scale_div);
            acc = fmaf(x, sgn > 0.f ? __fadd_rn(tw[k], nz) : __fsub_rn(tw[k], nz), acc);
        }
        acc = warp_sum(acc);
        if (lane == 0) {
            const float tb = p.theta32[(int64_t)par * p.D + p.b_start + j];
            const float nzb = mut_noise(p.sigma, __half2float(p.table[off + p.b_pstart + j]), p.scale, (int64_t)par * p.D + p.b_start + j, p.scale_div);
            s_logit[j] = acc + (sgn > 0.f ? __fadd_rn(tb, nzb) : __fsub_rn(tb, nzb));
        }
    }
    __syncthreads();
    float mx = -INFINITY;
    for (int j = threadIdx.x; j < p.V1; j += 256) mx = fmaxf(mx, s_logit[j]);
    mx = warp_max(mx);
    if (lane == 0) s_red[wq] = mx;
    __syncthreads();
    mx = s_red[0];
    for (int q = 1; q < 8; ++q) mx = fmaxf(mx, s_red[q]);
    __syncthreads();
    float sm = 0.f;
    for (int j = threadIdx.x; j < p.V1; j += 256) sm += expf(s_logit[j] - mx);
    sm = warp_sum(sm);
    if (lane == 0) s_red[wq] = sm;
    __syncthreads();
    sm = 0.f;
    for (int q = 0; q < 8; ++q) sm += s_red[q];
    const float lse = logf(sm);
    float best = -INFINITY; int idx = 0x7fffffff;
    for (int j = threadIdx.x; j < p.V1; j += 256) {
        const float l = __fsub_rn(__fsub_rn(s_logit[j], mx), lse);
        if (l > best) { best = l; idx = j; }              // ascending j per thread: the first index wins ties
    }
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
    }
    __syncthreads();
    if (lane == 0) { s_red[wq] = best; s_idx[wq] = idx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = 1; q < 8; ++q) if (s_red[q] > best || (s_red[q] == best && s_idx[q] < idx)) { best = s_red[q]; idx = s_idx[q]; }
        p.tok[grow] = idx; p.lp[grow] = best;
    }
}

}  // namespace nes
