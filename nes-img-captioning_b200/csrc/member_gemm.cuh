// Population GEMM for perturbed members, factored as  y_m = x_m theta^T + s_m sigma x_m eps_m^T
// (member m = theta + s_m*sigma*eps_m is never materialised).  One CTA computes a 128-row x BN-column
// output tile of ONE member: TMA streams the activation tile, the matching theta tile (split-fp16 shadow,
// L2 resident, shared by every member) and the member's eps tile straight out of the HBM noise table;
// theta and eps tiles sit back to back in shared memory so ONE tcgen05.mma of N = 2*BN (3*BN in EXACT
// mode) produces both partial products into adjacent TMEM column ranges; the epilogue combines them in
// fp32 with the exact scale factors and applies the fused tail (bias + maxout-LSTM cell, or online
// arg-max / log-sum-exp over the vocabulary tile, or a plain store).
//
// Reference math: captioning/nets.py:98-134 (LSTMCore.forward), :195-208 (img_embed, logit, log_softmax, max).
#pragma once
#include "common.cuh"
#include "tc_ptx.cuh"

namespace nes {

enum { EPI_PLAIN = 0, EPI_LSTM = 1, EPI_LOGIT = 2 };

struct alignas(64) SegMaps {
    CUtensorMap a_hi, a_lo;    // activations [rows_total, K] fp16 (x * 2^5 split hi/lo), box 64 x 128
    CUtensorMap w_hi, w_lo;    // theta shadow viewed as [n_parents*D_pad/K, K], box 64 x chunk_rows
    CUtensorMap eps;           // noise table viewed as [L/K, K], box 64 x chunk_rows
};
struct SegInfo {
    int32_t K;
    int32_t a_shared;          // 1: every member reads A rows [0, rows_pad) (shared fc batch)
    int64_t w_pstart;          // padded start of the weight tensor inside a noise row / theta shadow
    int64_t b_start, b_pstart; // bias tensor: flat start / padded start (-1: no bias)
};
// per (row, vocabulary tile): max logit + its first index, sum exp(logit - mx), and the EARLIEST index whose logit is
// within tie_delta of the tile max (the reference takes argmax over float32 LOG-PROBS, whose quantum near
// -log(V) is ~1e-6: logits closer than that tie and torch.max returns the lowest index -- see k_finalize)
struct alignas(16) LogitPartial { float mx; int32_t idx; float sum; float lo_val; int32_t lo_idx; int32_t pad[3]; };

struct GemmParams {
    SegMaps maps[2];
    SegInfo seg[2];
    int32_t nseg;
    int32_t rows, rows_pad, m_tiles, n_tiles, group;   // geometry / scheduling
    int32_t N;                 // valid output columns per chunk (R for LSTM, N otherwise)
    int32_t chunk_stride;      // weight rows between chunks (R for LSTM)
    int64_t D, D_pad;
    const int64_t* off; const float* sign; const int32_t* parent;   // per member (wave-local index)
    const float* theta32; const __half* table;
    float sigma;
    // PLAIN
    float* out_f32; int32_t ld_out;          // [member*rows_pad + row, ld_out] (nullable)
    __half* out_hi; __half* out_lo;          // next A operand [member*rows_pad + row, N] (nullable)
    // LSTM
    float* c_state;                          // [member*rows_pad + row, R] in place
    float* h_trace;                          // nullable, [member*rows + row, R]
    // LOGIT
    LogitPartial* partials;                  // [member*rows_pad + row, n_tiles]
    float* logit_trace; int64_t ld_trace;    // nullable, [member*rows + row, ld_trace]
    float tie_delta;                         // one float32 ulp of log(V+1)
};

template <int EPI, int PREC> struct GemmCfg {
    static constexpr int BM = 128, BK = 64;
    static constexpr int NCH = (EPI == EPI_LSTM) ? 5 : 1;           // weight-row chunks gathered per tile
    static constexpr int CR  = (EPI == EPI_LSTM) ? 16 : 128;        // rows per chunk
    static constexpr int BN  = NCH * CR;                            // 80 or 128
    static constexpr int NB  = (PREC == NES_PREC_EXACT) ? 3 : 2;    // B sub-tiles: [theta_lo] theta_hi eps
    static constexpr int NA  = (PREC == NES_PREC_EXACT) ? 2 : 1;    // A sub-tiles: x_hi [x_lo]
    static constexpr int A_BYTES = BM * BK * 2;                     // 16 KB
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = NA * A_BYTES + NB * B_BYTES;
    static constexpr int STAGES = (PREC == NES_PREC_EXACT) ? (EPI == EPI_LSTM ? 3 : 2) : (EPI == EPI_LSTM ? 3 : 2);
    static constexpr int TMEM_COLS = (NB * BN <= 256) ? 256 : 512;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 1024 /*barriers, bias*/;
    static constexpr int THREADS = 192;                             // warp0 TMA, warp1 MMA, warps 2-5 epilogue
};

__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ void split_half_dev(float x, float scale, __half& hi, __half& lo) {
    float xs = x * scale;
    hi = __float2half_rn(xs);
    lo = __float2half_rn(xs - __half2float(hi));
}

template <int EPI, int PREC>
__global__ void __launch_bounds__(192) k_member_gemm(const __grid_constant__ GemmParams p) {
    using C = GemmCfg<EPI, PREC>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* ctrl = smem + C::STAGES * C::STAGE_BYTES;
    uint64_t* full_bar  = reinterpret_cast<uint64_t*>(ctrl);
    uint64_t* empty_bar = full_bar + C::STAGES;
    uint64_t* acc_bar   = empty_bar + C::STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
    float* s_bias = reinterpret_cast<float*>(ctrl + 128);            // BN floats (<= 512 B)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- tile coordinates: CTAs of one weight-sharing group are adjacent (L2 reuse of theta / eps tiles)
    const int per_group = p.group * p.m_tiles;
    const int64_t gid = blockIdx.x / ((int64_t)p.n_tiles * per_group);
    const int rem = (int)(blockIdx.x - gid * (int64_t)p.n_tiles * per_group);
    const int nt = rem / per_group;
    const int sub = rem % per_group;
    const int64_t member = gid * p.group + sub / p.m_tiles;
    const int mt = sub % p.m_tiles;
    const int64_t off = p.off[member];
    const float sgn = p.sign[member];
    const int par = p.parent ? p.parent[member] : 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        tc::mbar_init(acc_bar, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // =========================== TMA producer ===========================
        if (lane == 0) {
            for (int s = 0; s < p.nseg; ++s) {
                tc::prefetch_tmap(&p.maps[s].a_hi); tc::prefetch_tmap(&p.maps[s].w_hi); tc::prefetch_tmap(&p.maps[s].eps);
            }
            int stage = 0; uint32_t phase = 0;
            for (int s = 0; s < p.nseg; ++s) {
                const SegMaps& mp = p.maps[s];
                const SegInfo& sg = p.seg[s];
                const int32_t a_row = (int32_t)((sg.a_shared ? 0 : member * p.rows_pad) + mt * C::BM);
                const int32_t w_row0 = (int32_t)(((int64_t)par * p.D_pad + sg.w_pstart) / sg.K) + nt * C::CR;
                const int32_t e_row0 = (int32_t)((off + sg.w_pstart) / sg.K) + nt * C::CR;
                for (int kb = 0; kb < sg.K / C::BK; ++kb) {
                    tc::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = smem + stage * C::STAGE_BYTES;
                    tc::mbar_expect_tx(&full_bar[stage], C::STAGE_BYTES);
                    tc::tma_load_2d(st, &mp.a_hi, &full_bar[stage], kb * C::BK, a_row);
                    if (PREC == NES_PREC_EXACT) tc::tma_load_2d(st + C::A_BYTES, &mp.a_lo, &full_bar[stage], kb * C::BK, a_row);
                    uint8_t* sb = st + C::NA * C::A_BYTES;
#pragma unroll
                    for (int c = 0; c < C::NCH; ++c) {
                        const int32_t dr = c * p.chunk_stride;
                        int sub_tile = 0;
                        if (PREC == NES_PREC_EXACT) {
                            tc::tma_load_2d(sb + (sub_tile * C::BN + c * C::CR) * 128, &mp.w_lo, &full_bar[stage], kb * C::BK, w_row0 + dr);
                            ++sub_tile;
                        }
                        tc::tma_load_2d(sb + (sub_tile * C::BN + c * C::CR) * 128, &mp.w_hi, &full_bar[stage], kb * C::BK, w_row0 + dr);
                        ++sub_tile;
                        tc::tma_load_2d(sb + (sub_tile * C::BN + c * C::CR) * 128, &mp.eps, &full_bar[stage], kb * C::BK, e_row0 + dr);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // =========================== MMA issuer (one thread) ===========================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            bool first = true;
            for (int s = 0; s < p.nseg; ++s) {
                for (int kb = 0; kb < p.seg[s].K / C::BK; ++kb) {
                    tc::mbar_wait(&full_bar[stage], phase);
                    tc::tc_fence_after();
                    const uint32_t sa = tc::smem_u32(smem + stage * C::STAGE_BYTES);
                    const uint32_t sb = sa + C::NA * C::A_BYTES;
#pragma unroll
                    for (int k = 0; k < C::BK / 16; ++k) {
                        const uint64_t a_hi = tc::make_smem_desc_sw128(sa + k * 32);
                        const uint32_t acc = (first && k == 0) ? 0u : 1u;
                        // x_hi * [theta_lo | theta_hi | eps]  (FAST: [theta_hi | eps])
                        constexpr int NTOT = C::NB * C::BN;
#pragma unroll
                        for (int n0 = 0; n0 < NTOT; n0 += 256) {
                            const int nn = (NTOT - n0) < 256 ? (NTOT - n0) : 256;
                            tc::umma_f16(tmem_base + n0, a_hi, tc::make_smem_desc_sw128(sb + n0 * 128 + k * 32),
                                         tc::make_idesc_f16(128, nn), acc);
                        }
                        if (PREC == NES_PREC_EXACT) {
                            // x_lo * [theta_hi | eps]  -> columns [BN, 3BN)
                            const uint64_t a_lo = tc::make_smem_desc_sw128(sa + C::A_BYTES + k * 32);
                            constexpr int N2 = 2 * C::BN;
                            tc::umma_f16(tmem_base + C::BN, a_lo, tc::make_smem_desc_sw128(sb + C::BN * 128 + k * 32),
                                         tc::make_idesc_f16(128, N2), 1u);
                        }
                    }
                    first = false;
                    tc::umma_commit(&empty_bar[stage]);      // frees the smem stage once these MMAs retire
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
            tc::umma_commit(acc_bar);                        // accumulators complete
        }
    } else {
        // =========================== epilogue (warps 2..5 = TMEM lane quarters 2,3,0,1) ===========================
        const int q = warp & 3;
        const int et = (warp - 2) * 32 + lane;               // 0..127 among epilogue threads
        const int row = q * 32 + lane;                       // accumulator row = TMEM lane
        const int m_local = mt * C::BM + row;                // row within the member's batch
        const bool row_ok = m_local < p.rows;
        const int64_t grow = member * p.rows_pad + m_local;  // row in the wave buffers
        const float sW = 1.0f / (kThetaScale * kActScale);
        const float sE = sgn * p.sigma / kActScale;

        // member bias for this tile: (theta_b + s*sigma*eps_b), summed over segments (i2h + h2h for the LSTM)
        for (int j = et; j < C::BN; j += 128) {
            const int c = j / C::CR, u = j % C::CR;
            const int n = c * p.chunk_stride + nt * C::CR + u;          // index inside the bias tensor
            const bool ok = (nt * C::CR + u) < p.N;
            float b = 0.f;
            if (ok) {
                for (int s = 0; s < p.nseg; ++s) {
                    if (p.seg[s].b_start >= 0) {
                        const float tb = p.theta32[(int64_t)par * p.D + p.seg[s].b_start + n];
                        const float eb = __half2float(p.table[off + p.seg[s].b_pstart + n]);
                        const float nz = __fmul_rn(p.sigma, eb);
                        b += (sgn > 0.f) ? __fadd_rn(tb, nz) : __fsub_rn(tb, nz);
                    }
                }
            }
            s_bias[j] = b;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");

        tc::mbar_wait(acc_bar, 0);
        tc::tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
        constexpr int COL_HI = (PREC == NES_PREC_EXACT) ? C::BN : 0;
        constexpr int COL_EPS = (C::NB - 1) * C::BN;

        if (EPI == EPI_LSTM) {
            // columns: chunk c (0 in, 1 forget, 2 out, 3/4 maxout candidates) x 16 hidden units
            const int j0 = nt * 16;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                float pre[5][8];
#pragma unroll
                for (int c = 0; c < 5; ++c) {
                    float vh[8], ve[8];
                    tc::tmem_ld8(trow + COL_HI + c * 16 + half * 8, vh);
                    tc::tmem_ld8(trow + COL_EPS + c * 16 + half * 8, ve);
                    if (PREC == NES_PREC_EXACT) {
                        float vl[8];
                        tc::tmem_ld8(trow + c * 16 + half * 8, vl);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int u = 0; u < 8; ++u) vh[u] += vl[u];
                    } else {
                        tc::tmem_ld_wait();
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) pre[c][u] = vh[u] * sW + ve[u] * sE + s_bias[c * 16 + half * 8 + u];
                }
                if (row_ok) {
                    float* cp = p.c_state + grow * p.N + j0 + half * 8;
                    float cprev[8];
                    *reinterpret_cast<float4*>(&cprev[0]) = *reinterpret_cast<const float4*>(cp);
                    *reinterpret_cast<float4*>(&cprev[4]) = *reinterpret_cast<const float4*>(cp + 4);
                    float cn[8], hn[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const float ig = sigmoid_f(pre[0][u]), fg = sigmoid_f(pre[1][u]), og = sigmoid_f(pre[2][u]);
                        const float cand = fmaxf(pre[3][u], pre[4][u]);
                        cn[u] = __fadd_rn(__fmul_rn(fg, cprev[u]), __fmul_rn(ig, cand));
                        hn[u] = og * tanhf(cn[u]);
                    }
                    *reinterpret_cast<float4*>(cp) = *reinterpret_cast<float4*>(&cn[0]);
                    *reinterpret_cast<float4*>(cp + 4) = *reinterpret_cast<float4*>(&cn[4]);
                    __half hh[8], hl[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) split_half_dev(hn[u], kActScale, hh[u], hl[u]);
                    *reinterpret_cast<uint4*>(p.out_hi + grow * p.N + j0 + half * 8) = *reinterpret_cast<uint4*>(hh);
                    if (p.out_lo) *reinterpret_cast<uint4*>(p.out_lo + grow * p.N + j0 + half * 8) = *reinterpret_cast<uint4*>(hl);
                    if (p.h_trace) {
                        float* ht = p.h_trace + (member * p.rows + m_local) * (int64_t)p.N + j0 + half * 8;
#pragma unroll
                        for (int u = 0; u < 8; ++u) ht[u] = hn[u];
                    }
                }
            }
        } else {
            // PLAIN / LOGIT: 128 output columns, 16 at a time
            float run_max = -INFINITY, run_sum = 0.f, lo_val = -INFINITY; int run_idx = 0, lo_idx = 0;
            const int n_base = nt * C::BN;
#pragma unroll 1
            for (int c0 = 0; c0 < C::BN; c0 += 16) {
                float vh[16], ve[16];
                tc::tmem_ld16(trow + COL_HI + c0, vh);
                tc::tmem_ld16(trow + COL_EPS + c0, ve);
                if (PREC == NES_PREC_EXACT) {
                    float vl[16];
                    tc::tmem_ld16(trow + c0, vl);
                    tc::tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 16; ++u) vh[u] += vl[u];
                } else {
                    tc::tmem_ld_wait();
                }
                float y[16];
#pragma unroll
                for (int u = 0; u < 16; ++u) y[u] = vh[u] * sW + ve[u] * sE + s_bias[c0 + u];
                if (EPI == EPI_PLAIN) {
                    if (row_ok) {
#pragma unroll
                        for (int u = 0; u < 16; ++u) {
                            const int n = n_base + c0 + u;
                            if (n < p.N) {
                                if (p.out_f32) p.out_f32[grow * p.ld_out + n] = y[u];
                                if (p.out_hi) {
                                    __half a, b; split_half_dev(y[u], kActScale, a, b);
                                    p.out_hi[grow * p.N + n] = a;
                                    if (p.out_lo) p.out_lo[grow * p.N + n] = b;
                                }
                            }
                        }
                    }
                } else {
                    const float prev_max = run_max;
#pragma unroll
                    for (int u = 0; u < 16; ++u) {
                        const int n = n_base + c0 + u;
                        const float v = (n < p.N) ? y[u] : -INFINITY;
                        y[u] = v;
                        if (v > run_max) {                                   // strict >: first index wins exact ties (torch.max)
                            if (!(lo_val >= v - p.tie_delta)) { lo_val = v; lo_idx = n; }
                            run_max = v; run_idx = n;
                        }
                    }
                    if (p.logit_trace && row_ok) {
                        float* lt = p.logit_trace + (member * p.rows + m_local) * p.ld_trace;
#pragma unroll
                        for (int u = 0; u < 16; ++u) { const int n = n_base + c0 + u; if (n < p.N) lt[n] = y[u]; }
                    }
                    if (run_max > prev_max) run_sum = run_sum * __expf(prev_max - run_max);
                    if (run_max > -INFINITY) {
#pragma unroll
                        for (int u = 0; u < 16; ++u) run_sum += (PREC == NES_PREC_EXACT) ? expf(y[u] - run_max) : __expf(y[u] - run_max);
                    }
                }
            }
            if (EPI == EPI_LOGIT && row_ok) {
                LogitPartial lp; lp.mx = run_max; lp.idx = run_idx; lp.sum = run_sum; lp.lo_val = lo_val; lp.lo_idx = lo_idx;
                lp.pad[0] = lp.pad[1] = lp.pad[2] = 0;
                float4* dst = reinterpret_cast<float4*>(&p.partials[grow * p.n_tiles + nt]);
                dst[0] = reinterpret_cast<float4*>(&lp)[0]; dst[1] = reinterpret_cast<float4*>(&lp)[1];
            }
        }
        tc::tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) { tc::tc_fence_after(); tc::tmem_dealloc(tmem_base, C::TMEM_COLS); }
}

}  // namespace nes
