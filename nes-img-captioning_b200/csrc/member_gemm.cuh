// Population GEMM for perturbed members, factored as  y_m = x_m theta^T + s_m sigma x_m eps_m^T
// (member m = theta + s_m*sigma*eps_m is never materialised).
//
// Persistent, warp-specialised tcgen05 kernel, one CTA per SM.  A CTA owns whole "units" (one member x one
// 128-row block of its batch) and sweeps every weight tile j of the layer for that unit, so
//   * all CTAs walk j in lock step: the theta tile j (split-fp16 shadow) is fetched from HBM once and then hit
//     in L2 by every member of the same parent; the eps tile streams straight out of the HBM noise table and
//     is shared through L2 by the antithetic partner running on the neighbouring SM;
//   * the vocabulary arg-max / log-sum-exp of a row lives in ONE thread's registers across all tiles, so
//     neither logits nor per-tile partials ever reach HBM;
//   * TMEM holds two accumulator buffers: the epilogue of tile j overlaps the MMAs of tile j+1.
// Roles: warps 0 and 10 = TMA producers (alternate stages), warp 1 = MMA issuer (one thread), warps 2..9 =
// epilogue (TMEM lane quarter = warp & 3, two warps per quarter split the tile's columns).  theta and eps tiles sit back to back in shared memory so ONE tcgen05.mma of N = 2*BN
// produces both partial products into adjacent TMEM column ranges [theta-acc | eps-acc]; in EXACT mode the
// low-order products (x_hi*theta_lo, x_lo*theta_hi, x_lo*eps) accumulate into the same columns.  The
// epilogue combines the two ranges in fp32 with the exact scale factors and applies the fused tail
// (bias + maxout-LSTM cell, or online arg-max / log-sum-exp, or a plain store).
//
// Reference math: captioning/nets.py:98-134 (LSTMCore.forward), :195-208 (img_embed, logit, log_softmax, max).
#pragma once
#include "common.cuh"
#include "tc_ptx.cuh"

namespace nes {

enum { EPI_PLAIN = 0, EPI_LSTM = 1, EPI_LOGIT = 2 };

struct alignas(64) SegMaps {
    CUtensorMap a_hi, a_lo;    // activations [rows_total, K] fp16 (x * 2^5 split hi/lo), box BK x 128
    CUtensorMap w_hi, w_lo;    // theta shadow viewed as [n_parents*D_pad/K, K], box BK x chunk_rows
    CUtensorMap eps;           // noise table viewed as [L/K, K], box BK x chunk_rows
};
struct SegInfo {
    int32_t K;
    int32_t a_shared;          // 1: every member reads A rows [0, rows_pad) (shared fc batch)
    int64_t w_pstart;          // padded start of the weight tensor inside a noise row / theta shadow
    int64_t b_start, b_pstart; // bias tensor: flat start / padded start (-1: no bias)
};
// per row, over the WHOLE vocabulary: max logit + its first index, sum exp(logit - mx), and the EARLIEST index
// whose logit is within tie_delta of the max (the reference takes argmax over float32 LOG-PROBS, whose quantum
// near -log(V) is ~1e-6: logits closer than that tie and torch.max returns the lowest index -- see k_finalize)
struct alignas(16) LogitPartial { float mx; int32_t idx; float sum; float lo_val; int32_t lo_idx; int32_t pad[3]; };

struct GemmParams {
    SegMaps maps[2];
    SegInfo seg[2];
    int32_t nseg;
    int32_t rows, rows_pad, m_tiles, n_tiles;          // geometry
    int32_t n_units;                                   // n_members * m_tiles
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
    LogitPartial* partials;                  // [member*rows_pad + row][2]: one per column half (epilogue warp pair)
    float* logit_trace; int64_t ld_trace;    // nullable, [member*rows + row, ld_trace]
    float tie_delta;                         // one float32 ulp of log(V+1)
    const uint8_t* unit_mask;                // nullable: only units with a non-zero flag are computed (HYBRID fallback)
    // PLAIN with a batch that all members share: x theta_p^T + theta_b is the same for every member of parent p, so it is computed once
    // per parent slot (a launch of this kernel with one pseudo member per slot and sigma = 0, float32 out) and handed in as y0
    // [slot * rows_pad + row, N]; this launch then streams and multiplies the eps tiles only (2 of the 5 split-fp16 passes)
    const float* y0;
};

template <int EPI, int PREC> struct GemmCfg {
    static constexpr int BM = 128;
    static constexpr int BK = (PREC == NES_PREC_EXACT) ? 32 : 64;   // k-block: 64-byte / 128-byte swizzled rows
    static constexpr int SWZ = BK * 2;                              // swizzle span in bytes
    static constexpr int NCH = (EPI == EPI_LSTM) ? 5 : 1;           // weight-row chunks gathered per tile
    static constexpr int CR  = (EPI == EPI_LSTM) ? 16 : 128;        // rows per chunk
    static constexpr int BN  = NCH * CR;                            // 80 or 128
    static constexpr int NB  = (PREC == NES_PREC_EXACT) ? 3 : 2;    // B sub-tiles: theta_hi eps [theta_lo]
    static constexpr int NA  = (PREC == NES_PREC_EXACT) ? 2 : 1;    // A sub-tiles: x_hi [x_lo]
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = NA * A_BYTES + NB * B_BYTES;
    static constexpr int CTRL_BYTES = 2048;                         // barriers, tmem slot, bias[2][BN]
    static constexpr int SMEM_MAX = 227 * 1024;
    static constexpr int STAGES_FIT = (SMEM_MAX - 1024 - CTRL_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
    static constexpr int ACC_STRIDE = 256;                          // TMEM columns between the two accumulator buffers
    static constexpr int TMEM_COLS = 512;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + CTRL_BYTES;
    static constexpr int THREADS = 352;                             // warps 0 and 10 TMA (alternate stages), warp 1 MMA, warps 2-9 epilogue
    static_assert(STAGES >= 2, "pipeline needs at least two stages");
    static_assert(2 * BN <= ACC_STRIDE, "accumulator buffer too narrow");
    static_assert(STAGE_BYTES % 1024 == 0 && A_BYTES % 1024 == 0 && B_BYTES % 1024 == 0, "swizzle atoms need 1024-byte alignment");
};

__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ void split_half_dev(float x, float scale, __half& hi, __half& lo) {
    float xs = x * scale;
    hi = __float2half_rn(xs);
    lo = __float2half_rn(xs - __half2float(hi));
}

template <int EPI, int PREC>
__global__ void __launch_bounds__(352, 1) k_member_gemm(const __grid_constant__ GemmParams p) {
    using C = GemmCfg<EPI, PREC>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* ctrl = smem + C::STAGES * C::STAGE_BYTES;
    uint64_t* full_bar   = reinterpret_cast<uint64_t*>(ctrl);
    uint64_t* empty_bar  = full_bar + C::STAGES;
    uint64_t* tfull_bar  = empty_bar + C::STAGES;     // [2] accumulator buffer complete (MMA -> epilogue)
    uint64_t* tempty_bar = tfull_bar + 2;             // [2] accumulator buffer drained  (epilogue -> MMA)
    uint32_t* tmem_slot  = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* s_bias = reinterpret_cast<float*>(ctrl + 512);            // [2][BN] floats (<= 1 KB)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&tfull_bar[b], 1); tc::mbar_init(&tempty_bar[b], 8); }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 || warp == 10) {
        // ===== TMA producers: a stage costs one thread ~450 cycles of barrier + issue latency (measured, tools/tma_bw.cu),
        // ===== more than the MMAs of a small stage take, so two warps fill alternate stages
        const int pid = warp == 0 ? 0 : 1;
        if (lane == 0) {
            for (int s = 0; s < p.nseg; ++s) {
                tc::prefetch_tmap(&p.maps[s].a_hi); tc::prefetch_tmap(&p.maps[s].w_hi); tc::prefetch_tmap(&p.maps[s].eps);
                if (PREC == NES_PREC_EXACT) { tc::prefetch_tmap(&p.maps[s].a_lo); tc::prefetch_tmap(&p.maps[s].w_lo); }
            }
            int stage = 0; uint32_t phase = 0; uint32_t q = 0;
            for (int u = blockIdx.x; u < p.n_units; u += gridDim.x) {
                if (p.unit_mask && !p.unit_mask[u]) continue;
                const int64_t member = u / p.m_tiles;
                const int mt = u % p.m_tiles;
                const int64_t off = p.off[member];
                const int par = p.parent ? p.parent[member] : 0;
                for (int nt = 0; nt < p.n_tiles; ++nt) {
                    for (int s = 0; s < p.nseg; ++s) {
                        const SegMaps& mp = p.maps[s];
                        const SegInfo& sg = p.seg[s];
                        const int32_t a_row = (int32_t)((sg.a_shared ? 0 : member * p.rows_pad) + mt * C::BM);
                        const int32_t w_row0 = (int32_t)(((int64_t)par * p.D_pad + sg.w_pstart) / sg.K) + nt * C::CR;
                        const int32_t e_row0 = (int32_t)((off + sg.w_pstart) / sg.K) + nt * C::CR;
                        const int nkb = sg.K / C::BK;
                        for (int kb = 0; kb < nkb; ++kb) {
                            if (((q++) & 1) != pid) { if (++stage == C::STAGES) { stage = 0; phase ^= 1; } continue; }
                            tc::mbar_wait(&empty_bar[stage], phase ^ 1);
                            uint8_t* st = smem + stage * C::STAGE_BYTES;
                            const bool eps_only = EPI == EPI_PLAIN && PREC == NES_PREC_EXACT && p.y0 != nullptr;
                            tc::mbar_expect_tx(&full_bar[stage], eps_only ? C::NA * C::A_BYTES + C::B_BYTES : C::STAGE_BYTES);
                            tc::tma_load_2d(st, &mp.a_hi, &full_bar[stage], kb * C::BK, a_row);
                            if (PREC == NES_PREC_EXACT) tc::tma_load_2d(st + C::A_BYTES, &mp.a_lo, &full_bar[stage], kb * C::BK, a_row);
                            uint8_t* sb = st + C::NA * C::A_BYTES;      // [theta_hi | eps | theta_lo]
                            if (C::NCH > 1) {       // the gate blocks of this tile in one 3-D request per operand
                                tc::tma_load_3d(sb, &mp.w_hi, &full_bar[stage], kb * C::BK, w_row0, 0);
                                tc::tma_load_3d(sb + C::B_BYTES, &mp.eps, &full_bar[stage], kb * C::BK, e_row0, 0);
                                if (PREC == NES_PREC_EXACT) tc::tma_load_3d(sb + 2 * C::B_BYTES, &mp.w_lo, &full_bar[stage], kb * C::BK, w_row0, 0);
                            } else if (eps_only) {
                                tc::tma_load_2d(sb + C::B_BYTES, &mp.eps, &full_bar[stage], kb * C::BK, e_row0);
                            } else {
                                tc::tma_load_2d(sb, &mp.w_hi, &full_bar[stage], kb * C::BK, w_row0);
                                tc::tma_load_2d(sb + C::B_BYTES, &mp.eps, &full_bar[stage], kb * C::BK, e_row0);
                                if (PREC == NES_PREC_EXACT) tc::tma_load_2d(sb + 2 * C::B_BYTES, &mp.w_lo, &full_bar[stage], kb * C::BK, w_row0);
                            }
                            if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // =========================== MMA issuer (one thread) ===========================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            uint32_t it = 0;                                  // tile counter of this CTA
            constexpr uint32_t idesc2 = tc::make_idesc_f16(128, 2 * C::BN);
            constexpr uint32_t idesc1 = tc::make_idesc_f16(128, C::BN);
            for (int u = blockIdx.x; u < p.n_units; u += gridDim.x) {
                if (p.unit_mask && !p.unit_mask[u]) continue;
                for (int nt = 0; nt < p.n_tiles; ++nt, ++it) {
                    const uint32_t buf = it & 1;
                    tc::mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);     // epilogue has drained this buffer
                    tc::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * C::ACC_STRIDE;
                    bool first = true;
                    for (int s = 0; s < p.nseg; ++s) {
                        const int nkb = p.seg[s].K / C::BK;
                        for (int kb = 0; kb < nkb; ++kb) {
                            tc::mbar_wait(&full_bar[stage], phase);
                            tc::tc_fence_after();
                            const uint32_t sa = tc::smem_u32(smem + stage * C::STAGE_BYTES);
                            const uint32_t sb = sa + C::NA * C::A_BYTES;
#pragma unroll
                            for (int k = 0; k < C::BK / 16; ++k) {
                                const uint64_t a_hi = tc::make_smem_desc<C::SWZ>(sa + k * 32);
                                if (EPI == EPI_PLAIN && PREC == NES_PREC_EXACT && p.y0 != nullptr) {
                                    // (x_hi + x_lo) * eps -> eps-acc; the theta term comes in through y0
                                    const uint64_t b_eps = tc::make_smem_desc<C::SWZ>(sb + C::B_BYTES + k * 32);
                                    tc::umma_f16(d_tmem + C::BN, a_hi, b_eps, idesc1, (first && k == 0) ? 0u : 1u);
                                    tc::umma_f16(d_tmem + C::BN, tc::make_smem_desc<C::SWZ>(sa + C::A_BYTES + k * 32), b_eps, idesc1, 1u);
                                    continue;
                                }
                                // x_hi * [theta_hi | eps] -> [theta-acc | eps-acc]
                                tc::umma_f16(d_tmem, a_hi, tc::make_smem_desc<C::SWZ>(sb + k * 32), idesc2, (first && k == 0) ? 0u : 1u);
                                if (PREC == NES_PREC_EXACT) {
                                    // x_hi * theta_lo -> theta-acc ;  x_lo * [theta_hi | eps] -> both
                                    tc::umma_f16(d_tmem, a_hi, tc::make_smem_desc<C::SWZ>(sb + 2 * C::B_BYTES + k * 32), idesc1, 1u);
                                    const uint64_t a_lo = tc::make_smem_desc<C::SWZ>(sa + C::A_BYTES + k * 32);
                                    tc::umma_f16(d_tmem, a_lo, tc::make_smem_desc<C::SWZ>(sb + k * 32), idesc2, 1u);
                                }
                            }
                            first = false;
                            tc::umma_commit(&empty_bar[stage]);      // frees the smem stage once these MMAs retire
                            if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                        }
                    }
                    tc::umma_commit(&tfull_bar[buf]);                // accumulators of this tile complete
                }
            }
        }
    } else {
        // ======= epilogue: warps 2..9.  TMEM lane quarter = warp & 3 (rows), column half = (warp - 2) >> 2 =======
        const int q = warp & 3;
        const int hf = (warp - 2) >> 2;                      // which half of the tile's columns / hidden units
        const int et = (warp - 2) * 32 + lane;               // 0..255 among epilogue threads
        const int row = q * 32 + lane;                       // accumulator row = TMEM lane
        const float sW = 1.0f / (kThetaScale * kActScale);
        constexpr int COL_EPS = C::BN;
        constexpr float kLog2e = 1.4426950408889634f;
        // member bias of tile (u, nt), element et: theta_b + s*sigma*eps_b, summed over segments (i2h + h2h for the LSTM).
        // Loaded one tile AHEAD into a register so the dependent global loads never sit on the epilogue's critical path.
        auto bias_val = [&](int u, int nt) -> float {
            if (u >= p.n_units || et >= C::BN) return 0.f;
            const int64_t member = u / p.m_tiles;
            const int c = et / C::CR, uu = et % C::CR;
            if (nt * C::CR + uu >= p.N) return 0.f;
            const int n = c * p.chunk_stride + nt * C::CR + uu;          // index inside the bias tensor
            const int64_t off = p.off[member];
            const float sgn = p.sign[member];
            const int par = p.parent ? p.parent[member] : 0;
            float b = 0.f;
            for (int s = 0; s < p.nseg; ++s) {
                if (p.seg[s].b_start >= 0) {
                    const float tb = (EPI == EPI_PLAIN && p.y0) ? 0.f : p.theta32[(int64_t)par * p.D + p.seg[s].b_start + n];
                    const float nz = __fmul_rn(p.sigma, __half2float(p.table[off + p.seg[s].b_pstart + n]));
                    b += (sgn > 0.f) ? __fadd_rn(tb, nz) : __fsub_rn(tb, nz);
                }
            }
            return b;
        };
        auto next_active = [&](int u) -> int { while (p.unit_mask && u < p.n_units && !p.unit_mask[u]) u += gridDim.x; return u; };
        float b_cur = bias_val(next_active(blockIdx.x), 0);
        uint32_t it = 0;
        for (int u = blockIdx.x; u < p.n_units; u += gridDim.x) {
            if (p.unit_mask && !p.unit_mask[u]) continue;
            const int64_t member = u / p.m_tiles;
            const int mt = u % p.m_tiles;
            const float sgn = p.sign[member];
            const int m_local = mt * C::BM + row;                // row within the member's batch
            const bool row_ok = m_local < p.rows;
            const int64_t grow = member * p.rows_pad + m_local;  // row in the wave buffers
            const float sE = sgn * p.sigma / kActScale;
            // running arg-max / log-sum-exp of this row over this thread's half of every vocabulary tile (LOGIT)
            float run_max = -INFINITY, run_sum = 0.f, lo_val = -INFINITY; int run_idx = 0, lo_idx = 0;

            for (int nt = 0; nt < p.n_tiles; ++nt, ++it) {
                const uint32_t buf = it & 1;
                float* bias = s_bias + buf * C::BN;
                if (et < C::BN) bias[et] = b_cur;
                asm volatile("bar.sync 1, 256;" ::: "memory");
                b_cur = (nt + 1 < p.n_tiles) ? bias_val(u, nt + 1) : bias_val(next_active(u + gridDim.x), 0);
                float cprev[8];                                   // LSTM: this row's cell state, fetched while the MMAs run
                if (EPI == EPI_LSTM && row_ok) {
                    const float4* cp4 = reinterpret_cast<const float4*>(p.c_state + grow * p.N + nt * 16 + hf * 8);
                    *reinterpret_cast<float4*>(&cprev[0]) = cp4[0];
                    *reinterpret_cast<float4*>(&cprev[4]) = cp4[1];
                }
                tc::mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
                tc::tc_fence_after();
                const uint32_t trow = tmem_base + buf * C::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
                {
                if (EPI == EPI_LSTM) {
                    // columns: chunk c (0 in, 1 forget, 2 out, 3/4 maxout candidates) x 16 hidden units; this warp owns 8 of them
                    const int j0 = nt * 16 + hf * 8;
                    float pre[5][8];
#pragma unroll
                    for (int c = 0; c < 5; ++c) {
                        float vh[8], ve[8];
                        tc::tmem_ld8(trow + c * 16 + hf * 8, vh);
                        tc::tmem_ld8(trow + COL_EPS + c * 16 + hf * 8, ve);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int uu = 0; uu < 8; ++uu) pre[c][uu] = vh[uu] * sW + ve[uu] * sE + bias[c * 16 + hf * 8 + uu];
                    }
                    if (row_ok) {
                        float cn[8], hn[8];
#pragma unroll
                        for (int uu = 0; uu < 8; ++uu) {
                            const float ig = sigmoid_f(pre[0][uu]), fg = sigmoid_f(pre[1][uu]), og = sigmoid_f(pre[2][uu]);
                            const float cand = fmaxf(pre[3][uu], pre[4][uu]);
                            cn[uu] = __fadd_rn(__fmul_rn(fg, cprev[uu]), __fmul_rn(ig, cand));
                            hn[uu] = og * tanhf(cn[uu]);
                        }
                        float* cp = p.c_state + grow * p.N + j0;
                        *reinterpret_cast<float4*>(cp) = *reinterpret_cast<float4*>(&cn[0]);
                        *reinterpret_cast<float4*>(cp + 4) = *reinterpret_cast<float4*>(&cn[4]);
                        __half hh[8], hl[8];
#pragma unroll
                        for (int uu = 0; uu < 8; ++uu) split_half_dev(hn[uu], kActScale, hh[uu], hl[uu]);
                        *reinterpret_cast<uint4*>(p.out_hi + grow * p.N + j0) = *reinterpret_cast<uint4*>(hh);
                        if (p.out_lo) *reinterpret_cast<uint4*>(p.out_lo + grow * p.N + j0) = *reinterpret_cast<uint4*>(hl);
                        if (p.h_trace) {
                            float* ht = p.h_trace + (member * p.rows + m_local) * (int64_t)p.N + j0;
#pragma unroll
                            for (int uu = 0; uu < 8; ++uu) ht[uu] = hn[uu];
                        }
                    }
                } else {
                    // PLAIN / LOGIT: this warp owns 64 of the tile's 128 output columns, 32 at a time
                    const int n_base = nt * C::BN;
#pragma unroll 1
                    for (int c0 = hf * 64; c0 < hf * 64 + 64; c0 += 32) {
                        float y[32];
                        if (EPI == EPI_PLAIN && PREC == NES_PREC_EXACT && p.y0 != nullptr) {
                            float ve[32];
                            tc::tmem_ld32(trow + COL_EPS + c0, ve);
                            tc::tmem_ld_wait();
                            const int par = p.parent ? p.parent[member] : 0;
                            const float* y0 = p.y0 + ((int64_t)par * p.rows_pad + m_local) * p.N + n_base + c0;
                            float yp[32];
                            if (row_ok && n_base + c0 + 32 <= p.N && (p.N & 3) == 0) {
#pragma unroll
                                for (int v4 = 0; v4 < 8; ++v4) *reinterpret_cast<float4*>(&yp[v4 * 4]) = __ldg(reinterpret_cast<const float4*>(y0) + v4);
                            } else {
#pragma unroll
                                for (int uu = 0; uu < 32; ++uu) yp[uu] = (row_ok && n_base + c0 + uu < p.N) ? y0[uu] : 0.f;
                            }
#pragma unroll
                            for (int uu = 0; uu < 32; ++uu) y[uu] = ve[uu] * sE + bias[c0 + uu] + yp[uu];
                        } else {
                            float vh[32], ve[32];
                            tc::tmem_ld32(trow + c0, vh);
                            tc::tmem_ld32(trow + COL_EPS + c0, ve);
                            tc::tmem_ld_wait();
#pragma unroll
                            for (int uu = 0; uu < 32; ++uu) y[uu] = vh[uu] * sW + ve[uu] * sE + bias[c0 + uu];
                        }
                        if (EPI == EPI_PLAIN) {
                            // a thread owns 32 consecutive columns of its row: 16-byte stores when the whole chunk is inside N and the
                            // rows are 16-byte aligned (element-wise 2-byte stores from 32 lanes in 32 different rows made this epilogue,
                            // not the MMAs, the bound of img_embed: 65 us per tile whatever the population)
                            const bool vec_ok = n_base + c0 + 32 <= p.N && (p.N & 7) == 0 && (!p.out_f32 || (p.ld_out & 3) == 0);
                            if (row_ok && vec_ok) {
                                if (p.out_f32) {
                                    float4* o = reinterpret_cast<float4*>(p.out_f32 + grow * p.ld_out + n_base + c0);
#pragma unroll
                                    for (int v4 = 0; v4 < 8; ++v4) o[v4] = *reinterpret_cast<float4*>(&y[v4 * 4]);
                                }
                                if (p.out_hi) {
                                    __align__(16) __half hh[32];
                                    __align__(16) __half hl[32];
#pragma unroll
                                    for (int uu = 0; uu < 32; ++uu) split_half_dev(y[uu], kActScale, hh[uu], hl[uu]);
                                    uint4* oh = reinterpret_cast<uint4*>(p.out_hi + grow * p.N + n_base + c0);
#pragma unroll
                                    for (int v8 = 0; v8 < 4; ++v8) oh[v8] = *reinterpret_cast<uint4*>(&hh[v8 * 8]);
                                    if (p.out_lo) {
                                        uint4* ol = reinterpret_cast<uint4*>(p.out_lo + grow * p.N + n_base + c0);
#pragma unroll
                                        for (int v8 = 0; v8 < 4; ++v8) ol[v8] = *reinterpret_cast<uint4*>(&hl[v8 * 8]);
                                    }
                                }
                            } else if (row_ok) {
#pragma unroll
                                for (int uu = 0; uu < 32; ++uu) {
                                    const int n = n_base + c0 + uu;
                                    if (n < p.N) {
                                        if (p.out_f32) p.out_f32[grow * p.ld_out + n] = y[uu];
                                        if (p.out_hi) {
                                            __half a, b; split_half_dev(y[uu], kActScale, a, b);
                                            p.out_hi[grow * p.N + n] = a;
                                            if (p.out_lo) p.out_lo[grow * p.N + n] = b;
                                        }
                                    }
                                }
                            }
                        } else {
                            if (n_base + c0 + 32 > p.N) {                            // ragged last tile: mask the columns past V+1
#pragma unroll
                                for (int uu = 0; uu < 32; ++uu) if (n_base + c0 + uu >= p.N) y[uu] = -INFINITY;
                            }
                            if (p.logit_trace && row_ok) {
                                float* lt = p.logit_trace + (member * p.rows + m_local) * p.ld_trace;
#pragma unroll
                                for (int uu = 0; uu < 32; ++uu) { const int n = n_base + c0 + uu; if (n < p.N) lt[n] = y[uu]; }
                            }
                            float cm = y[0];
#pragma unroll
                            for (int uu = 1; uu < 32; ++uu) cm = fmaxf(cm, y[uu]);
                            if (cm > run_max) {                                      // rare after the first tiles: a new running max in this chunk
                                const float prev_max = run_max;
#pragma unroll
                                for (int uu = 0; uu < 32; ++uu) {
                                    const float v = y[uu];
                                    if (v > run_max) {                               // strict >: first index wins exact ties (torch.max)
                                        if (!(lo_val >= v - p.tie_delta)) {          // the tie candidate fell out of the window
                                            if (run_max >= v - p.tie_delta) { lo_val = run_max; lo_idx = run_idx; }   // previous max still inside, and earlier
                                            else { lo_val = v; lo_idx = n_base + c0 + uu; }
                                        }
                                        run_max = v; run_idx = n_base + c0 + uu;
                                    }
                                }
                                run_sum *= tc::ex2_approx((prev_max - run_max) * kLog2e);
                            }
                            if (run_max > -INFINITY) {
                                const float ml2 = run_max * kLog2e;
                                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                                for (int uu = 0; uu < 32; uu += 2) {
                                    s0 += tc::ex2_approx(fmaf(y[uu], kLog2e, -ml2));
                                    s1 += tc::ex2_approx(fmaf(y[uu + 1], kLog2e, -ml2));
                                }
                                run_sum += s0 + s1;
                            }
                        }
                    }
                }
                }
                // this accumulator buffer may be overwritten by the MMAs of tile it+2
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&tempty_bar[buf]);
            }
            if (EPI == EPI_LOGIT && row_ok) {
                LogitPartial lp; lp.mx = run_max; lp.idx = run_idx; lp.sum = run_sum; lp.lo_val = lo_val; lp.lo_idx = lo_idx;
                lp.pad[0] = lp.pad[1] = lp.pad[2] = 0;
                float4* dst = reinterpret_cast<float4*>(&p.partials[grow * 2 + hf]);
                dst[0] = reinterpret_cast<float4*>(&lp)[0]; dst[1] = reinterpret_cast<float4*>(&lp)[1];
            }
        }
        tc::tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) { tc::tc_fence_after(); tc::tmem_dealloc(tmem_base, C::TMEM_COLS); }
}

}  // namespace nes
