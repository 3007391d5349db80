// Member GEMMs on MATERIALISED member weights, for the two layers whose weights are reused by every decode step.
//
// The factored kernel (member_gemm.cuh) never forms theta +- sigma*eps; that is the right trade for a weight that is
// used once per generation (img_embed), but the LSTM weights are used 17 times and logit.weight 16 times per
// generation, and in split-fp16 the factored form costs 5 MMA passes per k-step (x_hi*[theta_hi|eps], x_hi*theta_lo,
// x_lo*[theta_hi|eps]) and three operand tiles per stage.  With 180 GB of HBM a wave of 1000 members can simply hold
//     W_m = fl32(theta_p +- fl32(sigma*eps_m))            (exactly the reference's member weights, algorithm/nets.py:113)
// as split fp16 (LSTM: W*2^8 = hi + lo, 10.8 MB per member) or single fp16 (vocabulary projection of the HYBRID
// scheme, 9.7 MB per member), written once per generation by k_materialise.  The decode GEMMs then are plain
//     LSTM   : acc = x_hi*W_hi + x_lo*W_hi + x_hi*W_lo   -> ONE 240-column accumulator (48 hidden units x 5 gates),
//              3 MMAs of N = 240 per k-step, TMEM double-buffered, maxout-LSTM cell fused in the epilogue;
//     LOGITC : acc = x_hi*W16, N = 256, the member's x block (128 x K fp16) stays RESIDENT in shared memory for all 38
//              vocabulary tiles (the factored kernel re-fetched it per tile: L2 -> SM traffic was its bound),
//              online max / sum-exp per row in registers, arg-max candidates appended to a per-row list in HBM.
// Same roles as the factored kernel: warps 0 and 10 TMA producers, warp 1 MMA issuer, warps 2..9 epilogue.
//
// Reference math: captioning/nets.py:98-134 (LSTMCore.forward), :202-208 (logit, log_softmax, max).
#pragma once
#include "member_gemm.cuh"

namespace nes {

enum { DEN_LSTM = 0, DEN_LOGITC = 1, DEN_PLAIN = 2 };
#ifndef NES_LSTM_TILE_UNITS
#define NES_LSTM_TILE_UNITS 48
#endif
constexpr int kLstmTileUnits = NES_LSTM_TILE_UNITS;   // hidden units per LSTM tile (multiple of 16); accumulator column = gate * units + unit
constexpr int kLstmHalfUnits = kLstmTileUnits / 2;    // per epilogue warp of a TMEM lane quarter
constexpr int kLstmSubBatches = kLstmHalfUnits / 8;
static_assert(kLstmTileUnits % 16 == 0 && 5 * kLstmTileUnits <= 256, "LSTM tile shape");
constexpr int kLstmTileCols = 5 * kLstmTileUnits;
constexpr int kCandCap = 64;                // candidate entries per list (append-only; overflow -> split-fp16 fallback)
#ifndef NES_LOGITC_WARPS
#define NES_LOGITC_WARPS 16                 // epilogue warps of the LOGITC kernel: 8 or 16 (A/B builds)
#endif
constexpr int kCandLists = NES_LOGITC_WARPS / 4;               // lists per row: one per LOGITC epilogue warp of the row's TMEM lane quarter (64 of a tile's 256 columns each)

struct alignas(16) CandHead { float mx; float sum; int32_t n; int32_t pad; };
struct alignas(8) CandEntry { float v; int32_t idx; };
static_assert(kCandCap >= 32, "refine_row reads the first 32 entries of every list unconditionally");

struct alignas(64) DenseParams {
    CUtensorMap a_hi[2], a_lo[2];           // activations of segment s: [rows_total, K_s] fp16, box BK x 128
    CUtensorMap b_hi[2], b_lo[2];           // materialised weights of segment s: [W * b_rows, K_s] fp16, box BK x BN
    int32_t K[2]; int32_t nseg;
    int32_t rows, rows_pad, m_tiles, n_tiles, n_units;
    int32_t N;                              // valid outputs: R (LSTM hidden units) or V+1 (LOGITC)
    int32_t b_rows;                         // weight rows per member in the materialised buffer
    int32_t a_shared;                       // 1: every member reads activation rows [0, rows_pad) (shared fc batch, PLAIN)
    const float* bias; int32_t bias_stride; // [W][bias_stride] member bias in accumulator-column order
    float* c_state; __half* out_hi; __half* out_lo; float* h_trace;       // LSTM (PLAIN: out_hi / out_lo and / or out_f32)
    float* out_f32; int32_t ld_out;                                       // PLAIN: [member*rows_pad + row, ld_out] float32 (nullable)
    CandHead* heads; CandEntry* entries; const float* margin;            // LOGITC
    // LOGITC with few units (the elite candidates of a NIC-ES generation): the vocabulary sweep of a unit is cut into n_split ranges of
    // tiles_per_split tiles, one work item each, so that 2 units keep more than 2 SMs busy; row `grow` then owns n_split * kCandLists lists
    int32_t n_split, tiles_per_split;
    // the row's |h|^2 in pieces, written by the cell kernel that produced h (LSTM: one piece per (tile, column half)) and summed by the
    // LOGITC epilogue into the HYBRID error margin (hybrid.cuh: k_row_margin's formula) -- saves a pass over h per decode step
    float* hsq; int32_t hsq_n;                                            // [row * hsq_n + piece]; LSTM writes, LOGITC reads (nullable)
    const int32_t* mg_parent; const float* mg_theta_rowmax; const float* mg_eps_rowmax; const float* mg_scale_max; float mg_sigma; float* margin_out;
};

template <int EPI> struct DenseCfg {
    static constexpr bool LSTM = EPI == DEN_LSTM;
    static constexpr bool SPLIT = EPI != DEN_LOGITC;                     // split-fp16 operands streamed per stage (LSTM, PLAIN)
    static constexpr int BM = 128;
    static constexpr int BK = SPLIT ? 32 : 64;
    static constexpr int SWZ = BK * 2;
    static constexpr int BN = LSTM ? kLstmTileCols : 256;
    static constexpr int NA = SPLIT ? 2 : 1;
    static constexpr int NB = SPLIT ? 2 : 1;
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int ARES_KB = SPLIT ? 0 : 8;                        // resident A k-blocks (LOGITC needs K <= 512)
    static constexpr int ARES_BYTES = ARES_KB * A_BYTES;
    static constexpr int STAGE_BYTES = (SPLIT ? NA * A_BYTES : 0) + NB * B_BYTES;
    static constexpr int CTRL_BYTES = 2048;                              // barriers [0,512), tmem slot, bias [512, 1536)
    static constexpr int SMEM_MAX = 227 * 1024;
    static constexpr int STAGES_FIT = (SMEM_MAX - 1024 - CTRL_BYTES - ARES_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 6 ? 6 : STAGES_FIT;
    static constexpr int ACC_STRIDE = 256;
    static constexpr int TMEM_COLS = 512;
    static constexpr int SMEM_BYTES = ARES_BYTES + STAGES * STAGE_BYTES + 1024 + CTRL_BYTES;
    // warps 0 and 10 TMA producers, warp 1 MMA issuer, the rest epilogue: 8 warps (2 per TMEM lane quarter), or 16 for LOGITC, whose
    // epilogue (scale + bias, running max, one exp per logit) is the bound once K is small (config 4: K = 128, 512 rows)
    static constexpr int EPI_WARPS = EPI == DEN_LOGITC ? NES_LOGITC_WARPS : 8;
    static constexpr int THREADS = (3 + EPI_WARPS) * 32;
    static_assert(STAGES >= 3, "pipeline too shallow");
    static_assert(BN <= ACC_STRIDE && BN % 16 == 0, "accumulator shape");
    static_assert(STAGE_BYTES % 1024 == 0 && A_BYTES % 1024 == 0 && B_BYTES % 1024 == 0, "swizzle atoms need 1024-byte alignment");
    static_assert(SMEM_BYTES <= SMEM_MAX, "shared memory budget");
};

template <int EPI>
__global__ void __launch_bounds__(DenseCfg<EPI>::THREADS, 1) k_member_dense(const __grid_constant__ DenseParams p) {
    using C = DenseCfg<EPI>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* ares = smem;                                             // LOGITC: resident x block, ARES_KB k-blocks
    uint8_t* ring = smem + C::ARES_BYTES;
    uint8_t* ctrl = ring + C::STAGES * C::STAGE_BYTES;
    uint64_t* full_bar   = reinterpret_cast<uint64_t*>(ctrl);
    uint64_t* empty_bar  = full_bar + C::STAGES;
    uint64_t* tfull_bar  = empty_bar + C::STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint64_t* afull_bar  = tempty_bar + 2;                            // [8] resident A k-block landed
    uint64_t* aempty_bar = afull_bar + 8;                             // [8] last MMA of the unit that reads it retired
    uint32_t* tmem_slot  = reinterpret_cast<uint32_t*>(aempty_bar + 8);
    float* s_bias = reinterpret_cast<float*>(ctrl + 512);             // [BN]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&tfull_bar[b], 1); tc::mbar_init(&tempty_bar[b], C::EPI_WARPS); }
        for (int k = 0; k < 8; ++k) { tc::mbar_init(&afull_bar[k], 1); tc::mbar_init(&aempty_bar[k], 1); }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // work items: LSTM = (unit, weight tile) pairs in tile-minor order, so the CTAs that run concurrently share a unit's
    // activation block through L2 (it is fetched from HBM once, not once per tile); LOGITC = whole units, because the
    // running arg-max / sum-exp of a row lives in one thread's registers across all vocabulary tiles
    const int n_items = C::SPLIT ? p.n_units * p.n_tiles : p.n_units * p.n_split;

    if (warp == 0 || warp == 10) {
        // =========================== TMA producers (two warps fill alternate stages) ===========================
        const int pid = warp == 0 ? 0 : 1;
        if (lane == 0) {
            for (int s = 0; s < p.nseg; ++s) {
                tc::prefetch_tmap(&p.a_hi[s]); tc::prefetch_tmap(&p.b_hi[s]);
                if (C::SPLIT) { tc::prefetch_tmap(&p.a_lo[s]); tc::prefetch_tmap(&p.b_lo[s]); }
            }
            int stage = 0; uint32_t phase = 0; uint32_t q = 0; uint32_t uc = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++uc) {
                const int u = C::SPLIT ? item / p.n_tiles : item / p.n_split;
                const int nt0 = C::SPLIT ? item - u * p.n_tiles : (item - u * p.n_split) * p.tiles_per_split;
                const int nt1 = C::SPLIT ? nt0 + 1 : min(p.n_tiles, nt0 + p.tiles_per_split);
                const int64_t member = u / p.m_tiles;
                const int mt = u % p.m_tiles;
                const int32_t a_row = (int32_t)((p.a_shared ? 0 : member * p.rows_pad) + mt * C::BM);
                const int32_t b_row0 = (int32_t)(member * p.b_rows);
                if (!C::SPLIT && pid == 0) {
                    // the unit's x block: k-block kb may be overwritten once the previous unit's last tile has consumed it
                    const int nkb = p.K[0] / C::BK;
                    for (int kb = 0; kb < nkb; ++kb) {
                        if (uc > 0) tc::mbar_wait(&aempty_bar[kb], (uc - 1) & 1);
                        tc::mbar_expect_tx(&afull_bar[kb], C::A_BYTES);
                        tc::tma_load_2d(ares + kb * C::A_BYTES, &p.a_hi[0], &afull_bar[kb], kb * C::BK, a_row);
                    }
                }
                for (int nt = nt0; nt < nt1; ++nt) {
                    for (int s = 0; s < p.nseg; ++s) {
                        const int nkb = p.K[s] / C::BK;
                        for (int kb = 0; kb < nkb; ++kb) {
                            if (((q++) & 1) != pid) { if (++stage == C::STAGES) { stage = 0; phase ^= 1; } continue; }
                            tc::mbar_wait(&empty_bar[stage], phase ^ 1);
                            uint8_t* st = ring + stage * C::STAGE_BYTES;
                            tc::mbar_expect_tx(&full_bar[stage], C::STAGE_BYTES);
                            if (C::SPLIT) {
                                tc::tma_load_2d(st, &p.a_hi[s], &full_bar[stage], kb * C::BK, a_row);
                                tc::tma_load_2d(st + C::A_BYTES, &p.a_lo[s], &full_bar[stage], kb * C::BK, a_row);
                                uint8_t* sb = st + 2 * C::A_BYTES;
                                tc::tma_load_2d(sb, &p.b_hi[s], &full_bar[stage], kb * C::BK, b_row0 + nt * C::BN);
                                tc::tma_load_2d(sb + C::B_BYTES, &p.b_lo[s], &full_bar[stage], kb * C::BK, b_row0 + nt * C::BN);
                            } else {
                                tc::tma_load_2d(st, &p.b_hi[s], &full_bar[stage], kb * C::BK, b_row0 + nt * C::BN);
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
            int stage = 0; uint32_t phase = 0; uint32_t it = 0; uint32_t uc = 0;
            constexpr uint32_t idesc = tc::make_idesc_f16(128, C::BN);
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++uc) {
                const int nt0 = C::SPLIT ? item % p.n_tiles : (item % p.n_split) * p.tiles_per_split;
                const int nt1 = C::SPLIT ? nt0 + 1 : min(p.n_tiles, nt0 + p.tiles_per_split);
                for (int nt = nt0; nt < nt1; ++nt, ++it) {
                    const uint32_t buf = it & 1;
                    tc::mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);
                    tc::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * C::ACC_STRIDE;
                    bool first = true;
                    for (int s = 0; s < p.nseg; ++s) {
                        const int nkb = p.K[s] / C::BK;
                        for (int kb = 0; kb < nkb; ++kb) {
                            if (!C::SPLIT && nt == nt0) tc::mbar_wait(&afull_bar[kb], uc & 1);
                            tc::mbar_wait(&full_bar[stage], phase);
                            tc::tc_fence_after();
                            const uint32_t st = tc::smem_u32(ring + stage * C::STAGE_BYTES);
#pragma unroll
                            for (int k = 0; k < C::BK / 16; ++k) {
                                if (C::SPLIT) {
                                    const uint64_t a_hi = tc::make_smem_desc<C::SWZ>(st + k * 32);
                                    const uint64_t a_lo = tc::make_smem_desc<C::SWZ>(st + C::A_BYTES + k * 32);
                                    const uint64_t b_hi = tc::make_smem_desc<C::SWZ>(st + 2 * C::A_BYTES + k * 32);
                                    const uint64_t b_lo = tc::make_smem_desc<C::SWZ>(st + 2 * C::A_BYTES + C::B_BYTES + k * 32);
                                    tc::umma_f16(d_tmem, a_hi, b_hi, idesc, (first && k == 0) ? 0u : 1u);
                                    tc::umma_f16(d_tmem, a_lo, b_hi, idesc, 1u);
                                    tc::umma_f16(d_tmem, a_hi, b_lo, idesc, 1u);
                                } else {
                                    const uint64_t a = tc::make_smem_desc<C::SWZ>(tc::smem_u32(ares + kb * C::A_BYTES) + k * 32);
                                    tc::umma_f16(d_tmem, a, tc::make_smem_desc<C::SWZ>(st + k * 32), idesc, (first && k == 0) ? 0u : 1u);
                                }
                            }
                            first = false;
                            tc::umma_commit(&empty_bar[stage]);
                            if (!C::SPLIT && nt == nt1 - 1) tc::umma_commit(&aempty_bar[kb]);
                            if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                        }
                    }
                    tc::umma_commit(&tfull_bar[buf]);
                }
            }
        }
    } else {
        // ======= epilogue: warps 2..9 (and 11..18 for LOGITC).  TMEM lane quarter = warp & 3 (rows); hf = which of the warps of that
        // quarter this is = which column half (quarter for LOGITC) of every tile it owns =======
        const int q = warp & 3;
        const int ew = warp < 10 ? warp - 2 : warp - 3;      // 0 .. EPI_WARPS-1 among epilogue warps
        const int hf = warp < 10 ? (warp - 2) >> 2 : 2 + ((warp - 11) >> 2);
        const int et = ew * 32 + lane;                       // index among epilogue threads
        const int row = q * 32 + lane;
        const float sW = 1.0f / (kThetaScale * kActScale);
        constexpr float kLog2e = 1.4426950408889634f;
        // member bias of tile (u, nt), accumulator column et; fetched one tile ahead
        auto bias_val = [&](int u, int nt) -> float {
            if (u >= p.n_units || et >= C::BN) return 0.f;
            const int64_t member = u / p.m_tiles;
            const int col = nt * C::BN + et;
            return col < p.bias_stride ? p.bias[member * p.bias_stride + col] : 0.f;
        };
        float b_cur = C::SPLIT ? bias_val(blockIdx.x / p.n_tiles, blockIdx.x % p.n_tiles)
                               : bias_val(blockIdx.x / p.n_split, (blockIdx.x % p.n_split) * p.tiles_per_split);
        uint32_t it = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int u = C::SPLIT ? item / p.n_tiles : item / p.n_split;
            const int sp = C::SPLIT ? 0 : item - u * p.n_split;
            const int nt0 = C::SPLIT ? item - u * p.n_tiles : sp * p.tiles_per_split;
            const int nt1 = C::SPLIT ? nt0 + 1 : min(p.n_tiles, nt0 + p.tiles_per_split);
            const int64_t member = u / p.m_tiles;
            const int mt = u % p.m_tiles;
            const int m_local = mt * C::BM + row;
            const bool row_ok = m_local < p.rows;
            const int64_t grow = member * p.rows_pad + m_local;
            float run_max = -INFINITY, run_sum = 0.f; int n_cand = 0;
            float margin = 0.f;
            if (EPI == DEN_LOGITC && row_ok) {
                if (p.hsq) {                                              // k_row_margin's bound from the pieces of |h|^2 (hybrid.cuh)
                    float ss = 0.f;
                    for (int i = 0; i < p.hsq_n; ++i) ss += p.hsq[grow * p.hsq_n + i];
                    const int par = p.mg_parent ? p.mg_parent[member] : 0;
                    const float smax = p.mg_scale_max ? p.mg_scale_max[par] : 1.f;
                    const float wsum = 2.f * (p.mg_theta_rowmax[par] + fabsf(p.mg_sigma) * p.mg_eps_rowmax[0] * smax);
                    margin = 2.f * 1.1f * 4.8828125e-4f * sqrtf(ss) * 1.0001f * (wsum + 1e-4f) + 1e-7f;
                    if (hf == 0) p.margin_out[grow] = margin;             // k_refine_finalize filters the lists with the same number
                } else margin = p.margin[grow];
            }
            const int64_t list = EPI != DEN_LOGITC ? 0 : (grow * p.n_split + sp) * kCandLists + hf;
            CandEntry* ent = EPI != DEN_LOGITC ? nullptr : p.entries + list * (int64_t)kCandCap;

            for (int nt = nt0; nt < nt1; ++nt, ++it) {
                const uint32_t buf = it & 1;
                asm volatile("bar.sync 1, %0;" ::"n"(C::EPI_WARPS * 32) : "memory");           // everyone is done with the previous tile's bias
                if (et < C::BN) s_bias[et] = b_cur;
                asm volatile("bar.sync 1, %0;" ::"n"(C::EPI_WARPS * 32) : "memory");
                {
                    const int nx = item + (int)gridDim.x;                  // first tile of this CTA's next work item
                    b_cur = (nt + 1 < nt1) ? bias_val(u, nt + 1)
                          : (C::SPLIT ? bias_val(nx / p.n_tiles, nx % p.n_tiles) : bias_val(nx / p.n_split, (nx % p.n_split) * p.tiles_per_split));
                }
                float cprev[kLstmHalfUnits];
                if (C::LSTM) {
#pragma unroll
                    for (int sb = 0; sb < kLstmSubBatches; ++sb) {
                        const int j0 = nt * kLstmTileUnits + hf * kLstmHalfUnits + sb * 8;
                        if (row_ok && j0 < p.N) {
                            const float4* cp4 = reinterpret_cast<const float4*>(p.c_state + grow * p.N + j0);
                            *reinterpret_cast<float4*>(&cprev[sb * 8]) = cp4[0];
                            *reinterpret_cast<float4*>(&cprev[sb * 8 + 4]) = cp4[1];
                        } else {
#pragma unroll
                            for (int uu = 0; uu < 8; ++uu) cprev[sb * 8 + uu] = 0.f;
                        }
                    }
                }
                tc::mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
                tc::tc_fence_after();
                const uint32_t trow = tmem_base + buf * C::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
                {
                if (C::LSTM) {
                    float hs = 0.f;
#pragma unroll
                    for (int sb = 0; sb < kLstmSubBatches; ++sb) {
                        const int cu = hf * kLstmHalfUnits + sb * 8;                 // first unit of this sub-batch inside the tile
                        const int j0 = nt * kLstmTileUnits + cu;        // hidden unit index
                        float pre[5][8];
#pragma unroll
                        for (int c = 0; c < 5; ++c) tc::tmem_ld8(trow + c * kLstmTileUnits + cu, pre[c]);
                        tc::tmem_ld_wait();
                        if (row_ok && j0 < p.N) {
                            float cn[8], hn[8];
#pragma unroll
                            for (int uu = 0; uu < 8; ++uu) {
                                float g[5];
#pragma unroll
                                for (int c = 0; c < 5; ++c) g[c] = pre[c][uu] * sW + s_bias[c * kLstmTileUnits + cu + uu];
                                const float ig = sigmoid_f(g[0]), fg = sigmoid_f(g[1]), og = sigmoid_f(g[2]);
                                const float cand = fmaxf(g[3], g[4]);
                                cn[uu] = __fadd_rn(__fmul_rn(fg, cprev[sb * 8 + uu]), __fmul_rn(ig, cand));
                                hn[uu] = og * tanhf(cn[uu]);
                            }
                            float* cp = p.c_state + grow * p.N + j0;
                            *reinterpret_cast<float4*>(cp) = *reinterpret_cast<float4*>(&cn[0]);
                            *reinterpret_cast<float4*>(cp + 4) = *reinterpret_cast<float4*>(&cn[4]);
                            __half hh[8], hl[8];
#pragma unroll
                            for (int uu = 0; uu < 8; ++uu) { split_half_dev(hn[uu], kActScale, hh[uu], hl[uu]); hs = fmaf(hn[uu], hn[uu], hs); }
                            *reinterpret_cast<uint4*>(p.out_hi + grow * p.N + j0) = *reinterpret_cast<uint4*>(hh);
                            *reinterpret_cast<uint4*>(p.out_lo + grow * p.N + j0) = *reinterpret_cast<uint4*>(hl);
                            if (p.h_trace) {
                                float* ht = p.h_trace + (member * p.rows + m_local) * (int64_t)p.N + j0;
#pragma unroll
                                for (int uu = 0; uu < 8; ++uu) ht[uu] = hn[uu];
                            }
                        }
                    }
                    if (p.hsq && row_ok) p.hsq[grow * p.hsq_n + nt * 2 + hf] = hs;
                } else if (EPI == DEN_PLAIN) {
                    // PLAIN: y = acc * 2^-13 + bias, written as the next split-fp16 A operand; this warp owns 128 of the 256 columns
                    const int n_base = nt * C::BN;
#pragma unroll 1
                    for (int c0 = hf * 128; c0 < hf * 128 + 128; c0 += 32) {
                        float y[32];
                        tc::tmem_ld32(trow + c0, y);
                        tc::tmem_ld_wait();
                        if (row_ok && n_base + c0 < p.N) {
#pragma unroll
                            for (int uu = 0; uu < 32; ++uu) y[uu] = fmaf(y[uu], sW, s_bias[c0 + uu]);
                            if (p.out_f32) {
#pragma unroll
                                for (int v4 = 0; v4 < 8; ++v4)
                                    if (n_base + c0 + v4 * 4 < p.N)
                                        *reinterpret_cast<float4*>(p.out_f32 + grow * p.ld_out + n_base + c0 + v4 * 4) = *reinterpret_cast<float4*>(&y[v4 * 4]);
                            }
                            __half hh[32], hl[32];
#pragma unroll
                            for (int uu = 0; uu < 32; ++uu) split_half_dev(y[uu], kActScale, hh[uu], hl[uu]);
#pragma unroll
                            for (int v4 = 0; v4 < 4; ++v4) {
                                if (p.out_hi && n_base + c0 + v4 * 8 < p.N) {
                                    *reinterpret_cast<uint4*>(p.out_hi + grow * p.N + n_base + c0 + v4 * 8) = *reinterpret_cast<uint4*>(&hh[v4 * 8]);
                                    *reinterpret_cast<uint4*>(p.out_lo + grow * p.N + n_base + c0 + v4 * 8) = *reinterpret_cast<uint4*>(&hl[v4 * 8]);
                                }
                            }
                        }
                    }
                } else {
                    // LOGITC: this warp owns 64 (128 with 8 epilogue warps) of the tile's 256 vocabulary columns, 32 at a time
                    const int n_base = nt * C::BN;
                    constexpr int kCols = 256 / kCandLists;
#pragma unroll 1
                    for (int c0 = hf * kCols; c0 < hf * kCols + kCols; c0 += 32) {
                        float y[32];
                        tc::tmem_ld32(trow + c0, y);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int uu = 0; uu < 32; ++uu) y[uu] = fmaf(y[uu], sW, s_bias[c0 + uu]);
                        if (n_base + c0 + 32 > p.N) {                            // ragged last tile: mask the columns past V+1
#pragma unroll
                            for (int uu = 0; uu < 32; ++uu) if (n_base + c0 + uu >= p.N) y[uu] = -INFINITY;
                        }
                        float cm = y[0];
#pragma unroll
                        for (int uu = 1; uu < 32; ++uu) cm = fmaxf(cm, y[uu]);
                        if (cm >= run_max - margin && cm > -INFINITY && row_ok) {
                            // some column of this chunk is within the error margin of the running max: append it.  The list
                            // is a superset of the final candidates (the threshold only rises); k_refine filters it.
                            const float prev_max = run_max;
                            run_max = fmaxf(run_max, cm);
                            const float thr = run_max - margin;
#pragma unroll
                            for (int uu = 0; uu < 32; ++uu) {
                                if (y[uu] >= thr) {
                                    if (n_cand < kCandCap) { CandEntry e; e.v = y[uu]; e.idx = n_base + c0 + uu; ent[n_cand] = e; }
                                    ++n_cand;
                                }
                            }
                            if (run_max > prev_max) run_sum *= tc::ex2_approx((prev_max - run_max) * kLog2e);
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
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&tempty_bar[buf]);
            }
            if (EPI == DEN_LOGITC && row_ok) {
                CandHead hd; hd.mx = run_max; hd.sum = run_sum; hd.n = n_cand; hd.pad = 0;
                *reinterpret_cast<float4*>(&p.heads[list]) = *reinterpret_cast<float4*>(&hd);
            }
        }
        tc::tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) { tc::tc_fence_after(); tc::tmem_dealloc(tmem_base, C::TMEM_COLS); }
}

// ---------------------------------------------------------------------------------------------
// k_materialise: member weights W = fl32(theta_p +- fl32(sigma * eps)) of one weight matrix, for every member of the
// wave, as fp16 operands scaled by 2^8 (hi, and lo = the rounding remainder when `lo` is given).
//   mode 0: output row j = weight row j                      (logit.weight [V+1, R])
//   mode 1: output row tile*240 + gate*48 + u = weight row gate*R + tile*48 + u, zero rows past R   (i2h / h2h [5R, K])
struct MatParams {
    const int64_t* off; const float* sign; const int32_t* parent;
    const float* theta32; const __half* table; int64_t D;
    const float* scale; int scale_div;      // nullable: safe-mutation scale [n_parents, D] (scale_div: divide by it)
    int64_t w_start, w_pstart;
    int32_t K, out_rows, mode, R, src_rows;
    float sigma;
    __half* hi; __half* lo;
    int64_t n_members;
};
constexpr int kMatRowsPerBlock = 64;
constexpr int kMatBatch = 4;                 // rows in flight per thread: all loads of a batch are issued before the first store
                                             // (8 rows at 2 CTAs per SM: 7.8 -> 8.8 ms per wave, 4 rows at 2 CTAs: 8.1 ms; same-box A/B)
__device__ __forceinline__ void st_stream_128(void* p, const uint4& v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_nc_128(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// grid = (ceil(out_rows / 64), ceil(members / kMatGroup)): one block writes 64 output rows of kMatGroup consecutive members; a row is
// covered by K/8 threads (16 bytes of eps each), so all index arithmetic is 32-bit and per row, not per element.  The rollout hands
// the members over grouped by parent (ESPopulation.sample sorts them; NIC-NES has one parent), so a thread keeps its kMatBatch rows of
// the parent's float32 theta (and safe-mutation scale) in registers across the group and only the eps slices and the operand tiles
// stream: theta would otherwise cross the L2 -> SM fabric once per member (4 of every 10 bytes this kernel moved).  All loads of a
// batch are issued before the first store; the operand tiles leave through streaming stores (they are next read a whole rollout phase later)
constexpr int kMatGroup = 8;
template <bool SCALED, bool HAS_LO>
__global__ void __launch_bounds__(256, SCALED ? 2 : 3) k_materialise(const MatParams p) {
    const int kv = p.K / 8;
    const int tpr = kv < 256 ? kv : 256;                 // threads per row
    const int rpi = 256 / tpr;                           // rows per iteration
    const int lr = threadIdx.x / tpr, kvi = threadIdx.x - lr * tpr;
    if (lr >= rpi) return;
    const int r_end = min(p.out_rows, ((int)blockIdx.x + 1) * kMatRowsPerBlock);
    for (int64_t g0 = (int64_t)blockIdx.y * kMatGroup; g0 < p.n_members; g0 += (int64_t)gridDim.y * kMatGroup) {
        const int64_t g1 = g0 + kMatGroup < p.n_members ? g0 + kMatGroup : p.n_members;
        for (int k = kvi * 8; k < p.K; k += tpr * 8) {
            for (int r0 = blockIdx.x * kMatRowsPerBlock + lr; r0 < r_end; r0 += rpi * kMatBatch) {
                int src[kMatBatch];
#pragma unroll
                for (int b = 0; b < kMatBatch; ++b) {
                    const int r = r0 + b * rpi;
                    src[b] = -2;                                     // past the block: nothing to write
                    if (r < r_end) {
                        src[b] = r;
                        if (p.mode == 1) {
                            const int tile = r / kLstmTileCols, rr = r - tile * kLstmTileCols;
                            const int c = rr / kLstmTileUnits, u = rr - c * kLstmTileUnits;
                            const int unit = tile * kLstmTileUnits + u;
                            src[b] = unit < p.R ? c * p.R + unit : -1;
                        } else if (r >= p.src_rows) src[b] = -1;
                    }
                }
                float tv[kMatBatch][8];
                float sv[SCALED ? kMatBatch : 1][8];
                int cur_par = -1;
                for (int64_t member = g0; member < g1; ++member) {
                    // W * 2^8 = fma(theta, 2^8, +-fl(2^8 sigma * eps)): scaling by a power of two commutes with every rounding, so this
                    // is 2^8 * fl(theta +- fl(sigma * eps)) -- the reference's member weight -- with one FFMA instead of FADD + FMUL
                    const float ssig = __fmul_rn(p.sign[member] > 0.f ? p.sigma : -p.sigma, kThetaScale);
                    const int par = p.parent ? p.parent[member] : 0;
                    const __half* __restrict__ ebase = p.table + p.off[member] + p.w_pstart;
                    uint4 ev[kMatBatch];
#pragma unroll
                    for (int b = 0; b < kMatBatch; ++b)
                        if (src[b] >= 0) ev[b] = ld_nc_128(ebase + (int64_t)src[b] * p.K + k);
                    if (par != cur_par) {
                        cur_par = par;
                        const int64_t pbase = (int64_t)par * p.D + p.w_start;
                        const float* __restrict__ tbase = p.theta32 + pbase;
                        const bool aligned = (reinterpret_cast<uintptr_t>(tbase) & 15) == 0;      // K is a multiple of 64, so every row start is too
#pragma unroll
                        for (int b = 0; b < kMatBatch; ++b) {
                            if (src[b] < 0) continue;
                            const float* tw = tbase + (int64_t)src[b] * p.K + k;
                            if (aligned) {
                                *reinterpret_cast<float4*>(&tv[b][0]) = __ldg(reinterpret_cast<const float4*>(tw));
                                *reinterpret_cast<float4*>(&tv[b][4]) = __ldg(reinterpret_cast<const float4*>(tw) + 1);
                            } else {
#pragma unroll
                                for (int e = 0; e < 8; ++e) tv[b][e] = __ldg(tw + e);
                            }
                            if (SCALED) {
                                const float* sc = p.scale + pbase + (int64_t)src[b] * p.K + k;
#pragma unroll
                                for (int e = 0; e < 8; ++e) sv[b][e] = __ldg(sc + e);
                            }
                        }
                    }
#pragma unroll
                    for (int b = 0; b < kMatBatch; ++b) {
                        if (src[b] == -2) continue;
                        __align__(16) __half2 hh[4];
                        __align__(16) __half2 hl[4];
                        if (src[b] >= 0) {
                            const __half2* eh = reinterpret_cast<const __half2*>(&ev[b]);
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 ef = __half22float2(eh[e]);
                                float n0 = __fmul_rn(ssig, ef.x), n1 = __fmul_rn(ssig, ef.y);
                                if (SCALED) { n0 = apply_scale(n0, sv[b][2 * e], p.scale_div); n1 = apply_scale(n1, sv[b][2 * e + 1], p.scale_div); }
                                const float x0 = __fmaf_rn(tv[b][2 * e], kThetaScale, n0), x1 = __fmaf_rn(tv[b][2 * e + 1], kThetaScale, n1);
                                hh[e] = __floats2half2_rn(x0, x1);
                                if (HAS_LO) {
                                    const float2 hf = __half22float2(hh[e]);
                                    hl[e] = __floats2half2_rn(__fsub_rn(x0, hf.x), __fsub_rn(x1, hf.y));
                                }
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e) { hh[e] = __floats2half2_rn(0.f, 0.f); hl[e] = hh[e]; }
                        }
                        const int64_t o = (member * p.out_rows + (r0 + b * rpi)) * (int64_t)p.K + k;
                        st_stream_128(p.hi + o, *reinterpret_cast<uint4*>(hh));
                        if (HAS_LO) st_stream_128(p.lo + o, *reinterpret_cast<uint4*>(hl));
                    }
                }
            }
        }
    }
}
// member biases in accumulator-column order: LSTM (mode 1) = (b_i2h +- sigma eps) + (b_h2h +- sigma eps); logit (mode 0)
struct MatBiasParams {
    const int64_t* off; const float* sign; const int32_t* parent;
    const float* theta32; const __half* table; int64_t D; const float* scale; int scale_div;
    int64_t b_start[2], b_pstart[2]; int32_t nb;
    int32_t out_cols, mode, R, src_n;
    float sigma; float* out; int64_t n_members;
    int64_t out_stride, out_off;            // out[member * out_stride + out_off + col] (out_stride 0: = out_cols)
};
__global__ void __launch_bounds__(256) k_materialise_bias(const MatBiasParams p) {
    const int64_t total = p.n_members * p.out_cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t member = i / p.out_cols;
        const int col = (int)(i - member * p.out_cols);
        int n = col;
        if (p.mode == 1) {
            const int tile = col / kLstmTileCols, c = (col % kLstmTileCols) / kLstmTileUnits, u = col % kLstmTileUnits;
            const int unit = tile * kLstmTileUnits + u;
            n = unit < p.R ? c * p.R + unit : -1;
        } else if (col >= p.src_n) n = -1;
        float b = 0.f;
        if (n >= 0) {
            const float sgn = p.sign[member];
            const int64_t par = p.parent ? p.parent[member] : 0;
            const int64_t off = p.off[member];
            for (int s = 0; s < p.nb; ++s) {
                const float tb = p.theta32[par * p.D + p.b_start[s] + n];
                const float nz = mut_noise(p.sigma, __half2float(p.table[off + p.b_pstart[s] + n]), p.scale, par * p.D + p.b_start[s] + n, p.scale_div);
                b += (sgn > 0.f) ? __fadd_rn(tb, nz) : __fsub_rn(tb, nz);
            }
        }
        p.out[member * (p.out_stride ? p.out_stride : (int64_t)p.out_cols) + p.out_off + col] = b;
    }
}

}  // namespace nes
