// LSTM step of the population on materialised member weights, as thread-block clusters of two CTAs that SHARE the activation block.
//
// k_member_dense<DEN_LSTM> (dense.cuh) gives each (member, 48-unit weight tile) work item to one CTA, which streams the item's
// 240 x K weight tile (hi + lo) AND re-fetches the member's whole 128 x K activation block (hi + lo, 512 KB at K = E + R = 1024) for
// it: 11 tiles per member => the activation block crosses the L2 -> SM fabric 11 times, 5.8 MB next to 10.5 MB of weights per member
// and step (ncu round 1: l1tex__m_xbar2l1tex_read_bytes 16.85 GB per launch against 11.7 GB of DRAM reads).  At the power-capped clock the
// L2 slices deliver about as many bytes per second as HBM does (6300 B / cycle, B300_MICROARCH.md), so those re-reads cost like DRAM
// traffic.  Here a cluster takes two consecutive work items; when both belong to the same member (10 of every 11 pairs at R = 512)
// CTA 0 fetches the hi half of every activation k-block and CTA 1 the lo half, each with a TMA MULTICAST into both CTAs' shared
// memory, so the block crosses the fabric once per pair instead of once per tile.  Weight tiles, MMAs (3 split-fp16 passes into
// one TMEM accumulator, double buffered) and the fused maxout-LSTM epilogue are per CTA and unchanged.
//
// Protocol per pipeline stage (same ring in both CTAs, both walk the same k-blocks in lock step):
//   producer   waits on ITS empty barrier, which counts the tcgen05.commit of BOTH CTAs' MMA threads (each commit is multicast to the
//              pair): nobody still reads the stage it is about to overwrite -- here or in the peer; arms its full barrier with the
//              bytes it will RECEIVE (own weight tiles + both activation halves, whoever sends them); issues its loads
//   MMA thread waits on its full barrier, issues the stage's MMAs, commits to the empty barrier of both CTAs
// A mixed pair (last tile of one member + first tile of the next) and the odd item at the very end run the same protocol with private,
// non-multicast activation loads (or none).  The cluster barrier at the end keeps either CTA's shared memory alive until its peer is done.
//
// Reference math: captioning/nets.py:98-134 (LSTMCore.forward).
#pragma once
#include "dense.cuh"

namespace nes {

namespace tc {
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_id_x() { uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t ncluster_id_x() { uint32_t r; asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 2-D tile load delivered to the same shared-memory offset (and signalled on the same mbarrier offset) in every CTA of cta_mask
__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int32_t c0, int32_t c1, uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%4, %5}], [%2], %3;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "h"(cta_mask), "r"(c0), "r"(c1)
        : "memory");
}
// arrive on the mbarrier at this offset in every CTA of cta_mask once all MMAs issued so far by this thread have completed
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(cta_mask) : "memory");
}
}  // namespace tc

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(352, 1) k_lstm_pair(const __grid_constant__ DenseParams p) {
    using C = DenseCfg<DEN_LSTM>;
    extern __shared__ uint8_t smem_raw[];
    // the ring must sit at the SAME shared-memory offset in both CTAs (multicast writes by offset): the dynamic segment starts at the
    // same address in every CTA of a kernel, so the same round-up applies
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* ring = smem;
    uint8_t* ctrl = ring + C::STAGES * C::STAGE_BYTES;
    uint64_t* full_bar   = reinterpret_cast<uint64_t*>(ctrl);
    uint64_t* empty_bar  = full_bar + C::STAGES;
    uint64_t* tfull_bar  = empty_bar + C::STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot  = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* s_bias = reinterpret_cast<float*>(ctrl + 512);             // [BN]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int cid = (int)tc::cluster_id_x(), ncl = (int)tc::ncluster_id_x();
    if (threadIdx.x == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 2); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&tfull_bar[b], 1); tc::mbar_init(&tempty_bar[b], 8); }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::cluster_sync_all();                                           // the peer's barriers exist before anything is sent to them
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_items = p.n_units * p.n_tiles;
    const int n_pairs = (n_items + 1) / 2;
    const int nkb_total = (p.K[0] + (p.nseg > 1 ? p.K[1] : 0)) / C::BK;
    (void)nkb_total;

    if (warp == 0 || warp == 10) {
        // =========================== TMA producers (two warps fill alternate stages) ===========================
        const int pid = warp == 0 ? 0 : 1;
        if (lane == 0) {
            for (int s = 0; s < p.nseg; ++s) {
                tc::prefetch_tmap(&p.a_hi[s]); tc::prefetch_tmap(&p.b_hi[s]); tc::prefetch_tmap(&p.a_lo[s]); tc::prefetch_tmap(&p.b_lo[s]);
            }
            int stage = 0; uint32_t phase = 0; uint32_t q = 0;
            for (int pq = cid; pq < n_pairs; pq += ncl) {
                const int item = 2 * pq + (int)rank;
                const bool valid = item < n_items;
                const bool shared = (2 * pq + 1 < n_items) && ((2 * pq) / p.n_tiles == (2 * pq + 1) / p.n_tiles);
                const int u = valid ? item / p.n_tiles : 0;
                const int nt = valid ? item - u * p.n_tiles : 0;
                const int64_t member = u / p.m_tiles;
                const int mt = u % p.m_tiles;
                const int32_t a_row = (int32_t)((p.a_shared ? 0 : member * p.rows_pad) + mt * C::BM);
                const int32_t b_row = (int32_t)(member * p.b_rows) + nt * C::BN;
                for (int s = 0; s < p.nseg; ++s) {
                    const int nkb = p.K[s] / C::BK;
                    for (int kb = 0; kb < nkb; ++kb) {
                        if (((q++) & 1) != (uint32_t)pid) { if (++stage == C::STAGES) { stage = 0; phase ^= 1; } continue; }
                        tc::mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* st = ring + stage * C::STAGE_BYTES;
                        tc::mbar_expect_tx(&full_bar[stage], valid ? C::STAGE_BYTES : 0);
                        if (shared) {
                            if (rank == 0) tc::tma_load_2d_mc(st, &p.a_hi[s], &full_bar[stage], kb * C::BK, a_row, (uint16_t)3);
                            else           tc::tma_load_2d_mc(st + C::A_BYTES, &p.a_lo[s], &full_bar[stage], kb * C::BK, a_row, (uint16_t)3);
                        } else if (valid) {
                            tc::tma_load_2d(st, &p.a_hi[s], &full_bar[stage], kb * C::BK, a_row);
                            tc::tma_load_2d(st + C::A_BYTES, &p.a_lo[s], &full_bar[stage], kb * C::BK, a_row);
                        }
                        if (valid) {
                            uint8_t* sb = st + 2 * C::A_BYTES;
                            tc::tma_load_2d(sb, &p.b_hi[s], &full_bar[stage], kb * C::BK, b_row);
                            tc::tma_load_2d(sb + C::B_BYTES, &p.b_lo[s], &full_bar[stage], kb * C::BK, b_row);
                        }
                        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // =========================== MMA issuer (one thread) ===========================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0; uint32_t it = 0;
            constexpr uint32_t idesc = tc::make_idesc_f16(128, C::BN);
            for (int pq = cid; pq < n_pairs; pq += ncl) {
                const bool valid = 2 * pq + (int)rank < n_items;
                const uint32_t buf = it & 1;
                if (valid) {
                    tc::mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);
                    tc::tc_fence_after();
                }
                const uint32_t d_tmem = tmem_base + buf * C::ACC_STRIDE;
                bool first = true;
                for (int s = 0; s < p.nseg; ++s) {
                    const int nkb = p.K[s] / C::BK;
                    for (int kb = 0; kb < nkb; ++kb) {
                        tc::mbar_wait(&full_bar[stage], phase);
                        tc::tc_fence_after();
                        if (valid) {
                            const uint32_t st = tc::smem_u32(ring + stage * C::STAGE_BYTES);
#pragma unroll
                            for (int k = 0; k < C::BK / 16; ++k) {
                                const uint64_t a_hi = tc::make_smem_desc<C::SWZ>(st + k * 32);
                                const uint64_t a_lo = tc::make_smem_desc<C::SWZ>(st + C::A_BYTES + k * 32);
                                const uint64_t b_hi = tc::make_smem_desc<C::SWZ>(st + 2 * C::A_BYTES + k * 32);
                                const uint64_t b_lo = tc::make_smem_desc<C::SWZ>(st + 2 * C::A_BYTES + C::B_BYTES + k * 32);
                                tc::umma_f16(d_tmem, a_hi, b_hi, idesc, (first && k == 0) ? 0u : 1u);
                                tc::umma_f16(d_tmem, a_lo, b_hi, idesc, 1u);
                                tc::umma_f16(d_tmem, a_hi, b_lo, idesc, 1u);
                            }
                            first = false;
                        }
                        tc::umma_commit_mc(&empty_bar[stage], (uint16_t)3);       // releases the stage here AND in the peer
                        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                    }
                }
                if (valid) { tc::umma_commit(&tfull_bar[buf]); ++it; }
            }
        }
    } else {
        // ======= epilogue: warps 2..9.  TMEM lane quarter = warp & 3 (rows), column half = (warp - 2) >> 2 =======
        const int q = warp & 3;
        const int hf = (warp - 2) >> 2;
        const int et = (warp - 2) * 32 + lane;               // 0..255 among epilogue threads
        const int row = q * 32 + lane;
        const float sW = 1.0f / (kThetaScale * kActScale);
        auto bias_val = [&](int item) -> float {              // member bias of the item's tile, accumulator column et; fetched one item ahead
            if (item >= n_items || et >= C::BN) return 0.f;
            const int u = item / p.n_tiles, nt = item - u * p.n_tiles;
            const int64_t member = u / p.m_tiles;
            const int col = nt * C::BN + et;
            return col < p.bias_stride ? p.bias[member * p.bias_stride + col] : 0.f;
        };
        float b_cur = bias_val(2 * cid + (int)rank);
        uint32_t it = 0;
        for (int pq = cid; pq < n_pairs; pq += ncl) {
            const int item = 2 * pq + (int)rank;
            if (item >= n_items) break;                       // only the very last pair can lack its second item
            const int u = item / p.n_tiles, nt = item - u * p.n_tiles;
            const int64_t member = u / p.m_tiles;
            const int mt = u % p.m_tiles;
            const int m_local = mt * C::BM + row;
            const bool row_ok = m_local < p.rows;
            const int64_t grow = member * p.rows_pad + m_local;
            const uint32_t buf = it & 1;
            asm volatile("bar.sync 1, 256;" ::: "memory");           // everyone is done with the previous tile's bias
            if (et < C::BN) s_bias[et] = b_cur;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            b_cur = bias_val(2 * (pq + ncl) + (int)rank);
            float cprev[kLstmHalfUnits];
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
            tc::mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
            tc::tc_fence_after();
            const uint32_t trow = tmem_base + buf * C::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
            float hs = 0.f;
#pragma unroll
            for (int sb = 0; sb < kLstmSubBatches; ++sb) {
                const int cu = hf * kLstmHalfUnits + sb * 8;                 // first unit of this sub-batch inside the tile
                const int j0 = nt * kLstmTileUnits + cu;                     // hidden unit index
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
            if (p.hsq && row_ok) p.hsq[grow * p.hsq_n + nt * 2 + hf] = hs;      // this (tile, half)'s piece of the row's |h|^2 (HYBRID margin)
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&tempty_bar[buf]);
            ++it;
        }
        tc::tc_fence_before();
    }
    __syncthreads();
    tc::cluster_sync_all();                                           // the peer may still be sending into / arriving on this CTA's shared memory
    if (warp == 1) { tc::tc_fence_after(); tc::tmem_dealloc(tmem_base, C::TMEM_COLS); }
}

}  // namespace nes
