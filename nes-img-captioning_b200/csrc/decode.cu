// Greedy decode of a wave of perturbed members: FCModel._sample(fc, greedy=True)
// (captioning/nets.py:183-245) for every member of the population at once.
//   t = 0      x0 = img_embed(fc)                      PLAIN member GEMM (K = Fd)
//              (h, c) = core(x0, 0)                    LSTM member GEMM  (K = E + R)
//   t = 1..T   x_t = embed(it_{t-1}) (it_0 = <bos> = 0)  k_finalize gather
//              (h, c) = core(x_t, (h, c))              LSTM member GEMM
//              logits = logit(h) -> argmax, log-prob    LOGIT member GEMM + k_finalize
// Finished-mask, early break and the seq / seq_logprobs writes follow :233-243.
#include "hybrid.cuh"
#include "lstm_pair.cuh"
#include <cstdlib>
#include <map>
#include <tuple>

namespace nes {

// ---------------------------------------------------------------------------------------------
// TMA descriptor creation through the driver entry point (no link-time libcuda dependency, so the
// library also loads on a CPU-only box)
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    }
    return fn;
}
// fp16 matrix viewed as [rows, K] row-major, box = bk columns x box_rows rows, swizzle span = one box row (128 B / 64 B).
// n_chunks > 1: 3-D view [n_chunks, rows - (n_chunks-1)*chunk_stride, K] whose outer dimension steps chunk_stride rows
// (the 5 gate blocks of an LSTM weight matrix), so one request gathers the box_rows-row slab of every gate.
static int make_map(CUtensorMap* m, const void* base, int64_t rows, int64_t K, int box_rows, int bk, int n_chunks = 1, int64_t chunk_stride = 0) {
    PFN_encodeTiled enc = get_encode();
    NES_CHECK(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
    NES_CHECK(K % 64 == 0, "inner dimension must be a multiple of 64");
    NES_CHECK(bk == 64 || bk == 32, "k-block must be 32 or 64 elements");
    const CUtensorMapSwizzle swz = bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    CUresult r;
    if (n_chunks <= 1) {
        NES_CHECK(rows >= 1, "empty tensor map");
        cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
        cuuint64_t gstr[1] = {(cuuint64_t)K * 2};
        cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)box_rows};
        cuuint32_t estr[2] = {1u, 1u};
        r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        const int64_t rows1 = rows - (int64_t)(n_chunks - 1) * chunk_stride;
        NES_CHECK(rows1 >= 1 && chunk_stride >= 1, "empty chunked tensor map");
        cuuint64_t gdim[3] = {(cuuint64_t)K, (cuuint64_t)rows1, (cuuint64_t)n_chunks};
        cuuint64_t gstr[2] = {(cuuint64_t)K * 2, (cuuint64_t)chunk_stride * K * 2};
        cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)box_rows, (cuuint32_t)n_chunks};
        cuuint32_t estr[3] = {1u, 1u, 1u};
        r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    NES_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed");
    return 0;
}

struct DecodeWorkspace {
    int W = 0, rows_cap = 0, rows_pad = 0;
    int E = 0, R = 0, Fd = 0, V1 = 0, T = 0, nt_logit = 0;
    // activations (fp16 split operands, scaled by kActScale)
    float* fc32 = nullptr; __half *fc_hi = nullptr, *fc_lo = nullptr; int64_t fc_members = 0;   // capacity in members (1 = shared only)
    __half *x_hi = nullptr, *x_lo = nullptr;
    __half *h_hi[2] = {nullptr, nullptr}, *h_lo[2] = {nullptr, nullptr};
    float* c_state = nullptr;
    float *pre_i = nullptr, *pre_h = nullptr;               // LayerNorm path
    float* y0 = nullptr; int64_t* pp_off = nullptr; float* pp_sign = nullptr; int32_t* pp_par = nullptr;   // img_embed theta term per parent slot (shared batch)
    float* x32 = nullptr; float* ebn_par = nullptr;         // vbn_e: the embedding before its batch norm [W * rows_pad, E]; member BN parameters [W][4E]
    LogitPartial* partials = nullptr;
    CandHead* heads = nullptr; CandEntry* entries = nullptr;                                       // HYBRID candidate lists
    // materialised member weights of the wave (dense.cuh): LSTM split fp16 per segment, vocabulary projection single fp16
    bool dense_lstm = false, dense_logit = false; int ntl = 0, ntv = 0;
    __half *wl_hi[2] = {nullptr, nullptr}, *wl_lo[2] = {nullptr, nullptr}; float* wl_bias = nullptr;
    __half* wv16 = nullptr; float* wv_bias = nullptr;
    __half *wi_hi = nullptr, *wi_lo = nullptr; float* wi_bias = nullptr;      // img_embed, only with a safe-mutation scale (lazy)
    bool dense_ln = false;                                                    // LayerNorm cell: dense PLAIN pre-activation GEMMs
    __half *wn_hi[2] = {nullptr, nullptr}, *wn_lo[2] = {nullptr, nullptr}; float* wn_bias[2] = {nullptr, nullptr}; float* ln_par = nullptr;
    float* hsq = nullptr; int hsq_n = 0;                     // |h|^2 pieces per row from the cell kernels (HYBRID margin without a pass over h)
    float* margin = nullptr; int32_t* pre_tok = nullptr; float* pre_lp = nullptr; uint8_t* unit_flag = nullptr; uint8_t* row_flag = nullptr;   // HYBRID
    unsigned long long* stats = nullptr;
    uint8_t* unfinished = nullptr; int32_t* unf_count = nullptr;
    // per call
    int64_t cap_members = 0;
    int64_t* d_off = nullptr; float* d_sign = nullptr; int32_t* d_parent = nullptr;
    int64_t* d_seq = nullptr; float* d_lp = nullptr; int64_t out_cap = 0;
    int64_t* d_forced = nullptr; int64_t forced_cap = 0;
    double* d_uniforms = nullptr; int64_t uniforms_cap = 0;      // sampled decode: the caller's draws [n_members, rows, T]
    float* samp_logits = nullptr;                                // sampled decode: all logits of one step of one wave [W * rows_pad, round_up(V1, 4)]
    __half* wv_lo = nullptr;                                     // sampled decode: low halves of the materialised logit.weight (split fp16)
    std::map<std::tuple<const void*, int64_t, int64_t, int, int, int, int64_t>, CUtensorMap> maps;
    bool attrs_set = false;
};

static int get_map(DecodeWorkspace* ws, CUtensorMap* out, const void* base, int64_t rows, int64_t K, int box_rows, int bk,
                   int n_chunks = 1, int64_t chunk_stride = 0) {
    auto key = std::make_tuple(base, rows, K, box_rows, bk, n_chunks, chunk_stride);
    auto it = ws->maps.find(key);
    if (it == ws->maps.end()) {
        CUtensorMap m;
        if (make_map(&m, base, rows, K, box_rows, bk, n_chunks, chunk_stride)) return 1;
        it = ws->maps.emplace(key, m).first;
    }
    *out = it->second;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// helper kernels
__global__ void __launch_bounds__(256) k_split_rows(const float* __restrict__ src, __half* __restrict__ hi, __half* __restrict__ lo,
                                                    int64_t n_blocks, int rows, int rows_pad, int K) {
    // src [n_blocks, rows, K] -> hi/lo [n_blocks, rows_pad, K] (padding rows zero)
    const int64_t total = n_blocks * rows_pad * (int64_t)K;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % K);
        const int64_t rr = i / K;
        const int r = (int)(rr % rows_pad);
        const int64_t b = rr / rows_pad;
        __half a = __float2half_rn(0.f), l = a;
        if (r < rows) split_half_dev(src[(b * rows + r) * K + k], kActScale, a, l);
        hi[i] = a; lo[i] = l;
    }
}

struct FinalizeParams {
    const LogitPartial* partials; int n_tiles;
    const int32_t* pre_tok; const float* pre_lp; const uint8_t* unit_flag; int m_tiles;   // HYBRID: refined arg-max unless the unit is flagged
    int pre_always;                                                                        // flagged units were recomputed into pre_tok too (k_exact_rows)
    int rows, rows_pad, T, E, t;                 // t in 1..T (0: <bos> gather only)
    int64_t n_members;
    const int64_t* off; const float* sign; const int32_t* parent;
    const float* theta32; const __half* table; int64_t D; const float* scale; int scale_div;
    int64_t emb_start, emb_pstart;
    float sigma;
    uint8_t* unfinished; int32_t* unf_count;     // [T+1, W]
    int W;
    int64_t* seq; float* lp;                     // wave-local [n_members, rows, T]
    const int64_t* forced;                       // nullable
    __half *x_hi, *x_lo;
    float* x32;                                  // vbn_e: the gathered embedding as float32 instead (k_embed_bn normalises it)
    int gather;                                  // write x for the next step
    const uint8_t* row_flag;                     // nullable: k_finalize handles only the rows flagged here
};
// what happens to a row once its arg-max token and log-prob are known: the finished-mask bookkeeping of captioning/nets.py:233-243, the
// seq / log-prob writes, and the gather of the next input embedding theta_emb[it] +- sigma eps_emb[it] (t = 0: the <bos> gather only)
// the two loads of finalize_row that do not depend on the token: issued before the candidate re-evaluation, not after it
struct FinalizePre { int unf_prev; int prev; float sgn; int64_t off; int par; };
__device__ __forceinline__ FinalizePre finalize_prefetch(const FinalizeParams& p, int64_t member, int64_t grow) {
    FinalizePre q; q.unf_prev = 1; q.prev = 1;
    if (p.t > 1) { q.unf_prev = (int)ldp_32(p.unf_count + (p.t - 1) * p.W + member); q.prev = p.unfinished[grow]; }
    q.sgn = __uint_as_float(ldp_32(p.sign + member));
    const uint2 o = ldp_64(p.off + member);
    q.off = (int64_t)(((uint64_t)o.y << 32) | o.x);
    q.par = p.parent ? (int)ldp_32(p.parent + member) : 0;
    return q;
}
template <bool SCALED>
__device__ __forceinline__ void finalize_row(const FinalizeParams& p, int64_t member, int r, int64_t grow, int lane, int idx, float best, const FinalizePre pre) {
    const float sgn = pre.sgn;
    int64_t tok = 0;
    if (p.t >= 1) {
        const float logp = best;                                  // = max log-prob
        const bool done = p.t > 1 && pre.unf_prev == 0;           // the reference broke out of the loop
        int64_t it = 0;
        if (!done) {
            const bool prev = p.t == 1 ? true : (pre.prev != 0);
            const bool unf = prev && idx > 0;
            it = unf ? idx : 0;
            if (lane == 0) {
                const int64_t o = (member * p.rows + r) * p.T + (p.t - 1);
                p.seq[o] = it; p.lp[o] = logp;
                p.unfinished[grow] = unf ? 1 : 0;
                if (unf) atomicAdd(&p.unf_count[p.t * p.W + member], 1);
            }
        }
        tok = it;
        if (p.forced) tok = p.forced[(member * p.rows + r) * p.T + (p.t - 1)];
    }
    if (!p.gather) return;
    const int par = pre.par;
    const float* te = p.theta32 + (int64_t)par * p.D + p.emb_start + tok * p.E;
    const __half* ee = p.table + pre.off + p.emb_pstart + tok * p.E;
    const float* se = SCALED ? p.scale + (int64_t)par * p.D + p.emb_start + tok * p.E : nullptr;
    // x * 2^5 = fma(theta, 2^5, +-fl(2^5 sigma eps)) = 2^5 fl(theta +- fl(sigma eps)): powers of two commute with the roundings
    const float ssig = __fmul_rn(sgn > 0.f ? p.sigma : -p.sigma, kActScale);
    const bool aligned = (reinterpret_cast<uintptr_t>(te) & 15) == 0;
    // ---- phase 3: the embedding row of the token, two chunks at a time with all their loads pinned ahead of the arithmetic
    for (int e0 = lane * 8; e0 < p.E; e0 += 512) {
      uint4 evs[2]; float tvs[2][8];
#pragma unroll
      for (int c2 = 0; c2 < 2; ++c2) {
          const int e = e0 + c2 * 256;
          const int es = e < p.E ? e : 0;
          evs[c2] = ldp_128(ee + es);
          if (aligned) {
              const uint4 t0 = ldp_128(te + es), t1 = ldp_128(te + es + 4);
              tvs[c2][0] = __uint_as_float(t0.x); tvs[c2][1] = __uint_as_float(t0.y); tvs[c2][2] = __uint_as_float(t0.z); tvs[c2][3] = __uint_as_float(t0.w);
              tvs[c2][4] = __uint_as_float(t1.x); tvs[c2][5] = __uint_as_float(t1.y); tvs[c2][6] = __uint_as_float(t1.z); tvs[c2][7] = __uint_as_float(t1.w);
          } else {
#pragma unroll
              for (int u = 0; u < 8; ++u) tvs[c2][u] = __uint_as_float(ldp_32(te + es + u));
          }
      }
#pragma unroll
      for (int c2 = 0; c2 < 2; ++c2) {
        const int e = e0 + c2 * 256;
        if (e >= p.E) break;
        const __half2* eh = reinterpret_cast<const __half2*>(&evs[c2]);
        const float* tv = tvs[c2];
        __align__(16) __half2 hh[4];
        __align__(16) __half2 hl[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float2 ef = __half22float2(eh[u]);
            float n0 = __fmul_rn(ssig, ef.x), n1 = __fmul_rn(ssig, ef.y);
            if (SCALED) { n0 = apply_scale(n0, se[e + 2 * u], p.scale_div); n1 = apply_scale(n1, se[e + 2 * u + 1], p.scale_div); }
            const float x0 = __fmaf_rn(tv[2 * u], kActScale, n0), x1 = __fmaf_rn(tv[2 * u + 1], kActScale, n1);
            hh[u] = __floats2half2_rn(x0, x1);
            const float2 hf = __half22float2(hh[u]);
            hl[u] = __floats2half2_rn(__fsub_rn(x0, hf.x), __fsub_rn(x1, hf.y));
        }
        if (p.x32) {                                 // x = (x * 2^5) / 2^5, exact
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                float n0 = __fmul_rn(ssig, __half22float2(eh[u]).x), n1 = __fmul_rn(ssig, __half22float2(eh[u]).y);
                if (SCALED) { n0 = apply_scale(n0, se[e + 2 * u], p.scale_div); n1 = apply_scale(n1, se[e + 2 * u + 1], p.scale_div); }
                p.x32[grow * p.E + e + 2 * u] = __fmaf_rn(tv[2 * u], kActScale, n0) * (1.0f / kActScale);
                p.x32[grow * p.E + e + 2 * u + 1] = __fmaf_rn(tv[2 * u + 1], kActScale, n1) * (1.0f / kActScale);
            }
            continue;
        }
        *reinterpret_cast<uint4*>(p.x_hi + grow * p.E + e) = *reinterpret_cast<uint4*>(hh);
        *reinterpret_cast<uint4*>(p.x_lo + grow * p.E + e) = *reinterpret_cast<uint4*>(hl);
      }
    }
}


// one warp per (member, row): merge the per-tile (max, argmax, sum-exp) partials (or take the refined arg-max), then finalize_row.
// row_flag given: only the rows flagged there (the ones k_refine_finalize left to the fallback) are handled.
__global__ void __launch_bounds__(128) k_finalize(const FinalizeParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (wid >= p.n_members * p.rows) return;
    const int64_t member = wid / p.rows;
    const int r = (int)(wid % p.rows);
    const int64_t grow = member * p.rows_pad + r;
    if (p.row_flag && !p.row_flag[grow]) return;
    float best = -INFINITY; int idx = 0x7fffffff;
    if (p.t >= 1) {
        if (p.pre_tok && (p.pre_always || !p.unit_flag[member * p.m_tiles + r / 128])) { idx = p.pre_tok[grow]; best = p.pre_lp[grow]; }
        else {
        // global max and sum-exp over the vocabulary tiles
        float mx = -INFINITY;
        for (int j = lane; j < p.n_tiles; j += 32) mx = fmaxf(mx, p.partials[grow * p.n_tiles + j].mx);
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float s = 0.f;
        for (int j = lane; j < p.n_tiles; j += 32) {
            const LogitPartial q = p.partials[grow * p.n_tiles + j];
            s += q.sum * expf(q.mx - mx);
        }
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float lse = logf(s);
        // the reference's argmax runs over float32 log-probs, log_softmax(x) = (x - max) - log(sum) (captioning/nets.py:202-208):
        // logits closer than the float32 quantum of log(sum) tie there and torch.max returns the LOWEST index.
        for (int j = lane; j < p.n_tiles; j += 32) {
            const LogitPartial q = p.partials[grow * p.n_tiles + j];
            const float l1 = __fsub_rn(__fsub_rn(q.mx, mx), lse), l2 = __fsub_rn(__fsub_rn(q.lo_val, mx), lse);
            if (l1 > best || (l1 == best && q.idx < idx)) { best = l1; idx = q.idx; }
            if (l2 > best || (l2 == best && q.lo_idx < idx)) { best = l2; idx = q.lo_idx; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
            if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
        }
        }
    }
    if (p.scale) finalize_row<true>(p, member, r, grow, lane, idx, best, finalize_prefetch(p, member, grow));
    else finalize_row<false>(p, member, r, grow, lane, idx, best, finalize_prefetch(p, member, grow));
}

// HYBRID: the fp32 re-evaluation of a row's arg-max candidates (refine_row, hybrid.cuh) and everything that follows from the token
// (finalize_row) in one pass, one warp per (member, row).  A row whose candidate lists overflowed is flagged instead (row_flag, and its
// unit for the fallback projection); k_finalize with row_flag set picks those up after the fallback.
template <bool SCALED>
__global__ void __launch_bounds__(128, 6) k_refine_finalize(const RefineParams rp, const FinalizeParams fp, uint8_t* __restrict__ row_flag) {
    __shared__ int s_idx[4][32];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int wid = (int)blockIdx.x * 4 + wq;                         // members * rows of a wave fit 31 bits (checked at launch)
    if (wid >= (int)rp.n_members * rp.rows) return;
    const int64_t member = wid / rp.rows;
    const int r = wid % rp.rows;
    const int64_t grow = member * rp.rows_pad + r;
    int idx = 0; float best = 0.f;
    const FinalizePre pre = finalize_prefetch(fp, member, grow);
    const bool ok = refine_row<SCALED>(rp, member, r, grow, lane, wq, s_idx, idx, best);
    if (lane == 0) row_flag[grow] = ok ? 0 : 1;
    if (ok) finalize_row<SCALED>(fp, member, r, grow, lane, idx, best, pre);
}

// vbn_e (captioning/nets.py:156-165): nn.BatchNorm1d(E, track_running_stats=False) on the output of img_embed / embed -- every feature
// normalised by its mean and biased variance over the ROWS of the member's batch (eval mode too), then the member's affine pair.
// grid (ceil(E / 32), members), 256 threads: lane = feature, warp = row subset; float32 in, split-fp16 GEMM operand out.
struct EmbedBnParams { const float* x32; __half *x_hi, *x_lo; const float* par; int site; int rows, rows_pad, E; };
__global__ void __launch_bounds__(256) k_embed_bn(const EmbedBnParams p) {
    __shared__ float s_red[8][32];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t member = blockIdx.y;
    const int j = blockIdx.x * 32 + lane;
    const bool live = j < p.E;
    const float* x = p.x32 + member * (int64_t)p.rows_pad * p.E;
    const float inv_n = 1.0f / (float)p.rows;
    auto reduce = [&](float v) { __syncthreads(); s_red[wq][lane] = v; __syncthreads(); float a = 0.f; for (int w = 0; w < 8; ++w) a += s_red[w][lane]; return a; };
    float acc = 0.f;
    if (live) for (int r = wq; r < p.rows; r += 8) acc += x[(int64_t)r * p.E + j];
    const float mean = reduce(acc) * inv_n;
    acc = 0.f;
    if (live) for (int r = wq; r < p.rows; r += 8) { const float d = x[(int64_t)r * p.E + j] - mean; acc = fmaf(d, d, acc); }
    const float rstd = 1.0f / sqrtf(reduce(acc) * inv_n + 1e-5f);
    const float w = (p.par && live) ? p.par[member * 4 * p.E + (2 * p.site) * p.E + j] : 1.f;
    const float b = (p.par && live) ? p.par[member * 4 * p.E + (2 * p.site + 1) * p.E + j] : 0.f;
    if (live) for (int r = wq; r < p.rows; r += 8) {
        const float y = (x[(int64_t)r * p.E + j] - mean) * rstd * w + b;
        __half hi, lo; split_half_dev(y, kActScale, hi, lo);
        const int64_t o = (member * p.rows_pad + r) * (int64_t)p.E + j;
        p.x_hi[o] = hi; p.x_lo[o] = lo;
    }
}

struct SampleParams {
    const float* logits;                         // [n_members * rows, V1] of this step
    const LogitPartial* partials; int n_tiles;   // nullable: the row's max / sum-exp are then taken from the logits themselves
    int64_t ld; int row_stride;                  // logits of (member, row) start at (member * row_stride + row) * ld
    const double* u;                             // [n_members, rows, T]
    int rows, rows_pad, T, t, V1; int64_t n_members;
    int32_t* tok; float* lp;                     // [n_members * rows_pad]
};
// Sampled decode (captioning/nets.py:210-231): one CTA per (member, row).  prob = exp(log_softmax(logits)) in float32, divided by its
// 1-norm, accumulated into a float64 cdf that is normalised by its last entry; the token is the first index whose cdf exceeds the
// caller's uniform draw (numpy's legacy choice: cdf.searchsorted(u, side='right')).
__global__ void __launch_bounds__(256) k_sample(const SampleParams p) {
    extern __shared__ float s_p[];               // V1 probabilities
    __shared__ float s_f[8];
    __shared__ double s_d[8];
    __shared__ int s_i[8];
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int64_t member = blockIdx.x / p.rows;
    const int r = (int)(blockIdx.x % p.rows);
    const int64_t grow = member * p.rows_pad + r;
    const float* lg = p.logits + (member * p.row_stride + r) * p.ld;
    float mx = -INFINITY, se = 0.f;
    if (p.partials) {
        for (int j = 0; j < p.n_tiles; ++j) mx = fmaxf(mx, p.partials[grow * p.n_tiles + j].mx);
        for (int j = 0; j < p.n_tiles; ++j) { const LogitPartial q = p.partials[grow * p.n_tiles + j]; se += q.sum * expf(q.mx - mx); }
    } else {                                     // log_softmax(x) = (x - max) - log(sum exp(x - max)) (captioning/nets.py:202)
        for (int j = tid; j < p.V1; j += 256) mx = fmaxf(mx, lg[j]);
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) s_f[wq] = mx;
        __syncthreads();
        mx = s_f[0];
        for (int w = 1; w < 8; ++w) mx = fmaxf(mx, s_f[w]);
        __syncthreads();
        for (int j = tid; j < p.V1; j += 256) se += expf(lg[j] - mx);
        for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
        if (lane == 0) s_f[wq] = se;
        __syncthreads();
        se = 0.f;
        for (int w = 0; w < 8; ++w) se += s_f[w];
        __syncthreads();
    }
    const float lse = logf(se);
    float acc = 0.f;
    for (int j = tid; j < p.V1; j += 256) { const float e = expf(__fsub_rn(__fsub_rn(lg[j], mx), lse)); s_p[j] = e; acc += e; }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) s_f[wq] = acc;
    __syncthreads();
    float norm1 = 0.f;
    for (int w = 0; w < 8; ++w) norm1 += s_f[w];
    // float64 cdf over contiguous chunks
    const int C = (p.V1 + 255) / 256, j0 = tid * C, j1 = min(p.V1, j0 + C);
    double loc = 0.0;
    for (int j = j0; j < j1; ++j) { const float nj = __fdiv_rn(s_p[j], norm1); s_p[j] = nj; loc += (double)nj; }
    double inc = loc;                            // inclusive scan over the 256 chunk sums
    for (int o = 1; o < 32; o <<= 1) { const double v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) s_d[wq] = inc;
    __syncthreads();
    double base = 0.0, total = 0.0;
    for (int w = 0; w < 8; ++w) { if (w < wq) base += s_d[w]; total += s_d[w]; }
    const double u = p.u[(member * p.rows + r) * p.T + (p.t - 1)];
    double run = base + (inc - loc);
    int first = 0x7fffffff;
    // cdf_j / cdf_last > u, with the float64 division done only where the cheap product test says it can be true
    const double pre = u * total * (1.0 - 1e-15);
    for (int j = j0; j < j1; ++j) { run += (double)s_p[j]; if (first == 0x7fffffff && run >= pre && run / total > u) first = j; }
    for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    if (lane == 0) s_i[wq] = first;
    __syncthreads();
    if (tid == 0) {
        int idx = 0x7fffffff;
        for (int w = 0; w < 8; ++w) idx = min(idx, s_i[w]);
        if (idx >= p.V1) idx = p.V1 - 1;         // u within rounding of 1: the last entry, as searchsorted would clamp through choice's own check
        p.tok[grow] = idx;
        p.lp[grow] = __fsub_rn(__fsub_rn(lg[idx], mx), lse);
    }
}

// The same sampler for 16-byte aligned logit rows of at most 1024 * NV entries (the dense projection pads ld to a multiple of 4): the row is
// read from global memory ONCE, 128 bits at a time with every load in flight together, and lives in registers through the max / sum-exp /
// probability passes (k_sample walks the row three times with a few loads in flight: 4 ms per step of 1000 members at V = 9487).
template <int NV>
__global__ void __launch_bounds__(256) k_sample_reg(const SampleParams p) {
    extern __shared__ float s_p[];               // V1 probabilities, padded to a multiple of 4
    __shared__ float s_f[8];
    __shared__ double s_d[8];
    __shared__ int s_i[8];
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int64_t member = blockIdx.x / p.rows;
    const int r = (int)(blockIdx.x % p.rows);
    const int64_t grow = member * p.rows_pad + r;
    const float* lg = p.logits + (member * p.row_stride + r) * p.ld;
    const float4* lg4 = reinterpret_cast<const float4*>(lg);
    const int n4 = (p.V1 + 3) / 4;
    float v[NV][4];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int q = tid + 256 * i;
        float4 x = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        if (q < n4) x = lg4[q];
        v[i][0] = x.x; v[i][1] = x.y; v[i][2] = x.z; v[i][3] = x.w;
    }
    {   // the padding columns of the last 128-bit word hold nothing meaningful
        const int q = (n4 - 1) & 255, i_last = (n4 - 1) >> 8;
        if (tid == q) {
#pragma unroll
            for (int i = 0; i < NV; ++i) if (i == i_last) {
#pragma unroll
                for (int e = 0; e < 4; ++e) if (4 * (n4 - 1) + e >= p.V1) v[i][e] = -INFINITY;
            }
        }
    }
    auto block_sum = [&](float x) { for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o); __syncthreads(); if (lane == 0) s_f[wq] = x; __syncthreads();
                                    float a = 0.f; for (int w = 0; w < 8; ++w) a += s_f[w]; return a; };
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < NV; ++i) mx = fmaxf(fmaxf(fmaxf(v[i][0], v[i][1]), fmaxf(v[i][2], v[i][3])), mx);
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_f[wq] = mx;
    __syncthreads();
    mx = s_f[0];
    for (int w = 1; w < 8; ++w) mx = fmaxf(mx, s_f[w]);
    float se = 0.f;                              // log_softmax(x) = (x - max) - log(sum exp(x - max)) (captioning/nets.py:202)
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
        for (int e = 0; e < 4; ++e) se += expf(v[i][e] - mx);
    }
    se = block_sum(se);
    const float lse = logf(se);
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int q = tid + 256 * i;
        float4 pr;
        pr.x = expf(__fsub_rn(__fsub_rn(v[i][0], mx), lse)); pr.y = expf(__fsub_rn(__fsub_rn(v[i][1], mx), lse));
        pr.z = expf(__fsub_rn(__fsub_rn(v[i][2], mx), lse)); pr.w = expf(__fsub_rn(__fsub_rn(v[i][3], mx), lse));
        acc += (pr.x + pr.y) + (pr.z + pr.w);
        if (q < n4) reinterpret_cast<float4*>(s_p)[q] = pr;
    }
    const float norm1 = block_sum(acc);          // its second barrier also publishes s_p
    // float64 cdf over contiguous chunks
    const int C = (p.V1 + 255) / 256, j0 = tid * C, j1 = min(p.V1, j0 + C);
    double loc = 0.0;
    for (int j = j0; j < j1; ++j) { const float nj = __fdiv_rn(s_p[j], norm1); s_p[j] = nj; loc += (double)nj; }
    double inc = loc;                            // inclusive scan over the 256 chunk sums
    for (int o = 1; o < 32; o <<= 1) { const double x = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += x; }
    if (lane == 31) s_d[wq] = inc;
    __syncthreads();
    double base = 0.0, total = 0.0;
    for (int w = 0; w < 8; ++w) { if (w < wq) base += s_d[w]; total += s_d[w]; }
    const double u = p.u[(member * p.rows + r) * p.T + (p.t - 1)];
    double run = base + (inc - loc);
    int first = 0x7fffffff;
    const double pre = u * total * (1.0 - 1e-15);
    // the crossing lies in the first chunk whose end reaches the threshold: chunks that end below it are skipped, and so are chunks that
    // START well above it (one of their predecessors holds it) -- one or two threads of the block walk their 38 entries
    if (base + inc >= pre && run <= u * total * (1.0 + 1e-12))
        for (int j = j0; j < j1; ++j) { run += (double)s_p[j]; if (first == 0x7fffffff && run >= pre && run / total > u) first = j; }
    for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    if (lane == 0) s_i[wq] = first;
    __syncthreads();
    if (tid == 0) {
        int idx = 0x7fffffff;
        for (int w = 0; w < 8; ++w) idx = min(idx, s_i[w]);
        if (idx >= p.V1) idx = p.V1 - 1;
        p.tok[grow] = idx;
        p.lp[grow] = __fsub_rn(__fsub_rn(lg[idx], mx), lse);
    }
}

struct LnParams {
    const float *pre_i, *pre_h; float* c_state; __half *h_hi, *h_lo; float* h_trace;
    int rows, rows_pad, R; int64_t n_members;
    float* hsq; int hsq_n;     // nullable: the row's |h|^2 (in hsq_n pieces; the batch-norm cell writes one per block of 32 units) for the HYBRID margin
    const float* ln_par;       // nullable (no affine): member LayerNorm parameters [n_members][4*5R + 2R]:
                               // i2h_ln.w, i2h_ln.b, h2h_ln.w, h2h_ln.b [5R each], c_ln.w, c_ln.b [R each] (k_materialise_bias)
};
__device__ __forceinline__ float warp_sum_f(float v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
// LayerNorm variant of the cell (captioning/nets.py:102-104,126-127): LN over 5R of each addend, LN over R of c'.
// One warp per (member, row); the normalised gate sums of the row live in shared memory.
__global__ void __launch_bounds__(128) k_ln_gates(const LnParams p) {
    extern __shared__ float s_all[];                     // [4 warps][5R + R]
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t wid = (int64_t)blockIdx.x * 4 + wq;
    if (wid >= p.n_members * p.rows) return;
    const int64_t member = wid / p.rows;
    const int r = (int)(wid % p.rows);
    const int64_t grow = member * p.rows_pad + r;
    const int R = p.R, N = 5 * p.R;
    float* s_sum = s_all + wq * (N + R);
    float* s_c = s_sum + N;
    const float* par = p.ln_par ? p.ln_par + member * (int64_t)(4 * N + 2 * R) : nullptr;
    const float* src[2] = {p.pre_i + grow * N, p.pre_h + grow * N};
    for (int a = 0; a < 2; ++a) {
        float s = 0.f;
        for (int j = lane; j < N; j += 32) s += src[a][j];
        const float mean = warp_sum_f(s) / (float)N;
        float v = 0.f;
        for (int j = lane; j < N; j += 32) { const float d = src[a][j] - mean; v += d * d; }
        const float rstd = 1.0f / sqrtf(warp_sum_f(v) / (float)N + 1e-5f);
        for (int j = lane; j < N; j += 32) {
            const float w = par ? par[(2 * a) * N + j] : 1.f, b = par ? par[(2 * a + 1) * N + j] : 0.f;
            const float y = (src[a][j] - mean) * rstd * w + b;
            s_sum[j] = a == 0 ? y : s_sum[j] + y;
        }
    }
    __syncwarp();
    float* cs = p.c_state + grow * R;
    float cs_sum = 0.f;
    for (int j = lane; j < R; j += 32) {
        const float ig = sigmoid_f(s_sum[j]), fg = sigmoid_f(s_sum[R + j]);
        const float cand = fmaxf(s_sum[3 * R + j], s_sum[4 * R + j]);
        const float cn = __fadd_rn(__fmul_rn(fg, cs[j]), __fmul_rn(ig, cand));
        s_c[j] = cn; cs[j] = cn; cs_sum += cn;
    }
    const float mean = warp_sum_f(cs_sum) / (float)R;
    float v = 0.f;
    for (int j = lane; j < R; j += 32) { const float d = s_c[j] - mean; v += d * d; }
    const float rstd = 1.0f / sqrtf(warp_sum_f(v) / (float)R + 1e-5f);
    float hs = 0.f;
    for (int j = lane; j < R; j += 32) {
        const float w = par ? par[4 * N + j] : 1.f, b = par ? par[4 * N + R + j] : 0.f;
        const float og = sigmoid_f(s_sum[2 * R + j]);
        const float hn = og * tanhf((s_c[j] - mean) * rstd * w + b);
        hs = fmaf(hn, hn, hs);
        __half a, l; split_half_dev(hn, kActScale, a, l);
        p.h_hi[grow * R + j] = a; p.h_lo[grow * R + j] = l;
        if (p.h_trace) p.h_trace[(member * p.rows + r) * (int64_t)R + j] = hn;
    }
    hs = warp_sum_f(hs);
    if (p.hsq && lane == 0) p.hsq[grow * p.hsq_n] = hs;
}

// The same cell with the row held in registers (R <= 256, i.e. config 4's R = 128): a lane owns NV = 5R / 32 pre-activation values of each
// addend, so each is read from global memory ONCE (the kernel above re-reads them for the mean, the variance and the normalisation and
// was bound by instruction issue: 5.5 ms per step of 2500 members x 512 rows); the gate inputs of unit j = the lane's values at
// j, j + R, ... live in the lane that owns column j mod 32, so the cell needs no shared memory either.
template <int RV>                                          // RV = R / 32 hidden units per lane
__global__ void __launch_bounds__(128) k_ln_gates_reg(const LnParams p) {
    constexpr int NV = 5 * RV;
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t wid = (int64_t)blockIdx.x * 4 + wq;
    if (wid >= p.n_members * p.rows) return;
    const int64_t member = wid / p.rows;
    const int r = (int)(wid % p.rows);
    const int64_t grow = member * p.rows_pad + r;
    constexpr int R = 32 * RV, N = 5 * R;
    const float* par = p.ln_par ? p.ln_par + member * (int64_t)(4 * N + 2 * R) : nullptr;
    float sum[NV];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        const float* src = (a == 0 ? p.pre_i : p.pre_h) + grow * N;
        float v[NV];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) { v[i] = src[i * 32 + lane]; s += v[i]; }        // column i * 32 + lane: coalesced
        const float mean = warp_sum_f(s) / (float)N;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) { const float d = v[i] - mean; q += d * d; }
        const float rstd = 1.0f / sqrtf(warp_sum_f(q) / (float)N + 1e-5f);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int j = i * 32 + lane;
            const float w = par ? par[(2 * a) * N + j] : 1.f, b = par ? par[(2 * a + 1) * N + j] : 0.f;
            const float y = (v[i] - mean) * rstd * w + b;
            sum[i] = a == 0 ? y : sum[i] + y;
        }
    }
    // unit j = u * 32 + lane (u < RV): gate g of unit j is column g * R + j = (g * RV + u) * 32 + lane -> sum[g * RV + u]
    float* cs = p.c_state + grow * R;
    float cn[RV];
    float cs_sum = 0.f;
#pragma unroll
    for (int u = 0; u < RV; ++u) {
        const float ig = sigmoid_f(sum[u]), fg = sigmoid_f(sum[RV + u]);
        const float cand = fmaxf(sum[3 * RV + u], sum[4 * RV + u]);
        cn[u] = __fadd_rn(__fmul_rn(fg, cs[u * 32 + lane]), __fmul_rn(ig, cand));
        cs[u * 32 + lane] = cn[u];
        cs_sum += cn[u];
    }
    const float mean = warp_sum_f(cs_sum) / (float)R;
    float q = 0.f;
#pragma unroll
    for (int u = 0; u < RV; ++u) { const float d = cn[u] - mean; q += d * d; }
    const float rstd = 1.0f / sqrtf(warp_sum_f(q) / (float)R + 1e-5f);
    float hs = 0.f;
#pragma unroll
    for (int u = 0; u < RV; ++u) {
        const int j = u * 32 + lane;
        const float w = par ? par[4 * N + j] : 1.f, b = par ? par[4 * N + R + j] : 0.f;
        const float og = sigmoid_f(sum[2 * RV + u]);
        const float hn = og * tanhf((cn[u] - mean) * rstd * w + b);
        hs = fmaf(hn, hn, hs);
        __half hi, lo; split_half_dev(hn, kActScale, hi, lo);
        p.h_hi[grow * R + j] = hi; p.h_lo[grow * R + j] = lo;
        if (p.h_trace) p.h_trace[(member * p.rows + r) * (int64_t)R + j] = hn;
    }
    hs = warp_sum_f(hs);
    if (p.hsq && lane == 0) p.hsq[grow * p.hsq_n] = hs;
}

// Batch-norm variant of the cell -- the reference's "virtual batch norm" (captioning/nets.py:87-91,99-101,124-125):
// nn.BatchNorm1d(track_running_stats=False), i.e. every feature is normalised with the mean / biased variance over the ROWS of the
// batch at hand, in eval mode too.  grid (ceil(R/32), members): a block owns 32 units of one member (lane = unit, warp = row subset),
// which closes over everything the unit needs: its 5 gate columns of both addends and its cell value, over all rows.
__global__ void __launch_bounds__(256) k_bn_gates(const LnParams p) {
    __shared__ float s_red[8][11][32];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t member = blockIdx.y;
    const int R = p.R, N = 5 * p.R;
    const int j = blockIdx.x * 32 + lane;
    const bool live = j < R;
    const float* par = p.ln_par ? p.ln_par + member * (int64_t)(4 * N + 2 * R) : nullptr;
    const float* pi = p.pre_i + member * (int64_t)p.rows_pad * N;
    const float* ph = p.pre_h + member * (int64_t)p.rows_pad * N;
    float* cs = p.c_state + member * (int64_t)p.rows_pad * R;
    const float inv_n = 1.0f / (float)p.rows;
    // block-wide column reduction of `v[q]` (q < nq) over the warps; result in s_mean / s_rstd via the callback
    auto reduce = [&](const float (&v)[11], int nq, float (&out)[11]) {
        __syncthreads();
        for (int q = 0; q < nq; ++q) s_red[wq][q][lane] = v[q];
        __syncthreads();
        for (int q = 0; q < nq; ++q) { float a = 0.f; for (int w = 0; w < 8; ++w) a += s_red[w][q][lane]; out[q] = a; }
    };
    float acc[11], tot[11], mean[11], rstd[11];
    // ---- statistics of the 10 pre-activation columns (gate g of addend a: q = 2 g + a)
    for (int q = 0; q < 11; ++q) acc[q] = 0.f;
    if (live) for (int r = wq; r < p.rows; r += 8)
        for (int g = 0; g < 5; ++g) { acc[2 * g] += pi[(int64_t)r * N + g * R + j]; acc[2 * g + 1] += ph[(int64_t)r * N + g * R + j]; }
    reduce(acc, 10, tot);
    for (int q = 0; q < 10; ++q) { mean[q] = tot[q] * inv_n; acc[q] = 0.f; }
    if (live) for (int r = wq; r < p.rows; r += 8)
        for (int g = 0; g < 5; ++g) {
            const float a = pi[(int64_t)r * N + g * R + j] - mean[2 * g], b = ph[(int64_t)r * N + g * R + j] - mean[2 * g + 1];
            acc[2 * g] = fmaf(a, a, acc[2 * g]); acc[2 * g + 1] = fmaf(b, b, acc[2 * g + 1]);
        }
    reduce(acc, 10, tot);
    for (int q = 0; q < 10; ++q) rstd[q] = 1.0f / sqrtf(tot[q] * inv_n + 1e-5f);
    float wgt[10], bia[10];
    for (int g = 0; g < 5; ++g)
        for (int a = 0; a < 2; ++a) {
            wgt[2 * g + a] = (par && live) ? par[(2 * a) * N + g * R + j] : 1.f;
            bia[2 * g + a] = (par && live) ? par[(2 * a + 1) * N + g * R + j] : 0.f;
        }
    auto gate = [&](int r, int g) {
        const float a = (pi[(int64_t)r * N + g * R + j] - mean[2 * g]) * rstd[2 * g] * wgt[2 * g] + bia[2 * g];
        const float b = (ph[(int64_t)r * N + g * R + j] - mean[2 * g + 1]) * rstd[2 * g + 1] * wgt[2 * g + 1] + bia[2 * g + 1];
        return a + b;
    };
    // ---- c' = f c + i max(g3, g4), its statistics over the rows
    acc[10] = 0.f;
    if (live) for (int r = wq; r < p.rows; r += 8) {
        const float ig = sigmoid_f(gate(r, 0)), fg = sigmoid_f(gate(r, 1));
        const float cand = fmaxf(gate(r, 3), gate(r, 4));
        const float cn = __fadd_rn(__fmul_rn(fg, cs[(int64_t)r * R + j]), __fmul_rn(ig, cand));
        cs[(int64_t)r * R + j] = cn;
        acc[10] += cn;
    }
    {
        float v[11], o[11]; v[0] = acc[10];
        reduce(v, 1, o); mean[10] = o[0] * inv_n;
        float d2 = 0.f;
        if (live) for (int r = wq; r < p.rows; r += 8) { const float d = cs[(int64_t)r * R + j] - mean[10]; d2 = fmaf(d, d, d2); }
        v[0] = d2; reduce(v, 1, o); rstd[10] = 1.0f / sqrtf(o[0] * inv_n + 1e-5f);
    }
    const float cw = (par && live) ? par[4 * N + j] : 1.f, cb = (par && live) ? par[4 * N + R + j] : 0.f;
    // ---- h = o tanh(bn(c'))
    for (int r = wq; r < p.rows; r += 8) {
        const int64_t grow = member * p.rows_pad + r;
        float hn = 0.f;
        if (live) {
            const float og = sigmoid_f(gate(r, 2));
            hn = og * tanhf((cs[(int64_t)r * R + j] - mean[10]) * rstd[10] * cw + cb);
            __half a, l; split_half_dev(hn, kActScale, a, l);
            p.h_hi[grow * R + j] = a; p.h_lo[grow * R + j] = l;
            if (p.h_trace) p.h_trace[(member * p.rows + r) * (int64_t)R + j] = hn;
        }
        const float hs = warp_sum_f(hn * hn);                 // this block's 32 units of the row
        if (p.hsq && lane == 0) p.hsq[grow * p.hsq_n + blockIdx.x] = hs;
    }
}

__global__ void __launch_bounds__(256) k_logprob_fitness(const float* __restrict__ lp, int64_t per_member, double* __restrict__ out) {
    __shared__ double sh[8];
    const float* src = lp + blockIdx.x * per_member;
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < per_member; i += 256) s += (double)src[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { double a = 0; for (int q = 0; q < 8; ++q) a += sh[q]; out[blockIdx.x] = a; }
}

// ---------------------------------------------------------------------------------------------
// host side
template <typename T> static int dmalloc(T** p, int64_t n) {
    NES_CUDA(cudaMalloc((void**)p, sizeof(T) * (size_t)(n > 0 ? n : 1)));
    return 0;
}

int decode_create(Ctx* c, int wave_members, int rows) {
    NES_CHECK(c->is_fc, "decode needs an fc_caption context");
    NES_CHECK(wave_members >= 1 && rows >= 1, "wave_members and rows must be positive");
    const nes_fc_dims& d = c->dims;
    NES_CHECK(d.input_encoding_size % 64 == 0 && d.rnn_size % 64 == 0 && d.fc_feat_size % 64 == 0,
              "input_encoding_size, rnn_size and fc_feat_size must be multiples of 64 (UMMA K block)");
    DecodeWorkspace* ws = new DecodeWorkspace();
    c->ws = ws;
    ws->W = wave_members; ws->rows_cap = rows; ws->rows_pad = (rows + 127) / 128 * 128;
    ws->E = d.input_encoding_size; ws->R = d.rnn_size; ws->Fd = d.fc_feat_size; ws->V1 = d.vocab_size + 1; ws->T = d.seq_length;
    ws->nt_logit = (ws->V1 + 127) / 128;
    const int64_t WR = (int64_t)ws->W * ws->rows_pad;
    if (dmalloc(&ws->x_hi, WR * ws->E) || dmalloc(&ws->x_lo, WR * ws->E)) return 1;
    for (int b = 0; b < 2; ++b) if (dmalloc(&ws->h_hi[b], WR * ws->R) || dmalloc(&ws->h_lo[b], WR * ws->R)) return 1;
    if (dmalloc(&ws->c_state, WR * ws->R)) return 1;
    if (dmalloc(&ws->partials, 2 * WR)) return 1;
    {
        const char* e = getenv("NES_FACTORED");                       // experiments / tests: keep every layer in the factored form
        const bool factored = e && atoi(e) != 0;
        ws->dense_lstm = !factored && !d.layer_n;
        ws->dense_logit = !factored && ws->R <= 512;
        ws->ntl = (ws->R + kLstmTileUnits - 1) / kLstmTileUnits;
        ws->ntv = (ws->V1 + 255) / 256;
        if (ws->dense_lstm) {
            const int64_t rows_l = (int64_t)ws->W * ws->ntl * kLstmTileCols;
            if (dmalloc(&ws->wl_hi[0], rows_l * ws->E) || dmalloc(&ws->wl_lo[0], rows_l * ws->E) ||
                dmalloc(&ws->wl_hi[1], rows_l * ws->R) || dmalloc(&ws->wl_lo[1], rows_l * ws->R) || dmalloc(&ws->wl_bias, rows_l)) return 1;
        }
        ws->dense_ln = !factored && d.layer_n;
        if (ws->dense_ln) {
            const int64_t rows_n = (int64_t)ws->W * 5 * ws->R;
            if (dmalloc(&ws->wn_hi[0], rows_n * ws->E) || dmalloc(&ws->wn_lo[0], rows_n * ws->E) || dmalloc(&ws->wn_hi[1], rows_n * ws->R) ||
                dmalloc(&ws->wn_lo[1], rows_n * ws->R) || dmalloc(&ws->wn_bias[0], rows_n) || dmalloc(&ws->wn_bias[1], rows_n)) return 1;
        }
        if (d.layer_n && d.layer_n_affine && dmalloc(&ws->ln_par, (int64_t)ws->W * 22 * ws->R)) return 1;
        if (ws->dense_logit) {
            if (dmalloc(&ws->wv16, (int64_t)ws->W * ws->V1 * ws->R) || dmalloc(&ws->wv_bias, (int64_t)ws->W * ws->V1) ||
                dmalloc(&ws->heads, kCandLists * WR) || dmalloc(&ws->entries, (int64_t)kCandLists * WR * kCandCap)) return 1;
        }
    }
    ws->hsq_n = ws->dense_lstm ? 2 * ws->ntl : (d.layer_n == 2 ? (ws->R + 31) / 32 : 1);
    if (dmalloc(&ws->hsq, WR * ws->hsq_n)) return 1;
    if (dmalloc(&ws->margin, WR) || dmalloc(&ws->pre_tok, WR) || dmalloc(&ws->pre_lp, WR) || dmalloc(&ws->unit_flag, WR / 128) || dmalloc(&ws->row_flag, WR) || dmalloc(&ws->stats, 4)) return 1;
    NES_CUDA(cudaMemsetAsync(ws->stats, 0, 32, c->stream));
    if (dmalloc(&ws->unfinished, WR) || dmalloc(&ws->unf_count, (int64_t)(ws->T + 2) * ws->W)) return 1;
    if (d.layer_n) { if (dmalloc(&ws->pre_i, WR * 5 * ws->R) || dmalloc(&ws->pre_h, WR * 5 * ws->R)) return 1; }
    if (d.vbn_e) { if (dmalloc(&ws->x32, WR * ws->E) || (d.vbn_e_affine && dmalloc(&ws->ebn_par, (int64_t)ws->W * 4 * ws->E))) return 1; }
    // zero everything once: padding rows of the operand buffers must stay finite forever
    NES_CUDA(cudaMemsetAsync(ws->x_hi, 0, sizeof(__half) * WR * ws->E, c->stream));
    NES_CUDA(cudaMemsetAsync(ws->x_lo, 0, sizeof(__half) * WR * ws->E, c->stream));
    for (int b = 0; b < 2; ++b) {
        NES_CUDA(cudaMemsetAsync(ws->h_hi[b], 0, sizeof(__half) * WR * ws->R, c->stream));
        NES_CUDA(cudaMemsetAsync(ws->h_lo[b], 0, sizeof(__half) * WR * ws->R, c->stream));
    }
    NES_CUDA(cudaMemsetAsync(ws->c_state, 0, sizeof(float) * WR * ws->R, c->stream));
    return 0;
}

void decode_destroy(Ctx* c) {
    DecodeWorkspace* ws = c->ws;
    if (!ws) return;
    void* ptrs[] = {ws->fc32, ws->fc_hi, ws->fc_lo, ws->x_hi, ws->x_lo, ws->h_hi[0], ws->h_hi[1], ws->h_lo[0], ws->h_lo[1], ws->c_state,
                    ws->pre_i, ws->pre_h, ws->x32, ws->ebn_par, ws->y0, ws->pp_off, ws->pp_sign, ws->pp_par, ws->partials, ws->heads, ws->entries, ws->wl_hi[0], ws->wl_hi[1], ws->wl_lo[0], ws->wl_lo[1], ws->wl_bias, ws->wv16, ws->wv_bias, ws->wi_hi, ws->wi_lo, ws->wi_bias, ws->wn_hi[0], ws->wn_hi[1], ws->wn_lo[0], ws->wn_lo[1], ws->wn_bias[0], ws->wn_bias[1], ws->ln_par, ws->hsq, ws->margin, ws->pre_tok, ws->pre_lp, ws->unit_flag, ws->row_flag, ws->stats, ws->unfinished, ws->unf_count, ws->d_off, ws->d_sign, ws->d_parent, ws->d_seq,
                    ws->d_lp, ws->d_forced, ws->d_uniforms, ws->samp_logits, ws->wv_lo};
    for (void* p : ptrs) if (p) cudaFree(p);
    delete ws;
    c->ws = nullptr;
}

static int sm_count(Ctx* c) {
    static int n[64] = {0};
    if (!n[c->device]) { if (cudaDeviceGetAttribute(&n[c->device], cudaDevAttrMultiProcessorCount, c->device) != cudaSuccess || n[c->device] < 1) n[c->device] = 148; }
    return n[c->device];
}
static int env_debug();
template <int EPI, int PREC>
static int launch_gemm(Ctx* c, GemmParams& p, int64_t n_members) {
    using C = GemmCfg<EPI, PREC>;
    static bool attr_done = false;
    if (!attr_done) {
        NES_CUDA(cudaFuncSetAttribute(k_member_gemm<EPI, PREC>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        attr_done = true;
    }
    const int64_t units = n_members * p.m_tiles;
    NES_CHECK(units >= 1 && units < (1ll << 31), "bad unit count");
    p.n_units = (int32_t)units;
    // persistent: one CTA per SM, CTA b owns units b, b + grid, ... and sweeps every weight tile of each
    const int sms = sm_count(c);
    const unsigned grid = (unsigned)(units < sms ? units : sms);
    ProfScope prof(c, EPI == EPI_PLAIN ? PROF_PLAIN : (EPI == EPI_LSTM ? PROF_LSTM : (p.unit_mask ? PROF_FALLBACK : PROF_LOGIT)));
    k_member_gemm<EPI, PREC><<<grid, C::THREADS, C::SMEM_BYTES, c->stream>>>(p);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}
template <int EPI>
static int launch_gemm_prec(Ctx* c, GemmParams& p, int64_t n_members, int prec) {
    return prec == NES_PREC_EXACT ? launch_gemm<EPI, NES_PREC_EXACT>(c, p, n_members) : launch_gemm<EPI, NES_PREC_FAST>(c, p, n_members);
}

static int env_debug() { const char* e = getenv("NES_DEBUG"); return e ? atoi(e) : 0; }     // tests: bit 6 (64) sends every row through the HYBRID fallback
template <int EPI>
static int launch_dense(Ctx* c, DenseParams& p, int64_t n_members) {
    using C = DenseCfg<EPI>;
    static bool attr_done = false;
    if (!attr_done) {
        NES_CUDA(cudaFuncSetAttribute(k_member_dense<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        attr_done = true;
    }
    const int64_t units = n_members * p.m_tiles;
    NES_CHECK(units >= 1 && units < (1ll << 31), "bad unit count");
    p.n_units = (int32_t)units;
    const int sms = sm_count(c);
    if (p.n_split < 1) { p.n_split = 1; p.tiles_per_split = p.n_tiles; }
    const int64_t items = EPI != DEN_LOGITC ? units * p.n_tiles : units * p.n_split;   // see k_member_dense: LSTM / PLAIN work items are (unit, tile) pairs
    NES_CHECK(items < (1ll << 31), "too many work items");
    const unsigned grid = (unsigned)(items < sms ? items : sms);
    ProfScope prof(c, EPI == DEN_LSTM ? PROF_LSTM : (EPI == DEN_PLAIN ? PROF_PLAIN : PROF_LOGIT));
    k_member_dense<EPI><<<grid, C::THREADS, C::SMEM_BYTES, c->stream>>>(p);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

// LSTM step as clusters of two CTAs sharing the activation block (lstm_pair.cuh); NES_LSTM_PAIR=0 keeps the one-CTA-per-item kernel
static int lstm_pair_clusters(Ctx* c) {
    static int n[64] = {0};
    if (!n[c->device]) {
        using C = DenseCfg<DEN_LSTM>;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * 74); cfg.blockDim = dim3(C::THREADS); cfg.dynamicSmemBytes = C::SMEM_BYTES;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int m = 0;
        if (cudaOccupancyMaxActiveClusters(&m, k_lstm_pair, &cfg) != cudaSuccess || m < 1) { cudaGetLastError(); m = sm_count(c) / 2; }
        n[c->device] = m;
    }
    return n[c->device];
}
static bool lstm_pair_enabled() { static int v = -1; if (v < 0) { const char* e = getenv("NES_LSTM_PAIR"); v = (e && atoi(e) == 0) ? 0 : 1; } return v == 1; }
static int launch_lstm_pair(Ctx* c, DenseParams& p, int64_t n_members) {
    using C = DenseCfg<DEN_LSTM>;
    static bool attr_done = false;
    if (!attr_done) {
        NES_CUDA(cudaFuncSetAttribute(k_lstm_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        attr_done = true;
    }
    const int64_t units = n_members * p.m_tiles;
    NES_CHECK(units >= 1 && units * p.n_tiles < (1ll << 30), "bad unit count");
    p.n_units = (int32_t)units;
    const int64_t pairs = (units * p.n_tiles + 1) / 2;
    const int ncl = lstm_pair_clusters(c);
    const unsigned grid = 2u * (unsigned)(pairs < ncl ? pairs : ncl);
    ProfScope prof(c, PROF_LSTM);
    k_lstm_pair<<<grid, C::THREADS, C::SMEM_BYTES, c->stream>>>(p);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

// member weights of one wave, materialised once per rollout (dense.cuh)
static int materialise_wave(Ctx* c, int64_t nm, const int64_t* off, const float* sign, const int32_t* parent, float sigma, bool lstm, bool logit, bool img, bool ln, bool ln_par, bool ebn_par, bool logit_split = false) {
    DecodeWorkspace* ws = c->ws;
    const Layout& L = c->lay;
    ProfScope prof(c, PROF_MATERIALISE);
    auto mat = [&](int wt, int K, int out_rows, int mode, int src_rows, __half* hi, __half* lo) {
        MatParams q{};
        q.off = off; q.sign = sign; q.parent = parent; q.theta32 = c->theta32; q.table = c->table; q.D = L.D; q.scale = c->scale32; q.scale_div = c->scale_div;
        q.w_start = L.t[c->tix[wt]].start; q.w_pstart = L.t[c->tix[wt]].pstart; q.K = K; q.out_rows = out_rows; q.mode = mode; q.R = ws->R; q.src_rows = src_rows;
        q.sigma = sigma; q.hi = hi; q.lo = lo; q.n_members = nm;
        const int64_t groups = (nm + kMatGroup - 1) / kMatGroup;
        const dim3 grid((unsigned)((out_rows + kMatRowsPerBlock - 1) / kMatRowsPerBlock), (unsigned)(groups < 32768 ? groups : 32768));
        if (q.scale) { if (lo) k_materialise<true, true><<<grid, 256, 0, c->stream>>>(q); else k_materialise<true, false><<<grid, 256, 0, c->stream>>>(q); }
        else         { if (lo) k_materialise<false, true><<<grid, 256, 0, c->stream>>>(q); else k_materialise<false, false><<<grid, 256, 0, c->stream>>>(q); }
        c->launches++;
    };
    auto matb = [&](int b0, int b1, int out_cols, int mode, int src_n, float* out, int64_t out_stride = 0, int64_t out_off = 0) {
        MatBiasParams q{};
        q.out_stride = out_stride; q.out_off = out_off;
        q.off = off; q.sign = sign; q.parent = parent; q.theta32 = c->theta32; q.table = c->table; q.D = L.D; q.scale = c->scale32; q.scale_div = c->scale_div;
        q.b_start[0] = L.t[c->tix[b0]].start; q.b_pstart[0] = L.t[c->tix[b0]].pstart; q.nb = 1;
        if (b1 >= 0) { q.b_start[1] = L.t[c->tix[b1]].start; q.b_pstart[1] = L.t[c->tix[b1]].pstart; q.nb = 2; }
        q.out_cols = out_cols; q.mode = mode; q.R = ws->R; q.src_n = src_n; q.sigma = sigma; q.out = out; q.n_members = nm;
        k_materialise_bias<<<148 * 4, 256, 0, c->stream>>>(q);
        c->launches++;
    };
    if (lstm) {
        const int rows_l = ws->ntl * kLstmTileCols;
        mat(T_I2H_W, ws->E, rows_l, 1, 5 * ws->R, ws->wl_hi[0], ws->wl_lo[0]);
        mat(T_H2H_W, ws->R, rows_l, 1, 5 * ws->R, ws->wl_hi[1], ws->wl_lo[1]);
        matb(T_I2H_B, T_H2H_B, rows_l, 1, 5 * ws->R, ws->wl_bias);
    }
    if (ln) {
        const int N5 = 5 * ws->R;
        mat(T_I2H_W, ws->E, N5, 0, N5, ws->wn_hi[0], ws->wn_lo[0]);
        mat(T_H2H_W, ws->R, N5, 0, N5, ws->wn_hi[1], ws->wn_lo[1]);
        matb(T_I2H_B, -1, N5, 0, N5, ws->wn_bias[0]);
        matb(T_H2H_B, -1, N5, 0, N5, ws->wn_bias[1]);
    }
    if (ln_par) {        // member LayerNorm parameters (theta +- sigma eps of the six LN vectors) for k_ln_gates
        const int N5 = 5 * ws->R;
        const int64_t st = 22ll * ws->R;
        for (int i = 0; i < 4; ++i) matb(T_I2H_LN_W + i, -1, N5, 0, N5, ws->ln_par, st, (int64_t)i * N5);
        for (int i = 0; i < 2; ++i) matb(T_C_LN_W + i, -1, ws->R, 0, ws->R, ws->ln_par, st, 4ll * N5 + (int64_t)i * ws->R);
    }
    if (ebn_par) {       // member affine pairs of the two embedding batch norms (vbn_e): img_embed.1.{weight,bias}, embed.1.{weight,bias}
        const int roles[4] = {T_IMG_BN_W, T_IMG_BN_B, T_EMB_BN_W, T_EMB_BN_B};
        for (int i = 0; i < 4; ++i) matb(roles[i], -1, ws->E, 0, ws->E, ws->ebn_par, 4ll * ws->E, (int64_t)i * ws->E);
    }
    if (img) {
        mat(T_IMG_W, ws->Fd, ws->E, 0, ws->E, ws->wi_hi, ws->wi_lo);
        matb(T_IMG_B, -1, ws->E, 0, ws->E, ws->wi_bias);
    }
    if (logit) {
        mat(T_LOGIT_W, ws->R, ws->V1, 0, ws->V1, ws->wv16, logit_split ? ws->wv_lo : nullptr);
        matb(T_LOGIT_B, -1, ws->V1, 0, ws->V1, ws->wv_bias);
    }
    NES_CUDA(cudaGetLastError());
    return 0;
}

// fill one GEMM segment: activations A (hi/lo buffers with `a_rows` rows), weight tensor wt, bias tensor bt (-1 none)
static int fill_seg(Ctx* c, GemmParams& p, int s, const __half* a_hi, const __half* a_lo, int64_t a_rows, int K, int a_shared,
                    int wt, int bt, int box_rows, int prec, int n_chunks = 1, int64_t chunk_stride = 0) {
    const int bk = prec == NES_PREC_EXACT ? GemmCfg<EPI_PLAIN, NES_PREC_EXACT>::BK : GemmCfg<EPI_PLAIN, NES_PREC_FAST>::BK;
    DecodeWorkspace* ws = c->ws;
    const Layout& L = c->lay;
    if (get_map(ws, &p.maps[s].a_hi, a_hi, a_rows, K, 128, bk) || get_map(ws, &p.maps[s].a_lo, a_lo, a_rows, K, 128, bk)) return 1;
    const int64_t th_rows = (int64_t)c->n_parents * L.D_pad / K;
    if (get_map(ws, &p.maps[s].w_hi, c->theta_hi, th_rows, K, box_rows, bk, n_chunks, chunk_stride) ||
        get_map(ws, &p.maps[s].w_lo, c->theta_lo, th_rows, K, box_rows, bk, n_chunks, chunk_stride)) return 1;
    if (get_map(ws, &p.maps[s].eps, c->table, c->table_len / K, K, box_rows, bk, n_chunks, chunk_stride)) return 1;
    p.seg[s].K = K; p.seg[s].a_shared = a_shared; p.seg[s].w_pstart = L.t[c->tix[wt]].pstart;
    p.seg[s].b_start = bt >= 0 ? L.t[c->tix[bt]].start : -1;
    p.seg[s].b_pstart = bt >= 0 ? L.t[c->tix[bt]].pstart : -1;
    return 0;
}

static int ensure_call_buffers(Ctx* c, int64_t n_members, int rows, bool forced) {
    DecodeWorkspace* ws = c->ws;
    if (n_members > ws->cap_members) {
        if (ws->d_off) { cudaFree(ws->d_off); cudaFree(ws->d_sign); cudaFree(ws->d_parent); }
        if (dmalloc(&ws->d_off, n_members) || dmalloc(&ws->d_sign, n_members) || dmalloc(&ws->d_parent, n_members)) return 1;
        ws->cap_members = n_members;
    }
    const int64_t out_n = n_members * rows * ws->T;
    if (out_n > ws->out_cap) {
        if (ws->d_seq) { cudaFree(ws->d_seq); cudaFree(ws->d_lp); }
        if (dmalloc(&ws->d_seq, out_n) || dmalloc(&ws->d_lp, out_n)) return 1;
        ws->out_cap = out_n;
    }
    if (forced && out_n > ws->forced_cap) {
        if (ws->d_forced) cudaFree(ws->d_forced);
        if (dmalloc(&ws->d_forced, out_n)) return 1;
        ws->forced_cap = out_n;
    }
    return 0;
}

int decode_rollout(Ctx* c, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents,
                   const float* fc, int rows, int shared_batch, float sigma, int precision,
                   const int64_t* forced, int64_t* seq, float* logp, float* h_trace, float* logit_trace, const double* uniforms) {
    DecodeWorkspace* ws = c->ws;
    NES_CHECK(ws != nullptr, "context has no decode workspace");
    NES_CHECK(rows >= 1 && rows <= ws->rows_cap, "rows exceeds the workspace capacity given to nes_ctx_create_fc");
    NES_CHECK(precision == NES_PREC_FAST || precision == NES_PREC_EXACT || precision == NES_PREC_HYBRID, "unknown precision");
    NES_CHECK(n_members >= 0, "negative member count");
    const bool tracing = h_trace || logit_trace || forced;
    // HYBRID = split-fp16 everywhere except the vocabulary projection, which runs on fp16 operands with an exact
    // fp32 re-evaluation of the arg-max candidates (hybrid.cuh); a logit trace needs the accurate logits themselves
    const bool sampling = uniforms != nullptr;      // tokens are drawn from the full distribution: every accurate logit is needed
    NES_CHECK(!sampling || !logit_trace, "sampled decode and a logit trace exclude each other");
    const bool hybrid = precision == NES_PREC_HYBRID && !logit_trace && !sampling && ws->dense_logit;
    if (precision == NES_PREC_HYBRID) precision = NES_PREC_EXACT;
    const bool dense_lstm = ws->dense_lstm && precision == NES_PREC_EXACT;     // split-fp16 LSTM on materialised member weights
    const bool dense_ln = ws->dense_ln && precision == NES_PREC_EXACT;         // LayerNorm cell on dense pre-activation GEMMs
    const bool scaled = c->scale32 != nullptr;                                 // safe-mutation scale: every member weight is materialised
    // sampled decode needs every accurate logit: split-fp16 projection on the materialised logit.weight (hi + lo), float32 logits out
    const bool samp_dense = sampling && ws->dense_logit && (dense_lstm || dense_ln);
    if (scaled) {
        NES_CHECK((hybrid || samp_dense) && (dense_lstm || dense_ln), "a safe-mutation scale needs HYBRID precision on the dense kernels (no logit trace, R <= 512)");
        if (!ws->wi_hi) {
            if (dmalloc(&ws->wi_hi, (int64_t)ws->W * ws->E * ws->Fd) || dmalloc(&ws->wi_lo, (int64_t)ws->W * ws->E * ws->Fd) || dmalloc(&ws->wi_bias, (int64_t)ws->W * ws->E)) return 1;
        }
    }
    NES_CHECK(!tracing || n_members <= ws->W, "trace mode handles a single wave");
    if (n_members == 0) return 0;
    if (hybrid && !c->eps_rowmax_valid && launch_bound_norms(c, -1)) return 1;
    const Layout& L = c->lay;
    const int E = ws->E, R = ws->R, Fd = ws->Fd, V1 = ws->V1, T = ws->T;
    const int rows_pad = (rows + 127) / 128 * 128, m_tiles = rows_pad / 128;
    if (ensure_call_buffers(c, n_members, rows, forced != nullptr)) return 1;
    NES_CUDA(cudaMemcpyAsync(ws->d_off, offsets, sizeof(int64_t) * n_members, cudaMemcpyDefault, c->stream));
    NES_CUDA(cudaMemcpyAsync(ws->d_sign, signs, sizeof(float) * n_members, cudaMemcpyDefault, c->stream));
    if (parents) NES_CUDA(cudaMemcpyAsync(ws->d_parent, parents, sizeof(int32_t) * n_members, cudaMemcpyDefault, c->stream));
    else NES_CUDA(cudaMemsetAsync(ws->d_parent, 0, sizeof(int32_t) * n_members, c->stream));
    if (forced) NES_CUDA(cudaMemcpyAsync(ws->d_forced, forced, sizeof(int64_t) * n_members * rows * T, cudaMemcpyDefault, c->stream));
    if (sampling) {
        NES_CHECK(precision != NES_PREC_FAST, "sampled decode runs on the accurate logits (EXACT / HYBRID)");
        const int64_t un = n_members * rows * T;
        if (un > ws->uniforms_cap) {
            if (ws->d_uniforms) cudaFree(ws->d_uniforms);
            if (dmalloc(&ws->d_uniforms, un)) return 1;
            ws->uniforms_cap = un;
        }
        NES_CUDA(cudaMemcpyAsync(ws->d_uniforms, uniforms, sizeof(double) * un, cudaMemcpyDefault, c->stream));
        if (!ws->samp_logits && dmalloc(&ws->samp_logits, (int64_t)ws->W * ws->rows_pad * ((V1 + 3) / 4 * 4))) return 1;
        if (samp_dense && !ws->wv_lo && dmalloc(&ws->wv_lo, (int64_t)ws->W * V1 * ws->R)) return 1;
    }
    NES_CUDA(cudaMemsetAsync(ws->d_seq, 0, sizeof(int64_t) * n_members * rows * T, c->stream));
    NES_CUDA(cudaMemsetAsync(ws->d_lp, 0, sizeof(float) * n_members * rows * T, c->stream));

    // fc features -> split fp16 operand
    const int64_t fc_blocks = shared_batch ? 1 : n_members;
    const int64_t fc_need = shared_batch ? 1 : ws->W;
    if (fc_need > ws->fc_members || rows_pad > 0) {
        if (fc_need > ws->fc_members) {
            if (ws->fc32) { cudaFree(ws->fc32); cudaFree(ws->fc_hi); cudaFree(ws->fc_lo); }
            if (dmalloc(&ws->fc32, fc_need * ws->rows_cap * Fd) || dmalloc(&ws->fc_hi, fc_need * ws->rows_pad * Fd) ||
                dmalloc(&ws->fc_lo, fc_need * ws->rows_pad * Fd)) return 1;
            ws->fc_members = fc_need;
        }
    }
    if (c->rollout_timed) NES_CUDA(cudaEventRecord(c->ev_c, c->stream));

    for (int64_t m0 = 0; m0 < n_members; m0 += ws->W) {
        const int64_t nm = (n_members - m0) < ws->W ? (n_members - m0) : ws->W;
        const int64_t WR = nm * rows_pad;
        // ---- stage inputs / reset state
        const int64_t nb = shared_batch ? 1 : nm;
        if (!shared_batch || m0 == 0) {
            const float* src = fc + (shared_batch ? 0 : m0 * rows * (int64_t)Fd);
            NES_CUDA(cudaMemcpyAsync(ws->fc32, src, sizeof(float) * nb * rows * Fd, cudaMemcpyDefault, c->stream));
            k_split_rows<<<148 * 4, 256, 0, c->stream>>>(ws->fc32, ws->fc_hi, ws->fc_lo, nb, rows, rows_pad, Fd);
            c->launches++;
            NES_CUDA(cudaGetLastError());
        }
        (void)fc_blocks;
        NES_CUDA(cudaMemsetAsync(ws->c_state, 0, sizeof(float) * WR * R, c->stream));
        NES_CUDA(cudaMemsetAsync(ws->h_hi[0], 0, sizeof(__half) * WR * R, c->stream));
        NES_CUDA(cudaMemsetAsync(ws->h_lo[0], 0, sizeof(__half) * WR * R, c->stream));
        NES_CUDA(cudaMemsetAsync(ws->unf_count, 0, sizeof(int32_t) * (T + 2) * ws->W, c->stream));
        const bool ln_par = c->dims.layer_n && c->dims.layer_n_affine;
        const bool vbn_e = c->dims.vbn_e != 0, ebn_par = vbn_e && c->dims.vbn_e_affine;
        if ((dense_lstm || hybrid || samp_dense || dense_ln || ln_par || ebn_par) && materialise_wave(c, nm, ws->d_off + m0, ws->d_sign + m0, ws->d_parent + m0, sigma, dense_lstm, hybrid || samp_dense, scaled, dense_ln, ln_par, ebn_par, samp_dense)) return 1;
        auto embed_bn = [&](int site) -> int {          // batch norm of the embedding the wave just produced (vbn_e)
            EmbedBnParams q{};
            q.x32 = ws->x32; q.x_hi = ws->x_hi; q.x_lo = ws->x_lo; q.par = ebn_par ? ws->ebn_par : nullptr; q.site = site;
            q.rows = rows; q.rows_pad = rows_pad; q.E = E;
            ProfScope prof(c, PROF_FINALIZE);
            k_embed_bn<<<dim3((unsigned)((E + 31) / 32), (unsigned)nm), 256, 0, c->stream>>>(q);
            c->launches++;
            NES_CUDA(cudaGetLastError());
            return 0;
        };

        GemmParams base{};
        base.rows = rows; base.rows_pad = rows_pad; base.m_tiles = m_tiles;
        base.D = L.D; base.D_pad = L.D_pad;
        base.off = ws->d_off + m0; base.sign = ws->d_sign + m0; base.parent = ws->d_parent + m0;
        base.theta32 = c->theta32; base.table = c->table; base.sigma = sigma;
        // antithetic pairs (+eps, -eps adjacent members) run on neighbouring SMs at the same tile index and share
        // their eps tiles through L2

        const int64_t a_rows_w = (int64_t)ws->W * ws->rows_pad;
        // ---- t = 0 : x0 = img_embed(fc)
        if (scaled) {
            DenseParams p{};
            const int bk = DenseCfg<DEN_PLAIN>::BK;
            if (get_map(ws, &p.a_hi[0], ws->fc_hi, ws->fc_members * ws->rows_pad, Fd, 128, bk) || get_map(ws, &p.a_lo[0], ws->fc_lo, ws->fc_members * ws->rows_pad, Fd, 128, bk) ||
                get_map(ws, &p.b_hi[0], ws->wi_hi, (int64_t)ws->W * E, Fd, 256, bk) || get_map(ws, &p.b_lo[0], ws->wi_lo, (int64_t)ws->W * E, Fd, 256, bk)) return 1;
            p.a_hi[1] = p.a_hi[0]; p.a_lo[1] = p.a_lo[0]; p.b_hi[1] = p.b_hi[0]; p.b_lo[1] = p.b_lo[0];
            p.K[0] = Fd; p.K[1] = 0; p.nseg = 1; p.a_shared = shared_batch ? 1 : 0;
            p.rows = rows; p.rows_pad = rows_pad; p.m_tiles = m_tiles; p.n_tiles = (E + 255) / 256; p.N = E;
            p.b_rows = E; p.bias = ws->wi_bias; p.bias_stride = E;
            if (vbn_e) { p.out_f32 = ws->x32; p.ld_out = E; } else { p.out_hi = ws->x_hi; p.out_lo = ws->x_lo; }
            if (launch_dense<DEN_PLAIN>(c, p, nm)) return 1;
        } else {
            GemmParams p = base;
            p.nseg = 1; p.N = E; p.chunk_stride = 0; p.n_tiles = (E + 127) / 128;
            if (fill_seg(c, p, 0, ws->fc_hi, ws->fc_lo, ws->fc_members * ws->rows_pad, Fd, shared_batch ? 1 : 0, T_IMG_W, T_IMG_B, 128, precision)) return 1;
            if (shared_batch && precision == NES_PREC_EXACT && n_members >= 2 * (int64_t)c->n_parents) {     // it pays when members outnumber the slots
                // every member sees the same features: x theta_p^T + theta_b once per parent slot (sigma = 0 pseudo members), then the
                // members only add their sigma * x eps^T (2 of the 5 split-fp16 passes, no theta tiles through shared memory)
                const int np = c->n_parents;
                if (!ws->y0) {
                    if (dmalloc(&ws->y0, (int64_t)np * ws->rows_pad * E) || dmalloc(&ws->pp_off, np) || dmalloc(&ws->pp_sign, np) || dmalloc(&ws->pp_par, np)) return 1;
                    std::vector<float> ones(np, 1.0f); std::vector<int32_t> iota(np);
                    for (int i = 0; i < np; ++i) iota[i] = i;
                    NES_CUDA(cudaMemsetAsync(ws->pp_off, 0, sizeof(int64_t) * np, c->stream));
                    NES_CUDA(cudaMemcpyAsync(ws->pp_sign, ones.data(), sizeof(float) * np, cudaMemcpyHostToDevice, c->stream));
                    NES_CUDA(cudaMemcpyAsync(ws->pp_par, iota.data(), sizeof(int32_t) * np, cudaMemcpyHostToDevice, c->stream));
                    NES_CUDA(cudaStreamSynchronize(c->stream));
                }
                if (m0 == 0) {
                    GemmParams pt = p;
                    pt.off = ws->pp_off; pt.sign = ws->pp_sign; pt.parent = ws->pp_par; pt.sigma = 0.f;
                    pt.out_f32 = ws->y0; pt.ld_out = E; pt.out_hi = nullptr; pt.out_lo = nullptr; pt.y0 = nullptr;
                    if (launch_gemm_prec<EPI_PLAIN>(c, pt, np, precision)) return 1;
                }
                p.y0 = ws->y0;
            }
            if (vbn_e) { p.out_f32 = ws->x32; p.ld_out = E; } else { p.out_hi = ws->x_hi; p.out_lo = ws->x_lo; }
            if (launch_gemm_prec<EPI_PLAIN>(c, p, nm, precision)) return 1;
        }
        if (vbn_e && embed_bn(0)) return 1;
        auto make_fin = [&](int t) {
            FinalizeParams f{};
            f.partials = ws->partials; f.n_tiles = 2; f.rows = rows; f.rows_pad = rows_pad; f.T = T; f.E = E; f.t = t;
            f.n_members = nm; f.off = base.off; f.sign = base.sign; f.parent = base.parent; f.theta32 = c->theta32; f.table = c->table; f.scale = c->scale32; f.scale_div = c->scale_div;
            f.D = L.D; f.emb_start = L.t[c->tix[T_EMB_W]].start; f.emb_pstart = L.t[c->tix[T_EMB_W]].pstart; f.sigma = sigma;
            f.unfinished = ws->unfinished; f.unf_count = ws->unf_count; f.W = ws->W;
            f.seq = ws->d_seq + m0 * rows * T; f.lp = ws->d_lp + m0 * rows * T;
            f.forced = forced ? ws->d_forced + m0 * rows * T : nullptr;
            f.x_hi = ws->x_hi; f.x_lo = ws->x_lo; f.gather = t < T ? 1 : 0; f.x32 = vbn_e ? ws->x32 : nullptr;
            if (sampling && t >= 1) { f.pre_tok = ws->pre_tok; f.pre_lp = ws->pre_lp; f.unit_flag = ws->unit_flag; f.m_tiles = m_tiles; f.pre_always = 1; }
            if (hybrid && t >= 1) { f.pre_tok = ws->pre_tok; f.pre_lp = ws->pre_lp; f.unit_flag = ws->unit_flag; f.m_tiles = m_tiles; f.pre_always = scaled ? 1 : 0; }
            return f;
        };
        int hb = 0;   // h buffer holding the current state
        for (int t = 0; t <= T; ++t) {
            // ---- (h, c) = core(x_t, (h, c))
            float* ht = h_trace ? h_trace + (int64_t)t * n_members * rows * R : nullptr;
            if (dense_lstm) {
                DenseParams p{};
                const int bk = DenseCfg<DEN_LSTM>::BK;
                const int64_t b_rows_all = (int64_t)ws->W * ws->ntl * kLstmTileCols;
                if (get_map(ws, &p.a_hi[0], ws->x_hi, a_rows_w, E, 128, bk) || get_map(ws, &p.a_lo[0], ws->x_lo, a_rows_w, E, 128, bk) ||
                    get_map(ws, &p.a_hi[1], ws->h_hi[hb], a_rows_w, R, 128, bk) || get_map(ws, &p.a_lo[1], ws->h_lo[hb], a_rows_w, R, 128, bk) ||
                    get_map(ws, &p.b_hi[0], ws->wl_hi[0], b_rows_all, E, kLstmTileCols, bk) || get_map(ws, &p.b_lo[0], ws->wl_lo[0], b_rows_all, E, kLstmTileCols, bk) ||
                    get_map(ws, &p.b_hi[1], ws->wl_hi[1], b_rows_all, R, kLstmTileCols, bk) || get_map(ws, &p.b_lo[1], ws->wl_lo[1], b_rows_all, R, kLstmTileCols, bk)) return 1;
                p.K[0] = E; p.K[1] = R; p.nseg = t == 0 ? 1 : 2;      // t = 0: h = 0, the h2h product vanishes (its bias is in wl_bias)
                p.rows = rows; p.rows_pad = rows_pad; p.m_tiles = m_tiles; p.n_tiles = ws->ntl; p.N = R;
                p.b_rows = ws->ntl * kLstmTileCols; p.bias = ws->wl_bias; p.bias_stride = ws->ntl * kLstmTileCols;
                p.c_state = ws->c_state; p.out_hi = ws->h_hi[hb ^ 1]; p.out_lo = ws->h_lo[hb ^ 1]; p.h_trace = ht;
                p.hsq = hybrid ? ws->hsq : nullptr; p.hsq_n = ws->hsq_n;
                if (lstm_pair_enabled() ? launch_lstm_pair(c, p, nm) : launch_dense<DEN_LSTM>(c, p, nm)) return 1;
            } else if (!c->dims.layer_n) {
                GemmParams p = base;
                p.nseg = 2; p.N = R; p.chunk_stride = R; p.n_tiles = R / 16;
                if (fill_seg(c, p, 0, ws->x_hi, ws->x_lo, a_rows_w, E, 0, T_I2H_W, T_I2H_B, 16, precision, 5, R)) return 1;
                if (fill_seg(c, p, 1, ws->h_hi[hb], ws->h_lo[hb], a_rows_w, R, 0, T_H2H_W, T_H2H_B, 16, precision, 5, R)) return 1;
                p.c_state = ws->c_state; p.out_hi = ws->h_hi[hb ^ 1]; p.out_lo = ws->h_lo[hb ^ 1]; p.h_trace = ht;
                if (launch_gemm_prec<EPI_LSTM>(c, p, nm, precision)) return 1;
            } else {
                for (int s = 0; s < 2 && dense_ln; ++s) {
                    DenseParams p{};
                    const int bk = DenseCfg<DEN_PLAIN>::BK;
                    const int K = s == 0 ? E : R;
                    const __half* ah = s == 0 ? ws->x_hi : ws->h_hi[hb]; const __half* al = s == 0 ? ws->x_lo : ws->h_lo[hb];
                    if (get_map(ws, &p.a_hi[0], ah, a_rows_w, K, 128, bk) || get_map(ws, &p.a_lo[0], al, a_rows_w, K, 128, bk) ||
                        get_map(ws, &p.b_hi[0], ws->wn_hi[s], (int64_t)ws->W * 5 * R, K, 256, bk) || get_map(ws, &p.b_lo[0], ws->wn_lo[s], (int64_t)ws->W * 5 * R, K, 256, bk)) return 1;
                    p.a_hi[1] = p.a_hi[0]; p.a_lo[1] = p.a_lo[0]; p.b_hi[1] = p.b_hi[0]; p.b_lo[1] = p.b_lo[0];
                    p.K[0] = K; p.K[1] = 0; p.nseg = 1; p.a_shared = 0;
                    p.rows = rows; p.rows_pad = rows_pad; p.m_tiles = m_tiles; p.n_tiles = (5 * R + 255) / 256; p.N = 5 * R;
                    p.b_rows = 5 * R; p.bias = ws->wn_bias[s]; p.bias_stride = 5 * R;
                    p.out_f32 = s == 0 ? ws->pre_i : ws->pre_h; p.ld_out = 5 * R;
                    if (launch_dense<DEN_PLAIN>(c, p, nm)) return 1;
                }
                for (int s = 0; s < 2 && !dense_ln; ++s) {
                    GemmParams p = base;
                    p.nseg = 1; p.N = 5 * R; p.chunk_stride = 0; p.n_tiles = (5 * R + 127) / 128;
                    if (s == 0) { if (fill_seg(c, p, 0, ws->x_hi, ws->x_lo, a_rows_w, E, 0, T_I2H_W, T_I2H_B, 128, precision)) return 1; }
                    else        { if (fill_seg(c, p, 0, ws->h_hi[hb], ws->h_lo[hb], a_rows_w, R, 0, T_H2H_W, T_H2H_B, 128, precision)) return 1; }
                    p.out_f32 = s == 0 ? ws->pre_i : ws->pre_h; p.ld_out = 5 * R;
                    if (launch_gemm_prec<EPI_PLAIN>(c, p, nm, precision)) return 1;
                }
                LnParams q{};
                q.pre_i = ws->pre_i; q.pre_h = ws->pre_h; q.c_state = ws->c_state; q.h_hi = ws->h_hi[hb ^ 1]; q.h_lo = ws->h_lo[hb ^ 1];
                q.h_trace = ht; q.rows = rows; q.rows_pad = rows_pad; q.R = R; q.n_members = nm;
                q.ln_par = c->dims.layer_n_affine ? ws->ln_par : nullptr;
                q.hsq = hybrid ? ws->hsq : nullptr; q.hsq_n = ws->hsq_n;
                const size_t ln_smem = sizeof(float) * 4 * 6 * R;
                if (ln_smem > 48 * 1024) { static size_t set_to = 0; if (set_to < ln_smem) { NES_CUDA(cudaFuncSetAttribute(k_ln_gates, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ln_smem)); set_to = ln_smem; } }
                ProfScope prof(c, PROF_FINALIZE);
                if (c->dims.layer_n == 2) k_bn_gates<<<dim3((unsigned)((R + 31) / 32), (unsigned)nm), 256, 0, c->stream>>>(q);
                else if (R == 128) k_ln_gates_reg<4><<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(q);
                else if (R == 64) k_ln_gates_reg<2><<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(q);
                else if (R == 256) k_ln_gates_reg<8><<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(q);
                else k_ln_gates<<<(unsigned)((nm * rows + 3) / 4), 128, ln_smem, c->stream>>>(q);
                c->launches++;
                NES_CUDA(cudaGetLastError());
            }
            hb ^= 1;
            // ---- logits of step t (t >= 1), argmax / log-prob, next input
            if (t >= 1) {
                GemmParams p = base;
                p.nseg = 1; p.N = V1; p.chunk_stride = 0; p.n_tiles = ws->nt_logit;
                { int ex; frexpf(logf((float)V1), &ex); p.tie_delta = ldexpf(1.0f, ex - 1 - 23); }
                p.partials = ws->partials;
                if (hybrid) {
                    // (1) per-row error bound, (2) fp16-operand logits + candidate lists, (3) fp32 re-evaluation of the candidates
                    MarginParams mg{};
                    mg.h_hi = ws->h_hi[hb]; mg.h_lo = ws->h_lo[hb]; mg.R = R; mg.rows = rows; mg.rows_pad = rows_pad; mg.n_members = nm;
                    mg.parent = base.parent; mg.theta_rowmax = c->d_theta_rowmax; mg.eps_rowmax = c->d_eps_rowmax; mg.scale_max = c->d_scale_max; mg.sigma = sigma; mg.margin = ws->margin;
                    // the cell kernel that produced h left the pieces of |h|^2 behind: LOGITC forms the margin itself; only the factored
                    // LSTM kernel (NES_FACTORED) needs the separate pass over h
                    const bool hsq_ok = dense_lstm || c->dims.layer_n != 0;
                    if (!hsq_ok) {
                        ProfScope prof(c, PROF_FINALIZE);
                        k_row_margin<<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(mg); c->launches++;
                    }
                    NES_CUDA(cudaGetLastError());
                    int n_split = 1;
                    {
                        DenseParams pc{};
                        const int bk = DenseCfg<DEN_LOGITC>::BK;
                        if (get_map(ws, &pc.a_hi[0], ws->h_hi[hb], a_rows_w, R, 128, bk) ||
                            get_map(ws, &pc.b_hi[0], ws->wv16, (int64_t)ws->W * V1, R, 256, bk)) return 1;
                        pc.a_lo[0] = pc.a_hi[0]; pc.b_lo[0] = pc.b_hi[0];
                        pc.K[0] = R; pc.K[1] = 0; pc.nseg = 1;
                        pc.rows = rows; pc.rows_pad = rows_pad; pc.m_tiles = m_tiles; pc.n_tiles = ws->ntv; pc.N = V1;
                        pc.b_rows = V1; pc.bias = ws->wv_bias; pc.bias_stride = V1;
                        pc.heads = ws->heads; pc.entries = ws->entries; pc.margin = ws->margin;
                        if (hsq_ok) {
                            pc.hsq = ws->hsq; pc.hsq_n = ws->hsq_n; pc.mg_parent = base.parent; pc.mg_theta_rowmax = c->d_theta_rowmax;
                            pc.mg_eps_rowmax = c->d_eps_rowmax; pc.mg_scale_max = c->d_scale_max; pc.mg_sigma = sigma; pc.margin_out = ws->margin;
                        }
                        // a handful of units (the elite candidates of a NIC-ES generation): cut every unit's vocabulary sweep into ranges so
                        // that the launch spreads over the SMs (0.15 ms -> 0.02 ms per step for 2 members); the lists of a row multiply
                        const int64_t units = nm * m_tiles;
                        const int64_t cap = ((int64_t)ws->W * ws->rows_pad) / (nm * rows_pad);
                        int64_t sp = std::min<int64_t>(std::min<int64_t>(sm_count(c) / units, cap), ws->ntv);
                        const char* se = getenv("NES_LOGITC_SPLIT");                  // tests: 0 keeps few-member calls on whole-unit work items
                        if (se && atoi(se) == 0) sp = 1;
                        pc.n_split = 1; pc.tiles_per_split = ws->ntv;
                        if (sp >= 2) { pc.tiles_per_split = (int32_t)((ws->ntv + sp - 1) / sp); pc.n_split = (ws->ntv + pc.tiles_per_split - 1) / pc.tiles_per_split; }
                        n_split = pc.n_split;
                        if (launch_dense<DEN_LOGITC>(c, pc, nm)) return 1;
                    }
                    NES_CUDA(cudaMemsetAsync(ws->unit_flag, 0, (size_t)nm * m_tiles, c->stream));
                    RefineParams rf{};
                    rf.heads = ws->heads; rf.entries = ws->entries; rf.margin = ws->margin; rf.h_hi = ws->h_hi[hb]; rf.h_lo = ws->h_lo[hb]; rf.R = R; rf.V1 = V1;
                    rf.rows = rows; rf.rows_pad = rows_pad; rf.m_tiles = m_tiles; rf.n_members = nm;
                    rf.off = base.off; rf.sign = base.sign; rf.parent = base.parent; rf.theta32 = c->theta32; rf.table = c->table; rf.D = L.D; rf.scale = c->scale32; rf.scale_div = c->scale_div;
                    rf.w_start = L.t[c->tix[T_LOGIT_W]].start; rf.w_pstart = L.t[c->tix[T_LOGIT_W]].pstart; rf.b_start = L.t[c->tix[T_LOGIT_B]].start; rf.b_pstart = L.t[c->tix[T_LOGIT_B]].pstart;
                    rf.sigma = sigma; rf.tok = ws->pre_tok; rf.lp = ws->pre_lp; rf.unit_flag = ws->unit_flag; rf.stats = c->profiling ? ws->stats : nullptr;
                    rf.force_overflow = (env_debug() & 64) ? 1 : 0; rf.n_split = n_split;
                    { ProfScope prof(c, PROF_FINALIZE);       // (3) + the rows' bookkeeping and next-input gather, for every row that did not overflow
                      NES_CHECK(nm * rows < (1ll << 30), "too many rows in a wave");
                      if (c->scale32) k_refine_finalize<true><<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(rf, make_fin(t), ws->row_flag);
                      else k_refine_finalize<false><<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(rf, make_fin(t), ws->row_flag);
                      c->launches++; }
                    NES_CUDA(cudaGetLastError());
                    p.unit_mask = ws->unit_flag;      // (4) units whose candidate list overflowed: the split-fp16 kernel, everything else skips
                    if (scaled) {                     //     ... or, with scaled noise, every logit of those rows in fp32 (k_exact_rows)
                        ProfScope prof(c, PROF_FALLBACK);
                        k_exact_rows<<<(unsigned)(nm * rows), 256, sizeof(float) * V1, c->stream>>>(rf); c->launches++;
                        NES_CUDA(cudaGetLastError());
                    }
                }
                const int64_t samp_ld = (V1 + 3) / 4 * 4;
                if (samp_dense) {
                    DenseParams pd{};
                    const int bk = DenseCfg<DEN_PLAIN>::BK;
                    if (get_map(ws, &pd.a_hi[0], ws->h_hi[hb], a_rows_w, R, 128, bk) || get_map(ws, &pd.a_lo[0], ws->h_lo[hb], a_rows_w, R, 128, bk) ||
                        get_map(ws, &pd.b_hi[0], ws->wv16, (int64_t)ws->W * V1, R, 256, bk) || get_map(ws, &pd.b_lo[0], ws->wv_lo, (int64_t)ws->W * V1, R, 256, bk)) return 1;
                    pd.a_hi[1] = pd.a_hi[0]; pd.a_lo[1] = pd.a_lo[0]; pd.b_hi[1] = pd.b_hi[0]; pd.b_lo[1] = pd.b_lo[0];
                    pd.K[0] = R; pd.K[1] = 0; pd.nseg = 1; pd.a_shared = 0;
                    pd.rows = rows; pd.rows_pad = rows_pad; pd.m_tiles = m_tiles; pd.n_tiles = (V1 + 255) / 256; pd.N = V1;
                    pd.b_rows = V1; pd.bias = ws->wv_bias; pd.bias_stride = V1;
                    pd.out_f32 = ws->samp_logits; pd.ld_out = (int32_t)samp_ld;
                    if (launch_dense<DEN_PLAIN>(c, pd, nm)) return 1;
                }
                if (!(hybrid && scaled) && !samp_dense) {
                if (fill_seg(c, p, 0, ws->h_hi[hb], ws->h_lo[hb], a_rows_w, R, 0, T_LOGIT_W, T_LOGIT_B, 128, precision)) return 1;
                p.logit_trace = logit_trace ? logit_trace + (int64_t)(t - 1) * n_members * rows * V1 : nullptr; p.ld_trace = V1;
                if (sampling) p.logit_trace = ws->samp_logits;
                if (launch_gemm_prec<EPI_LOGIT>(c, p, nm, precision)) return 1;
                }
                if (sampling) {
                    SampleParams sp{};
                    sp.logits = ws->samp_logits; sp.partials = samp_dense ? nullptr : ws->partials; sp.n_tiles = 2; sp.u = ws->d_uniforms + m0 * rows * T;
                    sp.ld = samp_dense ? samp_ld : V1; sp.row_stride = samp_dense ? rows_pad : rows;
                    sp.rows = rows; sp.rows_pad = rows_pad; sp.T = T; sp.t = t; sp.V1 = V1; sp.n_members = nm; sp.tok = ws->pre_tok; sp.lp = ws->pre_lp;
                    const size_t smem = sizeof(float) * ((V1 + 3) / 4 * 4);
                    const int n4 = (V1 + 3) / 4;
                    const bool vec = sp.ld % 4 == 0 && (reinterpret_cast<uintptr_t>(sp.logits) & 15) == 0 && n4 <= 256 * 10;
                    ProfScope prof(c, PROF_FINALIZE);
                    if (vec && n4 <= 256 * 2) {
                        k_sample_reg<2><<<(unsigned)(nm * rows), 256, smem, c->stream>>>(sp);
                    } else if (vec) {
                        static bool attr_set = false;
                        if (!attr_set) { NES_CUDA(cudaFuncSetAttribute(k_sample_reg<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 256 * 10 * 4)); attr_set = true; }
                        k_sample_reg<10><<<(unsigned)(nm * rows), 256, smem, c->stream>>>(sp);
                    } else {
                        if (smem > 48 * 1024) { static size_t set_to = 0; if (set_to < smem) { NES_CUDA(cudaFuncSetAttribute(k_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set_to = smem; } }
                        k_sample<<<(unsigned)(nm * rows), 256, smem, c->stream>>>(sp);
                    }
                    c->launches++;
                    NES_CUDA(cudaGetLastError());
                }
            }
            if (t < T || t >= 1) {
                FinalizeParams f = make_fin(t);
                if (hybrid && t >= 1) f.row_flag = ws->row_flag;       // the rows k_refine_finalize left to the fallback (normally none)
                ProfScope prof(c, PROF_FINALIZE);
                k_finalize<<<(unsigned)((nm * rows + 3) / 4), 128, 0, c->stream>>>(f);
                c->launches++;
                NES_CUDA(cudaGetLastError());
                if (vbn_e && f.gather && embed_bn(1)) return 1;
            }
        }
    }
    if (c->rollout_timed) NES_CUDA(cudaEventRecord(c->ev_d, c->stream));
    if (seq)  NES_CUDA(cudaMemcpyAsync(seq, ws->d_seq, sizeof(int64_t) * n_members * rows * T, cudaMemcpyDefault, c->stream));
    if (logp) NES_CUDA(cudaMemcpyAsync(logp, ws->d_lp, sizeof(float) * n_members * rows * T, cudaMemcpyDefault, c->stream));
    return 0;
}

int launch_bound_norms(Ctx* c, int parent) {
    if (!c->is_fc) return 0;
    const int R = c->dims.rnn_size, V1 = c->dims.vocab_size + 1;
    if (parent >= 0) {
        NES_CUDA(cudaMemsetAsync(c->d_theta_rowmax + parent, 0, sizeof(float), c->stream));
        k_rowmax_theta<<<148, 256, 0, c->stream>>>(c->theta32 + (int64_t)parent * c->lay.D + c->lay.t[c->tix[T_LOGIT_W]].start, V1, R, c->d_theta_rowmax + parent);
    } else {
        NES_CUDA(cudaMemsetAsync(c->d_eps_rowmax, 0, sizeof(float), c->stream));
        k_rowmax_eps<<<148 * 8, 256, 0, c->stream>>>(c->table, c->table_len / R, R, c->d_eps_rowmax);
        c->eps_rowmax_valid = true;
    }
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

int launch_bound_norms_batch(Ctx* c, const int32_t* parents, int n) {
    // c->d_ranks still holds the packed (src | dst << 32) items of launch_evolve_batch; the dst slots are the parents to refresh
    if (!c->is_fc || n == 0) return 0;
    const int R = c->dims.rnn_size, V1 = c->dims.vocab_size + 1;
    (void)parents;
    k_zero_slots<<<(n + 255) / 256, 256, 0, c->stream>>>(c->d_ranks, n, c->d_theta_rowmax);
    k_rowmax_theta_batch<<<dim3(37, n), 256, 0, c->stream>>>(c->theta32, c->lay.D, c->lay.t[c->tix[T_LOGIT_W]].start, V1, R, c->d_ranks, c->d_theta_rowmax);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

int decode_hybrid_stats(Ctx* c, int64_t* out4, int reset) {
    DecodeWorkspace* ws = c->ws;
    NES_CHECK(ws != nullptr, "context has no decode workspace");
    NES_CUDA(cudaMemcpyAsync(out4, ws->stats, 32, cudaMemcpyDeviceToHost, c->stream));
    NES_CUDA(cudaStreamSynchronize(c->stream));
    if (reset) NES_CUDA(cudaMemsetAsync(ws->stats, 0, 32, c->stream));
    return 0;
}

const int64_t* decode_resident_seq(Ctx* c, int64_t n_members, int rows, int* T) {
    DecodeWorkspace* ws = c->ws;
    if (!ws || !ws->d_seq || n_members * rows * ws->T > ws->out_cap) return nullptr;
    *T = ws->T;
    return ws->d_seq;
}

const float* decode_resident_logprobs(Ctx* c, int64_t n_members, int rows) {
    DecodeWorkspace* ws = c->ws;
    if (!ws || !ws->d_lp || n_members * rows * ws->T > ws->out_cap) return nullptr;
    return ws->d_lp;
}

int decode_logprob_fitness(Ctx* c, int64_t n_members, int rows, double* out) {
    DecodeWorkspace* ws = c->ws;
    NES_CHECK(ws && ws->d_lp && n_members * rows * ws->T <= ws->out_cap, "no rollout results resident for that many members");
    if (n_members == 0) return 0;
    if (ensure_small(c, n_members)) return 1;
    k_logprob_fitness<<<(unsigned)n_members, 256, 0, c->stream>>>(ws->d_lp, (int64_t)rows * ws->T, c->d_fitness);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    NES_CUDA(cudaMemcpyAsync(out, c->d_fitness, sizeof(double) * n_members, cudaMemcpyDefault, c->stream));
    return 0;
}

}  // namespace nes

// ---------------------------------------------------------------------------------------------
// standalone self test of the tcgen05/TMA pipeline (PLAIN epilogue, one pseudo member)
namespace nes {
__global__ void k_split_scaled(const float* __restrict__ src, __half* __restrict__ hi, __half* __restrict__ lo, int64_t n, int64_t n_pad, float scale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
        __half a = __float2half_rn(0.f), l = a;
        if (i < n && src) {
            if (lo) split_half_dev(src[i], scale, a, l); else a = __float2half_rn(src[i] * scale);
        }
        hi[i] = a; if (lo) lo[i] = l;
    }
}
}  // namespace nes

extern "C" int nes_gemm_selftest(int device, int M, int N, int K, const float* A, const float* W, const float* E, float sigma,
                                 int precision, float* Cout) {
    using namespace nes;
    NES_CHECK(M >= 1 && N >= 1 && K >= 64 && K % 64 == 0, "selftest: need M,N >= 1 and K a multiple of 64");
    NES_CHECK(A && W && Cout, "selftest: null argument");
    NES_CUDA(cudaSetDevice(device));
    const int64_t Mp = (M + 127) / 128 * 128, Np = (N + 127) / 128 * 128;
    float *dA = nullptr, *dW = nullptr, *dE = nullptr, *dC = nullptr;
    __half *a_hi, *a_lo, *w_hi, *w_lo, *e16;
    int64_t* d_off; float* d_sign;
    NES_CUDA(cudaMalloc((void**)&dA, sizeof(float) * M * K)); NES_CUDA(cudaMalloc((void**)&dW, sizeof(float) * N * K));
    NES_CUDA(cudaMalloc((void**)&dC, sizeof(float) * M * N));
    NES_CUDA(cudaMemcpy(dA, A, sizeof(float) * M * K, cudaMemcpyDefault)); NES_CUDA(cudaMemcpy(dW, W, sizeof(float) * N * K, cudaMemcpyDefault));
    if (E) { NES_CUDA(cudaMalloc((void**)&dE, sizeof(float) * N * K)); NES_CUDA(cudaMemcpy(dE, E, sizeof(float) * N * K, cudaMemcpyDefault)); }
    NES_CUDA(cudaMalloc((void**)&a_hi, 2 * Mp * K)); NES_CUDA(cudaMalloc((void**)&a_lo, 2 * Mp * K));
    NES_CUDA(cudaMalloc((void**)&w_hi, 2 * Np * K)); NES_CUDA(cudaMalloc((void**)&w_lo, 2 * Np * K)); NES_CUDA(cudaMalloc((void**)&e16, 2 * Np * K));
    NES_CUDA(cudaMalloc((void**)&d_off, 8)); NES_CUDA(cudaMalloc((void**)&d_sign, 4));
    const int64_t zero = 0; const float one = 1.f;
    NES_CUDA(cudaMemcpy(d_off, &zero, 8, cudaMemcpyHostToDevice)); NES_CUDA(cudaMemcpy(d_sign, &one, 4, cudaMemcpyHostToDevice));
    k_split_scaled<<<256, 256>>>(dA, a_hi, a_lo, (int64_t)M * K, Mp * K, kActScale);
    k_split_scaled<<<256, 256>>>(dW, w_hi, w_lo, (int64_t)N * K, Np * K, kThetaScale);
    k_split_scaled<<<256, 256>>>(dE, e16, nullptr, (int64_t)N * K, Np * K, 1.0f);
    NES_CUDA(cudaGetLastError());
    GemmParams p{};
    const int bk = precision == NES_PREC_EXACT ? GemmCfg<EPI_PLAIN, NES_PREC_EXACT>::BK : GemmCfg<EPI_PLAIN, NES_PREC_FAST>::BK;
    if (make_map(&p.maps[0].a_hi, a_hi, Mp, K, 128, bk) || make_map(&p.maps[0].a_lo, a_lo, Mp, K, 128, bk) || make_map(&p.maps[0].w_hi, w_hi, Np, K, 128, bk) ||
        make_map(&p.maps[0].w_lo, w_lo, Np, K, 128, bk) || make_map(&p.maps[0].eps, e16, Np, K, 128, bk)) return 1;
    p.maps[1] = p.maps[0];
    p.nseg = 1; p.seg[0].K = K; p.seg[0].a_shared = 0; p.seg[0].w_pstart = 0; p.seg[0].b_start = -1; p.seg[0].b_pstart = -1;
    p.rows = M; p.rows_pad = (int)Mp; p.m_tiles = (int)(Mp / 128); p.n_tiles = (int)(Np / 128); p.N = N; p.chunk_stride = 0;
    p.D = 0; p.D_pad = 0; p.off = d_off; p.sign = d_sign; p.parent = nullptr; p.theta32 = nullptr; p.table = nullptr; p.sigma = sigma;
    p.out_f32 = dC; p.ld_out = N;
    Ctx tmp; tmp.stream = 0;
    int rc = launch_gemm_prec<EPI_PLAIN>(&tmp, p, 1, precision);
    if (!rc) { NES_CUDA(cudaDeviceSynchronize()); NES_CUDA(cudaMemcpy(Cout, dC, sizeof(float) * M * N, cudaMemcpyDefault)); }
    void* ptrs[] = {dA, dW, dE, dC, a_hi, a_lo, w_hi, w_lo, e16, d_off, d_sign};
    for (void* q : ptrs) if (q) cudaFree(q);
    return rc;
}
