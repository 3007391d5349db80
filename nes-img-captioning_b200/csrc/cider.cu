// CIDEr-D reward of decoded captions on the device (the reference's training fitness).
//
// Reference: CaptPolicy.compute_ciders (src/captioning/policies.py:145-193) -> CiderD.compute_score ->
// CiderScorer.compute_cider (cider/pyciderevalcap/ciderD/ciderD_scorer.py:120-199), fed with array_to_str strings
// (src/algorithm/tools/utils.py:34-40: a caption is the tokens up to and including the first 0).
//
// Everything that depends only on the batch (reference captions, document frequencies) is prepared once per batch by the
// host (nes-img-captioning_b200/cider.py, same float64 operation order as the reference) and uploaded: an idf hash table
// keyed by packed n-grams, and per reference caption its tf-idf entries, norms and length.  The per-member work -- 2P x
// rows hypotheses per iteration, the Python-dict hot spot of the reference -- is k_cider: one warp per caption, float64,
// every sum in the reference's order (n-gram length, then position; references in order), so scores are bit-identical.
#include "common.cuh"

namespace nes {

constexpr int kCiderMaxT = 32;                       // longest hypothesis (tokens)
constexpr int kCiderSlots = 4 * kCiderMaxT;          // (k, i) n-gram slots of one hypothesis
constexpr int kCiderPenalty = 128;                   // penalty[delta + 64]

struct CiderTables {
    uint64_t* idf_keys = nullptr; double* idf_vals = nullptr; int64_t idf_cap = 0; double idf_default = 0.0;
    int32_t* img_ref_ptr = nullptr; int32_t n_img = 0;
    int32_t* ref_ent_ptr = nullptr; int32_t n_refs = 0;
    uint64_t* ent_key = nullptr; double* ent_w = nullptr; double* ref_norm = nullptr; int32_t* ref_len = nullptr;
    double* penalty = nullptr;
    double* d_scores = nullptr; int64_t scores_cap = 0;
    int64_t* d_seq = nullptr; int64_t seq_cap = 0;
};

__host__ __device__ inline uint64_t mix64(uint64_t x) {      // splitmix64 finaliser
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31; return x;
}

__global__ void k_cider_hash_insert(const uint64_t* __restrict__ keys, const double* __restrict__ vals, int64_t n,
                                    uint64_t* __restrict__ tk, double* __restrict__ tv, int64_t cap) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t key = keys[i];
        int64_t h = (int64_t)(mix64(key) & (uint64_t)(cap - 1));
        for (;;) {
            const unsigned long long prev = atomicCAS(reinterpret_cast<unsigned long long*>(tk + h), 0ull, (unsigned long long)key);
            if (prev == 0ull || prev == key) { tv[h] = vals[i]; break; }
            h = (h + 1) & (cap - 1);
        }
    }
}

struct CiderParams {
    const int64_t* seq; int64_t n_members; int rows, T, rows_per_img;
    const uint64_t* idf_keys; const double* idf_vals; int64_t idf_cap; double idf_default;
    const int32_t* img_ref_ptr; const int32_t* ref_ent_ptr; const uint64_t* ent_key; const double* ent_w;
    const double* ref_norm; const int32_t* ref_len; const double* penalty;
    double* scores;
};

// one warp per caption (member, row)
__global__ void __launch_bounds__(128) k_cider(const CiderParams p) {
    __shared__ uint64_t s_key[4][kCiderSlots];
    __shared__ double s_w[4][kCiderSlots];
    __shared__ int s_tok[4][kCiderMaxT];
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    const int64_t wid = (int64_t)blockIdx.x * 4 + wq;
    if (wid >= p.n_members * p.rows) return;
    const int r = (int)(wid % p.rows);
    const int64_t* sq = p.seq + wid * p.T;
    // ---- caption = tokens up to and including the first 0 (array_to_str)
    int tok = lane < p.T ? (int)sq[lane] : -1;
    const unsigned zeros = __ballot_sync(0xffffffffu, lane < p.T && tok == 0);
    const int L = zeros ? __ffs(zeros) : p.T;                     // number of words
    s_tok[wq][lane] = tok;
    __syncwarp();
    // ---- n-gram slots in the reference's dict order: k = 1..4 outer, position inner.  key 0 = not a first occurrence.
    uint64_t* key = s_key[wq]; double* w = s_w[wq];
    for (int s = lane; s < kCiderSlots; s += 32) {
        const int k = s / kCiderMaxT + 1, i = s % kCiderMaxT;
        uint64_t kk = 0; double ww = 0.0;
        if (i + k <= L) {
            kk = (uint64_t)k;
            for (int j = 0; j < k; ++j) kk |= (uint64_t)(s_tok[wq][i + j] & 0x7fff) << (3 + 15 * j);
            int tf = 0; bool first = true;
            for (int i2 = 0; i2 + k <= L; ++i2) {
                bool same = true;
                for (int j = 0; j < k; ++j) same = same && s_tok[wq][i2 + j] == s_tok[wq][i + j];
                if (same) { ++tf; if (i2 < i) first = false; }
            }
            if (first) {
                double idf = p.idf_default;
                if (p.idf_cap) {
                    int64_t h = (int64_t)(mix64(kk) & (uint64_t)(p.idf_cap - 1));
                    for (;;) {
                        const uint64_t t = p.idf_keys[h];
                        if (t == kk) { idf = p.idf_vals[h]; break; }
                        if (t == 0) break;
                        h = (h + 1) & (p.idf_cap - 1);
                    }
                }
                ww = (double)tf * idf;                             // float(term_freq) * (ref_len - df)
            } else kk = 0;
        }
        key[s] = kk; w[s] = ww;
    }
    __syncwarp();
    // ---- hypothesis norms (lane k-1 sums its n-gram length in position order) and length (= number of bigrams)
    double norm_h = 0.0;
    if (lane < 4) {
        for (int i = 0; i < kCiderMaxT; ++i) if (key[lane * kCiderMaxT + i]) norm_h = __dadd_rn(norm_h, __dmul_rn(w[lane * kCiderMaxT + i], w[lane * kCiderMaxT + i]));   // no FMA: pow(x, 2) then +=
        norm_h = sqrt(norm_h);
    }
    const int len_h = L >= 2 ? L - 1 : 0;
    // ---- similarity against every reference of the image: lane = (ref j within a group of 8, n-gram length k)
    const int img = r / p.rows_per_img;
    const int ref0 = p.img_ref_ptr[img], nref = p.img_ref_ptr[img + 1] - ref0;
    const int k = lane & 3, jj = lane >> 2;
    const double nh = __shfl_sync(0xffffffffu, norm_h, k);
    double score = 0.0;                                            // lanes 0..3: running sum over references, in order
    for (int jb = 0; jb < nref; jb += 8) {
        double val = 0.0;
        const int j = jb + jj;
        if (j < nref) {
            const int ref = ref0 + j;
            const int e0 = p.ref_ent_ptr[ref], e1 = p.ref_ent_ptr[ref + 1];
            for (int i = 0; i < kCiderMaxT; ++i) {
                const uint64_t hk = key[k * kCiderMaxT + i];
                if (!hk) continue;
                double rw = 0.0;
                for (int e = e0; e < e1; ++e) if (p.ent_key[e] == hk) { rw = p.ent_w[e]; break; }
                val = __dadd_rn(val, __dmul_rn(fmin(w[k * kCiderMaxT + i], rw), rw));            // no FMA contraction
            }
            const double nr = p.ref_norm[ref * 4 + k];
            if (nh != 0.0 && nr != 0.0) val /= (nh * nr);
            int d = len_h - p.ref_len[ref] + 64;
            d = d < 0 ? 0 : (d > kCiderPenalty - 1 ? kCiderPenalty - 1 : d);
            val *= p.penalty[d];
        }
        for (int q = 0; q < 8 && jb + q < nref; ++q) {             // score += sim(...), reference by reference
            const double v = __shfl_sync(0xffffffffu, val, q * 4 + k);
            if (lane < 4) score += v;
        }
    }
    // ---- np.mean over the four n-gram lengths, / number of references, * 10
    const double s1 = __shfl_sync(0xffffffffu, score, 1), s2 = __shfl_sync(0xffffffffu, score, 2), s3 = __shfl_sync(0xffffffffu, score, 3);
    if (lane == 0) {
        double m = (((score + s1) + s2) + s3) / 4.0;
        m /= (double)nref;
        m *= 10.0;
        p.scores[wid] = m;
    }
}

// ---------------------------------------------------------------------------------------------
// log-prob-shaped fitness criteria (src/captioning/fitness.py:43-132 as called at captioning/policies.py:119-123):
//   out = sum_{r,t} f(logprob[r,t]) * reward[r] * mask[r,t] / sum mask,   mask[r,0] = 1, mask[r,t] = seq[r,t-1] > 0
// element values in float32 like the reference's torch ops, the sum in float64 (the reference sums in float32: results agree
// to float32 round-off, tested at 1e-5 relative).  One block per member.
__global__ void __launch_bounds__(256) k_shaped_fitness(int kind, const float* __restrict__ lp, const int64_t* __restrict__ seq,
                                                        const double* __restrict__ scores, int rows, int T, double* __restrict__ out) {
    __shared__ double sh[2][8];
    const int64_t m = blockIdx.x;
    const float* l = lp + m * rows * T;
    const int64_t* s = seq + m * rows * T;
    const float log10_9 = (float)0.95424250943932487;               // np.log10(10 - 1) added to a float32 tensor
    double acc = 0.0, msum = 0.0;
    for (int i = threadIdx.x; i < rows * T; i += 256) {
        const int r = i / T, t = i - r * T;
        const float mask = (t == 0 || s[i - 1] > 0) ? 1.f : 0.f;
        const float reward = (float)scores[m * rows + r];
        const float x = l[i];
        float v;
        if (kind == NES_FIT_SC_LOSS) v = __fmul_rn(__fmul_rn(-x, reward), mask);
        else if (kind == NES_FIT_GREEDY_LOGPROB || kind == NES_FIT_GREEDY_AVGPROB) {
            const float pf = __fadd_rn(log10f(__fadd_rn(expf(x), (float)(1.0 / 9.0))), log10_9);
            if (kind == NES_FIT_GREEDY_LOGPROB) v = __fmul_rn(__fmul_rn(pf, reward), mask);
            else v = __fadd_rn(__fmul_rn(__fmul_rn(0.5f, reward), mask), __fmul_rn(__fmul_rn(__fmul_rn(0.5f, pf), reward), mask));
        } else if (kind == NES_FIT_GREEDY_EXPPROB) v = __fmul_rn(__fmul_rn(__fdiv_rn(__fsub_rn(expf(expf(x)), 1.f), (float)(2.718281828459045 - 1.0)), reward), mask);
        else v = __fmul_rn(__fmul_rn(expf(x), reward), mask);          // NES_FIT_GREEDY_LINPROB
        acc += (double)v; msum += (double)mask;
    }
    for (int o = 16; o > 0; o >>= 1) { acc += __shfl_xor_sync(0xffffffffu, acc, o); msum += __shfl_xor_sync(0xffffffffu, msum, o); }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = acc; sh[1][threadIdx.x >> 5] = msum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0;
        for (int q = 0; q < 8; ++q) { a += sh[0][q]; b += sh[1][q]; }
        out[m] = (double)((float)a / (float)b);
    }
}

int cider_shaped_fitness(Ctx* c, int kind, int64_t n_members, int rows, const double* scores, double* out) {
    CiderTables* t = c->cider;
    NES_CHECK(kind >= NES_FIT_SC_LOSS && kind <= NES_FIT_GREEDY_LINPROB, "unknown fitness kind");
    NES_CHECK(n_members >= 0 && out, "bad arguments");
    if (n_members == 0) return 0;
    int T = 0;
    const int64_t* d_seq = decode_resident_seq(c, n_members, rows, &T);
    const float* d_lp = decode_resident_logprobs(c, n_members, rows);
    NES_CHECK(d_seq && d_lp, "no rollout results resident for that many members");
    const double* d_scores = nullptr;
    double* tmp = nullptr;
    if (scores) {
        cudaPointerAttributes at{};
        const bool on_dev = cudaPointerGetAttributes(&at, scores) == cudaSuccess && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged);
        cudaGetLastError();
        if (on_dev) d_scores = scores;
        else {
            NES_CUDA(cudaMallocAsync((void**)&tmp, sizeof(double) * n_members * rows, c->stream));
            NES_CUDA(cudaMemcpyAsync(tmp, scores, sizeof(double) * n_members * rows, cudaMemcpyDefault, c->stream));
            d_scores = tmp;
        }
    } else {
        NES_CHECK(t && t->d_scores && t->scores_cap >= n_members * rows, "no CIDEr-D scores resident: call nes_cider_scores first or pass scores");
        d_scores = t->d_scores;
    }
    if (ensure_small(c, n_members)) return 1;
    k_shaped_fitness<<<(unsigned)n_members, 256, 0, c->stream>>>(kind, d_lp, d_seq, d_scores, rows, T, c->d_fitness);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    if (tmp) NES_CUDA(cudaFreeAsync(tmp, c->stream));
    NES_CUDA(cudaMemcpyAsync(out, c->d_fitness, sizeof(double) * n_members, cudaMemcpyDefault, c->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
static void cider_free(CiderTables* t) {
    void* ptrs[] = {t->idf_keys, t->idf_vals, t->img_ref_ptr, t->ref_ent_ptr, t->ent_key, t->ent_w, t->ref_norm, t->ref_len, t->penalty};
    for (void* q : ptrs) if (q) cudaFree(q);
    t->idf_keys = nullptr; t->idf_vals = nullptr; t->img_ref_ptr = nullptr; t->ref_ent_ptr = nullptr; t->ent_key = nullptr; t->ent_w = nullptr;
    t->ref_norm = nullptr; t->ref_len = nullptr; t->penalty = nullptr;
}
void cider_destroy(Ctx* c) {
    if (!c->cider) return;
    cider_free(c->cider);
    if (c->cider->d_scores) cudaFree(c->cider->d_scores);
    if (c->cider->d_seq) cudaFree(c->cider->d_seq);
    delete c->cider;
    c->cider = nullptr;
}
template <typename T> static int upload(T** dst, const T* src, int64_t n, cudaStream_t st) {
    NES_CUDA(cudaMalloc((void**)dst, sizeof(T) * (size_t)(n > 0 ? n : 1)));
    if (n > 0) NES_CUDA(cudaMemcpyAsync(*dst, src, sizeof(T) * n, cudaMemcpyDefault, st));
    return 0;
}

int cider_set_tables(Ctx* c, const uint64_t* idf_keys, const double* idf_vals, int64_t n_idf, double idf_default,
                     const int32_t* img_ref_ptr, int32_t n_img, const int32_t* ref_ent_ptr, int32_t n_refs,
                     const uint64_t* ent_key, const double* ent_w, const double* ref_norm, const int32_t* ref_len, const double* penalty) {
    NES_CHECK(n_idf >= 0 && n_img >= 1 && n_refs >= 1 && img_ref_ptr && ref_ent_ptr && ref_norm && ref_len && penalty, "bad CIDEr tables");
    if (!c->cider) c->cider = new CiderTables();
    CiderTables* t = c->cider;
    NES_CUDA(cudaStreamSynchronize(c->stream));
    cider_free(t);
    t->idf_default = idf_default; t->n_img = n_img; t->n_refs = n_refs;
    t->idf_cap = 0;
    if (n_idf > 0) {
        int64_t cap = 64; while (cap < 2 * n_idf) cap <<= 1;
        t->idf_cap = cap;
        NES_CUDA(cudaMalloc((void**)&t->idf_keys, sizeof(uint64_t) * cap));
        NES_CUDA(cudaMalloc((void**)&t->idf_vals, sizeof(double) * cap));
        NES_CUDA(cudaMemsetAsync(t->idf_keys, 0, sizeof(uint64_t) * cap, c->stream));
        uint64_t* dk = nullptr; double* dv = nullptr;
        if (upload(&dk, idf_keys, n_idf, c->stream) || upload(&dv, idf_vals, n_idf, c->stream)) return 1;
        k_cider_hash_insert<<<(unsigned)((n_idf + 255) / 256 < 4096 ? (n_idf + 255) / 256 : 4096), 256, 0, c->stream>>>(dk, dv, n_idf, t->idf_keys, t->idf_vals, cap);
        c->launches++;
        NES_CUDA(cudaGetLastError());
        NES_CUDA(cudaStreamSynchronize(c->stream));
        cudaFree(dk); cudaFree(dv);
    }
    int32_t n_ent = 0;
    NES_CUDA(cudaMemcpy(&n_ent, ref_ent_ptr + n_refs, sizeof(int32_t), cudaMemcpyDefault));
    if (upload(&t->img_ref_ptr, img_ref_ptr, (int64_t)n_img + 1, c->stream) || upload(&t->ref_ent_ptr, ref_ent_ptr, (int64_t)n_refs + 1, c->stream) ||
        upload(&t->ent_key, ent_key, n_ent, c->stream) || upload(&t->ent_w, ent_w, n_ent, c->stream) ||
        upload(&t->ref_norm, ref_norm, (int64_t)n_refs * 4, c->stream) || upload(&t->ref_len, ref_len, n_refs, c->stream) ||
        upload(&t->penalty, penalty, kCiderPenalty, c->stream)) return 1;
    NES_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int cider_scores(Ctx* c, const int64_t* seq, int64_t n_members, int rows, int T, int rows_per_img, double* scores) {
    CiderTables* t = c->cider;
    NES_CHECK(t && t->img_ref_ptr, "nes_cider_set_tables has not been called");
    NES_CHECK(T >= 1 && T <= kCiderMaxT, "caption length must be 1..32 tokens");
    NES_CHECK(rows >= 1 && rows_per_img >= 1 && (rows + rows_per_img - 1) / rows_per_img <= t->n_img, "more rows than reference images");
    NES_CHECK(n_members >= 0 && scores, "bad arguments");
    if (n_members == 0) return 0;
    const int64_t n = n_members * rows;
    if (n > t->scores_cap) {
        if (t->d_scores) cudaFree(t->d_scores);
        NES_CUDA(cudaMalloc((void**)&t->d_scores, sizeof(double) * n));
        t->scores_cap = n;
    }
    const int64_t* d_seq = nullptr;
    if (seq) {
        cudaPointerAttributes at{};
        const bool on_dev = cudaPointerGetAttributes(&at, seq) == cudaSuccess && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged);
        cudaGetLastError();
        if (on_dev) d_seq = seq;
        else {
            if (n * T > t->seq_cap) {
                if (t->d_seq) cudaFree(t->d_seq);
                NES_CUDA(cudaMalloc((void**)&t->d_seq, sizeof(int64_t) * n * T));
                t->seq_cap = n * T;
            }
            NES_CUDA(cudaMemcpyAsync(t->d_seq, seq, sizeof(int64_t) * n * T, cudaMemcpyDefault, c->stream));
            d_seq = t->d_seq;
        }
    } else {
        int Tres = 0;
        d_seq = decode_resident_seq(c, n_members, rows, &Tres);
        NES_CHECK(d_seq != nullptr, "no rollout results resident for that many members");
        NES_CHECK(Tres == T, "T differs from the context's seq_length");
    }
    CiderParams p{};
    p.seq = d_seq; p.n_members = n_members; p.rows = rows; p.T = T; p.rows_per_img = rows_per_img;
    p.idf_keys = t->idf_keys; p.idf_vals = t->idf_vals; p.idf_cap = t->idf_cap; p.idf_default = t->idf_default;
    p.img_ref_ptr = t->img_ref_ptr; p.ref_ent_ptr = t->ref_ent_ptr; p.ent_key = t->ent_key; p.ent_w = t->ent_w;
    p.ref_norm = t->ref_norm; p.ref_len = t->ref_len; p.penalty = t->penalty; p.scores = t->d_scores;
    k_cider<<<(unsigned)((n + 3) / 4), 128, 0, c->stream>>>(p);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    NES_CUDA(cudaMemcpyAsync(scores, t->d_scores, sizeof(double) * n, cudaMemcpyDefault, c->stream));
    return 0;
}

}  // namespace nes
