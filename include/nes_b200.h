/* nes_b200.h -- C ABI of the B200-native NIC-ES / NIC-NES population-evaluation path.
 *
 * The reference (rubencart/NES-img-captioning, pure Python) has no FFI; its
 * boundary for this path is the Python Policy / PolicyNet / NESMaster /
 * Optimizer objects.  Each entry point below names the reference method it
 * replaces (paths relative to /root/reference/src).  The Python mirror of those
 * classes lives in nes-img-captioning_b200/{nets,policies,master}.py and calls
 * nothing but these symbols (ctypes).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; the message is
 *     available from nes_last_error() (thread-local).
 *   - every pointer argument may be a HOST or a DEVICE pointer (UVA); data is
 *     staged through context-owned device buffers with cudaMemcpyAsync on the
 *     context stream.  The caller owns every buffer it passes; the context keeps
 *     no alias to caller memory after return (contrast: PolicyNet.set_from_vector,
 *     algorithm/nets.py:163-165, aliases the caller's vector).
 *   - one context per device, not thread-safe, all work on one CUDA stream.
 *   - calls are asynchronous w.r.t. the host unless an output pointer is a host
 *     pointer (then the call synchronises the stream before returning).
 */
#ifndef NES_B200_H
#define NES_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nes_ctx nes_ctx;

/* fc_caption dimensions: ModelOptions of algorithm/policies.py:36-41 as used by
 * captioning/nets.py:138-176 (FCModel) and :73-96 (LSTMCore). */
typedef struct nes_fc_dims {
    int32_t vocab_size;            /* V; logits have V+1 columns */
    int32_t input_encoding_size;   /* E */
    int32_t rnn_size;              /* R */
    int32_t fc_feat_size;          /* Fd */
    int32_t seq_length;            /* T = 16 (captioning/nets.py:147) */
    int32_t layer_n;               /* 1: LayerNorm on i2h / h2h / c (captioning/nets.py:93-96); 2: the reference's "virtual batch norm",
                                      BatchNorm1d over the rows of the batch without running statistics (captioning/nets.py:87-91) */
    int32_t layer_n_affine;        /* elementwise affine parameters of either (layer_n_affine / vbn_affine) */
    int32_t vbn_e;                 /* 1: batch norm over the rows of the batch on the outputs of img_embed and embed as well
                                      (captioning/nets.py:156-165: Sequential(layer, BatchNorm1d(track_running_stats=False, affine=vbn_affine)));
                                      vbn_e_affine (the reference's vbn_affine) adds img_embed.1.{weight,bias} and embed.1.{weight,bias} */
    int32_t vbn_e_affine;
} nes_fc_dims;

/* precision of the population GEMMs */
#define NES_PREC_FAST  0   /* fp16 operands (theta_hi, x_hi, eps exact), fp32 accumulate          */
#define NES_PREC_EXACT 1   /* split-fp16 (hi+lo) theta and activations: ~fp32 products, 5 MMAs    */
#define NES_PREC_HYBRID 2  /* EXACT, except the vocabulary projection: fp16-operand logits with a  */
                           /* rigorous error bound + exact fp32 re-evaluation of the arg-max      */
                           /* candidates (same tokens as EXACT; log-probs within ~1e-5)           */

/* optimizer kinds (algorithm/nic_nes/optimizers.py) */
#define NES_OPT_SGD  0
#define NES_OPT_ADAM 1

const char* nes_last_error(void);
int  nes_version(void);

/* ---- context ------------------------------------------------------------------------------- */
/* fc_caption context: flat-theta layout = FCModel.named_parameters() order (captioning/nets.py:150-153; with vbn_e the
 * Sequential wrappers put img_embed.1.* right behind img_embed.0.* and embed.1.* behind embed.0.weight).
 * table_len: number of fp16 elements of the shared noise table (multiple of 4).
 * n_parents: how many theta vectors the context holds (1 for NIC-NES; mu for NIC-ES).
 * wave_members / rows: capacity of the decode workspace (members evaluated per wave, rows per rollout). */
int nes_ctx_create_fc(nes_ctx** out, int device, const nes_fc_dims* dims, uint64_t seed,
                      int64_t table_len, int n_parents, int wave_members, int rows);
/* generic flat context (master-side math only: ranks, gradient, optimizer; no decode), e.g. MnistNet
 * (classification/nets.py) or the config-5 sweep.  counts[n_tensors] in flat order; granule = slot alignment. */
int nes_ctx_create_flat(nes_ctx** out, int device, const int64_t* counts, int n_tensors, int64_t granule,
                        uint64_t seed, int64_t table_len, int n_parents);
int nes_ctx_destroy(nes_ctx* ctx);
int nes_set_stream(nes_ctx* ctx, void* cuda_stream);
int nes_sync(nes_ctx* ctx);
int64_t nes_num_params(const nes_ctx* ctx);   /* D      = PolicyNet.nb_learnable_params (algorithm/nets.py:46-50) */
int64_t nes_padded_params(const nes_ctx* ctx);/* D_pad  (noise-row length in table elements) */
int64_t nes_granule(const nes_ctx* ctx);      /* G */
int64_t nes_table_len(const nes_ctx* ctx);
int64_t nes_gpu_launches(const nes_ctx* ctx); /* kernels launched by this context so far */

/* ---- noise table (replaces the per-worker torch normal_ of algorithm/nets.py:92-102) -------- */
int nes_fill_table(nes_ctx* ctx);                                     /* K1: Philox4x32-10 -> Box-Muller -> fp16 */
int nes_table_slice(nes_ctx* ctx, int64_t start, int64_t count, void* out_f16);
int nes_sample_offsets(nes_ctx* ctx, uint64_t iteration, int64_t F, int64_t* out);
int nes_sample_parents(nes_ctx* ctx, uint64_t iteration, int64_t F, int32_t n_parents, int32_t* out);
/* ESWorker.fitness with selection == 'tournament' (algorithm/nic_es/nic_es_worker.py:150-158): per offspring, tournament_size distinct
 * parent indices drawn without replacement, the lowest one (parents are sorted best first) wins.  The reference draws from an unseeded
 * numpy RandomState; here a uniformly random subset by Floyd's algorithm on the context's Philox stream (spec: oracle/noise.py).
 * 1 <= n_parents <= 4096; tournament_size is clipped to n_parents; iteration < 2^48. */
int nes_sample_parents_tournament(nes_ctx* ctx, uint64_t iteration, int64_t F, int32_t n_parents, int32_t tournament_size, int32_t* out);

/* ---- theta (Policy.set_from_parameter_vector / parameter_vector, algorithm/policies.py:102-111) */
int nes_set_theta(nes_ctx* ctx, int parent, const void* theta, int is_f64);
int nes_get_theta(nes_ctx* ctx, int parent, void* out, int as_f64);
/* PolicyNet.evolve (algorithm/nets.py:83-119) against table row `offset`: noise = f32(sigma)*eps,
 * member = theta + sign*noise.  Either output may be NULL.  If dst_parent >= 0 the member also becomes
 * parent `dst_parent` (NIC-ES survivor materialisation, nic_es_worker.py:188 without the .pth file). */
int nes_evolve(nes_ctx* ctx, int parent, int64_t offset, float sign, float sigma,
               float* member_out, float* noise_out, int dst_parent);

/* n evolve calls with dst_parent set, as one launch (the survivors of a NIC-ES generation, nic_es_master.py:155-167 +
 * nic_es_worker.py:188): parent dst[i] = theta(src[i]) + sigmas[i] * eps(offsets[i]) (sigma 0 copies an elite).  HOST arrays;
 * the destination slots must be distinct and disjoint from the source slots. */
int nes_evolve_batch(nes_ctx* ctx, int n, const int32_t* src_parent, const int64_t* offsets, const float* sigmas,
                     const int32_t* dst_parent);

/* ESMaster.selection (algorithm/nic_es/nic_es_master.py:155-167) + the parent files ESIteration.record_parents keeps
 * (algorithm/nic_es/iteration.py:83-99, written by the workers at nic_es_worker.py:188), without a host round trip: the n fitnesses
 * are ranked like Python's stable `sort(key=fitness, reverse=True)` (equal fitnesses keep their submission order; NaN ranks last) and
 * the n_keep best offspring are re-derived into parent slots dst_base + r <- theta[parents[i_r]] + sigma * eps(offsets[i_r]), r = rank.
 * fitness float64 [n], offsets int64 [n], parents int32 [n] (absolute slot indices; NULL = slot 0): host or device; order_out
 * (nullable) int64 [n_keep] = i_r.  The destination slots must not be read by any member (a wild or overlapping index falls back to a
 * slot outside the destination range instead of faulting). */
int nes_es_select_evolve(nes_ctx* ctx, const double* fitness, const int64_t* offsets, const int32_t* parents, int64_t n, int n_keep,
                         float sigma, int dst_base, int64_t* order_out);
/* parent slot dst <- parent slot src in stream order, with every derived copy (an elite entering the podium is a file copy in the
 * reference: algorithm/tools/podium.py:43-62; so is ESIteration.set_next_elites_to_evaluate, nic_es/iteration.py:113-129) */
int nes_copy_parent(nes_ctx* ctx, int src_parent, int dst_parent);

/* Safe mutations (PolicyNet.evolve, algorithm/nets.py:104-111).  NES_MUT_PROPORTIONAL (the reference's "SM-PROPORTIONAL", shipped in
 * experiments/mscoco_es.json): noise = sigma * eps (.) |theta_p|, exact zeros of theta_p replaced by mean|theta_p|.  Applies to
 * nes_evolve and nes_rollout (HYBRID precision); K7 scales its result by the scale of parent 0.
 * NES_MUT_SENSITIVITY ("SM-VECTOR", and "SM-G-SUM" / "SM-G-ABS" with sensitivities computed by the caller's autograd,
 * algorithm/nets.py:104-105, safe_mutations.py:27-32): noise = sigma * eps / sens_p, a true float32 division like the reference's;
 * sens_p is 1 until nes_set_sensitivity supplies it.
 * NES_MUT_SCALE: noise = sigma * eps (.) s_p with a caller-supplied multiplier s_p (nes_set_sensitivity supplies it; 1 until then).
 * With theta_p = 0 (LayerNorm weights 1), s_p = the Xavier std of each weight (0 for biases) and sigma = 1 a member is a freshly
 * initialised model (PolicyNet.initialize_params, algorithm/nets.py:52-81): the `parent is None` first generation of NIC-ES
 * (nic_es_worker.py:166-169) as noise-table rows.
 * A NES_MUT_SENSITIVITY slot that nes_evolve_batch / nes_es_select_evolve overwrite has its divisors set to NaN until
 * nes_set_sensitivity supplies the new parent's (the reference recomputes them per (task, parent), nic_es_worker.py:172-173). */
enum { NES_MUT_DEFAULT = 0, NES_MUT_PROPORTIONAL = 1, NES_MUT_SENSITIVITY = 2, NES_MUT_SCALE = 3 };
int nes_set_mutation(nes_ctx* ctx, int kind);
/* sens: D positive float32 divisors, host or device, as Sensitivity.set_sensitivity leaves them (clamped below at the underflow value,
 * divided by their minimum); parent < 0 sets every parent slot.  Under NES_MUT_SCALE the vector holds the multipliers instead. */
int nes_set_sensitivity(nes_ctx* ctx, int parent, const float* sens);

/* ---- MNIST classifier population rollout (BASELINE configs[0], experiments/mnist_es.json): replaces, for n_members perturbed
 *      members at once, ESWorker.fitness nic_es_worker.py:142-192 -> ClfPolicy.rollout classification/policies.py:16-48 ->
 *      MnistNet.forward classification/nets.py:25-37 (and ClfPolicy.accuracy_on :50-86 through `correct`) ------------------------
 * The context is a flat one (nes_ctx_create_flat) with MnistNet's six tensors in parameter_vector order: conv1.weight 250,
 * conv1.bias 10, conv2.weight 5000, conv2.bias 20, fc1.weight 3200, fc1.bias 10.  member m = theta[parents[m]] +
 * signs[m]*sigma*eps(offsets[m]) under the mutation set by nes_set_mutation; parents may be NULL (all 0); sigma = 0 evaluates theta.
 * images: float32 [rows, 1, 28, 28] if shared_batch else [n_members, rows, 1, 28, 28]; labels: int64 [rows] / [n_members, rows].
 * fitness: float64 [n_members] = -CrossEntropyLoss (mean over rows); correct (nullable): int32 [n_members] rows whose arg-max equals
 * the label; logits (nullable): float32 [n_members, rows, 10].  Host or device pointers. */
int nes_mnist_rollout(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents,
                      const float* images, const int64_t* labels, int rows, int shared_batch, float sigma,
                      double* fitness, int32_t* correct, float* logits);

/* ---- population rollout (NESWorker.fitness nic_nes_worker.py:131-154 / ESWorker.fitness
 *      nic_es_worker.py:142-175 -> CaptPolicy.rollout captioning/policies.py:86-116 ->
 *      FCModel._sample(greedy) captioning/nets.py:183-245), CIDEr scoring excluded --------------- */
/* member m = theta[parents[m]] + signs[m]*sigma*eps(offsets[m]); parents may be NULL (all 0).
 * fc: float32 [rows, Fd] if shared_batch else [n_members, rows, Fd].
 * seq: int64 [n_members, rows, T]; logprobs: float32 [n_members, rows, T]. */
int nes_rollout(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs,
                const int32_t* parents, const float* fc, int rows, int shared_batch, float sigma,
                int precision, int64_t* seq, float* logprobs);
/* FCModel._sample(fc_feats, greedy=False) (src/captioning/nets.py:210-231) for every member: the token of (member, row, step t) is
 * numpy's legacy choice(p = exp(logprobs) / |.|_1) evaluated at the caller's uniform uniforms[member][row][t-1] in [0, 1) (float64,
 * host or device) -- the reference draws them from the unseeded global generator, one per row per step.  logprobs holds the
 * log-probability of the sampled token.  Runs on the accurate (split-fp16) logits; finished-row bookkeeping as nes_rollout. */
int nes_rollout_sample(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents,
                       const float* fc, int rows, int shared_batch, float sigma, const double* uniforms, int64_t* seq,
                       float* logprobs);

/* test hook: teacher-forced decode of ONE wave with per-step dumps.  forced_tokens int64 [n_members, rows, T]
 * (NULL = greedy feedback); h_trace float32 [T+1, n_members, rows, R]; logit_trace float32 [T, n_members, rows, V+1]
 * (either may be NULL). */
int nes_rollout_trace(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs,
                      const int32_t* parents, const float* fc, int rows, int shared_batch, float sigma,
                      int precision, const int64_t* forced_tokens, int64_t* seq, float* logprobs,
                      float* h_trace, float* logit_trace);
/* sum over time of seq_logprobs per member -> float64 [n_members] (device-side stand-in fitness for benches) */
int nes_logprob_fitness(nes_ctx* ctx, int64_t n_members, int rows, double* out);

/* ---- master math ------------------------------------------------------------------------------ */
/* K6: NESMaster.compute_centered_ranks + weights (nic_nes_master.py:178-179,184-205); stable tie rule.
 * fitness float64 [F,2]; centered float64 [F,2] (nullable); weights float32 [F] (nullable; kept in ctx). */
int nes_rank_weights(nes_ctx* ctx, const double* fitness, int64_t F, double* centered, float* weights);
/* nes_rank_weights for callers that cannot allocate (code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Same arithmetic and kernels as nes_rank_weights,
 * with the int64 rank scratch in a fixed device array of the library (2 * NES_STATIC_RANK_MAX_F entries = 512 KB, held by every GPU
 * that loads the library until exit) instead of the context's growable buffer.  The call allocates and frees nothing, copies
 * nothing and does not synchronise: it enqueues two kernels on the context stream (nes_set_stream) and returns.
 *   fitness  float64 [F,2], DEVICE pointer;  centered float64 [F,2] / weights float32 [F]: DEVICE pointers, either may be NULL.
 *   F > NES_STATIC_RANK_MAX_F (or a host pointer) is refused with an error; the context's resident weights (the ones
 *   nes_gradient(weights = NULL) uses) are NOT updated -- pass `weights` on.
 *   All calls share the one scratch array per GPU: calls that may overlap on the GPU (different streams, two contexts on one device)
 *   race on it; calls in stream order are safe (every entry that is read was written by the same call). */
#define NES_STATIC_RANK_MAX_F 32768
int nes_rank_weights_static(nes_ctx* ctx, const double* fitness, int64_t F, double* centered, float* weights);
/* ranks of a flat float64 vector (compute_ranks, nic_nes_master.py:196-205) -> int64 [n] */
int nes_compute_ranks(nes_ctx* ctx, const double* x, int64_t n, int64_t* ranks);
/* K7: NESMaster.gradient_estimate / batched_weighted_sum (nic_nes_master.py:170-182,207-221):
 * g = f32(sigma) * sum_i w_i eps_i / (2F) over table rows; weights NULL = the ctx weights of the last
 * nes_rank_weights.  g_out float32 [D] nullable (the gradient stays resident in the ctx). */
int nes_gradient(nes_ctx* ctx, const int64_t* offsets, const float* weights, int64_t F, float sigma, float* g_out);

/* K7 sharded over the parameter axis for one process per GPU: shard `shard` of `n_shards` sums ALL F rows for its contiguous slice
 * of elements (same arithmetic as nes_gradient) and leaves zeros elsewhere, so an all-reduce(sum) of the n_shards vectors equals
 * nes_gradient's result bit for bit on every rank (replaces the master-only np.dot of nic_nes_master.py:207-221). */
int nes_gradient_shard(nes_ctx* ctx, const int64_t* offsets, const float* weights, int64_t F, float sigma, int shard, int n_shards,
                       float* g_out);
/* K8: G = -g + l2*theta; Optimizer.update (optimizers.py:15-22) with SGD (:44-47) or Adam (:78-83);
 * float64 state, float64 master theta, float32 + split-fp16 evaluation copies refreshed in the same pass.
 * grad NULL = the resident gradient.  ratio_out float64 (nullable) = |step| / |theta|. */
int nes_opt_init(nes_ctx* ctx, int kind, double stepsize, double momentum_or_beta1, double beta2, double epsilon);
int nes_opt_set_stepsize(nes_ctx* ctx, double stepsize);
int nes_opt_update(nes_ctx* ctx, const float* grad, double l2coeff, double* ratio_out);
/* Optimizer.update(globalg) verbatim (optimizers.py:15-22): globalg float32 or float64 [D]; numpy keeps
 * (1-beta)*globalg in float32 when globalg is float32, and so does this. */
int nes_opt_update_direct(nes_ctx* ctx, const void* globalg, int is_f64, double* ratio_out);
int nes_opt_get_state(nes_ctx* ctx, double* m, double* v, int64_t* t);   /* Optimizer.save_to_file payload */
int nes_opt_set_state(nes_ctx* ctx, const double* m, const double* v, int64_t t);
/* one whole master step: ranks -> weights -> gradient -> update (run_master nic_nes_master.py:123-136) */
int nes_master_step(nes_ctx* ctx, const double* fitness, const int64_t* offsets, int64_t F, float sigma,
                    double l2coeff, double* ratio_out);

/* timing of the last K7 launch on the context stream, CUDA events (ms); bench.py roofline */
int nes_last_gradient_ms(nes_ctx* ctx, float* ms);
int nes_last_rollout_ms(nes_ctx* ctx, float* ms);
/* per-kernel timing of the decode launches (CUDA events on the context stream).  kinds: 0 plain member GEMM
 * (img_embed / LayerNorm pre-activations), 1 LSTM member GEMM, 2 vocabulary-projection member GEMM, 3 finalize / margin /
 * refine, 4 masked split-fp16 fallback of the HYBRID vocabulary projection, 5 materialisation of the wave's member weights.
 * nes_profile_read synchronises, returns the summed durations [6] and launch counts [6] since the last read. */
int nes_profile_enable(nes_ctx* ctx, int on);
int nes_profile_read(nes_ctx* ctx, float* ms_by_kind, int32_t* count_by_kind);
/* HYBRID bookkeeping, counted while profiling is enabled: out4 = {rows refined, candidate logits re-evaluated in fp32,
 * rows whose candidate list overflowed (their unit took the split-fp16 fallback), 0}. */
int nes_hybrid_stats(nes_ctx* ctx, int64_t* out4, int reset);

/* CIDEr-D reward of decoded captions.  Replaces the per-rollout Python scorer call of
 * CaptPolicy.compute_ciders (src/captioning/policies.py:145-193 -> cider/pyciderevalcap/ciderD/ciderD_scorer.py:120-213).
 * nes_cider_set_tables uploads what depends only on the batch (prepared by the host in the reference's float64 operation
 * order, nes-img-captioning_b200/cider.py): idf[ngram] = ref_len - log(max(1, df)) for the n-grams with df > 1 (packed keys:
 * n | t0<<3 | t1<<18 | t2<<33 | t3<<48, tokens < 32768; absent n-grams take idf_default), and per image its reference
 * captions as CSR lists of (key, tf*idf) in dict order with their four norms and length; penalty[d + 64] =
 * e**(-(d*d)/(2 sigma^2)) for d = len(hyp) - len(ref).  All pointers host memory.
 * nes_cider_scores scores n_members x rows captions (seq int64 [n_members, rows, T] host or device; NULL = the tokens of
 * the last nes_rollout, resident on the device); row r is scored against image r / rows_per_img; scores float64
 * [n_members, rows] = the per-caption `scores` array of CiderD.compute_score, bit-identical. */
int nes_cider_set_tables(nes_ctx* ctx, const uint64_t* idf_keys, const double* idf_vals, int64_t n_idf, double idf_default,
                         const int32_t* img_ref_ptr, int32_t n_img, const int32_t* ref_ent_ptr, int32_t n_refs,
                         const uint64_t* ent_key, const double* ent_w, const double* ref_norm, const int32_t* ref_len,
                         const double* penalty);
int nes_cider_scores(nes_ctx* ctx, const int64_t* seq, int64_t n_members, int rows, int T, int rows_per_img, double* scores);

/* Log-prob-shaped fitness criteria on the resident rollout results (src/captioning/fitness.py:43-132 called at
 * src/captioning/policies.py:119-123): out[m] = sum f(logprob) * reward * mask / sum mask over the member's rows x T.
 * scores float64 [n_members, rows] = per-caption CIDEr-D (host or device; NULL = the result of the last nes_cider_scores). */
enum { NES_FIT_SC_LOSS = 0,          /* LogFitnessCriterion     -logprob                                  */
       NES_FIT_GREEDY_LOGPROB = 1,   /* AltLogFitnessCriterion  log10(p + 1/9) + log10(9)                 */
       NES_FIT_GREEDY_AVGPROB = 2,   /* AvgLogFitnessCriterion  0.5 + 0.5 * (log10(p + 1/9) + log10(9))   */
       NES_FIT_GREEDY_EXPPROB = 3,   /* ExpFitnessCriterion     (exp(p) - 1) / (e - 1)                    */
       NES_FIT_GREEDY_LINPROB = 4 }; /* LinFitnessCriterion     p                                         */
int nes_shaped_fitness(nes_ctx* ctx, int kind, int64_t n_members, int rows, const double* scores, double* out);

/* standalone tensor-core self test through the same tcgen05/TMA kernel the decode uses:
 * C[M,N] = A[M,K] W[N,K]^T + sigma * A[M,K] f16(E[N,K])^T, float32 in/out (host or device pointers). */
int nes_gemm_selftest(int device, int M, int N, int K, const float* A, const float* W, const float* E,
                      float sigma, int precision, float* C);

#ifdef __cplusplus
}
#endif
#endif
