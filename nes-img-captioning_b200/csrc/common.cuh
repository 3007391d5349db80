// Shared declarations for the B200-native NIC-ES/NIC-NES hot path.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/nes_b200.h"

namespace nes {

// ---------------------------------------------------------------------------------------------
// errors
void set_error(const std::string& msg);
#define NES_CHECK(cond, msg)                                                        \
    do { if (!(cond)) { ::nes::set_error(std::string(msg) + " [" #cond "] at " __FILE__ ":" + std::to_string(__LINE__)); return 1; } } while (0)
#define NES_CUDA(call)                                                              \
    do { cudaError_t e__ = (call); if (e__ != cudaSuccess) {                        \
        ::nes::set_error(std::string("CUDA error: ") + cudaGetErrorString(e__) + " in " #call " at " __FILE__ ":" + std::to_string(__LINE__)); \
        return 1; } } while (0)

// ---------------------------------------------------------------------------------------------
// flat / padded layout of theta and of one noise row (spec: oracle/noise.py header)
constexpr int kMaxTensors = 20;
struct TensorDesc {
    int64_t start;   // flat start (elements)
    int64_t count;   // elements
    int64_t pstart;  // start inside a padded noise row / padded theta shadow (multiple of G)
    int32_t rows;    // [rows, cols] for matrices, rows = count, cols = 1 for vectors
    int32_t cols;
};
struct Layout {
    int n;
    TensorDesc t[kMaxTensors];
    int64_t D, D_pad, G;
};
// fc_caption tensors by ROLE; Ctx::tix[role] is the tensor's position in the flat (named_parameters()) order, -1 if the model has none
enum { T_IMG_W = 0, T_IMG_B, T_EMB_W, T_LOGIT_W, T_LOGIT_B, T_I2H_W, T_I2H_B, T_H2H_W, T_H2H_B,
       T_I2H_LN_W, T_I2H_LN_B, T_H2H_LN_W, T_H2H_LN_B, T_C_LN_W, T_C_LN_B,
       T_IMG_BN_W, T_IMG_BN_B, T_EMB_BN_W, T_EMB_BN_B, T_ROLES };

constexpr int kAbsBlocks = 296;
// the perturbation of one parameter: f32(sigma) * eps, times the safe-mutation scale when one is set (algorithm/nets.py:101-111)
// the perturbation of one parameter: sigma*eps, times the safe-mutation scale (SM-PROPORTIONAL, algorithm/nets.py:107-111) or divided by
// the sensitivity (SM-VECTOR / SM-G-*, :104-105: noise /= sensitivity) -- a true division, as the reference's
__device__ __forceinline__ float apply_scale(float nz, float s, int div) { return div ? __fdiv_rn(nz, s) : __fmul_rn(nz, s); }
__device__ __forceinline__ float mut_noise(float sigma, float eps, const float* __restrict__ scale, int64_t idx, int div) {
    const float nz = __fmul_rn(sigma, eps);
    return scale ? apply_scale(nz, scale[idx], div) : nz;
}

// theta / activation scaling of the split-fp16 operands (exact powers of two)
constexpr float kThetaScale = 256.0f;     // theta_hi + theta_lo = theta * 2^8
constexpr float kActScale   = 32.0f;      // x_hi + x_lo = x * 2^5

// ---------------------------------------------------------------------------------------------
struct DecodeWorkspace;   // decode.cu
struct CiderTables;       // cider.cu

struct Ctx {
    int device = 0;
    cudaStream_t stream = 0;
    bool is_fc = false;
    nes_fc_dims dims{};
    Layout lay{};
    int tix[T_ROLES];                 // role -> index into lay.t (fc_caption contexts)
    uint64_t seed = 0;
    int64_t table_len = 0;
    int n_parents = 1;
    int64_t launches = 0;

    __half* table = nullptr;          // [table_len]
    double* theta64 = nullptr;        // [n_parents, D]  master copy
    float*  theta32 = nullptr;        // [n_parents, D]  evaluation copy (what the reference's workers load)
    __half* theta_hi = nullptr;       // [n_parents, D_pad] padded layout, theta*2^8 split hi/lo
    __half* theta_lo = nullptr;
    uint8_t* theta_is_f32 = nullptr;  // host flags [n_parents]: master theta still has float32 provenance
    uint8_t* theta64_stale = nullptr; // host flags [n_parents]: theta64 not yet refreshed from theta32 (NIC-ES survivors never read it)
    Layout* d_lay = nullptr;          // device copy of lay
    float* d_theta_rowmax = nullptr;  // [n_parents] max row 2-norm of logit.weight (HYBRID error bound)
    float* d_eps_rowmax = nullptr;    // [1] max 2-norm over the rows of the [L/R, R] view of the noise table
    bool eps_rowmax_valid = false;

    // safe mutations (algorithm/nets.py:104-111).  NES_MUT_PROPORTIONAL: noise = sigma*eps (.) s_p with s_p = |theta_p|, exact zeros
    // replaced by mean|theta_p|; s_p is refreshed whenever parent p changes (launch_theta_shadow)
    int mut_kind = 0;
    float* scale32 = nullptr;         // [n_parents, D] (nullptr: default mutation)
    int scale_div = 0;                // 1: scale32 holds sensitivities, the noise is divided by them (NES_MUT_SENSITIVITY)
    float* d_scale_max = nullptr;     // [n_parents] max of s_p (HYBRID error bound)
    uint8_t* scale_stale = nullptr;   // host flags [n_parents]: NES_MUT_SENSITIVITY divisors of this slot predate its current parent
    double* d_abs_partials = nullptr; // [kAbsBlocks]

    // master math
    float*  grad = nullptr;           // [D]
    double* opt_m = nullptr;          // [D]
    double* opt_v = nullptr;          // [D]
    int64_t opt_t = 0;
    int opt_kind = -1;
    double opt_stepsize = 0, opt_b1 = 0, opt_b2 = 0, opt_eps = 0;
    double* d_norm_partials = nullptr; int norm_blocks = 0;
    double* d_norms = nullptr;        // [2]
    double* d_G64 = nullptr;          // [D] staging for a float64 globalg (lazy)

    // scratch for per-iteration small arrays
    int64_t* d_offsets = nullptr; float* d_weights = nullptr; double* d_fitness = nullptr;
    double* d_centered = nullptr; int64_t* d_ranks = nullptr; int64_t small_cap = 0;
    int64_t* d_off_sorted = nullptr; float* d_w_sorted = nullptr;          // K7: the rows in table order (launch_gradient)

    cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr, ev_d = nullptr;
    bool grad_timed = false, rollout_timed = false;

    // optional per-kernel timing (bench.py roofline): event pairs around every decode kernel launch
    bool profiling = false;
    struct ProfRec { int kind; cudaEvent_t a, b; };
    std::vector<ProfRec> prof_recs;        // records of the current profiling window
    std::vector<ProfRec> prof_pool;        // recycled events

    DecodeWorkspace* ws = nullptr;
    CiderTables* cider = nullptr;
};
enum { PROF_PLAIN = 0, PROF_LSTM = 1, PROF_LOGIT = 2, PROF_FINALIZE = 3, PROF_FALLBACK = 4, PROF_MATERIALISE = 5, PROF_KINDS = 6 };
struct ProfScope {                          // RAII: record an event pair around one launch when profiling is on
    Ctx* c; int idx = -1;
    ProfScope(Ctx* c_, int kind) : c(c_) {
        if (!c->profiling) return;
        Ctx::ProfRec r;
        if (!c->prof_pool.empty()) { r = c->prof_pool.back(); c->prof_pool.pop_back(); }
        else { cudaEventCreate(&r.a); cudaEventCreate(&r.b); }
        r.kind = kind;
        cudaEventRecord(r.a, c->stream);
        c->prof_recs.push_back(r); idx = (int)c->prof_recs.size() - 1;
    }
    ~ProfScope() { if (idx >= 0) cudaEventRecord(c->prof_recs[idx].b, c->stream); }
};

int ensure_small(Ctx* c, int64_t n);   // (re)allocate the small per-iteration scratch arrays for n members

// noise.cu
int launch_fill_table(Ctx* c);
int launch_sample_offsets(Ctx* c, uint64_t iteration, int64_t F, int64_t* d_out);
int launch_sample_parents(Ctx* c, uint64_t iteration, int64_t F, int32_t n_parents, int32_t* d_out);
int launch_sample_parents_tournament(Ctx* c, uint64_t iteration, int64_t F, int32_t n_parents, int32_t tournament_size, int32_t* d_out);
int launch_bound_norms(Ctx* c, int parent);       // parent >= 0: theta_rowmax[parent]; parent < 0: eps_rowmax (decode.cu)
int launch_theta_shadow(Ctx* c, int parent);      // theta32 -> theta_hi/lo (padded)
int launch_prop_scale(Ctx* c, int parent);        // theta32 -> scale32 (SM-PROPORTIONAL)
int set_sensitivity(Ctx* c, int parent, const float* sens);   // host or device vector of D divisors (parent < 0: all parents)
int set_mutation(Ctx* c, int kind);
int launch_set_theta(Ctx* c, int parent, const void* d_src, bool is_f64);
int launch_evolve(Ctx* c, int parent, int64_t offset, float sign, float sigma, float* d_member, float* d_noise, int dst_parent);
int refresh_theta64(Ctx* c, int parent);          // theta64 <- (double)theta32 if the slot is stale
int launch_evolve_batch(Ctx* c, int n, const int32_t* src, const int64_t* offsets, const float* sigmas, const int32_t* dst);   // host arrays
int launch_es_select_evolve(Ctx* c, const double* d_fit, const int64_t* d_off, const int32_t* d_par, int64_t n, int n_keep, float sigma,
                            int dst_base, int64_t* d_order);   // device arrays; d_order [n_keep] device
int copy_parent(Ctx* c, int src, int dst);
int launch_bound_norms_batch(Ctx* c, const int32_t* parents, int n);   // theta_rowmax of several parents in one launch (decode.cu)

// master.cu
int launch_ranks(Ctx* c, const double* d_x, int64_t n, int64_t* d_ranks);
int launch_rank_weights(Ctx* c, const double* d_fit, int64_t F, double* d_centered, float* d_w);
constexpr int64_t kStaticRankCap = 2 * NES_STATIC_RANK_MAX_F;      // entries of the static rank scratch (include/nes_b200.h)
int launch_rank_weights_static(Ctx* c, const double* d_fit, int64_t F, double* d_centered, float* d_w);
int launch_gradient(Ctx* c, const int64_t* d_off, const float* d_w, int64_t F, float sigma, float* d_g, int shard = 0, int n_shards = 1);
int launch_opt_update(Ctx* c, const float* d_g, const double* d_G64, int direct, double l2, double* ratio_host_or_null);

// decode.cu
int decode_create(Ctx* c, int wave_members, int rows);
void decode_destroy(Ctx* c);
int decode_rollout(Ctx* c, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents,
                   const float* fc, int rows, int shared_batch, float sigma, int precision,
                   const int64_t* forced, int64_t* seq, float* logp, float* h_trace, float* logit_trace, const double* uniforms = nullptr);
int decode_logprob_fitness(Ctx* c, int64_t n_members, int rows, double* out);
int decode_hybrid_stats(Ctx* c, int64_t* out4, int reset);
const int64_t* decode_resident_seq(Ctx* c, int64_t n_members, int rows, int* T);
const float* decode_resident_logprobs(Ctx* c, int64_t n_members, int rows);   // device seq of the last rollout (nullptr if not resident)

// mnist.cu
int mnist_rollout(Ctx* c, int64_t n_members, const int64_t* d_off, const float* d_sign, const int32_t* d_parent, const float* d_images,
                  const int64_t* d_labels, int rows, int shared_batch, float sigma, float* d_loss, int32_t* d_pred,
                  double* d_fitness, int32_t* d_correct, float* d_logits);

// cider.cu
int cider_set_tables(Ctx* c, const uint64_t* idf_keys, const double* idf_vals, int64_t n_idf, double idf_default,
                     const int32_t* img_ref_ptr, int32_t n_img, const int32_t* ref_ent_ptr, int32_t n_refs,
                     const uint64_t* ent_key, const double* ent_w, const double* ref_norm, const int32_t* ref_len, const double* penalty);
int cider_scores(Ctx* c, const int64_t* seq, int64_t n_members, int rows, int T, int rows_per_img, double* scores);
void cider_destroy(Ctx* c);
int cider_shaped_fitness(Ctx* c, int kind, int64_t n_members, int rows, const double* scores, double* out);

}  // namespace nes

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11) -- identical to oracle/noise.py
namespace nes {
__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
constexpr uint32_t kOffsetDomain = 0x4F464653u;  // "OFFS"
constexpr uint32_t kParentDomain = 0x50415245u;  // "PARE"
constexpr uint32_t kTournDomain  = 0x544F5552u;  // "TOUR"
}  // namespace nes
