// extern "C" surface declared in include/nes_b200.h.
#include "common.cuh"
#include <cstring>
#include <numeric>

namespace nes {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }

static int64_t lcm64(int64_t a, int64_t b) { return a / std::gcd(a, b) * b; }

static bool is_device_ptr(const void* p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
// an output that lands in host memory must be complete when the call returns
static int finish_output(Ctx* c, const void* out) {
    if (out && !is_device_ptr(out)) NES_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

static void build_layout(Layout& L, const int64_t* counts, const int32_t* rows, const int32_t* cols, int n, int64_t G) {
    L.n = n; L.G = G;
    int64_t pos = 0, p = 0;
    for (int k = 0; k < n; ++k) {
        L.t[k].start = pos; L.t[k].count = counts[k]; L.t[k].pstart = p;
        L.t[k].rows = rows ? rows[k] : (int32_t)counts[k]; L.t[k].cols = cols ? cols[k] : 1;
        pos += counts[k];
        p = (p + counts[k] + G - 1) / G * G;
    }
    L.D = pos; L.D_pad = p;
}

int ensure_small(Ctx* c, int64_t n) {
    if (n <= c->small_cap) return 0;
    void* olds[] = {c->d_offsets, c->d_weights, c->d_fitness, c->d_centered, c->d_ranks, c->d_off_sorted, c->d_w_sorted};
    for (void* o : olds) if (o) cudaFree(o);
    int64_t cap = n < 1024 ? 1024 : n;
    NES_CUDA(cudaMalloc((void**)&c->d_offsets, sizeof(int64_t) * cap));
    NES_CUDA(cudaMalloc((void**)&c->d_weights, sizeof(float) * cap));
    NES_CUDA(cudaMalloc((void**)&c->d_fitness, sizeof(double) * 2 * cap));
    NES_CUDA(cudaMalloc((void**)&c->d_centered, sizeof(double) * 2 * cap));
    NES_CUDA(cudaMalloc((void**)&c->d_ranks, sizeof(int64_t) * 2 * cap));
    NES_CUDA(cudaMalloc((void**)&c->d_off_sorted, sizeof(int64_t) * cap));
    NES_CUDA(cudaMalloc((void**)&c->d_w_sorted, sizeof(float) * cap));
    c->small_cap = cap;
    return 0;
}

static int ctx_common_init(Ctx* c, int device, uint64_t seed, int64_t table_len, int n_parents) {
    NES_CHECK(table_len > 0 && table_len % 4 == 0, "table_len must be a positive multiple of 4");
    NES_CHECK(n_parents >= 1, "n_parents must be >= 1");
    NES_CHECK(c->lay.D_pad <= table_len, "noise table shorter than one padded noise row");
    int n_dev = 0;
    NES_CUDA(cudaGetDeviceCount(&n_dev));
    NES_CHECK(device >= 0 && device < n_dev, "no such CUDA device (this library has no CPU fallback)");
    NES_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    NES_CUDA(cudaGetDeviceProperties(&prop, device));
    NES_CHECK(prop.major == 10, "this library is built for sm_100a (B200) only");
    c->device = device; c->seed = seed; c->table_len = table_len; c->n_parents = n_parents;
    const int64_t D = c->lay.D, Dp = c->lay.D_pad;
    NES_CUDA(cudaMalloc((void**)&c->table, sizeof(__half) * table_len));
    NES_CUDA(cudaMalloc((void**)&c->theta64, sizeof(double) * n_parents * D));
    NES_CUDA(cudaMalloc((void**)&c->theta32, sizeof(float) * n_parents * D));
    NES_CUDA(cudaMalloc((void**)&c->theta_hi, sizeof(__half) * n_parents * Dp));
    NES_CUDA(cudaMalloc((void**)&c->theta_lo, sizeof(__half) * n_parents * Dp));
    NES_CUDA(cudaMalloc((void**)&c->grad, sizeof(float) * D));
    NES_CUDA(cudaMalloc((void**)&c->opt_m, sizeof(double) * D));
    NES_CUDA(cudaMalloc((void**)&c->opt_v, sizeof(double) * D));
    NES_CUDA(cudaMalloc((void**)&c->d_lay, sizeof(Layout)));
    NES_CUDA(cudaMalloc((void**)&c->d_theta_rowmax, sizeof(float) * n_parents));
    NES_CUDA(cudaMalloc((void**)&c->d_eps_rowmax, sizeof(float)));
    NES_CUDA(cudaMemset(c->d_theta_rowmax, 0, sizeof(float) * n_parents));
    NES_CUDA(cudaMemset(c->d_eps_rowmax, 0, sizeof(float)));
    c->norm_blocks = 148 * 8;
    NES_CUDA(cudaMalloc((void**)&c->d_norm_partials, sizeof(double) * 2 * c->norm_blocks));
    NES_CUDA(cudaMalloc((void**)&c->d_norms, sizeof(double) * 2));
    NES_CUDA(cudaMemcpy(c->d_lay, &c->lay, sizeof(Layout), cudaMemcpyHostToDevice));
    NES_CUDA(cudaMemset(c->theta64, 0, sizeof(double) * n_parents * D));
    NES_CUDA(cudaMemset(c->theta32, 0, sizeof(float) * n_parents * D));
    NES_CUDA(cudaMemset(c->theta_hi, 0, sizeof(__half) * n_parents * Dp));
    NES_CUDA(cudaMemset(c->theta_lo, 0, sizeof(__half) * n_parents * Dp));
    NES_CUDA(cudaMemset(c->grad, 0, sizeof(float) * D));
    NES_CUDA(cudaMemset(c->opt_m, 0, sizeof(double) * D));
    NES_CUDA(cudaMemset(c->opt_v, 0, sizeof(double) * D));
    c->theta_is_f32 = new uint8_t[n_parents];
    for (int i = 0; i < n_parents; ++i) c->theta_is_f32[i] = 1;
    c->theta64_stale = new uint8_t[n_parents];
    for (int i = 0; i < n_parents; ++i) c->theta64_stale[i] = 0;
    c->scale_stale = new uint8_t[n_parents];
    for (int i = 0; i < n_parents; ++i) c->scale_stale[i] = 0;
    NES_CUDA(cudaEventCreate(&c->ev_a)); NES_CUDA(cudaEventCreate(&c->ev_b));
    NES_CUDA(cudaEventCreate(&c->ev_c)); NES_CUDA(cudaEventCreate(&c->ev_d));
    return ensure_small(c, 1024);
}

}  // namespace nes

using namespace nes;
#define CTX(c) reinterpret_cast<Ctx*>(c)
#define NEED_CTX(c) NES_CHECK((c) != nullptr, "null context")

extern "C" {

const char* nes_last_error(void) { return g_err.c_str(); }
int nes_version(void) { return 100; }

int nes_ctx_create_fc(nes_ctx** out, int device, const nes_fc_dims* dims, uint64_t seed, int64_t table_len, int n_parents,
                      int wave_members, int rows) {
    NES_CHECK(out && dims, "null argument");
    NES_CHECK(dims->vocab_size >= 1 && dims->input_encoding_size >= 1 && dims->rnn_size >= 1 && dims->fc_feat_size >= 1 && dims->seq_length >= 1,
              "bad fc_caption dimensions");
    Ctx* c = new Ctx();
    c->is_fc = true; c->dims = *dims;
    const int64_t V1 = dims->vocab_size + 1, E = dims->input_encoding_size, R = dims->rnn_size, Fd = dims->fc_feat_size;
    int64_t counts[kMaxTensors]; int32_t rws[kMaxTensors], cls[kMaxTensors]; int n = 0;
    for (int r = 0; r < T_ROLES; ++r) c->tix[r] = -1;
    auto add = [&](int role, int64_t r, int64_t cl) { counts[n] = r * cl; rws[n] = (int32_t)r; cls[n] = (int32_t)cl; c->tix[role] = n; ++n; };
    const bool ebn_affine = dims->vbn_e && dims->vbn_e_affine;
    add(T_IMG_W, E, Fd); add(T_IMG_B, E, 1);
    if (ebn_affine) { add(T_IMG_BN_W, E, 1); add(T_IMG_BN_B, E, 1); }
    add(T_EMB_W, V1, E);
    if (ebn_affine) { add(T_EMB_BN_W, E, 1); add(T_EMB_BN_B, E, 1); }
    add(T_LOGIT_W, V1, R); add(T_LOGIT_B, V1, 1); add(T_I2H_W, 5 * R, E); add(T_I2H_B, 5 * R, 1); add(T_H2H_W, 5 * R, R); add(T_H2H_B, 5 * R, 1);
    if (dims->layer_n && dims->layer_n_affine) {
        add(T_I2H_LN_W, 5 * R, 1); add(T_I2H_LN_B, 5 * R, 1); add(T_H2H_LN_W, 5 * R, 1); add(T_H2H_LN_B, 5 * R, 1); add(T_C_LN_W, R, 1); add(T_C_LN_B, R, 1);
    }
    build_layout(c->lay, counts, rws, cls, n, lcm64(lcm64(Fd, E), R));
    if (ctx_common_init(c, device, seed, table_len, n_parents)) { delete c; return 1; }
    if (wave_members > 0 && decode_create(c, wave_members, rows)) { nes_ctx_destroy(reinterpret_cast<nes_ctx*>(c)); return 1; }
    *out = reinterpret_cast<nes_ctx*>(c);
    return 0;
}

int nes_ctx_create_flat(nes_ctx** out, int device, const int64_t* counts, int n_tensors, int64_t granule, uint64_t seed,
                        int64_t table_len, int n_parents) {
    NES_CHECK(out && counts, "null argument");
    NES_CHECK(n_tensors >= 1 && n_tensors <= kMaxTensors, "between 1 and 16 tensors");
    NES_CHECK(granule >= 8 && granule % 8 == 0, "granule must be a positive multiple of 8 elements (128-bit table loads)");
    for (int k = 0; k < n_tensors; ++k) NES_CHECK(counts[k] >= 1, "empty tensor");
    Ctx* c = new Ctx();
    build_layout(c->lay, counts, nullptr, nullptr, n_tensors, granule);
    if (ctx_common_init(c, device, seed, table_len, n_parents)) { delete c; return 1; }
    *out = reinterpret_cast<nes_ctx*>(c);
    return 0;
}

int nes_ctx_destroy(nes_ctx* ctx) {
    if (!ctx) return 0;
    Ctx* c = CTX(ctx);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    decode_destroy(c);
    cider_destroy(c);
    void* ptrs[] = {c->table, c->theta64, c->theta32, c->theta_hi, c->theta_lo, c->grad, c->opt_m, c->opt_v, c->d_lay, c->d_norm_partials,
                    c->d_norms, c->d_G64, c->d_theta_rowmax, c->d_eps_rowmax, c->scale32, c->d_scale_max, c->d_abs_partials, c->d_offsets, c->d_weights, c->d_fitness, c->d_centered, c->d_ranks, c->d_off_sorted, c->d_w_sorted};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (c->ev_a) cudaEventDestroy(c->ev_a);
    if (c->ev_b) cudaEventDestroy(c->ev_b);
    if (c->ev_c) cudaEventDestroy(c->ev_c);
    if (c->ev_d) cudaEventDestroy(c->ev_d);
    for (auto& r : c->prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    for (auto& r : c->prof_pool) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    delete[] c->theta_is_f32;
    delete[] c->theta64_stale;
    delete[] c->scale_stale;
    delete c;
    return 0;
}

int nes_set_stream(nes_ctx* ctx, void* s) { NEED_CTX(ctx); CTX(ctx)->stream = reinterpret_cast<cudaStream_t>(s); return 0; }
int nes_sync(nes_ctx* ctx) { NEED_CTX(ctx); NES_CUDA(cudaStreamSynchronize(CTX(ctx)->stream)); return 0; }
int64_t nes_num_params(const nes_ctx* ctx) { return ctx ? reinterpret_cast<const Ctx*>(ctx)->lay.D : -1; }
int64_t nes_padded_params(const nes_ctx* ctx) { return ctx ? reinterpret_cast<const Ctx*>(ctx)->lay.D_pad : -1; }
int64_t nes_granule(const nes_ctx* ctx) { return ctx ? reinterpret_cast<const Ctx*>(ctx)->lay.G : -1; }
int64_t nes_table_len(const nes_ctx* ctx) { return ctx ? reinterpret_cast<const Ctx*>(ctx)->table_len : -1; }
int64_t nes_gpu_launches(const nes_ctx* ctx) { return ctx ? reinterpret_cast<const Ctx*>(ctx)->launches : -1; }

int nes_fill_table(nes_ctx* ctx) { NEED_CTX(ctx); return launch_fill_table(CTX(ctx)); }

int nes_table_slice(nes_ctx* ctx, int64_t start, int64_t count, void* out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(start >= 0 && count >= 0 && start + count <= c->table_len, "slice outside the table");
    if (count == 0) return 0;
    NES_CUDA(cudaMemcpyAsync(out, c->table + start, sizeof(__half) * count, cudaMemcpyDefault, c->stream));
    return finish_output(c, out);
}

int nes_sample_offsets(nes_ctx* ctx, uint64_t iteration, int64_t F, int64_t* out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(F >= 0, "negative F");
    if (F == 0) return 0;
    if (ensure_small(c, F)) return 1;
    if (launch_sample_offsets(c, iteration, F, c->d_offsets)) return 1;
    if (out) { NES_CUDA(cudaMemcpyAsync(out, c->d_offsets, sizeof(int64_t) * F, cudaMemcpyDefault, c->stream)); return finish_output(c, out); }
    return 0;
}

int nes_sample_parents(nes_ctx* ctx, uint64_t iteration, int64_t F, int32_t n_parents, int32_t* out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(F >= 0 && out, "bad arguments");
    if (F == 0) return 0;
    if (ensure_small(c, F)) return 1;
    int32_t* tmp = reinterpret_cast<int32_t*>(c->d_ranks);
    if (launch_sample_parents(c, iteration, F, n_parents, tmp)) return 1;
    NES_CUDA(cudaMemcpyAsync(out, tmp, sizeof(int32_t) * F, cudaMemcpyDefault, c->stream));
    return finish_output(c, out);
}

int nes_sample_parents_tournament(nes_ctx* ctx, uint64_t iteration, int64_t F, int32_t n_parents, int32_t tournament_size, int32_t* out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(F >= 0 && out, "bad arguments");
    if (F == 0) return 0;
    if (ensure_small(c, F)) return 1;
    int32_t* tmp = reinterpret_cast<int32_t*>(c->d_ranks);
    if (launch_sample_parents_tournament(c, iteration, F, n_parents, tournament_size, tmp)) return 1;
    NES_CUDA(cudaMemcpyAsync(out, tmp, sizeof(int32_t) * F, cudaMemcpyDefault, c->stream));
    return finish_output(c, out);
}

int nes_set_theta(nes_ctx* ctx, int parent, const void* theta, int is_f64) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(theta, "null theta");
    NES_CHECK(parent >= 0 && parent < c->n_parents, "parent index out of range");
    const int64_t D = c->lay.D;
    const void* src = theta;
    void* tmp = nullptr;
    if (!is_device_ptr(theta)) {
        const size_t bytes = (is_f64 ? sizeof(double) : sizeof(float)) * D;
        NES_CUDA(cudaMallocAsync(&tmp, bytes, c->stream));
        NES_CUDA(cudaMemcpyAsync(tmp, theta, bytes, cudaMemcpyHostToDevice, c->stream));
        src = tmp;
    }
    int rc = launch_set_theta(c, parent, src, is_f64 != 0);
    if (tmp) { cudaFreeAsync(tmp, c->stream); NES_CUDA(cudaStreamSynchronize(c->stream)); }
    return rc;
}

int nes_get_theta(nes_ctx* ctx, int parent, void* out, int as_f64) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(out, "null output");
    NES_CHECK(parent >= 0 && parent < c->n_parents, "parent index out of range");
    const int64_t D = c->lay.D;
    if (as_f64 && refresh_theta64(c, parent)) return 1;
    if (as_f64) NES_CUDA(cudaMemcpyAsync(out, c->theta64 + parent * D, sizeof(double) * D, cudaMemcpyDefault, c->stream));
    else        NES_CUDA(cudaMemcpyAsync(out, c->theta32 + parent * D, sizeof(float) * D, cudaMemcpyDefault, c->stream));
    return finish_output(c, out);
}

int nes_evolve(nes_ctx* ctx, int parent, int64_t offset, float sign, float sigma, float* member_out, float* noise_out, int dst_parent) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(parent >= 0 && parent < c->n_parents && dst_parent < c->n_parents, "parent index out of range");
    NES_CHECK(sign == 1.0f || sign == -1.0f, "sign must be +1 or -1");
    const int64_t D = c->lay.D;
    float *dm = member_out, *dn = noise_out;
    float *tm = nullptr, *tn = nullptr;
    if (member_out && !is_device_ptr(member_out)) { NES_CUDA(cudaMallocAsync((void**)&tm, sizeof(float) * D, c->stream)); dm = tm; }
    if (noise_out && !is_device_ptr(noise_out)) { NES_CUDA(cudaMallocAsync((void**)&tn, sizeof(float) * D, c->stream)); dn = tn; }
    int rc = launch_evolve(c, parent, offset, sign, sigma, dm, dn, dst_parent);
    if (!rc && tm) NES_CUDA(cudaMemcpyAsync(member_out, tm, sizeof(float) * D, cudaMemcpyDeviceToHost, c->stream));
    if (!rc && tn) NES_CUDA(cudaMemcpyAsync(noise_out, tn, sizeof(float) * D, cudaMemcpyDeviceToHost, c->stream));
    if (tm) cudaFreeAsync(tm, c->stream);
    if (tn) cudaFreeAsync(tn, c->stream);
    if (tm || tn) NES_CUDA(cudaStreamSynchronize(c->stream));
    return rc;
}

int nes_evolve_batch(nes_ctx* ctx, int n, const int32_t* src_parent, const int64_t* offsets, const float* sigmas, const int32_t* dst_parent) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(n == 0 || (src_parent && offsets && sigmas && dst_parent), "null argument");
    return launch_evolve_batch(c, n, src_parent, offsets, sigmas, dst_parent);
}

int nes_es_select_evolve(nes_ctx* ctx, const double* fitness, const int64_t* offsets, const int32_t* parents, int64_t n, int n_keep,
                         float sigma, int dst_base, int64_t* order_out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(n >= 1 && fitness && offsets, "bad arguments");
    // inputs that live on the host are staged in stream order; device inputs are used where they are (no host round trip)
    std::vector<void*> temps;
    auto stage = [&](const void* src, size_t bytes, const void** out) -> int {
        if (!src || is_device_ptr(src)) { *out = src; return 0; }
        void* d = nullptr;
        NES_CUDA(cudaMallocAsync(&d, bytes, c->stream));
        temps.push_back(d);
        NES_CUDA(cudaMemcpyAsync(d, src, bytes, cudaMemcpyHostToDevice, c->stream));
        *out = d; return 0;
    };
    const void *d_fit, *d_off, *d_par;
    int64_t* d_order = nullptr;
    int rc = stage(fitness, sizeof(double) * n, &d_fit) || stage(offsets, sizeof(int64_t) * n, &d_off) || stage(parents, sizeof(int32_t) * n, &d_par);
    if (!rc) {
        if (order_out && is_device_ptr(order_out)) d_order = order_out;
        else { rc = cudaMallocAsync((void**)&d_order, sizeof(int64_t) * (n_keep > 0 ? n_keep : 1), c->stream) != cudaSuccess; if (!rc) temps.push_back(d_order); }
    }
    if (!rc) rc = launch_es_select_evolve(c, (const double*)d_fit, (const int64_t*)d_off, (const int32_t*)d_par, n, n_keep, sigma, dst_base, d_order);
    if (!rc && order_out && d_order != order_out && n_keep > 0) rc = cudaMemcpyAsync(order_out, d_order, sizeof(int64_t) * n_keep, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess;
    const bool host_io = !temps.empty();
    for (void* t : temps) cudaFreeAsync(t, c->stream);
    if (!rc && host_io) NES_CUDA(cudaStreamSynchronize(c->stream));          // pageable host inputs / a host output must be complete on return
    if (rc && g_err.empty()) set_error("nes_es_select_evolve: CUDA call failed");
    return rc ? 1 : 0;
}

int nes_copy_parent(nes_ctx* ctx, int src_parent, int dst_parent) {
    NEED_CTX(ctx);
    return copy_parent(CTX(ctx), src_parent, dst_parent);
}

int nes_rollout_trace(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents,
                      const float* fc, int rows, int shared_batch, float sigma, int precision, const int64_t* forced,
                      int64_t* seq, float* logprobs, float* h_trace, float* logit_trace) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(c->is_fc && c->ws, "not an fc_caption context with a decode workspace");
    NES_CHECK(n_members >= 0, "negative member count");
    if (n_members == 0) return 0;
    NES_CHECK(offsets && signs && fc, "null argument");
    const int64_t T = c->dims.seq_length, R = c->dims.rnn_size, V1 = c->dims.vocab_size + 1;
    float *dh = h_trace, *dl = logit_trace, *th = nullptr, *tl = nullptr;
    const size_t hb = sizeof(float) * (T + 1) * n_members * rows * R, lb = sizeof(float) * T * n_members * rows * V1;
    if (h_trace && !is_device_ptr(h_trace)) { NES_CUDA(cudaMalloc((void**)&th, hb)); NES_CUDA(cudaMemsetAsync(th, 0, hb, c->stream)); dh = th; }
    if (logit_trace && !is_device_ptr(logit_trace)) { NES_CUDA(cudaMalloc((void**)&tl, lb)); NES_CUDA(cudaMemsetAsync(tl, 0, lb, c->stream)); dl = tl; }
    int rc = decode_rollout(c, n_members, offsets, signs, parents, fc, rows, shared_batch, sigma, precision, forced, seq, logprobs, dh, dl);
    if (!rc && th) NES_CUDA(cudaMemcpyAsync(h_trace, th, hb, cudaMemcpyDeviceToHost, c->stream));
    if (!rc && tl) NES_CUDA(cudaMemcpyAsync(logit_trace, tl, lb, cudaMemcpyDeviceToHost, c->stream));
    if (th || tl) { cudaStreamSynchronize(c->stream); if (th) cudaFree(th); if (tl) cudaFree(tl); }
    if (rc) return rc;
    if (finish_output(c, seq) || finish_output(c, logprobs)) return 1;
    return 0;
}

int nes_rollout_sample(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents, const float* fc,
                       int rows, int shared_batch, float sigma, const double* uniforms, int64_t* seq, float* logprobs) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(c->is_fc && c->ws, "not an fc_caption context with a decode workspace");
    NES_CHECK(n_members >= 0, "negative member count");
    if (n_members == 0) return 0;
    NES_CHECK(offsets && signs && fc && uniforms, "null argument");
    if (decode_rollout(c, n_members, offsets, signs, parents, fc, rows, shared_batch, sigma, NES_PREC_EXACT, nullptr, seq, logprobs, nullptr, nullptr, uniforms)) return 1;
    if (finish_output(c, seq) || finish_output(c, logprobs)) return 1;
    return 0;
}

int nes_rollout(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents, const float* fc,
                int rows, int shared_batch, float sigma, int precision, int64_t* seq, float* logprobs) {
    return nes_rollout_trace(ctx, n_members, offsets, signs, parents, fc, rows, shared_batch, sigma, precision, nullptr, seq, logprobs,
                             nullptr, nullptr);
}

int nes_logprob_fitness(nes_ctx* ctx, int64_t n_members, int rows, double* out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(out, "null output");
    if (decode_logprob_fitness(c, n_members, rows, out)) return 1;
    return finish_output(c, out);
}

int nes_cider_set_tables(nes_ctx* ctx, const uint64_t* idf_keys, const double* idf_vals, int64_t n_idf, double idf_default,
                         const int32_t* img_ref_ptr, int32_t n_img, const int32_t* ref_ent_ptr, int32_t n_refs,
                         const uint64_t* ent_key, const double* ent_w, const double* ref_norm, const int32_t* ref_len, const double* penalty) {
    NEED_CTX(ctx);
    return cider_set_tables(CTX(ctx), idf_keys, idf_vals, n_idf, idf_default, img_ref_ptr, n_img, ref_ent_ptr, n_refs, ent_key, ent_w, ref_norm, ref_len, penalty);
}

int nes_cider_scores(nes_ctx* ctx, const int64_t* seq, int64_t n_members, int rows, int T, int rows_per_img, double* scores) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (cider_scores(c, seq, n_members, rows, T, rows_per_img, scores)) return 1;
    return finish_output(c, scores);
}

// ---- MNIST classifier population rollout (mnist.cu)
int nes_mnist_rollout(nes_ctx* ctx, int64_t n_members, const int64_t* offsets, const float* signs, const int32_t* parents,
                      const float* images, const int64_t* labels, int rows, int shared_batch, float sigma,
                      double* fitness, int32_t* correct, float* logits) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(n_members >= 0 && rows >= 1, "bad geometry");
    if (n_members == 0) return 0;
    NES_CHECK(offsets && signs && images && labels && fitness, "null argument");
    std::vector<void*> temps;
    auto release = [&]() { for (void* t : temps) cudaFreeAsync(t, c->stream); };
    auto dev_in = [&](const void* src, size_t bytes, const void** out) -> int {
        if (!src || is_device_ptr(src)) { *out = src; return 0; }
        void* d = nullptr;
        NES_CUDA(cudaMallocAsync(&d, bytes, c->stream));
        temps.push_back(d);
        NES_CUDA(cudaMemcpyAsync(d, src, bytes, cudaMemcpyHostToDevice, c->stream));
        *out = d; return 0;
    };
    auto dev_out = [&](void* dst, size_t bytes, void** out) -> int {
        if (dst && is_device_ptr(dst)) { *out = dst; return 0; }
        void* d = nullptr;
        NES_CUDA(cudaMallocAsync(&d, bytes, c->stream));
        temps.push_back(d);
        *out = d; return 0;
    };
    if (!is_device_ptr(offsets))
        for (int64_t i = 0; i < n_members; ++i) {
            NES_CHECK(offsets[i] >= 0 && offsets[i] + c->lay.D_pad <= c->table_len && offsets[i] % c->lay.G == 0, "offset outside the noise table or not slot aligned");
            NES_CHECK(!parents || is_device_ptr(parents) || (parents[i] >= 0 && parents[i] < c->n_parents), "parent index out of range");
            NES_CHECK(is_device_ptr(signs) || signs[i] == 1.0f || signs[i] == -1.0f, "sign must be +1 or -1");
        }
    const int64_t n_img = (shared_batch ? 1 : n_members) * (int64_t)rows, n_out = n_members * (int64_t)rows;
    const void *d_off, *d_sign, *d_par, *d_img, *d_lab;
    void *d_loss, *d_pred, *d_fit, *d_cor, *d_log = nullptr;
    int rc = dev_in(offsets, sizeof(int64_t) * n_members, &d_off) || dev_in(signs, sizeof(float) * n_members, &d_sign) ||
             dev_in(parents, sizeof(int32_t) * n_members, &d_par) || dev_in(images, sizeof(float) * n_img * 784, &d_img) ||
             dev_in(labels, sizeof(int64_t) * n_img, &d_lab) || dev_out(nullptr, sizeof(float) * n_out, &d_loss) ||
             dev_out(nullptr, sizeof(int32_t) * n_out, &d_pred) || dev_out(fitness, sizeof(double) * n_members, &d_fit) ||
             dev_out(correct, sizeof(int32_t) * n_members, &d_cor) || (logits ? dev_out(logits, sizeof(float) * n_out * 10, &d_log) : 0);
    if (!rc) rc = mnist_rollout(c, n_members, (const int64_t*)d_off, (const float*)d_sign, (const int32_t*)d_par, (const float*)d_img, (const int64_t*)d_lab,
                                rows, shared_batch, sigma, (float*)d_loss, (int32_t*)d_pred, (double*)d_fit, (int32_t*)d_cor, (float*)d_log);
    if (!rc && d_fit != fitness) rc = cudaMemcpyAsync(fitness, d_fit, sizeof(double) * n_members, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess;
    if (!rc && correct && d_cor != correct) rc = cudaMemcpyAsync(correct, d_cor, sizeof(int32_t) * n_members, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess;
    if (!rc && logits && d_log != logits) rc = cudaMemcpyAsync(logits, d_log, sizeof(float) * n_out * 10, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess;
    release();
    if (!rc && (!is_device_ptr(fitness) || (correct && !is_device_ptr(correct)) || (logits && !is_device_ptr(logits)))) NES_CUDA(cudaStreamSynchronize(c->stream));
    if (rc && g_err.empty()) set_error("nes_mnist_rollout: CUDA copy failed");
    return rc ? 1 : 0;
}

int nes_set_mutation(nes_ctx* ctx, int kind) {
    NEED_CTX(ctx);
    return set_mutation(CTX(ctx), kind);
}

int nes_set_sensitivity(nes_ctx* ctx, int parent, const float* sens) {
    NEED_CTX(ctx);
    return set_sensitivity(CTX(ctx), parent, sens);
}

int nes_shaped_fitness(nes_ctx* ctx, int kind, int64_t n_members, int rows, const double* scores, double* out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (cider_shaped_fitness(c, kind, n_members, rows, scores, out)) return 1;
    return finish_output(c, out);
}

int nes_compute_ranks(nes_ctx* ctx, const double* x, int64_t n, int64_t* ranks) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(n >= 0 && (n == 0 || (x && ranks)), "bad arguments");
    if (n == 0) return 0;
    if (ensure_small(c, (n + 1) / 2)) return 1;
    NES_CUDA(cudaMemcpyAsync(c->d_fitness, x, sizeof(double) * n, cudaMemcpyDefault, c->stream));
    if (launch_ranks(c, c->d_fitness, n, c->d_ranks)) return 1;
    NES_CUDA(cudaMemcpyAsync(ranks, c->d_ranks, sizeof(int64_t) * n, cudaMemcpyDefault, c->stream));
    return finish_output(c, ranks);
}

int nes_rank_weights(nes_ctx* ctx, const double* fitness, int64_t F, double* centered, float* weights) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(F >= 0 && (F == 0 || fitness), "bad arguments");
    if (F == 0) return 0;
    if (ensure_small(c, F)) return 1;
    NES_CUDA(cudaMemcpyAsync(c->d_fitness, fitness, sizeof(double) * 2 * F, cudaMemcpyDefault, c->stream));
    if (launch_rank_weights(c, c->d_fitness, F, c->d_centered, c->d_weights)) return 1;
    if (centered) NES_CUDA(cudaMemcpyAsync(centered, c->d_centered, sizeof(double) * 2 * F, cudaMemcpyDefault, c->stream));
    if (weights) NES_CUDA(cudaMemcpyAsync(weights, c->d_weights, sizeof(float) * F, cudaMemcpyDefault, c->stream));
    if (finish_output(c, centered) || finish_output(c, weights)) return 1;
    return 0;
}

// No counterpart in the reference: added for callers that cannot allocate (include/nes_b200.h).
// Enqueues two kernels on the context stream and returns; no cudaMalloc / cudaFree / copy, the context's own buffers are not touched.
int nes_rank_weights_static(nes_ctx* ctx, const double* fitness, int64_t F, double* centered, float* weights) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(F >= 0, "negative F");
    NES_CHECK(F <= NES_STATIC_RANK_MAX_F, "nes_rank_weights_static: F exceeds NES_STATIC_RANK_MAX_F (32768 antithetic pairs); use nes_rank_weights");
    if (F == 0) return 0;
    NES_CHECK(fitness && (centered || weights), "nes_rank_weights_static: null fitness, or neither output given");
    NES_CHECK(is_device_ptr(fitness) && (!centered || is_device_ptr(centered)) && (!weights || is_device_ptr(weights)),
              "nes_rank_weights_static takes device pointers only (it stages nothing)");
    return launch_rank_weights_static(c, fitness, F, centered, weights);
}

int nes_gradient(nes_ctx* ctx, const int64_t* offsets, const float* weights, int64_t F, float sigma, float* g_out) {
    return nes_gradient_shard(ctx, offsets, weights, F, sigma, 0, 1, g_out);
}

int nes_gradient_shard(nes_ctx* ctx, const int64_t* offsets, const float* weights, int64_t F, float sigma, int shard, int n_shards, float* g_out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(F >= 0, "negative F");
    if (ensure_small(c, F)) return 1;
    if (F > 0) {
        if (offsets) NES_CUDA(cudaMemcpyAsync(c->d_offsets, offsets, sizeof(int64_t) * F, cudaMemcpyDefault, c->stream));
        if (weights) NES_CUDA(cudaMemcpyAsync(c->d_weights, weights, sizeof(float) * F, cudaMemcpyDefault, c->stream));
    }
    if (launch_gradient(c, c->d_offsets, c->d_weights, F, sigma, c->grad, shard, n_shards)) return 1;
    if (g_out) { NES_CUDA(cudaMemcpyAsync(g_out, c->grad, sizeof(float) * c->lay.D, cudaMemcpyDefault, c->stream)); return finish_output(c, g_out); }
    return 0;
}

int nes_opt_init(nes_ctx* ctx, int kind, double stepsize, double b1, double b2, double eps) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(kind == NES_OPT_SGD || kind == NES_OPT_ADAM, "unknown optimizer kind");
    c->opt_kind = kind; c->opt_stepsize = stepsize; c->opt_b1 = b1; c->opt_b2 = b2; c->opt_eps = eps; c->opt_t = 0;
    NES_CUDA(cudaMemsetAsync(c->opt_m, 0, sizeof(double) * c->lay.D, c->stream));
    NES_CUDA(cudaMemsetAsync(c->opt_v, 0, sizeof(double) * c->lay.D, c->stream));
    return 0;
}
int nes_opt_set_stepsize(nes_ctx* ctx, double stepsize) { NEED_CTX(ctx); CTX(ctx)->opt_stepsize = stepsize; return 0; }

int nes_opt_update(nes_ctx* ctx, const float* grad, double l2coeff, double* ratio_out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (grad) NES_CUDA(cudaMemcpyAsync(c->grad, grad, sizeof(float) * c->lay.D, cudaMemcpyDefault, c->stream));
    return launch_opt_update(c, c->grad, nullptr, 0, l2coeff, ratio_out);
}

int nes_opt_update_direct(nes_ctx* ctx, const void* globalg, int is_f64, double* ratio_out) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(globalg, "null globalg");
    if (is_f64) {
        if (!c->d_G64) NES_CUDA(cudaMalloc((void**)&c->d_G64, sizeof(double) * c->lay.D));
        NES_CUDA(cudaMemcpyAsync(c->d_G64, globalg, sizeof(double) * c->lay.D, cudaMemcpyDefault, c->stream));
        return launch_opt_update(c, nullptr, c->d_G64, 2, 0.0, ratio_out);
    }
    NES_CUDA(cudaMemcpyAsync(c->grad, globalg, sizeof(float) * c->lay.D, cudaMemcpyDefault, c->stream));
    return launch_opt_update(c, c->grad, nullptr, 1, 0.0, ratio_out);
}

int nes_opt_get_state(nes_ctx* ctx, double* m, double* v, int64_t* t) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (m) NES_CUDA(cudaMemcpyAsync(m, c->opt_m, sizeof(double) * c->lay.D, cudaMemcpyDefault, c->stream));
    if (v) NES_CUDA(cudaMemcpyAsync(v, c->opt_v, sizeof(double) * c->lay.D, cudaMemcpyDefault, c->stream));
    if (t) *t = c->opt_t;
    NES_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
int nes_opt_set_state(nes_ctx* ctx, const double* m, const double* v, int64_t t) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (m) NES_CUDA(cudaMemcpyAsync(c->opt_m, m, sizeof(double) * c->lay.D, cudaMemcpyDefault, c->stream));
    if (v) NES_CUDA(cudaMemcpyAsync(c->opt_v, v, sizeof(double) * c->lay.D, cudaMemcpyDefault, c->stream));
    c->opt_t = t;
    NES_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int nes_master_step(nes_ctx* ctx, const double* fitness, const int64_t* offsets, int64_t F, float sigma, double l2coeff, double* ratio_out) {
    NEED_CTX(ctx);
    NES_CHECK(F >= 1 && fitness, "master step needs at least one evaluated pair");
    if (nes_rank_weights(ctx, fitness, F, nullptr, nullptr)) return 1;
    if (nes_gradient(ctx, offsets, nullptr, F, sigma, nullptr)) return 1;
    return nes_opt_update(ctx, nullptr, l2coeff, ratio_out);
}

int nes_profile_enable(nes_ctx* ctx, int on) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    c->profiling = on != 0;
    return 0;
}
int nes_profile_read(nes_ctx* ctx, float* ms_by_kind, int32_t* count_by_kind) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    NES_CHECK(ms_by_kind && count_by_kind, "null output");
    NES_CUDA(cudaStreamSynchronize(c->stream));
    for (int k = 0; k < PROF_KINDS; ++k) { ms_by_kind[k] = 0.f; count_by_kind[k] = 0; }
    for (auto& r : c->prof_recs) {
        float ms = 0.f;
        NES_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
        ms_by_kind[r.kind] += ms; count_by_kind[r.kind] += 1;
        c->prof_pool.push_back(r);
    }
    c->prof_recs.clear();
    return 0;
}

int nes_hybrid_stats(nes_ctx* ctx, int64_t* out4, int reset) {
    NEED_CTX(ctx);
    NES_CHECK(out4, "null output");
    return decode_hybrid_stats(CTX(ctx), out4, reset);
}

int nes_last_gradient_ms(nes_ctx* ctx, float* ms) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (!c->grad_timed) { c->grad_timed = true; *ms = -1.f; return 0; }   // first call arms the timer
    NES_CUDA(cudaEventSynchronize(c->ev_b));
    NES_CUDA(cudaEventElapsedTime(ms, c->ev_a, c->ev_b));
    return 0;
}
int nes_last_rollout_ms(nes_ctx* ctx, float* ms) {
    NEED_CTX(ctx); Ctx* c = CTX(ctx);
    if (!c->rollout_timed) { c->rollout_timed = true; *ms = -1.f; return 0; }
    NES_CUDA(cudaEventSynchronize(c->ev_d));
    NES_CUDA(cudaEventElapsedTime(ms, c->ev_c, c->ev_d));
    return 0;
}

}  // extern "C"
