// Population rollout of the reference's MNIST classifier (BASELINE configs[0], experiments/mnist_es.json):
//   MnistNet.forward (src/classification/nets.py:25-37): conv 1->10 5x5, relu, maxpool 2, conv 10->20 5x5, relu, maxpool 2, fc 320->10
//   ClfPolicy.rollout (src/classification/policies.py:16-48): fitness = -CrossEntropyLoss(outputs, labels) (mean over the batch)
//   ClfPolicy.accuracy_on (:50-86): prediction = argmax(outputs, 1), first index on ties
// for members theta_p + s*sigma*eps(offset) [safe-mutation scaled] that are never written to HBM: a CTA rebuilds its member's 8490
// parameters in shared memory straight from theta32 and the noise-table row (the arithmetic of k_evolve, so they equal nes_evolve's
// bit for bit) and runs kImagesPerBlock images through the net with every intermediate in shared memory.  FP32 CUDA-core work: the
// whole net is 0.47 MFLOP-pairs per image, and a member's parameters (34 KB) are read once per CTA -- nothing here is GEMM shaped.
#include "common.cuh"

namespace nes {

namespace {
constexpr int kW1 = 0, kB1 = 250, kW2 = 260, kB2 = 5260, kW3 = 5280, kB3 = 8480, kParams = 8490;   // flat starts (parameter_vector order)
constexpr int kImagesPerBlock = 8;
constexpr int kThreads = 256;

struct MnistSmem {
    float w[kParams + 6];          // member parameters
    float img[28 * 28];
    float p1[10 * 12 * 12];        // after conv1 + relu + pool
    float p2[320];                 // after conv2 + relu + pool, flattened like x.view(-1, 4*4*20): [c][y][x]
    float logits[10];
};

__global__ void __launch_bounds__(kThreads) k_mnist_rollout(const float* __restrict__ theta32, const __half* __restrict__ table, const Layout* __restrict__ lay,
                                                            const int64_t* __restrict__ offs, const float* __restrict__ signs, const int32_t* __restrict__ parents,
                                                            float sigma, const float* __restrict__ scale, int scale_div,
                                                            const float* __restrict__ images, const int64_t* __restrict__ labels, int rows, int shared_batch,
                                                            float* __restrict__ loss, int32_t* __restrict__ pred, float* __restrict__ logits_out) {
    extern __shared__ uint8_t smem_raw[];
    MnistSmem& s = *reinterpret_cast<MnistSmem*>(smem_raw);
    const int m = blockIdx.y, tid = threadIdx.x;
    const int par = parents ? parents[m] : 0;
    const int64_t D = lay->D;
    const float* th = theta32 + (int64_t)par * D;
    const float* sc = scale ? scale + (int64_t)par * D : nullptr;
    const __half* row = table + offs[m];
    const float sign = signs[m];
    for (int k = 0; k < lay->n; ++k) {
        const TensorDesc t = lay->t[k];
        for (int j = tid; j < (int)t.count; j += kThreads) {
            const float th0 = th[t.start + j];
            float v = th0;
            if (sigma != 0.f) {
                const float nz = mut_noise(sigma, __half2float(row[t.pstart + j]), sc, t.start + j, scale_div);
                v = sign > 0.f ? __fadd_rn(th0, nz) : __fsub_rn(th0, nz);
            }
            s.w[t.start + j] = v;
        }
    }
    const float* img_base = images + (shared_batch ? 0 : (int64_t)m * rows * 784);
    const int r0 = blockIdx.x * kImagesPerBlock;
    for (int r = r0; r < min(r0 + kImagesPerBlock, rows); ++r) {
        __syncthreads();                                   // parameters ready / previous image done
        for (int j = tid; j < 784; j += kThreads) s.img[j] = img_base[(int64_t)r * 784 + j];
        __syncthreads();
        // conv1 + bias + relu + maxpool: one pooled output per item, its 6x6 input patch in registers
        for (int o = tid; o < 1440; o += kThreads) {
            const int c = o / 144, py = (o % 144) / 12, px = o % 12;
            const float* w = s.w + kW1 + c * 25;
            float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;
#pragma unroll
            for (int ky = 0; ky < 5; ++ky) {
                const float* i0 = s.img + (2 * py + ky) * 28 + 2 * px;
                const float* i1 = i0 + 28;
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) {
                    const float wv = w[ky * 5 + kx];
                    a00 = fmaf(wv, i0[kx], a00); a01 = fmaf(wv, i0[kx + 1], a01);
                    a10 = fmaf(wv, i1[kx], a10); a11 = fmaf(wv, i1[kx + 1], a11);
                }
            }
            const float mx = fmaxf(fmaxf(a00, a01), fmaxf(a10, a11)) + s.w[kB1 + c];
            s.p1[o] = fmaxf(mx, 0.f);
        }
        __syncthreads();
        // conv2 + bias + relu + maxpool
        for (int o = tid; o < 320; o += kThreads) {
            const int c = o / 16, py = (o % 16) / 4, px = o % 4;
            float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;
            for (int ci = 0; ci < 10; ++ci) {
                const float* w = s.w + kW2 + (c * 10 + ci) * 25;
                const float* pin = s.p1 + ci * 144;
#pragma unroll
                for (int ky = 0; ky < 5; ++ky) {
                    const float* i0 = pin + (2 * py + ky) * 12 + 2 * px;
                    const float* i1 = i0 + 12;
#pragma unroll
                    for (int kx = 0; kx < 5; ++kx) {
                        const float wv = w[ky * 5 + kx];
                        a00 = fmaf(wv, i0[kx], a00); a01 = fmaf(wv, i0[kx + 1], a01);
                        a10 = fmaf(wv, i1[kx], a10); a11 = fmaf(wv, i1[kx + 1], a11);
                    }
                }
            }
            const float mx = fmaxf(fmaxf(a00, a01), fmaxf(a10, a11)) + s.w[kB2 + c];
            s.p2[o] = fmaxf(mx, 0.f);
        }
        __syncthreads();
        // fc1: warp w computes outputs w and w + 8
        const int warp = tid >> 5, lane = tid & 31;
        for (int o = warp; o < 10; o += kThreads / 32) {
            const float* w = s.w + kW3 + o * 320;
            float a = 0.f;
            for (int j = lane; j < 320; j += 32) a = fmaf(w[j], s.p2[j], a);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
            if (lane == 0) s.logits[o] = a + s.w[kB3 + o];
        }
        __syncthreads();
        if (tid == 0) {
            // CrossEntropyLoss = log_softmax + nll (policies.py:39-40); argmax with the first index on ties (:72)
            float mx = s.logits[0]; int am = 0;
            for (int o = 1; o < 10; ++o) if (s.logits[o] > mx) { mx = s.logits[o]; am = o; }
            float se = 0.f;
            for (int o = 0; o < 10; ++o) se += expf(s.logits[o] - mx);
            const int64_t lab = labels[(shared_batch ? 0 : (int64_t)m * rows) + r];
            const float lp = (s.logits[lab] - mx) - logf(se);
            loss[(int64_t)m * rows + r] = -lp;
            pred[(int64_t)m * rows + r] = am;
        }
        if (logits_out && tid < 10) logits_out[((int64_t)m * rows + r) * 10 + tid] = s.logits[tid];
    }
}

// fitness = -mean loss (float64 accumulation in row order), correct = number of rows whose prediction equals the label
__global__ void __launch_bounds__(128) k_mnist_reduce(const float* __restrict__ loss, const int32_t* __restrict__ pred, const int64_t* __restrict__ labels,
                                                      int rows, int shared_batch, int64_t n_members, double* __restrict__ fitness, int32_t* __restrict__ correct) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_members) return;
    double acc = 0.0; int ok = 0;
    for (int r = 0; r < rows; ++r) {
        acc += (double)loss[m * rows + r];
        ok += pred[m * rows + r] == (int32_t)labels[(shared_batch ? 0 : m * rows) + r];
    }
    fitness[m] = -(acc / rows);
    if (correct) correct[m] = ok;
}
}  // namespace

// all pointers are device pointers here (capi.cu stages host buffers)
int mnist_rollout(Ctx* c, int64_t n_members, const int64_t* d_off, const float* d_sign, const int32_t* d_parent, const float* d_images,
                  const int64_t* d_labels, int rows, int shared_batch, float sigma, float* d_loss, int32_t* d_pred,
                  double* d_fitness, int32_t* d_correct, float* d_logits) {
    const Layout& L = c->lay;
    static const int64_t want[6] = {250, 10, 5000, 20, 3200, 10};
    NES_CHECK(!c->is_fc && L.n == 6, "the MNIST rollout needs a flat context with the 6 MnistNet tensors (conv1, conv2, fc1: weight, bias)");
    for (int k = 0; k < 6; ++k) NES_CHECK(L.t[k].count == want[k], "tensor sizes are not MnistNet's (250, 10, 5000, 20, 3200, 10)");
    NES_CHECK(rows >= 1 && n_members >= 0, "bad geometry");
    if (n_members == 0) return 0;
    NES_CHECK(n_members <= 65535, "at most 65535 members per call");
    static bool attr_set = false;
    if (!attr_set) { NES_CUDA(cudaFuncSetAttribute(k_mnist_rollout, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MnistSmem))); attr_set = true; }
    dim3 grid((rows + kImagesPerBlock - 1) / kImagesPerBlock, (unsigned)n_members);
    k_mnist_rollout<<<grid, kThreads, sizeof(MnistSmem), c->stream>>>(c->theta32, c->table, c->d_lay, d_off, d_sign, d_parent, sigma, c->scale32, c->scale_div,
                                                                      d_images, d_labels, rows, shared_batch, d_loss, d_pred, d_logits);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    k_mnist_reduce<<<(unsigned)((n_members + 127) / 128), 128, 0, c->stream>>>(d_loss, d_pred, d_labels, rows, shared_batch, n_members, d_fitness, d_correct);
    c->launches++;
    NES_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace nes
