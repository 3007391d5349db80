// Population rollout of the reference's MNIST classifier (BASELINE configs[0], experiments/mnist_es.json):
//   MnistNet.forward (src/classification/nets.py:25-37): conv 1->10 5x5, relu, maxpool 2, conv 10->20 5x5, relu, maxpool 2, fc 320->10
//   ClfPolicy.rollout (src/classification/policies.py:16-48): fitness = -CrossEntropyLoss(outputs, labels) (mean over the batch)
//   ClfPolicy.accuracy_on (:50-86): prediction = argmax(outputs, 1), first index on ties
// for members theta_p + s*sigma*eps(offset) [safe-mutation scaled] that are never written to HBM: a CTA rebuilds its member's 8490
// parameters in shared memory straight from theta32 and the noise-table row (the arithmetic of k_evolve, so they equal nes_evolve's
// bit for bit) and runs kImagesPerBlock images through the net stage by stage with every intermediate in shared memory; the
// convolutions are register tiled (5 output channels x the 2x2 pooling window per thread) so that the FMA pipe, not the
// shared-memory pipe, is the limit.  FP32 CUDA-core work: the
// whole net is 0.47 M multiply-adds per image, and a member's parameters (34 KB) are read once per CTA -- nothing here is GEMM shaped.
// ncu (profiles/r1_full_v18_mnist_summary.txt): FMA pipe 62 % active, issue slots 76 %, 6.8 ms for 5000 members x 64 images.
#include "common.cuh"

namespace nes {

namespace {
constexpr int kImagesPerBlock = 8;
constexpr int kThreads = 256;
// shared-memory parameter layout: every 5x5 filter padded to 28 floats so that its taps load as 7 broadcast LDS.128
constexpr int kFilt = 28;
constexpr int sW1 = 0, sW2 = sW1 + 10 * kFilt, sW3 = sW2 + 200 * kFilt, sB1 = sW3 + 3200, sB2 = sB1 + 12, sB3 = sB2 + 20, sParams = sB3 + 12;

struct MnistSmem {
    float w[sParams];                               // member parameters (36.5 KB)
    float img[kImagesPerBlock][28 * 28];            // 25 KB; dead after conv1, p2 lives on top of it
    float p1[kImagesPerBlock][10 * 12 * 12];        // after conv1 + relu + pool (46 KB)
    float logits[kImagesPerBlock][10];
};
static_assert(sizeof(MnistSmem) <= 112 * 1024, "two CTAs per SM");
static_assert(sW2 % 4 == 0 && sW3 % 4 == 0 && kFilt % 4 == 0, "filters must be 16-byte aligned");

// flat parameter index of tensor k, element j -> position in MnistSmem::w
__device__ __forceinline__ int smem_pos(int k, int j) {
    switch (k) {
        case 0: return sW1 + (j / 25) * kFilt + j % 25;
        case 1: return sB1 + j;
        case 2: return sW2 + (j / 25) * kFilt + j % 25;
        case 3: return sB2 + j;
        case 4: return sW3 + j;
        default: return sB3 + j;
    }
}

// CO output channels of one pooled position: acc[c][0..3] = the four conv outputs under the 2x2 pooling window, accumulated over one
// input plane.  in: the plane's 6x6 patch origin (row stride LD, 8-byte aligned); wf: CO filters, stride WS floats (16-byte aligned).
// Loads per plane: 18 LDS.64 of inputs (conflict free) + 7*CO broadcast LDS.128 of taps for 100*CO FMAs -- the first version of this
// kernel issued 0.85 shared-memory wavefronts per FMA and sat at 97 % of the LSU pipe (profiles/r1_full_v17_mnist_summary.txt)
template <int CO, int LD, int WS>
__device__ __forceinline__ void conv_plane(const float* __restrict__ in, const float* __restrict__ wf, float (&acc)[CO][4]) {
    float x[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float2 v = *reinterpret_cast<const float2*>(in + r * LD + 2 * q);
            x[r][2 * q] = v.x; x[r][2 * q + 1] = v.y;
        }
#pragma unroll
    for (int c = 0; c < CO; ++c) {
        float w[28];
#pragma unroll
        for (int q = 0; q < 7; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(wf + c * WS + 4 * q);
            w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                const float wv = w[ky * 5 + kx];
                acc[c][0] = fmaf(wv, x[ky][kx], acc[c][0]);     acc[c][1] = fmaf(wv, x[ky][kx + 1], acc[c][1]);
                acc[c][2] = fmaf(wv, x[ky + 1][kx], acc[c][2]); acc[c][3] = fmaf(wv, x[ky + 1][kx + 1], acc[c][3]);
            }
    }
}

__global__ void __launch_bounds__(kThreads, 2) k_mnist_rollout(const float* __restrict__ theta32, const __half* __restrict__ table, const Layout* __restrict__ lay,
                                                               const int64_t* __restrict__ offs, const float* __restrict__ signs, const int32_t* __restrict__ parents,
                                                               float sigma, const float* __restrict__ scale, int scale_div,
                                                               const float* __restrict__ images, const int64_t* __restrict__ labels, int rows, int shared_batch,
                                                               float* __restrict__ loss, int32_t* __restrict__ pred, float* __restrict__ logits_out) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    MnistSmem& s = *reinterpret_cast<MnistSmem*>(smem_raw);
    float (*p2)[320] = reinterpret_cast<float (*)[320]>(&s.img[0][0]);      // after conv2 + relu + pool, [c][y][x] like x.view(-1, 320)
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
            s.w[smem_pos(k, j)] = v;
        }
    }
    const float* img_base = images + (shared_batch ? 0 : (int64_t)m * rows * 784);
    const int r0 = blockIdx.x * kImagesPerBlock;
    const int n_img = min(kImagesPerBlock, rows - r0);
    for (int j = tid; j < n_img * 784; j += kThreads) (&s.img[0][0])[j] = img_base[(int64_t)r0 * 784 + j];
    __syncthreads();
    // conv1 + bias + relu + maxpool: item = (image, half of the channels, pooled position), 5 channels per item
    for (int o = tid; o < n_img * 2 * 144; o += kThreads) {
        const int im = o / 288, cg = (o % 288) / 144, pos = o % 144, py = pos / 12, px = pos % 12;
        float acc[5][4] = {};
        conv_plane<5, 28, kFilt>(&s.img[im][(2 * py) * 28 + 2 * px], s.w + sW1 + cg * 5 * kFilt, acc);
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            const float mx = fmaxf(fmaxf(acc[c][0], acc[c][1]), fmaxf(acc[c][2], acc[c][3])) + s.w[sB1 + cg * 5 + c];
            s.p1[im][(cg * 5 + c) * 144 + pos] = fmaxf(mx, 0.f);
        }
    }
    __syncthreads();
    // conv2 + bias + relu + maxpool: item = (image, group of 5 output channels, pooled position): 8 x 4 x 16 = 512 items = two full rounds
    for (int o = tid; o < n_img * 4 * 16; o += kThreads) {
        const int im = o / 64, cg = (o % 64) / 16, pos = o % 16, py = pos / 4, px = pos % 4;
        float acc[5][4] = {};
        for (int ci = 0; ci < 10; ++ci)
            conv_plane<5, 12, 10 * kFilt>(&s.p1[im][ci * 144 + (2 * py) * 12 + 2 * px], s.w + sW2 + (cg * 5 * 10 + ci) * kFilt, acc);
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            const float mx = fmaxf(fmaxf(acc[c][0], acc[c][1]), fmaxf(acc[c][2], acc[c][3])) + s.w[sB2 + cg * 5 + c];
            p2[im][(cg * 5 + c) * 16 + pos] = fmaxf(mx, 0.f);
        }
    }
    __syncthreads();
    // fc1: one warp per (image, output)
    const int warp = tid >> 5, lane = tid & 31;
    for (int o = warp; o < n_img * 10; o += kThreads / 32) {
        const int im = o / 10, k = o % 10;
        const float* w = s.w + sW3 + k * 320;
        float a = 0.f;
        for (int j = lane; j < 320; j += 32) a = fmaf(w[j], p2[im][j], a);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
        if (lane == 0) s.logits[im][k] = a + s.w[sB3 + k];
    }
    __syncthreads();
    if (tid < n_img) {
        // CrossEntropyLoss = log_softmax + nll (policies.py:39-40); argmax with the first index on ties (:72)
        const float* lg = s.logits[tid];
        const int r = r0 + tid;
        float mx = lg[0]; int am = 0;
        for (int o = 1; o < 10; ++o) if (lg[o] > mx) { mx = lg[o]; am = o; }
        float se = 0.f;
        for (int o = 0; o < 10; ++o) se += expf(lg[o] - mx);
        const int64_t lab = labels[(shared_batch ? 0 : (int64_t)m * rows) + r];
        const float lp = (lg[lab] - mx) - logf(se);
        loss[(int64_t)m * rows + r] = -lp;
        pred[(int64_t)m * rows + r] = am;
    }
    if (logits_out && tid < n_img * 10) logits_out[((int64_t)m * rows + r0) * 10 + tid] = (&s.logits[0][0])[tid];
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
    NES_CUDA(cudaFuncSetAttribute(k_mnist_rollout, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MnistSmem)));   // per device: cheap enough to repeat
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
