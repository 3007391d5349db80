// Microbenchmark: how fast can TMA fill shared memory on B200?  (design input for member_gemm.cuh)
//   ./tma_bw <buffer_MB> <box_rows> <bk(32|64)> <stages> <ctas> <cluster(1|2|4)> <iters>
// Every CTA streams boxes [box_rows x bk] fp16 from a [rows, 512] fp16 matrix; with cluster > 1 each CTA loads
// 1/cluster of every box and multicasts it to the whole cluster (so every CTA still receives full boxes).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include "../nes-img-captioning_b200/csrc/tc_ptx.cuh"
using namespace nes;

__device__ __forceinline__ void tma_load_2d_mc(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(tc::smem_u32(dst)), "l"((uint64_t)m), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() { asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// remote arrive on the same barrier offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(tc::smem_u32(bar)), "r"(cta));
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}

struct P { CUtensorMap map; int box_rows, bk, stages, iters, cluster, rows_total, nbox; };

__global__ void __launch_bounds__(64, 1) k_bw(const __grid_constant__ P p, unsigned long long* cyc) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    const int one_box = p.box_rows * p.bk * 2;
    const int box_bytes = one_box * p.nbox;
    uint64_t* full = (uint64_t*)(smem + p.stages * box_bytes);
    uint64_t* empty = full + p.stages;
    const uint32_t rank = p.cluster > 1 ? cluster_rank() : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], p.cluster); }
        tc::fence_barrier_init();
    }
    __syncthreads();
    if (p.cluster > 1) cluster_sync_all();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkb = 512 / p.bk;
    // each cluster streams its own row range so nothing is shared between clusters
    const int cl = blockIdx.x / p.cluster;
    const int ncl = gridDim.x / p.cluster;
    const int rows_per = (p.rows_total / ncl) / p.box_rows * p.box_rows;
    long long t0 = clock64();
    if (warp == 0 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        long long tw = 0, ti = 0;
        for (int i = 0; i < p.iters; ++i) {
            long long a0 = clock64();
            const int row = cl * rows_per + ((i / nkb) * p.box_rows) % rows_per;
            const int kb = i % nkb;
            tc::mbar_wait(&empty[stage], phase ^ 1);
            long long a1 = clock64();
            tc::mbar_expect_tx(&full[stage], box_bytes);
            if (p.cluster == 1) { for (int b = 0; b < p.nbox; ++b) tc::tma_load_2d(smem + stage * box_bytes + b * one_box, &p.map, &full[stage], kb * p.bk, row + b * p.box_rows); }
            else {
                const int sub = p.box_rows / p.cluster;
                tma_load_2d_mc(smem + stage * box_bytes + rank * sub * p.bk * 2, &p.map, &full[stage], kb * p.bk, row + rank * sub,
                               (uint16_t)((1u << p.cluster) - 1));
            }
            long long a2 = clock64(); tw += a1 - a0; ti += a2 - a1;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (blockIdx.x == 0) { cyc[1024] = tw; cyc[1025] = ti; }
    } else if (warp == 1 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        long long cw = 0, ca = 0;
        for (int i = 0; i < p.iters; ++i) {
            long long a0 = clock64();
            tc::mbar_wait(&full[stage], phase);
            long long a1 = clock64();
            if (p.cluster == 1) tc::mbar_arrive(&empty[stage]);
            else for (int c = 0; c < p.cluster; ++c) mbar_arrive_remote(&empty[stage], c);
            long long a2 = clock64(); cw += a1 - a0; ca += a2 - a1;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (blockIdx.x == 0) { cyc[1026] = cw; cyc[1027] = ca; }
    }
    __syncthreads();
    long long t1 = clock64();
    if (p.cluster > 1) cluster_sync_all();
    if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                            const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    if (argc < 8) { printf("usage\n"); return 1; }
    const long mb = atol(argv[1]); P p{}; p.box_rows = atoi(argv[2]); p.bk = atoi(argv[3]); p.stages = atoi(argv[4]);
    const int ctas = atoi(argv[5]); p.cluster = atoi(argv[6]); p.iters = atoi(argv[7]); p.nbox = argc > 8 ? atoi(argv[8]) : 1;
    const long rows = mb * 1024 * 1024 / 1024;     // rows of 512 fp16 = 1 KB
    p.rows_total = (int)rows;
    void* buf; cudaMalloc(&buf, rows * 1024); cudaMemset(buf, 0, rows * 1024);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    cuuint64_t gdim[2] = {512, (cuuint64_t)rows}; cuuint64_t gstr[1] = {1024};
    cuuint32_t box[2] = {(cuuint32_t)p.bk, (cuuint32_t)(p.box_rows / p.cluster)}; cuuint32_t es[2] = {1, 1};
    CUresult r = ((PFN_enc)fn)(&p.map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, buf, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               p.bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    unsigned long long* cyc; cudaMalloc(&cyc, 8 * 1100);
    const int smem = p.stages * p.nbox * p.box_rows * p.bk * 2 + 1024 + 256;
    cudaFuncSetAttribute(k_bw, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = p.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        cudaError_t e = cudaLaunchKernelEx(&cfg, k_bw, p, cyc);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep > 0 && ms < best) best = ms;
    }
    unsigned long long h[1100]; cudaMemcpy(h, cyc, 8 * 1100, cudaMemcpyDeviceToHost);
    printf("  per iter cycles: producer wait %.0f issue %.0f | consumer wait %.0f arrive %.0f | total %.0f\n", (double)h[1024]/p.iters, (double)h[1025]/p.iters, (double)h[1026]/p.iters, (double)h[1027]/p.iters, (double)h[0]/p.iters);
    double bytes = (double)ctas * p.iters * p.nbox * p.box_rows * p.bk * 2;
    printf("nbox %d buf %ld MB box %dx%d stages %d ctas %d cluster %d: %.3f ms  delivered %.2f TB/s  (%.1f B/cycle/SM by clock64, L2 reads %.2f TB/s)\n", p.nbox, mb, p.box_rows, p.bk,
           p.stages, ctas, p.cluster, best, bytes / best / 1e9, (double)p.iters * p.nbox * p.box_rows * p.bk * 2 / (double)h[0], bytes / p.cluster / best / 1e9);
    return 0;
}
