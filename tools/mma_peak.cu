// Tensor-pipe ceiling probe: tcgen05.mma issued back to back on operands already in shared memory (no TMA, no
// epilogue), for kind::i8 and kind::f16, cta_group::1 (128 x 256) and cta_group::2 (256 x 256).  The numbers
// are the denominators the GroupGEMM kernels are compared with.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I mxmoe_b200/csrc tools/mma_peak.cu -o tools/mma_peak
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"

using namespace mxb::ptx;

template <bool PAIR, bool I8>
__global__ void __launch_bounds__(128, 1) peak_kernel(int iters, int n_cols) {
  extern __shared__ uint8_t raw[];
  uint32_t a = smem_u32(raw);
  uint8_t* base = raw + ((1024u - (a & 1023u)) & 1023u);
  uint64_t* bar = reinterpret_cast<uint64_t*>(base + 4 * 49152);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 4 * 49152 / 4; i += 128) reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u * (i & 3);
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (warp == 0) {
    if (PAIR) { tmem_alloc_pair(tptr, 512); tmem_relinquish_pair(); }
    else      { tmem_alloc(tptr, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tptr;
  const bool issuer = warp == 1 && lane == 0 && (!PAIR || cluster_ctarank() == 0);
  if (issuer) {
    const uint32_t cf = I8 ? 2u : 1u, ab = I8 ? 1u : 0u;
    const uint32_t idesc = (cf << 4) | (ab << 7) | (ab << 10) | ((uint32_t(n_cols) >> 3) << 17) | (((PAIR ? 256u : 128u) >> 4) << 24);
    for (int it = 0; it < iters; ++it) {
      const uint32_t st = smem_u32(base + (it & 3) * 49152);
      const uint64_t da = umma_desc_sw128(st), db = umma_desc_sw128(st + 16384);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (PAIR) { if (I8) umma_i8_pair(tmem + (it & 1) * 256, da + 2 * k, db + 2 * k, idesc, 1); else umma_f16_pair(tmem + (it & 1) * 256, da + 2 * k, db + 2 * k, idesc, 1); }
        else      { if (I8) umma_i8(tmem + (it & 1) * 256, da + 2 * k, db + 2 * k, idesc, 1); else umma_f16(tmem + (it & 1) * 256, da + 2 * k, db + 2 * k, idesc, 1); }
      }
    }
    if (PAIR) umma_commit_pair(bar, 1); else umma_commit(bar);
    mbar_wait(bar, 0);
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  if (warp == 0) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem, 512); else tmem_dealloc(tmem, 512);
  }
}

template <bool PAIR, bool I8>
void run(const char* name, int sms, int n_cols) {
  const int smem = 4 * 49152 + 2048;
  auto k = peak_kernel<PAIR, I8>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 20000;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sms & ~1);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = PAIR ? 2 : 1;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaLaunchKernelEx(&cfg, k, 200, n_cols);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  cudaLaunchKernelEx(&cfg, k, iters, n_cols);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  // per iteration and CTA: 4 MMAs of 128 rows x n_cols x (32 B of K: 32 int8 / 16 fp16)
  const double kk = I8 ? 32.0 : 16.0;
  const double ops = 2.0 * 128 * n_cols * kk * 4 * double(iters) * (sms & ~1);
  printf("%-28s N=%3d  %8.3f ms  %8.1f T%sOPS  (%s)\n", name, n_cols, ms, ops / ms / 1e9, I8 ? "" : "FL", cudaGetErrorString(e));
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
  for (int n : {256, 128, 64}) {
    run<false, true>("i8  cta_group::1 M=128", p.multiProcessorCount, n);
    run<true, true>("i8  cta_group::2 M=256", p.multiProcessorCount, n);
    run<false, false>("f16 cta_group::1 M=128", p.multiProcessorCount, n);
    run<true, false>("f16 cta_group::2 M=256", p.multiProcessorCount, n);
  }
  return 0;
}
