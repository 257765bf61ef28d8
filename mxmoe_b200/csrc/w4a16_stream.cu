// Piece (4) of the hot path: weight-only w4a16 for LOW-COUNT experts (decode, M of a few rows) — HBM streaming.
//
// Reference: cta_gemm_multistage_qb_v2 / cta_gemm_wxa16g128 (mxmoe/kernels/src/include/cta_gemm.cuh:113-421) with
// 16-row tiles, dequant in registers (quantize.cuh:134-214), used for the memory-bound experts of the 512-token
// and decode workloads.  For M << 128 the tcgen05 path would spend a full 128-row UMMA tile and a shared-memory
// unpack per weight tile; here every 4-bit weight is read from HBM exactly once, dequantized in registers with
// the reference's arithmetic (LOP3 0x6400 magic, hsub, w = hfma(q, scale, zp): one fp16 rounding) and fed to
// mma.sync.m16n8k16 against the few activation rows, fp32 accumulation, fp16 output.
//
// Mapping: a warp owns 32 weight rows (4 groups of 8) of one problem over the whole K; lane (g = lane/4,
// t = lane%4) streams 16 bytes (32 k) of row g of every group per 128-k chunk.  The dot product does not care in
// which order k is visited, so the 8 values of one 32-bit word are used as this lane's k-slots of two MMAs and
// the activation fragment is loaded with the SAME k permutation (one 16-byte load per row and word).  A fragments
// are reused across the 4 row groups.  Grid: one warp task per (problem, 32-row slab); empty experts cost nothing.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "gg_kernel.cuh"

namespace mxb {
namespace {

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 ldg_stream(const void* p) {  // weights are read once: do not pollute L1
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

constexpr int WS_WARPS = 4;       // warps per CTA
constexpr int WS_ROWS = 32;       // weight rows per warp task

// one 32-bit word (8 unsigned 4-bit codes, nibble order k = 0,2,4,6,1,3,5,7) -> 4 half2 = k (0,1) (2,3) (4,5) (6,7)
__device__ __forceinline__ void dequant_word(uint32_t w, __half2 bias, __half2 sc, __half2 zp, uint32_t (&o)[4]) {
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    uint32_t t = ((w >> (4 * s)) & 0x000F000Fu) | 0x64006400u;
    __half2 h = __hfma2(__hsub2(*reinterpret_cast<__half2*>(&t), bias), sc, zp);
    o[s] = *reinterpret_cast<uint32_t*>(&h);
  }
}

// Problems are routed here by DevProblem::path == 1 (few rows per expert).  16 activation rows are held at a time
// (more rows re-stream the slab, which then comes from L2); weights AND activation fragments of chunk kc+128 are
// in flight while chunk kc is dequantized and multiplied.
__global__ void __launch_bounds__(WS_WARPS * 32, 3) w4a16_stream_kernel(const DevProblem* __restrict__ probs, int P,
                                                                     const int32_t* __restrict__ task_prefix /*[P+1] slabs*/) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int task = blockIdx.x * WS_WARPS + (threadIdx.x >> 5);
  if (task >= task_prefix[P]) return;
  int p = 0;
  while (task >= task_prefix[p + 1]) ++p;
  const DevProblem& d = probs[p];
  if (d.path != 1) return;
  int M = d.M, row0 = 0;
  if (d.m_range != nullptr) {
    row0 = d.m_range[0];
    M = d.m_range[1] - row0;
  }
  if (M <= 0) return;
  const int N = d.N, K = d.K;
  const int n0 = (task - task_prefix[p]) * WS_ROWS;
  const int64_t lda = d.lda, ldc = d.ldc;
  const __half* A = static_cast<const __half*>(d.A_ptr) + static_cast<int64_t>(row0) * lda;
  const uint8_t* B = static_cast<const uint8_t*>(d.B_ptr);
  const __half* sb = d.scale_b;
  const bool sym = d.sym != 0;
  const int gsz = d.gsize > 0 ? d.gsize : K;  // 128 or per-channel
  const __half2 bias = __float2half2_rn(sym ? 1031.f : 1024.f);
  __half* C = d.C + static_cast<int64_t>(row0) * ldc;

  for (int m0 = 0; m0 < M; m0 += 16) {  // normally one pass
    float acc[4][4];
#pragma unroll
    for (int b = 0; b < 4; ++b)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[b][c] = 0.f;
    const int r0 = m0 + g, r1 = r0 + 8;
    const __half* a0p = A + static_cast<int64_t>(r0) * lda + t * 32;
    const __half* a1p = A + static_cast<int64_t>(r1) * lda + t * 32;
    uint4 wnext[4], alo_n[4], ahi_n[4];
    auto fetch = [&](int kc) {
#pragma unroll
      for (int ng = 0; ng < 4; ++ng) {
        const int n = n0 + ng * 8 + g;
        wnext[ng] = make_uint4(0, 0, 0, 0);
        if (n < N && kc < K) wnext[ng] = ldg_stream(B + static_cast<int64_t>(n) * (K >> 1) + (kc >> 1) + t * 16);
      }
#pragma unroll
      for (int wd = 0; wd < 4; ++wd) {
        alo_n[wd] = ahi_n[wd] = make_uint4(0, 0, 0, 0);
        if (kc < K) {
          if (r0 < M) alo_n[wd] = *reinterpret_cast<const uint4*>(a0p + kc + wd * 8);
          if (r1 < M) ahi_n[wd] = *reinterpret_cast<const uint4*>(a1p + kc + wd * 8);
        }
      }
    };
    fetch(0);
    __half2 sc[4], zp[4];
    for (int kc = 0; kc < K; kc += 128) {
      uint4 w[4], alo[4], ahi[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        w[i] = wnext[i];
        alo[i] = alo_n[i];
        ahi[i] = ahi_n[i];
      }
      fetch(kc + 128);
      if (kc % gsz == 0) {
        const int gi = kc / gsz;
#pragma unroll
        for (int ng = 0; ng < 4; ++ng) {
          const int n = n0 + ng * 8 + g;
          sc[ng] = zp[ng] = __float2half2_rn(0.f);
          if (n < N) {
            if (sym) sc[ng] = __half2half2(sb[static_cast<int64_t>(gi) * N + n]);
            else {
              const __half2 v = *reinterpret_cast<const __half2*>(sb + (static_cast<int64_t>(gi) * N + n) * 2);
              sc[ng] = __half2half2(__low2half(v));
              zp[ng] = __half2half2(__high2half(v));
            }
          }
        }
      }
      const uint32_t* ww[4] = {&w[0].x, &w[1].x, &w[2].x, &w[3].x};
#pragma unroll
      for (int wd = 0; wd < 4; ++wd) {  // word wd of the 16 bytes: k = kc + t*32 + wd*8 .. +8
#pragma unroll
        for (int ng = 0; ng < 4; ++ng) {
          uint32_t bq[4];
          dequant_word(ww[ng][wd], bias, sc[ng], zp[ng], bq);
          mma16816(acc[ng], alo[wd].x, ahi[wd].x, alo[wd].y, ahi[wd].y, bq[0], bq[1]);
          mma16816(acc[ng], alo[wd].z, ahi[wd].z, alo[wd].w, ahi[wd].w, bq[2], bq[3]);
        }
      }
    }
    // ---- fp16 output: rows g / g+8, columns n0 + ng*8 + 2t, +1
#pragma unroll
    for (int ng = 0; ng < 4; ++ng) {
      const int col = n0 + ng * 8 + 2 * t;
      if (col < N) {
        if (r0 < M) *reinterpret_cast<__half2*>(C + static_cast<int64_t>(r0) * ldc + col) = __floats2half2_rn(acc[ng][0], acc[ng][1]);
        if (r1 < M) *reinterpret_cast<__half2*>(C + static_cast<int64_t>(r1) * ldc + col) = __floats2half2_rn(acc[ng][2], acc[ng][3]);
      }
    }
  }
}

}  // namespace

void w4a16_stream_launch(const DevProblem* probs, int P, const int32_t* task_prefix_dev, int total_tasks, int classes, cudaStream_t st) {
  if (total_tasks <= 0) return;
  const int grid = (total_tasks + WS_WARPS - 1) / WS_WARPS;
  (void)classes;
  w4a16_stream_kernel<<<grid, WS_WARPS * 32, 0, st>>>(probs, P, task_prefix_dev);
}

int w4a16_stream_rows_per_task() { return WS_ROWS; }

}  // namespace mxb
