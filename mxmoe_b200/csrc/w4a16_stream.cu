// Piece (4) of the hot path: weight-only w4a16 for LOW-COUNT experts (decode, M of a few rows) — HBM streaming.
//
// Reference: cta_gemm_multistage_qb_v2 / cta_gemm_wxa16g128 (mxmoe/kernels/src/include/cta_gemm.cuh:113-421) with
// 16-row tiles, dequant in registers (quantize.cuh:134-214), used for the memory-bound experts of the 512-token
// and decode workloads.  For M << 128 the job is to move every 4-bit weight from HBM exactly once with as few
// instructions per byte as possible; the tensor-core path would spend a 128-row UMMA tile and a shared-memory
// unpack per weight tile.
//
// B200 design:
//   * a CTA (12 warps, one per SM) takes 384 weight rows of ONE problem; the problem's <= 16 activation rows are
//     staged once in shared memory;
//   * every warp streams 16-row slabs of the packed weights through its OWN ring of shared-memory stages filled by
//     TMA (one 16-row x 128-byte box per stage, 128-byte swizzle, mbarrier transaction counts): no register
//     staging, no cross-warp synchronisation, up to 16 KB in flight per warp.  (A first version used one 256-byte
//     cp.async.bulk per row: the per-copy cost of the TMA unit capped it at ~1.2 TB/s.)
//   * the WEIGHTS are the m16 operand of mma.sync.m16n8k16 and the tokens the n8 operand, so one MMA consumes
//     8 weights per lane and the token fragment is ONE 16-byte shared load per 32-bit weight word;
//   * dequantisation is the reference's arithmetic (0x6400 magic, q - bias exactly, w = hfma(q, scale, zp): one fp16
//     rounding); the high nibbles are taken without a shift: (w & 0x00F000F0 | 0x6400) * 1/16 - (64 + bias) is exact;
//   * the dot product does not care in which order k is visited: a lane's 8 nibbles of a word are its k-slots of
//     two MMAs and the token fragment is read with the same k order.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "gg_kernel.cuh"
#include "ptx.cuh"

namespace mxb {
namespace {

using namespace ptx;

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
constexpr int WS_WARPS = 12;
constexpr int WS_SLAB = 16;                       // weight rows per slab (the m16 of the MMA)
constexpr int WS_SLABS_PER_CTA = 24;              // 384 weight rows per CTA task (two slabs per warp)
constexpr int WS_STAGE_K = 256;                   // k per stage: 128 bytes of every row = one 2 KB TMA box (16 rows)
constexpr int WS_ROW_PITCH = 128;                 // TMA SWIZZLE_128B: 16-byte chunk c of row r sits at chunk c ^ (r & 7)
constexpr int WS_STAGE_BYTES = WS_SLAB * WS_ROW_PITCH;
constexpr int WS_MAX_STAGES = 8;

// one 32-bit word (8 unsigned 4-bit codes, nibble order k = 0,2,4,6,1,3,5,7) -> 4 half2 = k (0,1) (2,3) (4,5) (6,7)
//   lo_bias = 1024 + bias, hi_bias = 64 + bias (bias 7 for symmetric codes, 0 for asymmetric)
__device__ __forceinline__ void dequant_word(uint32_t w, __half2 lo_bias, __half2 hi_bias, __half2 sc, __half2 zp, uint32_t (&o)[4]) {
  const __half2 sixteenth = __float2half2_rn(0.0625f);
  const uint32_t w8 = w >> 8;
  uint32_t t0 = (w & 0x000F000Fu) | 0x64006400u, t1 = (w & 0x00F000F0u) | 0x64006400u;
  uint32_t t2 = (w8 & 0x000F000Fu) | 0x64006400u, t3 = (w8 & 0x00F000F0u) | 0x64006400u;
  __half2 q0 = __hsub2(*reinterpret_cast<__half2*>(&t0), lo_bias);
  __half2 q1 = __hfma2(*reinterpret_cast<__half2*>(&t1), sixteenth, __hneg2(hi_bias));
  __half2 q2 = __hsub2(*reinterpret_cast<__half2*>(&t2), lo_bias);
  __half2 q3 = __hfma2(*reinterpret_cast<__half2*>(&t3), sixteenth, __hneg2(hi_bias));
  q0 = __hfma2(q0, sc, zp);
  q1 = __hfma2(q1, sc, zp);
  q2 = __hfma2(q2, sc, zp);
  q3 = __hfma2(q3, sc, zp);
  o[0] = *reinterpret_cast<uint32_t*>(&q0);
  o[1] = *reinterpret_cast<uint32_t*>(&q1);
  o[2] = *reinterpret_cast<uint32_t*>(&q2);
  o[3] = *reinterpret_cast<uint32_t*>(&q3);
}

}  // namespace

namespace {

// Problems are routed here by DevProblem::path == 1 (few rows per expert).  16 activation rows are resident at a time
// (more rows re-stream the weights, which then come from L2).
__global__ void __launch_bounds__(WS_WARPS * 32, 1) w4a16_stream_kernel(const DevProblem* __restrict__ probs, int P,
                                                                     const int32_t* __restrict__ task_prefix /*[P+1] CTA tasks*/,
                                                                     int stages_in) {
  extern __shared__ __align__(1024) uint8_t ws_smem_raw[];
  uint8_t* ws_smem = ws_smem_raw + ((1024u - (smem_u32(ws_smem_raw) & 1023u)) & 1023u);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int stages = stages_in;
  const int task = blockIdx.x;
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
  const int64_t lda = d.lda, ldc = d.ldc;
  const __half* A = static_cast<const __half*>(d.A_ptr) + static_cast<int64_t>(row0) * lda;
  const __half* sb = d.scale_b;
  const bool sym = d.sym != 0;
  const int gsz = d.gsize > 0 ? d.gsize : K;  // 128 or per-channel
  const __half2 lo_bias = __float2half2_rn(sym ? 1031.f : 1024.f), hi_bias = __float2half2_rn(sym ? 71.f : 64.f);
  __half* C = d.C + static_cast<int64_t>(row0) * ldc;

  const int pitchT = K * 2 + 64;                                 // token row pitch: consecutive rows 4 bank groups apart
  uint8_t* tok = ws_smem;
  const int tokb = (16 * pitchT + 1023) & ~1023;                 // the ring is 1024-byte aligned (swizzle atom)
  uint8_t* ring = ws_smem + tokb + warp * stages * WS_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(ws_smem + tokb + WS_WARPS * stages * WS_STAGE_BYTES) + warp * WS_MAX_STAGES;
  if (lane < stages) mbar_init(&full[lane], 1);
  fence_mbar_init();
  __syncwarp();

  // this warp's slabs: slab0 + warp, + WS_WARPS, ... inside the CTA's slab window
  const int slab0 = (task - task_prefix[p]) * WS_SLABS_PER_CTA;
  const int slab_end = min(slab0 + WS_SLABS_PER_CTA, N / WS_SLAB);
  const int nst = (K + WS_STAGE_K - 1) / WS_STAGE_K;               // stages per slab
  const int my_slabs = slab0 + warp < slab_end ? (slab_end - slab0 - warp + WS_WARPS - 1) / WS_WARPS : 0;
  const int total = my_slabs * nst;
  uint32_t used = 0;             // stage uses so far (ring position and mbarrier parity)

  auto issue = [&](int i) {      // stage i of this pass: one TMA box of 16 rows x 128 B (bytes past K/2 are zero-filled)
    if (lane != 0) return;
    const int sl = i / nst, ks = i - sl * nst;
    const int n = (slab0 + warp + sl * WS_WARPS) * WS_SLAB;
    const int slot = (used + i) % stages;  // `used` is the value at the start of the pass
    {
      mbar_expect_tx(&full[slot], WS_STAGE_BYTES);
      tma_load_2d(ring + slot * WS_STAGE_BYTES, &d.tmB, &full[slot], ks * (WS_STAGE_K / 2), n);
    }
  };

  for (int m0 = 0; m0 < M; m0 += 16) {  // normally one pass
    const int mp = min(16, M - m0);
    const bool two = mp > 8;
    __syncthreads();  // previous pass done with the token rows
    const int fill_rows = two ? 16 : 8;  // rows mp..fill_rows-1 are zero; rows 8..15 are never read when mp <= 8
    for (int i = threadIdx.x; i < fill_rows * (K >> 3); i += WS_WARPS * 32) {
      const int r = i / (K >> 3), c = i - r * (K >> 3);
      uint4 v = make_uint4(0, 0, 0, 0);
      if (r < mp) v = *reinterpret_cast<const uint4*>(A + static_cast<int64_t>(m0 + r) * lda + c * 8);
      *reinterpret_cast<uint4*>(tok + r * pitchT + ((c ^ ((c >> 3) & 3)) << 4)) = v;  // unit swizzle: lanes t = 0..3 read distinct bank groups
    }
    const uint32_t used0 = used;
    for (int i = 0; i < min(stages, total); ++i) issue(i);
    __syncthreads();  // token rows visible to every warp

    // four independent accumulator chains per token tile: 16 dependent MMAs per stage on one accumulator exposed the
    // mma.sync latency (ncu: "wait" stalls), the chains are summed when the slab is written
    float acc0[4][4], acc1[4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc0[c][e] = acc1[c][e] = 0.f;
    // scales of stage i are fetched while stage i-1 is multiplied (a global load on the critical path of every stage
    // showed up as long-scoreboard stalls); lane t needs the group of k = k0 + 64 t
    // The raw (scale, zero) words are kept as loaded and only unpacked at the top of the NEXT stage, so nothing
    // waits on the global load inside the stage that issues it.
    __half2 sc0, zp0, sc1, zp1;
    uint32_t nraw0 = 0u, nraw1 = 0u;  // next stage: asym (scale | zero << 16), sym (scale)
    auto load_scales = [&](int sl, int ks) {
      nraw0 = nraw1 = 0u;  // lanes past K (and stages past the end) multiply by zero
      const int kl = ks * WS_STAGE_K + t * 64;
      if (sl >= my_slabs || kl >= K) return;
      const int64_t gi = static_cast<int64_t>(gsz == K ? 0 : kl >> 7) * N + (slab0 + warp + sl * WS_WARPS) * WS_SLAB + g;
      if (sym) {
        nraw0 = __half_as_ushort(sb[gi]);
        nraw1 = __half_as_ushort(sb[gi + 8]);
      } else {
        nraw0 = *reinterpret_cast<const uint32_t*>(sb + gi * 2);
        nraw1 = *reinterpret_cast<const uint32_t*>(sb + (gi + 8) * 2);
      }
    };
    load_scales(0, 0);
    int sl = 0, ks = 0;  // slab of this warp and k-stage inside it (no divisions in the loop)
    for (int i = 0; i < total; ++i) {
      const int n0 = (slab0 + warp + sl * WS_WARPS) * WS_SLAB;
      const int k0 = ks * WS_STAGE_K;
      {
        const __half2 r0 = *reinterpret_cast<const __half2*>(&nraw0), r1 = *reinterpret_cast<const __half2*>(&nraw1);
        sc0 = __half2half2(__low2half(r0));
        sc1 = __half2half2(__low2half(r1));
        zp0 = __half2half2(__high2half(r0));  // zero for symmetric codes (upper half of the raw word is 0)
        zp1 = __half2half2(__high2half(r1));
      }
      const int ks_n = ks + 1 == nst ? 0 : ks + 1, sl_n = ks + 1 == nst ? sl + 1 : sl;
      load_scales(sl_n, ks_n);
      const uint32_t u = used0 + i;
      const int slot = u % stages;
      mbar_wait(&full[slot], (u / stages) & 1);
      const uint8_t* st = ring + slot * WS_STAGE_BYTES;
      // lane t owns the 16-byte chunks 2t, 2t+1 of rows g and g+8 (k = k0 + 64 t .. + 64): under the XOR swizzle the
      // eight lanes of a quarter-warp then hit eight distinct bank groups.  K % 128 == 0, so a lane's 64 k are all
      // inside K or all in the zero-filled tail of the box; tail lanes run with zero scales on in-range token addresses.
      const int kl = min(k0 + t * 64, K - 64);
      const uint8_t* tk = tok + g * pitchT;
      const int u0 = kl >> 3;  // first 16-byte token unit of this lane
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int pc = ((2 * t + jj) ^ g) << 4;  // physical chunk of logical chunk 2t+jj in rows g and g+8
        const uint4 wl = *reinterpret_cast<const uint4*>(st + g * WS_ROW_PITCH + pc);
        const uint4 wh = *reinterpret_cast<const uint4*>(st + (g + 8) * WS_ROW_PITCH + pc);
        const uint32_t wlw[4] = {wl.x, wl.y, wl.z, wl.w}, whw[4] = {wh.x, wh.y, wh.z, wh.w};
#pragma unroll
        for (int wd = 0; wd < 4; ++wd) {  // word wd: k = kl + jj*32 + wd*8 .. +8
          uint32_t lo[4], hi[4];
          dequant_word(wlw[wd], lo_bias, hi_bias, sc0, zp0, lo);
          dequant_word(whw[wd], lo_bias, hi_bias, sc1, zp1, hi);
          const int un = u0 + jj * 4 + wd;
          const int off = (un ^ ((un >> 3) & 3)) << 4;  // token rows are stored with the same unit swizzle (see the fill)
          const uint4 b = *reinterpret_cast<const uint4*>(tk + off);
          mma16816(acc0[wd], lo[0], hi[0], lo[1], hi[1], b.x, b.y);
          mma16816(acc0[(wd + 2) & 3], lo[2], hi[2], lo[3], hi[3], b.z, b.w);
          if (two) {
            const uint4 b2 = *reinterpret_cast<const uint4*>(tk + 8 * pitchT + off);
            mma16816(acc1[wd], lo[0], hi[0], lo[1], hi[1], b2.x, b2.y);
            mma16816(acc1[(wd + 2) & 3], lo[2], hi[2], lo[3], hi[3], b2.z, b2.w);
          }
        }
      }
      __syncwarp();  // every lane is done reading the slot
      if (i + stages < total) issue(i + stages);
      if (ks == nst - 1) {
        // ---- fp16 output of this slab: weight rows n0+g / n0+g+8, tokens 2t, 2t+1 (+8)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float a[4];
#pragma unroll
          for (int e = 0; e < 4; ++e)
            a[e] = h ? (acc1[0][e] + acc1[1][e]) + (acc1[2][e] + acc1[3][e]) : (acc0[0][e] + acc0[1][e]) + (acc0[2][e] + acc0[3][e]);
          const int tk0 = h * 8 + 2 * t;
          if (tk0 < mp) {
            C[static_cast<int64_t>(m0 + tk0) * ldc + n0 + g] = __float2half_rn(a[0]);
            C[static_cast<int64_t>(m0 + tk0) * ldc + n0 + g + 8] = __float2half_rn(a[2]);
          }
          if (tk0 + 1 < mp) {
            C[static_cast<int64_t>(m0 + tk0 + 1) * ldc + n0 + g] = __float2half_rn(a[1]);
            C[static_cast<int64_t>(m0 + tk0 + 1) * ldc + n0 + g + 8] = __float2half_rn(a[3]);
          }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc0[c][e] = acc1[c][e] = 0.f;
      }
      ks = ks_n;
      sl = sl_n;
    }
    used = used0 + total;
  }
}

int ws_smem_bytes(int K, int* stages_out) {
  const int tokb = (16 * (K * 2 + 64) + 1023) & ~1023;
  int stages = (227 * 1024 - 2048 - tokb - WS_WARPS * WS_MAX_STAGES * 8) / (WS_WARPS * WS_STAGE_BYTES);
  if (stages > WS_MAX_STAGES) stages = WS_MAX_STAGES;
  *stages_out = stages;
  return tokb + WS_WARPS * stages * WS_STAGE_BYTES + WS_WARPS * WS_MAX_STAGES * 8 + 1024;
}

}  // namespace

// max_k: largest K among the problems routed here (sizes the token staging and the ring depth)
void w4a16_stream_launch(const DevProblem* probs, int P, const int32_t* task_prefix_dev, int total_tasks, int max_k, cudaStream_t st) {
  if (total_tasks <= 0) return;
  int stages = 0;
  const int smem = ws_smem_bytes(max_k, &stages);
  static int configured = 0;
  if (configured < smem) {
    cudaFuncSetAttribute(w4a16_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    configured = 227 * 1024;
  }
  w4a16_stream_kernel<<<total_tasks, WS_WARPS * 32, smem, st>>>(probs, P, task_prefix_dev, stages);
}

int w4a16_stream_rows_per_task() { return WS_SLABS_PER_CTA * WS_SLAB; }
int w4a16_stream_max_k() { return 4096; }

}  // namespace mxb
