// Piece (4) of the hot path: weight-only w4a16 for LOW-COUNT experts (decode, M of a few rows) — HBM streaming.
//
// Reference: cta_gemm_multistage_qb_v2 / cta_gemm_wxa16g128 (mxmoe/kernels/src/include/cta_gemm.cuh:113-421) with
// 16-row tiles, dequant in registers (quantize.cuh:134-214), used for the memory-bound experts of the 512-token
// and decode workloads.  For M << 128 the job is to move every 4-bit weight from HBM exactly once with as few
// instructions per byte as possible; the tensor-core path would spend a 128-row UMMA tile and a shared-memory
// unpack per weight tile.
//
// B200 design:
//   * PERSISTENT: one CTA per SM (16 warps) owns a CONTIGUOUS run of the launch's (expert, 16-token pass, 16-row
//     slab) work items, cut at slab granularity in a cost space (K-stages per slab) every CTA derives from the
//     device-resident row counts, so a CTA mostly stays inside one expert: its <= 16 activation rows are staged in
//     shared memory once per (expert, pass) with cp.async, and the weight pipeline never drains between slabs, passes
//     or experts;
//   * every warp streams 16-row slabs of the packed weights through its OWN ring of shared-memory stages filled by
//     TMA (one 16-row x 128-byte box per stage, 128-byte swizzle, mbarrier transaction counts): no register
//     staging, no cross-warp synchronisation.  Lane 0 of the warp is its producer and runs `stages` boxes ahead of
//     the warp, across slab / pass / expert boundaries, from a small list of (expert, row-group range) entries the
//     CTA builds in shared memory.  (A first version used one 256-byte cp.async.bulk per row: the per-copy cost of
//     the TMA unit capped it at ~1.2 TB/s; the second one launched one CTA per 384 rows and paid the token fill, the
//     problem search and the pipeline fill 4.7 times per SM.)
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
#include "quant_common.cuh"

namespace mxb {
namespace {

using namespace ptx;

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

constexpr int WS_WARPS = 16;
constexpr int WS_SLAB = 16;                       // weight rows per slab (the m16 of the MMA)
constexpr int WS_STAGE_K = 256;                   // k per stage: 128 bytes of every row = one 2 KB TMA box (16 rows)
constexpr int WS_ROW_PITCH = 128;                 // TMA SWIZZLE_128B: 16-byte chunk c of row r sits at chunk c ^ (r & 7)
constexpr int WS_STAGE_BYTES = WS_SLAB * WS_ROW_PITCH;
constexpr int WS_MAX_STAGES = 8;
constexpr int WS_MAX_P = 1024;                   // problems per launch (the unit prefix lives in shared memory)
constexpr int WS_MAX_ENT = 64;                    // (expert, group range) entries per list build

// one entry of the CTA's work list: slabs [sb, se) of problem p for the 16-token pass starting at row m0; warp w takes
// the slabs sb + ((w - woff) mod WS_WARPS) + j * WS_WARPS (woff continues the round-robin across entries)
struct __align__(16) WsEntry {
  int p, m0, M, row0;
  int sb, se, woff, nst;
};

// one 32-bit word (8 unsigned 4-bit codes, nibble order k = 0,2,4,6,1,3,5,7) -> 4 half2 = k (0,1) (2,3) (4,5) (6,7)
//   lo_bias = 1024 + bias, hi_bias = 64 + bias (bias 7 for symmetric codes, 0 for asymmetric)
//   magic = 0x64006400 arrives as a kernel argument: as a literal next to the literal mask, (w & mask) | magic costs two
//   LOP3 (one immediate per instruction); from a register it is one
__device__ __forceinline__ void dequant_word(uint32_t w, uint32_t magic, __half2 lo_bias, __half2 nhi_bias, __half2 sc, __half2 zp,
                                             uint32_t (&o)[4]) {
  const __half2 sixteenth = __float2half2_rn(0.0625f);
  const uint32_t w8 = w >> 8;
  uint32_t t0 = (w & 0x000F000Fu) | magic, t1 = (w & 0x00F000F0u) | magic;
  uint32_t t2 = (w8 & 0x000F000Fu) | magic, t3 = (w8 & 0x00F000F0u) | magic;
  __half2 q0 = __hsub2(*reinterpret_cast<__half2*>(&t0), lo_bias);
  __half2 q1 = __hfma2(*reinterpret_cast<__half2*>(&t1), sixteenth, nhi_bias);
  __half2 q2 = __hsub2(*reinterpret_cast<__half2*>(&t2), lo_bias);
  __half2 q3 = __hfma2(*reinterpret_cast<__half2*>(&t3), sixteenth, nhi_bias);
  q0 = __hfma2(q0, sc, zp);
  q1 = __hfma2(q1, sc, zp);
  q2 = __hfma2(q2, sc, zp);
  q3 = __hfma2(q3, sc, zp);
  o[0] = *reinterpret_cast<uint32_t*>(&q0);
  o[1] = *reinterpret_cast<uint32_t*>(&q1);
  o[2] = *reinterpret_cast<uint32_t*>(&q2);
  o[3] = *reinterpret_cast<uint32_t*>(&q3);
}

// cost weights of a problem's passes: a K-stage of a pass with more than 8 tokens (two n8 MMA tiles per weight fragment)
// takes WS_C2, one with at most 8 tokens WS_C1.  Measured per CTA (tools/ws_profile.py, c3): 1.8 us against 1.12 us per
// stage of 16 slabs.  Every pass but the last holds 16 tokens.
#ifndef WS_C1
#define WS_C1 5
#define WS_C2 8
#endif
__device__ __forceinline__ int ws_cost(int np, int M) { return np > 0 ? WS_C2 * (np - 1) + (M - 16 * (np - 1) > 8 ? WS_C2 : WS_C1) : 0; }
// first item (pass-major, slab-minor) that starts at or after cost offset c inside a problem
__device__ __forceinline__ int ws_item_of(int c, int M, int nslab, int nst) {
  const int np = (M + 15) >> 4;
  const int full = WS_C2 * (np - 1) * nslab * nst;  // cost of the 16-token passes before the last one
  if (c <= full) return (c + WS_C2 * nst - 1) / (WS_C2 * nst);
  const int wl = (M - 16 * (np - 1) > 8 ? WS_C2 : WS_C1) * nst;
  return (np - 1) * nslab + (c - full + wl - 1) / wl;
}
// The cost space is cut in units of WS_SPLIT slabs of one pass.  WS_SPLIT == WS_WARPS ("rounds": one slab for every warp)
// keeps every entry a whole number of rounds, so no warp idles at the barrier between two entries; WS_SPLIT == 1 balances
// the CTAs to one slab but an entry of s slabs then takes ceil(s / 16) slab times.
#ifndef WS_SPLIT
#define WS_SPLIT WS_WARPS
#endif
__device__ __forceinline__ int ws_split_units(int nslab) { return (nslab + WS_SPLIT - 1) / WS_SPLIT; }
__device__ __forceinline__ int ws_first_slab(const WsEntry& e, int warp) { return e.sb + ((warp - e.woff) & (WS_WARPS - 1)); }

// Producer cursor of one warp (only lane 0 uses it): the next stage to request.  The entry's constants ride along in
// registers, so a refill touches shared memory only when the cursor moves to another entry.
struct WsProd {
  int i, slab, ks;
  bool live;
  int desc_p;  // problem whose weight descriptor this lane acquired last (tma_desc_acquire, ptx.cuh)
  int p, se, nst;
};

// first entry at or after c.i this warp has a slab in
__device__ __forceinline__ void ws_prod_seek(WsProd& c, const WsEntry* ent, int n_ent, int warp) {
  for (; c.i < n_ent; ++c.i) {
    const WsEntry e = ent[c.i];
    const int first = ws_first_slab(e, warp);
    if (first < e.se) {
      c.slab = first;
      c.ks = 0;
      c.p = e.p;
      c.se = e.se;
      c.nst = e.nst;
      c.live = true;
      return;
    }
  }
  c.live = false;
}
__device__ __forceinline__ void ws_prod_advance(WsProd& c, const WsEntry* ent, int n_ent, int warp) {
  if (++c.ks < c.nst) return;
  c.ks = 0;
  c.slab += WS_WARPS;
  if (c.slab < c.se) return;
  ++c.i;
  ws_prod_seek(c, ent, n_ent, warp);
}
// request the cursor's stage into ring slot `slot` and move on
__device__ __forceinline__ void ws_prod_issue(WsProd& c, const DevProblem* probs, uint8_t* ring, uint64_t* full, uint32_t slot,
                                              const WsEntry* ent, int n_ent, int warp) {
  if (c.p != c.desc_p) {
    tma_desc_acquire(&probs[c.p].tmB);
    c.desc_p = c.p;
  }
  mbar_expect_tx(&full[slot], WS_STAGE_BYTES);
  tma_load_2d(ring + slot * WS_STAGE_BYTES, &probs[c.p].tmB, &full[slot], c.ks * (WS_STAGE_K / 2), c.slab * WS_SLAB);
  ws_prod_advance(c, ent, n_ent, warp);
}

struct WsProblem {  // per-entry constants of the consumer
  const __half* A;
  __half* C;
  const __half* sb;
  int64_t lda, ldc;
  int N, K, gsz;
  bool sym, epi_silu;
};

// One entry (16-token pass of one expert) for one warp: its slabs of [sb, se), K in 256-wide stages.
template <bool TWO>
__device__ __forceinline__ void ws_run_pass(const WsProblem& pb, const WsEntry& e, int mp, const uint8_t* tok, int pitchT,
                                            const uint8_t* ring, uint64_t* full, int stages, uint32_t& cslot, uint32_t& cphase, WsProd& prod,
                                            const DevProblem* probs, const WsEntry* ent, int n_ent, int warp, int lane, uint32_t magic) {
  const int g = lane >> 2, t = lane & 3;
  const int N = pb.N, K = pb.K;
  const __half2 lo_bias = __float2half2_rn(pb.sym ? 1031.f : 1024.f), nhi_bias = __float2half2_rn(pb.sym ? -71.f : -64.f);
  // loop-invariant lane offsets: weight chunks 2t, 2t+1 of rows g (and g+8: +1024) under the TMA 128-byte swizzle; token
  // units c = 0..7 of the lane's 128-byte k-block under the fill's unit swizzle (u ^ ((u >> 3) & 3), and (u >> 3) & 3 == t
  // for the unit block of lane t)
  const int wo0 = g * WS_ROW_PITCH + (((2 * t) ^ g) << 4), wo1 = g * WS_ROW_PITCH + (((2 * t + 1) ^ g) << 4);
  int x[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) x[c] = g * pitchT + ((c ^ t) << 4);
  const int pitch8 = 8 * pitchT;

  float acc0[4][4], acc1[TWO ? 4 : 1][4];
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      acc0[c][q] = 0.f;
      if (TWO) acc1[TWO ? c : 0][q] = 0.f;
    }
  // scales of the next stage are fetched while the current one is multiplied; lane t needs the group of k = k0 + 64 t.
  // The raw (scale, zero) words are kept as loaded and unpacked at the top of the NEXT stage.  `sp` walks the lane's
  // (scale, zero) words: + one stage (two groups) inside a slab, back to group t >> 1 of the warp's next slab after it.
  uint32_t nraw0 = 0u, nraw1 = 0u;  // asym (scale | zero << 16), sym (scale)
  const int esh = pb.sym ? 1 : 2;   // log2 bytes per (row, group) entry
  const int64_t st_stride = pb.gsz == K ? 0 : (static_cast<int64_t>(2 * N) << esh);
  const int64_t slab_step = (static_cast<int64_t>(WS_WARPS * WS_SLAB) << esh) - (e.nst - 1) * st_stride;
  const int slab0 = ws_first_slab(e, warp);
  const uint8_t* sp = reinterpret_cast<const uint8_t*>(pb.sb) +
                      ((static_cast<int64_t>(pb.gsz == K ? 0 : t >> 1) * N + slab0 * WS_SLAB + g) << esh);
  int kl_next = t * 64;  // first k of the lane's share of the stage `sp` points at
  auto load_scales = [&]() {  // (not __ldg: nvcc speculates it past the bounds this lambda is called under)
    nraw0 = nraw1 = 0u;       // lanes past K multiply by zero
    if (kl_next < K) {
      if (pb.sym) {
        nraw0 = ld_nc_u16(reinterpret_cast<const __half*>(sp));
        nraw1 = ld_nc_u16(reinterpret_cast<const __half*>(sp) + 8);
      } else {
        nraw0 = ld_nc_u32(reinterpret_cast<const uint32_t*>(sp));
        nraw1 = ld_nc_u32(reinterpret_cast<const uint32_t*>(sp) + 8);
      }
    }
  };
  if (slab0 < e.se) load_scales();
  for (int slab = slab0; slab < e.se; slab += WS_WARPS) {
    const int n0 = slab * WS_SLAB;
    for (int ks = 0; ks < e.nst; ++ks) {
      const __half2 r0 = *reinterpret_cast<const __half2*>(&nraw0), r1 = *reinterpret_cast<const __half2*>(&nraw1);
      const __half2 sc0 = __half2half2(__low2half(r0)), sc1 = __half2half2(__low2half(r1));
      const __half2 zp0 = __half2half2(__high2half(r0)), zp1 = __half2half2(__high2half(r1));  // zero for symmetric codes
      {
        const bool in_slab = ks + 1 < e.nst;
        sp += in_slab ? st_stride : slab_step;
        kl_next = in_slab ? kl_next + WS_STAGE_K : t * 64;
        if (in_slab || slab + WS_WARPS < e.se) load_scales();
      }
      mbar_wait(&full[cslot], cphase);
      const uint8_t* st = ring + cslot * WS_STAGE_BYTES;
      // lane t owns the 16-byte chunks 2t, 2t+1 of rows g and g+8 (k = k0 + 64 t .. + 64).  K % 128 == 0, so a lane's
      // 64 k are all inside K or all in the zero-filled tail of the box; tail lanes run with zero scales on in-range
      // token addresses.
      const int kl = min(ks * WS_STAGE_K + t * 64, K - 64);
      const uint8_t* tkb = tok + (kl << 1);
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const uint4 wl = *reinterpret_cast<const uint4*>(st + (jj ? wo1 : wo0));
        const uint4 wh = *reinterpret_cast<const uint4*>(st + (jj ? wo1 : wo0) + 8 * WS_ROW_PITCH);
        const uint32_t wlw[4] = {wl.x, wl.y, wl.z, wl.w}, whw[4] = {wh.x, wh.y, wh.z, wh.w};
#pragma unroll
        for (int wd = 0; wd < 4; ++wd) {  // word wd: k = kl + jj*32 + wd*8 .. +8
          uint32_t lo[4], hi[4];
          dequant_word(wlw[wd], magic, lo_bias, nhi_bias, sc0, zp0, lo);
          dequant_word(whw[wd], magic, lo_bias, nhi_bias, sc1, zp1, hi);
          const uint8_t* ta = tkb + x[wd] + jj * 64;
          const uint4 b = *reinterpret_cast<const uint4*>(ta);
          mma16816(acc0[wd], lo[0], hi[0], lo[1], hi[1], b.x, b.y);
          mma16816(acc0[(wd + 2) & 3], lo[2], hi[2], lo[3], hi[3], b.z, b.w);
          if (TWO) {
            const uint4 b2 = *reinterpret_cast<const uint4*>(ta + pitch8);
            mma16816(acc1[TWO ? wd : 0], lo[0], hi[0], lo[1], hi[1], b2.x, b2.y);
            mma16816(acc1[TWO ? (wd + 2) & 3 : 0], lo[2], hi[2], lo[3], hi[3], b2.z, b2.w);
          }
        }
      }
      __syncwarp();  // every lane is done reading the slot
      // refill the slot just freed: the producer's next slot IS this one (ring order; the slots requested so far are
      // exactly the ones consumed so far plus the ring depth)
      if (lane == 0 && prod.live) ws_prod_issue(prod, probs, const_cast<uint8_t*>(ring), full, cslot, ent, n_ent, warp);
      if (++cslot == static_cast<uint32_t>(stages)) {
        cslot = 0;
        cphase ^= 1u;
      }
    }
    // ---- fp16 output of this slab: weight rows n0+g / n0+g+8, tokens 2t, 2t+1 (+8)
    __half* C = pb.C + static_cast<int64_t>(e.m0) * pb.ldc + n0 + g;
#pragma unroll
    for (int h = 0; h < (TWO ? 2 : 1); ++h) {
      float a[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        a[q] = h ? (acc1[TWO ? 0 : 0][q] + acc1[TWO ? 1 : 0][q]) + (acc1[TWO ? 2 : 0][q] + acc1[TWO ? 3 : 0][q])
                 : (acc0[0][q] + acc0[1][q]) + (acc0[2][q] + acc0[3][q]);
      }
      const int tk0 = h * 8 + 2 * t;
      if (pb.epi_silu) {
        // MXB_EPI_SILU_MUL: weight rows are interleaved (row 2i = gate_i, row 2i+1 = up_i), so lane g (even) holds gate and
        // lane g+1 (four lanes up) the matching up values.  Both are rounded to fp16 first, like the two-step path that
        // stores [gate | up] and applies act_fn(gate) * up on the stored values; output column (n0 + g) / 2 (+4).
        float o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const __half mine = __float2half_rn(a[q]);
          const __half other = __shfl_xor_sync(0xffffffffu, mine, 4);
          o[q] = silu_mul_h(mine, other);
        }
        if ((g & 1) == 0) {
          __half* Ch = pb.C + static_cast<int64_t>(e.m0) * pb.ldc + ((n0 + g) >> 1);
          if (tk0 < mp) {
            Ch[static_cast<int64_t>(tk0) * pb.ldc] = __float2half_rn(o[0]);
            Ch[static_cast<int64_t>(tk0) * pb.ldc + 4] = __float2half_rn(o[2]);
          }
          if (tk0 + 1 < mp) {
            Ch[static_cast<int64_t>(tk0 + 1) * pb.ldc] = __float2half_rn(o[1]);
            Ch[static_cast<int64_t>(tk0 + 1) * pb.ldc + 4] = __float2half_rn(o[3]);
          }
        }
      } else {
        if (tk0 < mp) {
          C[static_cast<int64_t>(tk0) * pb.ldc] = __float2half_rn(a[0]);
          C[static_cast<int64_t>(tk0) * pb.ldc + 8] = __float2half_rn(a[2]);
        }
        if (tk0 + 1 < mp) {
          C[static_cast<int64_t>(tk0 + 1) * pb.ldc] = __float2half_rn(a[1]);
          C[static_cast<int64_t>(tk0 + 1) * pb.ldc + 8] = __float2half_rn(a[3]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        acc0[c][q] = 0.f;
        if (TWO) acc1[TWO ? c : 0][q] = 0.f;
      }
  }
}


// Problems are routed here by DevProblem::path == 1 (few rows per expert).  16 activation rows are resident at a time
// (more rows re-stream the CTA's slabs, which then come from L2).
//
// Work split: an item = (problem, 16-token pass, 16-row slab), items ordered problem-major, pass, slab; an item costs the
// problem's K-stages x 3, or x 4 when the pass holds more than 8 tokens (two n8 MMA tiles per weight fragment: measured
// 1.33x per stage).  Every CTA computes the same cost prefix over the problems from the DEVICE-resident row counts
// (like the tcgen05 kernels build their tile prefix) and takes the items of an equal slice of the total cost, so empty
// experts cost nothing, an expert with 40 rows weighs three passes, K = 2816 weighs twice K = 1408, and the static
// contiguous split stays balanced to one slab whatever the routing was.
#ifdef WS_PROFILE
// debug build only (make WS_PROFILE=1): per-CTA timeline of the last gate_up (slot 0, K == 2048) and down (slot 1) launch
__device__ unsigned long long ws_prof[2][160][8];
__device__ __forceinline__ unsigned long long ws_gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#endif

__global__ void __launch_bounds__(WS_WARPS * 32, 1) w4a16_stream_kernel(const DevProblem* __restrict__ probs, int P, int stages,
                                                                     int tok_bytes, uint32_t magic) {
  extern __shared__ __align__(1024) uint8_t ws_smem_raw[];
  uint8_t* ws_smem = ws_smem_raw + ((1024u - (smem_u32(ws_smem_raw) & 1023u)) & 1023u);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef WS_PROFILE
  const unsigned long long pf_t0 = ws_gtime();
  unsigned long long pf_fill = 0, pf_ent = 0, pf_two = 0, pf_later = 0, pf_list = 0;
#endif
  uint8_t* tok = ws_smem;
  uint8_t* ring = ws_smem + tok_bytes + warp * stages * WS_STAGE_BYTES;
  uint8_t* tail = ws_smem + tok_bytes + WS_WARPS * stages * WS_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(tail) + warp * WS_MAX_STAGES;
  WsEntry* ent = reinterpret_cast<WsEntry*>(tail + WS_WARPS * WS_MAX_STAGES * 8);
  int* ctl = reinterpret_cast<int*>(ent + WS_MAX_ENT);  // [0] entries in the list, [1] next problem, [2] list is the last one
  int* cost_prefix = ctl + 4;                           // [P+1]
  int* rows_of = cost_prefix + P + 1;                     // [P] rows of every problem (as read for the prefix)
  int* row0_of = rows_of + P;                             // [P] first row (m_range[0]); the list build reads no global memory
  int* shape_of = row0_of + P;                            // [P] slabs << 8 | K-stages
  if (lane < stages) mbar_init(&full[lane], 1);
  fence_mbar_init();

  for (int i = threadIdx.x; i < P; i += WS_WARPS * 32) {
    const DevProblem& d = probs[i];
    int M = d.M, r0 = 0;
    if (d.m_range != nullptr) {
      r0 = d.m_range[0];
      M = d.m_range[1] - r0;
    }
    const int np = (d.path == 1 && M > 0) ? (M + 15) >> 4 : 0;
    const int nst = (d.K + WS_STAGE_K - 1) / WS_STAGE_K;  // <= 16 (K <= 4096)
    rows_of[i] = np > 0 ? M : 0;
    row0_of[i] = r0;
    shape_of[i] = ((d.N / WS_SLAB) << 8) | nst;
    cost_prefix[i + 1] = ws_cost(np, M) * ws_split_units(d.N / WS_SLAB) * nst;
  }
  __syncthreads();
  if (warp == 0) {  // inclusive scan, 32 problems per step
    int carry = 0;
    for (int base = 0; base < P; base += 32) {
      int v = base + lane < P ? cost_prefix[base + lane + 1] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      if (base + lane < P) cost_prefix[base + lane + 1] = v + carry;
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
    if (lane == 0) cost_prefix[0] = 0;
  }
  __syncthreads();
  // this CTA's contiguous share of the cost
  const int U = cost_prefix[P];
  const int u_lo = static_cast<int>(static_cast<int64_t>(blockIdx.x) * U / gridDim.x);
  const int u_hi = static_cast<int>(static_cast<int64_t>(blockIdx.x + 1) * U / gridDim.x);
  if (u_lo >= u_hi) return;
#ifdef WS_PROFILE
  const unsigned long long pf_t1 = ws_gtime();
#endif

  int p_next = 0;
  {  // last problem whose cost range starts at or before u_lo
    int lo = 0, hi = P;  // invariant: cost_prefix[lo] <= u_lo < cost_prefix[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cost_prefix[mid] <= u_lo) lo = mid;
      else hi = mid;
    }
    p_next = lo;
  }
  uint32_t cslot = 0, cphase = 0;

  for (;;) {
    __syncthreads();  // the previous list is fully consumed
    if (warp == 0) {
      // One lane per problem: the items [ia, ib) of the problem inside this CTA's cost slice (a boundary inside an item
      // rounds up, identically in both neighbours), one entry per pass they touch.
      int n = 0, base = p_next, nxt = -1, slabs_before = 0;
      bool last = false;
      while (n < WS_MAX_ENT && !last) {
        const int p = base + lane;
        const bool inr = p < P && cost_prefix[min(p, P)] < u_hi;
        int ia = 0, ib = 0, nslab = 1, nslab_real = 1, nst = 1, M = 0, row0 = 0;  // nslab: split units per pass (WS_SPLIT slabs each)
        if (inr) {
          const int pre = cost_prefix[p], pre1 = cost_prefix[p + 1];
          if (pre1 > pre) {
            nst = shape_of[p] & 0xff;
            nslab_real = shape_of[p] >> 8;
            nslab = ws_split_units(nslab_real);
            M = rows_of[p];
            row0 = row0_of[p];
            // (DevProblem::c_row is not honoured here: api.cu keeps such problems off this path)
            ia = ws_item_of(max(u_lo, pre) - pre, M, nslab, nst);
            ib = ws_item_of(min(u_hi, pre1) - pre, M, nslab, nst);
          }
        }
        const bool has = ib > ia;
        const int pa = ia / nslab, pb = has ? (ib - 1) / nslab : pa;
        const int cnt = has ? pb - pa + 1 : 0;
        int nsl = 0;  // slabs of this problem's entries
        for (int ps = pa; has && ps <= pb; ++ps)
          nsl += min((ps == pb ? ib - pb * nslab : nslab) * WS_SPLIT, nslab_real) - (ps == pa ? ia - pa * nslab : 0) * WS_SPLIT;
        int pos = cnt, sl = nsl;  // exclusive scans of the entry counts and the slab counts
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int u = __shfl_up_sync(0xffffffffu, pos, o), v = __shfl_up_sync(0xffffffffu, sl, o);
          if (lane >= o) {
            pos += u;
            sl += v;
          }
        }
        const int tot_cnt = __shfl_sync(0xffffffffu, pos, 31), tot_sl = __shfl_sync(0xffffffffu, sl, 31);
        pos = n + pos - cnt;
        sl = slabs_before + sl - nsl;
        const bool fits = pos + cnt <= WS_MAX_ENT;
        if (has && fits) {
          for (int ps = pa; ps <= pb; ++ps) {
            WsEntry e;
            e.p = p;
            e.m0 = ps * 16;
            e.M = M;
            e.row0 = row0;
            e.sb = (ps == pa ? ia - pa * nslab : 0) * WS_SPLIT;
            e.se = min((ps == pb ? ib - pb * nslab : nslab) * WS_SPLIT, nslab_real);
            e.woff = sl & (WS_WARPS - 1);
            e.nst = nst;
            sl += e.se - e.sb;
            ent[pos++] = e;
          }
        }
        const unsigned over = __ballot_sync(0xffffffffu, has && !fits);
        const unsigned im = __ballot_sync(0xffffffffu, inr);
        if (over) {  // list full: resume at the first problem that did not fit (lanes after it did not write: pos is monotone)
          const int first = __ffs(over) - 1;
          nxt = base + first;
          n = __shfl_sync(0xffffffffu, pos, first);
          break;
        }
        n += tot_cnt;
        slabs_before += tot_sl;
        if (im != 0xffffffffu) last = true;
        else base += 32;
      }
      if (lane == 0) {
        ctl[0] = n;
        ctl[1] = nxt >= 0 ? nxt : base;  // `base` has moved past every problem looked at unless `last`
        ctl[2] = (last && nxt < 0) ? 1 : 0;
      }
    }
    __syncthreads();
    const int n_ent = ctl[0];
    const bool last_list = ctl[2] != 0;
    p_next = ctl[1];

    // producer: request the first `stages` boxes of this list
    WsProd prod = {0, 0, 0, false, -1, 0, 0, 0};
    if (lane == 0) {
      ws_prod_seek(prod, ent, n_ent, warp);
      uint32_t pslot = cslot;  // the ring is empty between two lists
      for (int s = 0; s < stages && prod.live; ++s) {
        ws_prod_issue(prod, probs, ring, full, pslot, ent, n_ent, warp);
        if (++pslot == static_cast<uint32_t>(stages)) pslot = 0;
      }
    }

    for (int i = 0; i < n_ent; ++i) {
      const WsEntry e = ent[i];
      const DevProblem& d = probs[e.p];
      WsProblem pb;
      pb.N = d.N;
      pb.K = d.K;
      pb.lda = d.lda;
      pb.ldc = d.ldc;
      const int32_t* a_rows = d.a_rows;
      const bool pro_silu = (d.epi_flags & 2) != 0;  // MXB_PRO_SILU_MUL: A = [gate | up], rows become silu(gate) * up
      pb.epi_silu = (d.epi_flags & 1) != 0;          // MXB_EPI_SILU_MUL: interleaved gate/up weight rows, C = act_fn(gate) * up
      pb.A = static_cast<const __half*>(d.A_ptr) + (a_rows ? 0 : static_cast<int64_t>(e.row0) * pb.lda);
      pb.C = d.C + static_cast<int64_t>(e.row0) * pb.ldc;
      pb.sb = d.scale_b;
      pb.sym = d.sym != 0;
      pb.gsz = d.gsize > 0 ? d.gsize : pb.K;
      const int K = pb.K;
      const int pitchT = K * 2 + 64;  // token row pitch: consecutive rows 4 bank groups apart
      const int mp = min(16, e.M - e.m0);
      const bool two = mp > 8;
      __syncthreads();  // previous entry done with the token rows
#ifdef WS_PROFILE
      const unsigned long long pf_f0 = ws_gtime();
#endif
      const int fill_rows = two ? 16 : 8;  // rows mp..fill_rows-1 are zero; rows 8..15 are never read when mp <= 8
      const int upr = K >> 3;              // 16-byte units per row
      for (int idx = threadIdx.x; idx < fill_rows * upr; idx += WS_WARPS * 32) {
        const int r = idx / upr, c = idx - r * upr;
        uint8_t* dst = tok + r * pitchT + ((c ^ ((c >> 3) & 3)) << 4);  // unit swizzle: lanes t = 0..3 read distinct bank groups
        if (r < mp) {
          // the token gather (a_rows) and act_fn(gate) * up (MXB_PRO_SILU_MUL) happen here, while the rows are staged
          const int64_t ar = a_rows ? a_rows[e.row0 + e.m0 + r] : e.m0 + r;
          const __half* srcp = pb.A + ar * pb.lda + c * 8;
          if (pro_silu) {
            const uint4 gq = ld_global_nc_v4(srcp), uq = ld_global_nc_v4(srcp + K);
            const __half* gh = reinterpret_cast<const __half*>(&gq);
            const __half* uh = reinterpret_cast<const __half*>(&uq);
            uint4 o;
            __half* oh = reinterpret_cast<__half*>(&o);
#pragma unroll
            for (int j = 0; j < 8; ++j) oh[j] = __float2half_rn(silu_mul_h(gh[j], uh[j]));
            *reinterpret_cast<uint4*>(dst) = o;
          } else {
            cp_async16(dst, srcp);
          }
        } else {
          *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
        }
      }
      cp_async_wait_all();
      __syncthreads();  // token rows visible to every warp
#ifdef WS_PROFILE
      pf_fill += ws_gtime() - pf_f0;
      pf_ent += 1;
      pf_two += two ? (e.se - e.sb) * e.nst : 0;
      pf_later += e.m0 > 0 ? (e.se - e.sb) * e.nst : 0;
#endif
      if (two) ws_run_pass<true>(pb, e, mp, tok, pitchT, ring, full, stages, cslot, cphase, prod, probs, ent, n_ent, warp, lane, magic);
      else ws_run_pass<false>(pb, e, mp, tok, pitchT, ring, full, stages, cslot, cphase, prod, probs, ent, n_ent, warp, lane, magic);
    }
#ifdef WS_PROFILE
    pf_list += 1;
#endif
    if (last_list) break;
  }
#ifdef WS_PROFILE
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long* o = ws_prof[probs[0].K == 2048 ? 0 : 1][blockIdx.x];
    o[0] = pf_t0;
    o[1] = pf_t1 - pf_t0;
    o[2] = ws_gtime() - pf_t0;
    o[3] = pf_fill;
    o[4] = pf_ent;
    o[5] = pf_two;
    o[6] = pf_later;
    o[7] = static_cast<unsigned long long>(u_hi - u_lo);
  }
#endif
}

int ws_smem_bytes(int K, int P, int* stages_out, int* tok_bytes_out) {
  const int tokb = (16 * (K * 2 + 64) + 1023) & ~1023;  // the ring is 1024-byte aligned (swizzle atom)
  const int fixed = WS_WARPS * WS_MAX_STAGES * 8 + WS_MAX_ENT * static_cast<int>(sizeof(WsEntry)) + 16 + (4 * P + 1) * 4 + 64;
  int stages = (227 * 1024 - 1024 - tokb - fixed) / (WS_WARPS * WS_STAGE_BYTES);
  if (stages > WS_MAX_STAGES) stages = WS_MAX_STAGES;
  *stages_out = stages;
  *tok_bytes_out = tokb;
  return 1024 + tokb + WS_WARPS * stages * WS_STAGE_BYTES + fixed;
}

}  // namespace

// est_units: host-side estimate of the work units (sizes the grid only; the kernel counts them from device-resident M)
// max_k: largest K among the problems routed here (sizes the token staging and the ring depth)
void w4a16_stream_launch(const DevProblem* probs, int P, int est_units, int max_k, int sm_count, cudaStream_t st) {
  if (est_units <= 0) return;
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(w4a16_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    configured = true;
  }
  const int grid = est_units < sm_count ? est_units : sm_count;
  for (int off = 0; off < P; off += WS_MAX_P) {  // the prefix of one launch lives in shared memory
    const int n = P - off < WS_MAX_P ? P - off : WS_MAX_P;
    int stages = 0, tokb = 0;
    const int smem = ws_smem_bytes(max_k, n, &stages, &tokb);
    w4a16_stream_kernel<<<grid, WS_WARPS * 32, smem, st>>>(probs + off, n, stages, tokb, 0x64006400u);
  }
}

#ifdef WS_PROFILE
extern "C" int mxb_debug_ws_profile(unsigned long long* out) {
  return static_cast<int>(cudaMemcpyFromSymbol(out, ws_prof, sizeof(ws_prof)));
}
#endif

int w4a16_stream_rows_per_group() { return WS_WARPS * WS_SLAB; }
int w4a16_stream_max_k() { return 4096; }

}  // namespace mxb
