// Piece (3) of the hot path: ONE persistent launch over a problem list that mixes precisions.
//
// What it replaces in the reference: the generated `groupgemm_hz_fused_i_impl` persistent kernel
// (mxmoe/kernels/compose_kernel.py:153-224) with its five mma.sync mainloops
// (mxmoe/kernels/src/include/cta_gemm.cuh:8,113,292,424,613), the linear-scan TileScheduler
// (tile_scheduler.cuh:5-50) and the GlobalTileC epilogue (mm_tile.cuh:445-496,610-662).
//
// B200 design (not a translation):
//   * one CTA per SM, 12 warps, warp-specialised:
//       warp 0      TMA producer        (cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx)
//       warp 1      MMA issuer          (one thread, tcgen05.mma kind::f16 / kind::i8, accumulators in TMEM)
//       warp 2      TMEM allocator
//       warps 4-7   epilogue            (tcgen05.ld -> dequant -> fp16 -> smem transpose -> 64B row stores)
//       warps 8-11  operand converters  (4-bit weights unpacked in shared memory into the UMMA layout)
//   * every problem carries its own precision, tile N, tensor maps and instruction descriptor at run
//     time — no code generation, no per-qconfig recompilation;
//   * the tile prefix sum is computed on the device from M values that may themselves live in device
//     memory (DevProblem::m_range), so routing never needs a host round trip;
//   * TMEM holds two 256-column accumulators so the epilogue of tile i overlaps the mainloop of tile i+1.
//
// Numerics follow the reference epilogue: int32 accumulator -> float, multiplied by the fp16-rounded
// product sa*sb (mm_tile.cuh:490), accumulated in fp32 per K-group, one rounding to fp16 (mm_tile.cuh:643).
#include <type_traits>
#include "gg_kernel.cuh"
#include "ptx.cuh"

namespace mxb {
using namespace ptx;

namespace {

struct SmemLayout {
  uint8_t* stages;       // GG_STAGES x (A | B)
  uint8_t* epi_stage;    // 4 warps x 32 rows x 64 B
  uint64_t* full;        // [GG_STAGES] TMA landed
  uint64_t* empty;       // [GG_STAGES] MMA finished reading
  uint64_t* conv;        // [GG_STAGES] converters finished writing B
  uint64_t* cfull;       // [GG_STAGES] packed B landed (used ONLY by converted tiles, so converter warps see every phase)
  uint64_t* tmem_full;   // [2]
  uint64_t* tmem_empty;  // [2]
  uint32_t* tmem_ptr;
  int32_t* prefix;       // [P+1] tile prefix sum
  int32_t* m_arr;        // [P]
  int32_t* row0;         // [P]
};

__device__ __forceinline__ SmemLayout carve(uint8_t* raw) {
  SmemLayout s;
  uint32_t a = smem_u32(raw);
  uint8_t* base = raw + ((1024u - (a & 1023u)) & 1023u);
  s.stages = base;
  s.epi_stage = base + GG_STAGES * GG_STAGE_BYTES;
  uint8_t* p = s.epi_stage + GG_EPI_STAGE_BYTES;
  s.full = reinterpret_cast<uint64_t*>(p);
  s.empty = s.full + GG_STAGES;
  s.conv = s.empty + GG_STAGES;
  s.cfull = s.conv + GG_STAGES;
  s.tmem_full = s.cfull + GG_STAGES;
  s.tmem_empty = s.tmem_full + 2;
  s.tmem_ptr = reinterpret_cast<uint32_t*>(s.tmem_empty + 2);
  s.prefix = reinterpret_cast<int32_t*>(p + 256);
  s.m_arr = s.prefix + GG_MAXP + 1;
  s.row0 = s.m_arr + GG_MAXP;
  return s;
}

// Walks this CTA's tiles: tile = blockIdx.x + i * gridDim.x over the flattened (problem, m, n) list,
// N fastest inside a problem (same rasterisation as tile_scheduler.cuh:39-49).
struct TileWalk {
  const int32_t* prefix;
  int P, tile, total, p;
  __device__ TileWalk(const int32_t* pre, int P_) : prefix(pre), P(P_), tile(blockIdx.x), total(pre[P_]), p(0) {}
  __device__ bool valid() {
    if (tile >= total) return false;
    while (tile >= prefix[p + 1]) ++p;
    return true;
  }
  __device__ int local() const { return tile - prefix[p]; }
  __device__ void advance() { tile += gridDim.x; }
};

__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// ---------------------------------------------------------------- epilogue helpers
// Writes 32 fp16 values of this lane's row into the warp's staging tile (swizzled against bank
// conflicts), then the warp stores 8 rows x 64 B per instruction (full 32-byte sectors).
__device__ __forceinline__ void stage_and_store(uint8_t* stg, const uint32_t (&h)[16], int lane, __half* C, int64_t ldc,
                                                int row_base /*global row of lane 0*/, int rows_valid /*rows of this warp in range*/,
                                                int col0, int N) {
  const int sw = (lane >> 1) & 3;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint4 v = make_uint4(h[4 * c], h[4 * c + 1], h[4 * c + 2], h[4 * c + 3]);
    *reinterpret_cast<uint4*>(stg + lane * 64 + ((c ^ sw) << 4)) = v;
  }
  __syncwarp();
  const int cc = lane & 3;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int r = i * 8 + (lane >> 2);
    uint4 v = *reinterpret_cast<const uint4*>(stg + r * 64 + ((cc ^ ((r >> 1) & 3)) << 4));
    int col = col0 + cc * 8;
    if (r < rows_valid && col < N) {
      st_global_v4(C + static_cast<int64_t>(row_base + r) * ldc + col, v);
    }
  }
  __syncwarp();
}

}  // namespace

__global__ void __launch_bounds__(GG_THREADS, 1) gg_persistent_kernel(const DevProblem* __restrict__ probs, int P) {
  extern __shared__ uint8_t smem_raw[];
  SmemLayout S = carve(smem_raw);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // ---- device-side tile table (replaces the host prefix sum + cudaMalloc/Memcpy of kernel_sketch.py:87-104)
  for (int p = threadIdx.x; p < P; p += GG_THREADS) {
    const DevProblem& d = probs[p];
    int m = d.M, r0 = 0;
    if (d.m_range != nullptr) {
      r0 = d.m_range[0];
      m = d.m_range[1] - r0;
      if (m < 0) m = 0;
    }
    S.m_arr[p] = m;
    S.row0[p] = r0;
    // per-problem tile count, scanned below (problems routed to the streaming kernel contribute none)
    S.prefix[p + 1] = d.path == 0 ? ((m + GG_BM - 1) / GG_BM) * d.n_tiles : 0;
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < GG_STAGES; ++i) {
      mbar_init(&S.full[i], 1);
      mbar_init(&S.empty[i], 1);
      mbar_init(&S.conv[i], 4);
      mbar_init(&S.cfull[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&S.tmem_full[i], 1);
      mbar_init(&S.tmem_empty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(S.tmem_ptr, 512);
    tmem_relinquish();
  }
  __syncthreads();
  if (warp == 1) {  // inclusive scan of tile counts, 32 at a time
    int carry = 0;
    if (lane == 0) S.prefix[0] = 0;
    for (int base = 0; base < P; base += 32) {
      int idx = base + lane;
      int v = idx < P ? S.prefix[idx + 1] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      if (idx < P) S.prefix[idx + 1] = v + carry;
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *S.tmem_ptr;

  if (warp == 0) {
    // ============================================================ TMA producer
    if (lane == 0) {
      int stage = 0, desc_p = -1;
      uint32_t phase = 0;
      for (TileWalk w(S.prefix, P); w.valid(); w.advance()) {
        const DevProblem& d = probs[w.p];
        const int loc = w.local();
        const int m_blk = loc / d.n_tiles, n_blk = loc - m_blk * d.n_tiles;
        const int a_row = S.row0[w.p] + m_blk * GG_BM;
        const int b_row = n_blk * d.bn;
        const int kblocks = d.kblocks, a_kstep = d.a_kstep, b_kstep = d.b_kstep;
        const uint32_t tx = d.tx_bytes;
        const bool has_conv = d.conv != CONV_NONE;
        const int b_off = has_conv ? GG_PACKED_OFF : 0;
        if (w.p != desc_p) {  // first tile of this problem on this CTA: drop any stale cached descriptor (ptx.cuh)
          tma_desc_acquire(&d.tmA);
          tma_desc_acquire(&d.tmB);
          desc_p = w.p;
        }
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&S.empty[stage], phase ^ 1);
          uint8_t* sa = S.stages + stage * GG_STAGE_BYTES;
          // converted tiles: A lands on `full`, the packed B on `cfull` (a barrier whose every phase the
          // converter warps observe; waiting on `full` from a role that skips tiles would alias phases)
          uint64_t* bbar = has_conv ? &S.cfull[stage] : &S.full[stage];
          if (has_conv) {
            mbar_expect_tx(&S.full[stage], GG_A_BYTES);
            mbar_expect_tx(bbar, tx - GG_A_BYTES);
          } else {
            mbar_expect_tx(&S.full[stage], tx);
          }
          tma_load_2d(sa, &d.tmA, &S.full[stage], kb * a_kstep, a_row);
          tma_load_2d(sa + GG_A_BYTES + b_off, &d.tmB, bbar, kb * b_kstep, b_row);
          if (++stage == GG_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ============================================================ MMA issuer (single thread)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0, conv_phase_bits = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (TileWalk w(S.prefix, P); w.valid(); w.advance()) {
        const DevProblem& d = probs[w.p];
        const int kblocks = d.kblocks;
        const int kpg = d.epi == EPI_INT_GROUP ? d.kb_per_group : kblocks;
        const uint32_t idesc = d.idesc;
        const bool is_i8 = d.mma_kind == MMA_I8, is_f8 = d.mma_kind == MMA_F8;
        const bool has_conv = d.conv != CONV_NONE;
        for (int kb0 = 0; kb0 < kblocks; kb0 += kpg) {
          mbar_wait(&S.tmem_empty[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * 256;
          const int kend = min(kb0 + kpg, kblocks);
          // one copy of the k-block loop per MMA kind (the kind is an opcode): no kind test per MMA on the issuing thread
          auto run_group = [&](auto kind) {
            for (int kb = kb0; kb < kend; ++kb) {
              mbar_wait(&S.full[stage], phase);
              if (has_conv) {
                mbar_wait(&S.conv[stage], (conv_phase_bits >> stage) & 1);
                conv_phase_bits ^= 1u << stage;
              }
              tc_fence_after();
              const uint32_t a_addr = smem_u32(S.stages + stage * GG_STAGE_BYTES);
              const uint64_t da = umma_desc_sw128(a_addr);
              const uint64_t db = umma_desc_sw128(a_addr + GG_A_BYTES);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint32_t accum = (kb > kb0 || k > 0) ? 1u : 0u;
                if constexpr (decltype(kind)::value == MMA_I8) umma_i8(d_tmem, da + 2 * k, db + 2 * k, idesc, accum);
                else if constexpr (decltype(kind)::value == MMA_F8) umma_f8(d_tmem, da + 2 * k, db + 2 * k, idesc, accum);
                else umma_f16(d_tmem, da + 2 * k, db + 2 * k, idesc, accum);
              }
              umma_commit(&S.empty[stage]);
              if (++stage == GG_STAGES) { stage = 0; phase ^= 1; }
            }
          };
          if (is_i8) run_group(std::integral_constant<int, MMA_I8>{});
          else if (is_f8) run_group(std::integral_constant<int, MMA_F8>{});
          else run_group(std::integral_constant<int, MMA_F16>{});
          umma_commit(&S.tmem_full[acc]);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ============================================================ epilogue
    const int q = warp - 4;  // TMEM lane quadrant == warp % 4
    uint8_t* stg = S.epi_stage + q * 2048;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (TileWalk w(S.prefix, P); w.valid(); w.advance()) {
      const DevProblem& d = probs[w.p];
      const int loc = w.local();
      const int m_blk = loc / d.n_tiles, n_blk = loc - m_blk * d.n_tiles;
      const int M = S.m_arr[w.p], N = d.N, bn = d.bn;
      const int row_in_prob = m_blk * GG_BM + q * 32 + lane;
      const int grow = S.row0[w.p] + row_in_prob;           // global row of this lane
      const int crow = (d.c_row != nullptr ? static_cast<int>(ld_nc_u32(d.c_row)) : S.row0[w.p]) + row_in_prob;  // its row in C (DevProblem::c_row)
      const bool row_ok = row_in_prob < M;
      const int rows_valid = M - (m_blk * GG_BM + q * 32);  // rows of this warp inside the problem
      const int n0 = n_blk * bn;
      const int epi = d.epi;
      const int64_t ldc = d.ldc;
      __half* C = d.C;
      const __half* sb = d.scale_b;
      const int shift = d.acc_shift;
      const bool facc = d.mma_kind == MMA_F8;  // fp32 accumulators (kind::f8f6f4)
      const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);

      if (epi != EPI_INT_GROUP) {
        __half sa = __float2half(1.f);
        if (epi == EPI_INT_ROWCOL && row_ok) sa = d.scale_a[static_cast<int64_t>(grow) * d.scale_a_ld];
        mbar_wait(&S.tmem_full[acc], acc_phase);
        tc_fence_after();
        for (int c = 0; c < bn; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32(lane_addr + acc * 256 + c, v);
          tmem_ld_wait();
          uint32_t h[16];
          const int col0 = n0 + c;
          if (epi == EPI_F32) {
#pragma unroll
            for (int j = 0; j < 16; ++j) h[j] = pack_half2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
          } else {
            if (d.acc_out != nullptr && row_ok) {
              int32_t* ao = d.acc_out + static_cast<int64_t>(row_in_prob) * N + col0;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j < N) ao[j] = static_cast<int32_t>(v[j]) >> shift;
            }
            const __half2 sa2 = __half2half2(sa);
            if (!facc) {  // integer accumulators become floats in place (one branch per tile, not a select per element)
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__int2float_rn(static_cast<int32_t>(v[j]) >> shift));
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 s4 = make_uint4(0, 0, 0, 0);
              if (col0 + g * 8 < N) s4 = ld_nc_v4(sb + col0 + g * 8);
              const uint32_t sw[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                __half2 prod = __hmul2(sa2, *reinterpret_cast<const __half2*>(&sw[j]));
                float2 pf = __half22float2(prod);
                h[g * 4 + j] = pack_half2(__uint_as_float(v[g * 8 + 2 * j]) * pf.x, __uint_as_float(v[g * 8 + 2 * j + 1]) * pf.y);
              }
            }
          }
          stage_and_store(stg, h, lane, C, ldc, crow - lane, rows_valid, col0, N);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&S.tmem_empty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      } else {
        // group-wise: one accumulator epoch per K group, fp32 running sum in registers (bn <= 64)
        float run[64];
#pragma unroll
        for (int j = 0; j < 64; ++j) run[j] = 0.f;
        const int kblocks = d.kblocks, kpg = d.kb_per_group;
        const int sald = d.scale_a_ld;
        int g = 0;
        for (int kb0 = 0; kb0 < kblocks; kb0 += kpg, ++g) {
          __half sa = __float2half(0.f);
          if (row_ok) sa = d.scale_a[static_cast<int64_t>(grow) * sald + g];
          const __half2 sa2 = __half2half2(sa);
          const __half* sbg = sb + static_cast<int64_t>(g) * N;
          mbar_wait(&S.tmem_full[acc], acc_phase);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < 64; c += 32) {
            if (c < bn) {
              uint32_t v[32];
              tmem_ld_32x32(lane_addr + acc * 256 + c, v);
              tmem_ld_wait();
              const int col0 = n0 + c;
              if (d.acc_out != nullptr && row_ok) {
                int32_t* ao = d.acc_out + (static_cast<int64_t>(g) * M + row_in_prob) * N + col0;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col0 + j < N) ao[j] = static_cast<int32_t>(v[j]) >> shift;
              }
#pragma unroll
              for (int gg = 0; gg < 4; ++gg) {
                uint4 s4 = make_uint4(0, 0, 0, 0);
                if (col0 + gg * 8 < N) s4 = ld_nc_v4(sbg + col0 + gg * 8);
                const uint32_t sw[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  __half2 prod = __hmul2(sa2, *reinterpret_cast<const __half2*>(&sw[j]));
                  float2 pf = __half22float2(prod);
                  int a0 = static_cast<int32_t>(v[gg * 8 + 2 * j]) >> shift;
                  int a1 = static_cast<int32_t>(v[gg * 8 + 2 * j + 1]) >> shift;
                  run[c + gg * 8 + 2 * j] = fmaf(__int2float_rn(a0), pf.x, run[c + gg * 8 + 2 * j]);
                  run[c + gg * 8 + 2 * j + 1] = fmaf(__int2float_rn(a1), pf.y, run[c + gg * 8 + 2 * j + 1]);
                }
              }
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&S.tmem_empty[acc]);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
#pragma unroll
        for (int c = 0; c < 64; c += 32) {
          if (c < bn) {
            uint32_t h[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) h[j] = pack_half2(run[c + 2 * j], run[c + 2 * j + 1]);
            stage_and_store(stg, h, lane, C, ldc, crow - lane, rows_valid, n0 + c, N);
          }
        }
      }
    }
  } else if (warp >= 8) {
    // ============================================================ operand converters (4 warps)
    const int ct = threadIdx.x - 256;  // 0..127
    int stage = 0;
    uint32_t cphase_bits = 0;  // per-stage parity of `cfull` (advances only on converted tiles)
    for (TileWalk w(S.prefix, P); w.valid(); w.advance()) {
      const DevProblem& d = probs[w.p];
      const int kblocks = d.kblocks;
      const int conv = d.conv;
      if (conv == CONV_NONE) {  // nothing to do: just keep the ring position in step
        stage = (stage + kblocks) % GG_STAGES;
        continue;
      }
      const int bn = d.bn;
      const int loc = w.local();
      const int n_blk = loc % d.n_tiles;
      const int n0 = n_blk * bn;
      if (conv == CONV_W4_I8) {
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&S.cfull[stage], (cphase_bits >> stage) & 1);
          cphase_bits ^= 1u << stage;
          uint8_t* bfin = S.stages + stage * GG_STAGE_BYTES + GG_A_BYTES;
          const uint8_t* bpk = bfin + GG_PACKED_OFF;
          // packed tile: bn rows x 64 B (128 k as nibbles); 16-byte item = 32 k: low nibbles k..k+15, high k+16..k+31.
          // bn/32 items per thread; all loads are issued before the first store (ILP against the smem latency).
          const int iters = bn >> 5;
          uint4 w4[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < iters) w4[j] = *reinterpret_cast<const uint4*>(bpk + (ct + j * 128) * 16);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j < iters) {
              const int i = ct + j * 128;
              const int row = i >> 2, c = i & 3;
              uint4 lo, hi;
              lo.x = (w4[j].x << 4) & 0xF0F0F0F0u; hi.x = w4[j].x & 0xF0F0F0F0u;
              lo.y = (w4[j].y << 4) & 0xF0F0F0F0u; hi.y = w4[j].y & 0xF0F0F0F0u;
              lo.z = (w4[j].z << 4) & 0xF0F0F0F0u; hi.z = w4[j].z & 0xF0F0F0F0u;
              lo.w = (w4[j].w << 4) & 0xF0F0F0F0u; hi.w = w4[j].w & 0xF0F0F0F0u;
              const int sw = row & 7;
              *reinterpret_cast<uint4*>(bfin + row * 128 + (((2 * c) ^ sw) << 4)) = lo;
              *reinterpret_cast<uint4*>(bfin + row * 128 + (((2 * c + 1) ^ sw) << 4)) = hi;
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&S.conv[stage]);
          if (++stage == GG_STAGES) stage = 0;
        }
      } else {
        // W-only: unsigned codes -> fp16 via the 0x6400 magic, then w = hfma(q, scale, zp) (quantize.cuh:189-190)
        const bool w4 = conv == CONV_W4_F16;
        const bool sym = d.sym != 0;
        const int N = d.N;
        const int kpg = d.kb_per_group;  // k-blocks (64 k each) per scale group; >= kblocks means per-channel
        const __half* sb = d.scale_b;
        const int rows_per_thread = (bn + 63) / 64;  // thread owns rows ct/2 + 64*j
        const int r_base = ct >> 1, hf = ct & 1;
        const float bias = w4 ? (sym ? 1031.f : 1024.f) : (sym ? 1151.f : 1024.f);
        const __half2 bias2 = __float2half2_rn(bias);
        __half2 sc[2], zp[2], sc_n[2], zp_n[2];
        // scales of group g+1 are fetched while group g is converted (a global load on the critical path of every
        // k-block made the converters, not the tensor cores, the bottleneck of the weight-only path)
        auto load_scales = [&](int g) {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int n = n0 + r_base + 64 * j;
            sc_n[j] = __float2half2_rn(0.f);
            zp_n[j] = __float2half2_rn(0.f);
            if (j < rows_per_thread && n < N && g * kpg < kblocks) {
              if (sym) {
                sc_n[j] = __half2half2(sb[static_cast<int64_t>(g) * N + n]);
              } else {
                __half2 t = *reinterpret_cast<const __half2*>(sb + (static_cast<int64_t>(g) * N + n) * 2);
                sc_n[j] = __half2half2(__low2half(t));
                zp_n[j] = __half2half2(__high2half(t));
              }
            }
          }
        };
        load_scales(0);
        for (int kb = 0; kb < kblocks; ++kb) {
          if (kb % kpg == 0) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              sc[j] = sc_n[j];
              zp[j] = zp_n[j];
            }
            load_scales(kb / kpg + 1);
          }
          mbar_wait(&S.cfull[stage], (cphase_bits >> stage) & 1);
          cphase_bits ^= 1u << stage;
          uint8_t* bfin = S.stages + stage * GG_STAGE_BYTES + GG_A_BYTES;
          const uint8_t* bpk = bfin + GG_PACKED_OFF;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int row = r_base + 64 * j;
            if (row < bn) {
              const int sw = row & 7;
              if (w4) {
                // packed row = 32 B (64 k); this thread converts 16 B = 32 k -> four 16-byte chunks
                uint4 p4 = *reinterpret_cast<const uint4*>(bpk + row * 32 + hf * 16);
                const uint32_t pw[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                  uint32_t o[4];
#pragma unroll
                  for (int s = 0; s < 4; ++s) {
                    uint32_t t = ((pw[c] >> (4 * s)) & 0x000F000Fu) | 0x64006400u;
                    __half2 hv = __hsub2(*reinterpret_cast<__half2*>(&t), bias2);
                    hv = __hfma2(hv, sc[j], zp[j]);
                    o[s] = *reinterpret_cast<uint32_t*>(&hv);
                  }
                  *reinterpret_cast<uint4*>(bfin + row * 128 + (((hf * 4 + c) ^ sw) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
                }
              } else {
                // 8-bit: packed row = 64 B (64 k); this thread converts 32 B -> four 16-byte chunks
                const uint4* pp = reinterpret_cast<const uint4*>(bpk + row * 64 + hf * 32);
#pragma unroll
                for (int c2 = 0; c2 < 2; ++c2) {
                  uint4 p4 = pp[c2];
                  const uint32_t pw[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
                  for (int c = 0; c < 2; ++c) {
                    uint32_t o[4];
#pragma unroll
                    for (int s = 0; s < 2; ++s) {
                      const uint32_t wv = pw[c * 2 + s];
                      uint32_t t0 = __byte_perm(wv, 0x64646464u, 0x4140);  // bytes 0,1 -> two halves 0x64xx
                      uint32_t t1 = __byte_perm(wv, 0x64646464u, 0x4342);  // bytes 2,3
                      __half2 h0 = __hfma2(__hsub2(*reinterpret_cast<__half2*>(&t0), bias2), sc[j], zp[j]);
                      __half2 h1 = __hfma2(__hsub2(*reinterpret_cast<__half2*>(&t1), bias2), sc[j], zp[j]);
                      o[2 * s] = *reinterpret_cast<uint32_t*>(&h0);
                      o[2 * s + 1] = *reinterpret_cast<uint32_t*>(&h1);
                    }
                    *reinterpret_cast<uint4*>(bfin + row * 128 + (((hf * 4 + c2 * 2 + c) ^ sw) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
                  }
                }
              }
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&S.conv[stage]);
          if (++stage == GG_STAGES) stage = 0;
        }
      }
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

int gg_configure() {
  return static_cast<int>(cudaFuncSetAttribute(gg_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GG_SMEM_BYTES));
}

void gg_launch(const DevProblem* probs, int P, int grid, cudaStream_t stream) {
  gg_persistent_kernel<<<grid, GG_THREADS, GG_SMEM_BYTES, stream>>>(probs, P);
}

}  // namespace mxb
