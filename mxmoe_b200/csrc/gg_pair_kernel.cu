// Piece (3), compute-bound form: the persistent grouped GEMM on CTA PAIRS (tcgen05 cta_group::2).
//
// Same contract as gg_kernel.cu (one launch over a problem list that mixes precisions; it replaces the
// reference's generated `groupgemm_hz_fused_i_impl`, mxmoe/kernels/compose_kernel.py:153-224, its
// mainloops cta_gemm.cuh:8,424 and the GlobalTileC epilogue mm_tile.cuh:445-496,610-662), but laid out
// for what bounds a B200 at bs=8192: the L2 -> SM operand traffic and the 128-row token-tile waste of
// MoE problems whose M is a few hundred rows.
//
//   * A cluster of 2 CTAs (one TPC) computes one 256 (weight rows) x NT (tokens) tile with
//     `tcgen05.mma.cta_group::2`, UMMA M = 256.  The WEIGHTS are the UMMA "A" operand (128 rows per CTA)
//     and the TOKENS the UMMA "B" operand (NT/2 rows per CTA): C^T = W * X^T.  With the operands swapped
//     the token extent of a tile is the UMMA N, which is any multiple of 16 up to 256, so an expert
//     with M rows is cut into ceil(M/256) equal tiles of ~M/ceil(M/256) tokens (<= 15 padded rows per
//     tile) instead of 128-row tiles (the reference pads to BM=128/256, tile_config.py).
//   * Per 128-byte K slab each CTA stages 16 KB of weights (8 KB when they are 4-bit) + <= 16 KB of
//     tokens for a 128 x 256 slice of the output: 2/3 of the L2 traffic of the single-CTA 128 x 256 tile,
//     and 32 KB stages leave room for a 6-deep ring (the single-CTA kernel has 3).
//   * 4-bit weights land packed in the upper half of their own stage and are expanded IN PLACE by the
//     converter warps (registers in between, one named barrier), so staging costs no extra smem.
//   * Barriers: TMA bytes of both CTAs are counted on the leader's `full`; the leader's single MMA
//     thread commits with a cluster multicast to `empty` / `tmem_full` of both CTAs; the epilogue and
//     converter warps of the peer arrive remotely on the leader's `tmem_empty` / `conv`.
//   * Epilogue: TMEM lane = weight row n, column = token.  The per-token scale is staged in smem, the
//     per-channel scale lives in one register; the arithmetic (int32 -> float, times the fp16-rounded
//     product sa*sb, one rounding to fp16) is the reference's, mm_tile.cuh:490,643.
//   * Orientation is chosen PER TILE (tile_tokm): a token chunk of >= 160 rows over a full 256-row weight tile
//     is computed the other way round — the same two smem regions with the operand descriptors swapped, tokens on
//     UMMA M — so that a thread of the epilogue owns an output ROW: the per-token scale is a register, the
//     per-channel scales are staged, gate/up pairs of the fused silu*mul sit in adjacent registers (no
//     shuffles) and the chunk leaves as 16-byte vectors after a warp-private shared-memory transpose (8 rows x
//     64 contiguous bytes per store instruction): a third of the weight-major drain's instructions, the same bits.
//
// Supported here: fp16, w8a8 and w4a4 with per-token / per-channel scales (g-1).  Lists containing
// group-wise or weight-only problems run on gg_kernel.cu.
#include <type_traits>
#include "gg_kernel.cuh"
#include "ptx.cuh"
#include "quant_common.cuh"

namespace mxb {
using namespace ptx;

namespace {

#ifndef GP_STAGES_N
#define GP_STAGES_N 6
#endif
constexpr int GP_STAGES = GP_STAGES_N;  // (-DGP_STAGES_N=4/5: ring-depth sensitivity runs, tools/r02_ab.sh)
constexpr int GP_THREADS = 512;  // 16 warps: TMA tokens, MMA, TMEM alloc, TMA weights, 4 epilogue, 4 converter, 4 epilogue
constexpr int GP_W_BYTES = 128 * 128;  // this CTA's 128 weight rows x 128 B
constexpr int GP_T_BYTES = 128 * 128;  // this CTA's half of the token tile, <= 128 rows x 128 B
constexpr int GP_STAGE_BYTES = GP_W_BYTES + GP_T_BYTES;
constexpr int GP_PACKED_OFF = GP_W_BYTES / 2;  // 4-bit weights land in the upper half of the W region
constexpr int GP_TBOX_ROWS = 32;               // token TMA box
constexpr int GP_TBOX_BYTES = GP_TBOX_ROWS * 128;
constexpr int GP_BT = 256;                     // max tokens per tile (UMMA N)
constexpr int GP_BN = 256;                     // weight rows per tile (UMMA M across the pair)

struct Smem {
  uint8_t* stages;
  uint64_t* full;        // [S] leader: token (and unconverted weight) bytes of both CTAs landed
  uint64_t* empty;       // [S] each CTA: the pair's MMAs finished reading this stage (multicast commit)
  uint64_t* conv;        // [S] leader: converter warps of both CTAs finished (8 arrivals)
  uint64_t* cfull;       // [S] each CTA: its own packed weights landed
  uint64_t* tmem_full;   // [2] each CTA (multicast commit)
  uint64_t* tmem_empty;  // [2] leader: epilogue warps of both CTAs drained the accumulator (16 arrivals)
  uint32_t* tmem_ptr;
  __half* sa_stage;      // [2][256] per-token scales of the tile being drained
  uint8_t* c_stage;      // [2 groups][32 tokens][GP_CS_PITCH B]: output chunk transposed to token-major before it leaves
  int32_t* prefix;       // [P+1]
  int32_t* m_arr;        // [P]
  int32_t* row0;         // [P]
};

constexpr int GP_BAR_BYTES = 512;
constexpr int GP_SA_BYTES = 2 * GP_BT * 2;
constexpr int GP_CS_PITCH = 128 * 2 + 16;          // a token's 128 outputs of this CTA + 16 B: conflict-free 16-byte row reads
constexpr int GP_CS_BYTES = 2 * 32 * GP_CS_PITCH;  // one 32-token chunk per epilogue group
constexpr int GP_SMEM_BYTES = 1024 + GP_STAGES * GP_STAGE_BYTES + GP_BAR_BYTES + GP_SA_BYTES + GP_CS_BYTES + (3 * GG_MAXP + 4) * 4;

__device__ __forceinline__ Smem carve(uint8_t* raw) {
  Smem s;
  uint32_t a = smem_u32(raw);
  uint8_t* base = raw + ((1024u - (a & 1023u)) & 1023u);
  s.stages = base;
  uint8_t* p = base + GP_STAGES * GP_STAGE_BYTES;
  s.full = reinterpret_cast<uint64_t*>(p);
  s.empty = s.full + GP_STAGES;
  s.conv = s.empty + GP_STAGES;
  s.cfull = s.conv + GP_STAGES;
  s.tmem_full = s.cfull + GP_STAGES;
  s.tmem_empty = s.tmem_full + 2;
  s.tmem_ptr = reinterpret_cast<uint32_t*>(s.tmem_empty + 2);
  s.sa_stage = reinterpret_cast<__half*>(p + GP_BAR_BYTES);
  s.c_stage = p + GP_BAR_BYTES + GP_SA_BYTES;
  s.prefix = reinterpret_cast<int32_t*>(p + GP_BAR_BYTES + GP_SA_BYTES + GP_CS_BYTES);
  s.m_arr = s.prefix + GG_MAXP + 1;
  s.row0 = s.m_arr + GG_MAXP;
  return s;
}

// One tile of the walk: weight rows [n0, n0+256) x tokens [t0, t0+nt) of problem p.
struct PairTile {
  int p, n0, t0, nt;
};

// Static round-robin over the flattened (problem, n chunk, token chunk) list, token chunks fastest so that
// consecutive tiles (running on neighbouring pairs at the same time) share one weight slab in L2.
struct PairWalk {
  const int32_t* prefix;
  const int32_t* m_arr;
  int tile, total, p, step;
  __device__ PairWalk(const Smem& S, int P) : prefix(S.prefix), m_arr(S.m_arr), tile(cluster_id_x()), total(S.prefix[P]), p(0),
                                              step(n_clusters_x()) {}
  __device__ bool next(PairTile* t) {
    if (tile >= total) return false;
    while (tile >= prefix[p + 1]) ++p;
    const int loc = tile - prefix[p];
    const int M = m_arr[p];
    const int tch = (M + GP_BT - 1) / GP_BT;               // token chunks
    const int per = (((M + tch - 1) / tch) + 15) & ~15;    // tokens per chunk, equalised, multiple of 16
    const int n_blk = loc / tch, t_blk = loc - n_blk * tch;
    t->p = p;
    t->n0 = n_blk * GP_BN;
    t->t0 = t_blk * per;
    t->nt = min(per, M - t->t0);
    tile += step;
    return true;
  }
};

#ifdef GP_TRACE
// -DGP_TRACE builds only (tools/gg_trace.py): nanosecond stamps of the first cluster's leader CTA, [role][event]
//   0 token producer: stage free seen      1 MMA: full seen   2 MMA: conv seen   3 MMA: k-block issued
//   4 converter warp 8: packed slab landed  5 converter: arrived on conv
//   6 epilogue warp 4: tmem_full seen (per tile)   7 epilogue: tile stored (per tile)   8 MMA: accumulator free seen (per tile)
constexpr int GP_TR_N = 2048;
__device__ unsigned long long gp_trace_buf[9][GP_TR_N];
__device__ __forceinline__ void gp_tr(int role, int idx) {
  if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && idx < GP_TR_N) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    gp_trace_buf[role][idx] = t;
  }
}
#define GP_TR(role, idx) gp_tr(role, idx)
#else
#define GP_TR(role, idx) do {} while (0)
#endif

// Orientation of a tile's accumulator: token-major — tokens on UMMA M, a thread drains one output ROW — for
// (nearly) full token chunks over a full 256-row weight tile; weight-major (tokens on UMMA N, sized to the chunk) for the ragged
// ones, where a 256-row M would waste tensor time on rows that do not exist.
#ifdef GP_X_NOSTG  // timing experiment: everything but the global stores of the token-major drain (never-true extra condition)
#define GP_X_STG_COND &&t.nt > 100000
#else
#define GP_X_STG_COND
#endif
// Measured (same box, c2 GEMM pair): 0.490 ms all weight-major, 0.479 ms with chunks of >= 160 tokens token-major (the extra
// tensor time of a 256-row M over 192 tokens is free: the tensor pipe is not what this kernel waits for);
// -DGP_TOKM_MIN=257 builds the all-weight-major kernel.
#ifndef GP_TOKM_MIN
#define GP_TOKM_MIN 160
#endif
__device__ __forceinline__ bool tile_tokm(const DevProblem& d, const PairTile& t) {
  // (not into a peer GPU's buffer, c_row: there the weight-major drain's transposed 256-byte row segments suit NVLink better than
  // 64-byte ones — expert-parallel c2 at 2 GPUs 0.868 ms against 0.894)
  return t.nt >= GP_TOKM_MIN && d.N - t.n0 >= GP_BN && d.acc_out == nullptr && d.c_row == nullptr;
}

__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GP_THREADS, 1) gg_pair_kernel(const DevProblem* __restrict__ probs, int P) {
  extern __shared__ uint8_t smem_raw[];
  Smem S = carve(smem_raw);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  // ---- device-side tile table (M may live in device memory: DevProblem::m_range)
  for (int p = threadIdx.x; p < P; p += GP_THREADS) {
    const DevProblem& d = probs[p];
    int m = d.M, r0 = 0;
    if (d.m_range != nullptr) {
      r0 = d.m_range[0];
      m = d.m_range[1] - r0;
      if (m < 0) m = 0;
    }
    S.m_arr[p] = m;
    S.row0[p] = r0;
    S.prefix[p + 1] = ((m + GP_BT - 1) / GP_BT) * d.n_tiles;
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < GP_STAGES; ++i) {
      mbar_init(&S.full[i], 1);
      mbar_init(&S.empty[i], 1);
      mbar_init(&S.conv[i], 8);
      mbar_init(&S.cfull[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&S.tmem_full[i], 1);
      mbar_init(&S.tmem_empty[i], 16);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(S.tmem_ptr, 512);
    tmem_relinquish_pair();
  }
  __syncthreads();
  if (warp == 1) {
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
  cluster_sync_all();  // barrier inits of both CTAs are visible before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *S.tmem_ptr;

#ifdef GP_MMA_W3  // the MMA issuer on scheduler 3, whose epilogue warps idle on token-major tiles of <= 192 tokens
  constexpr int kMmaWarp = 3, kWeightTmaWarp = 1;
#else
  constexpr int kMmaWarp = 1, kWeightTmaWarp = 3;
#endif
  if (warp == 0 || warp == kWeightTmaWarp) {
    // ============================================================ TMA producers (in EACH CTA): warp 0 tokens, warp 3 weights.
    // The whole warp walks the tiles converged and one elected lane issues: with a single-thread region the compiler
    // wraps every UTMALDG in a uniform-register "waterfall" loop, and one thread issuing 5 loads per stage was the
    // limiter of the first version of this kernel (ncu: producer busy 64 % of the time, converters starved).
    const bool tokens = warp == 0;
    int stage = 0, desc_p = -1;
    uint32_t phase = 0;
#ifdef GP_TRACE
    int tr_kb = 0;
#endif
    PairTile t;
    for (PairWalk w(S, P); w.next(&t);) {
      const DevProblem& d = probs[t.p];
      const int nt16 = (t.nt + 15) & ~15;
      const int half = nt16 >> 1;  // token rows this CTA stages
      // one 128-row box when (nearly) all of it is needed, else 32-row boxes: a ragged tile moves only its own rows
      const bool big = half > 96;
      const int boxes = big ? 1 : (half + GP_TBOX_ROWS - 1) / GP_TBOX_ROWS;
      const uint32_t tok_bytes = big ? uint32_t(GP_T_BYTES) : uint32_t(boxes) * GP_TBOX_BYTES;
      const int trow = S.row0[t.p] + t.t0 + int(rank) * half;
      const int wrow = t.n0 + int(rank) * 128;
      const int kblocks = d.kblocks, a_kstep = d.a_kstep, b_kstep = d.b_kstep;
      const bool has_conv = d.conv != CONV_NONE;
      // bytes counted on the leader's `full`: the token boxes of both CTAs (+ both weight halves when unconverted)
      const uint32_t tx_full = 2u * tok_bytes + (has_conv ? 0u : 2u * GP_W_BYTES);
      const CUtensorMap* tmT = big ? &d.tmA2 : &d.tmA;
      if (t.p != desc_p) {  // first tile of this problem on this CTA: drop any stale cached descriptor (ptx.cuh)
        if (lane < 3) tma_desc_acquire(lane == 0 ? &d.tmA : lane == 1 ? &d.tmA2 : &d.tmB);
        desc_p = t.p;
        __syncwarp();  // elect_one() may pick any lane to issue: ordered after the three fences
      }
      for (int kb = 0; kb < kblocks; ++kb) {
        mbar_wait(&S.empty[stage], phase ^ 1);
#ifdef GP_TRACE
        if (tokens) GP_TR(0, tr_kb++);
#endif
        uint8_t* sw = S.stages + stage * GP_STAGE_BYTES;
        const uint32_t full_leader = mapa_u32(smem_u32(&S.full[stage]), 0);
        if (elect_one()) {
          if (tokens) {
#ifdef GP_X_HALFTOK  // timing experiment (wrong results): token boxes only on even k-blocks
            const bool skip_tok = (kb & 1) != 0;
            if (leader) mbar_expect_tx(&S.full[stage], skip_tok ? tx_full - 2u * tok_bytes : tx_full);
            if (skip_tok) { /* nothing to load */ } else
#else
            if (leader) mbar_expect_tx(&S.full[stage], tx_full);
#endif
            {
            uint8_t* st = sw + GP_W_BYTES;
            if (big) {
              tma_load_2d_pair(st, tmT, full_leader, kb * a_kstep, trow);
            } else {
              for (int b = 0; b < boxes; ++b)
                tma_load_2d_pair(st + b * GP_TBOX_BYTES, tmT, full_leader, kb * a_kstep, trow + b * GP_TBOX_ROWS);
            }
            }
          } else if (has_conv) {
            mbar_expect_tx(&S.cfull[stage], GP_W_BYTES / 2);
            tma_load_2d(sw + GP_PACKED_OFF, &d.tmB, &S.cfull[stage], kb * b_kstep, wrow);
          } else {
            tma_load_2d_pair(sw, &d.tmB, full_leader, kb * b_kstep, wrow);
          }
        }
        __syncwarp();
        if (++stage == GP_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == kMmaWarp) {
    // ============================================================ MMA issuer (leader CTA; warp converged, one elected lane issues)
    if (leader) {
      int stage = 0;
      uint32_t phase = 0, conv_phase_bits = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      // Operand descriptors of the current stage as running values: a stage is 32 KB further, i.e. +2048 in the 14-bit
      // address field (no carry out of it: the ring ends below 227 KB), so a k-block costs two adds instead of building both
      // descriptors from the shared address (this lane's instruction count per k-block paces the tensor pipe).
      const uint64_t da0 = umma_desc_sw128(smem_u32(S.stages)), db0 = umma_desc_sw128(smem_u32(S.stages) + GP_W_BYTES);
      uint64_t da = da0, db = db0;
#ifdef GP_TRACE
      int tr_kb = 0, tr_tile = 0;
#endif
      PairTile t;
      for (PairWalk w(S, P); w.next(&t);) {
        const DevProblem& d = probs[t.p];
        const int kblocks = d.kblocks;
        // token-major tile: tokens on UMMA M (A operand, 128 rows of the chunk per CTA; rows past the chunk hold stale bytes whose
        // products only reach accumulator lanes nobody reads), the 256 weight rows on UMMA N
        const bool tokm = tile_tokm(d, t);
        const int nt16 = (t.nt + 15) & ~15;
        const uint32_t idesc = d.idesc | ((uint32_t(tokm ? GP_BN : nt16) >> 3) << 17);
        const bool is_i8 = d.mma_kind == MMA_I8, is_f8 = d.mma_kind == MMA_F8;
        const bool has_conv = d.conv != CONV_NONE;
#ifdef GP_MMA_LANE0
        if (lane != 0) continue;  // one lane runs the whole issue loop (no elect / reconvergence per k-block)
#endif
        mbar_wait(&S.tmem_empty[acc], acc_phase ^ 1);
#ifdef GP_TRACE
        GP_TR(8, tr_tile++);
#endif
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        // The MMA kind is an opcode, not an operand: one copy of the k-block loop per kind, chosen once per tile.  (A kind
        // test per MMA inside the loop costs 4 % of the whole GEMM: this one lane's issue rate is the kernel's pace.)
        const uint64_t dx = tokm ? db0 - da0 : 0;  // db - da is the same for every stage
        auto run_tile = [&](auto kind) {
          for (int kb = 0; kb < kblocks; ++kb) {
#ifdef GP_CONV_FULL
            if (!has_conv)
#endif
            mbar_wait(&S.full[stage], phase);
#ifdef GP_TRACE
            GP_TR(1, tr_kb);
#endif
            if (has_conv) {
              mbar_wait(&S.conv[stage], (conv_phase_bits >> stage) & 1);
              conv_phase_bits ^= 1u << stage;
            }
#ifdef GP_TRACE
            GP_TR(2, tr_kb);
#endif
            tc_fence_after();
#ifdef GP_MMA_LANE0
            {
#else
            if (elect_one()) {
#endif
#ifdef GP_X_QUARTERMMA  // timing experiment (wrong results): one of the four MMAs of a k-block
#define GP_KSTEPS 1
#else
#define GP_KSTEPS 4
#endif
#pragma unroll
              for (int k = 0; k < GP_KSTEPS; ++k) {
                const uint32_t accum = (kb > 0 || k > 0) ? 1u : 0u;
                const uint64_t ka = da + dx + 2 * k, kb_ = db - dx + 2 * k;  // token-major: A = token region, B = weight region
                if constexpr (decltype(kind)::value == MMA_I8) umma_i8_pair(d_tmem, ka, kb_, idesc, accum);
                else if constexpr (decltype(kind)::value == MMA_F8) umma_f8_pair(d_tmem, ka, kb_, idesc, accum);
                else umma_f16_pair(d_tmem, ka, kb_, idesc, accum);
              }
              umma_commit_pair(&S.empty[stage], 3);
              if (kb == kblocks - 1) umma_commit_pair(&S.tmem_full[acc], 3);
            }
#ifndef GP_MMA_LANE0
            __syncwarp();
#endif
#ifdef GP_TRACE
            GP_TR(3, tr_kb++);
#endif
            da += GP_STAGE_BYTES >> 4;
            db += GP_STAGE_BYTES >> 4;
            if (++stage == GP_STAGES) { stage = 0; phase ^= 1; da = da0; db = db0; }
          }
        };
        if (is_i8) run_tile(std::integral_constant<int, MMA_I8>{});
        else if (is_f8) run_tile(std::integral_constant<int, MMA_F8>{});
        else run_tile(std::integral_constant<int, MMA_F16>{});
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if ((warp >= 4 && warp < 8) || warp >= 12) {
    // ============================================================ epilogue (8 warps in EACH CTA; lane = weight row)
    // Two warps per TMEM lane quadrant (warps q+4 and q+12, both on scheduler q), taking alternate
    // 32-token column chunks: one warp per scheduler could not hide the latency of its own dependent chains.
    const int q = warp & 3;
    const int grp = warp >= 12 ? 1 : 0;
    const int et = grp * 128 + q * 32 + lane;  // 0..255
    int acc = 0;
    uint32_t acc_phase = 0;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tmem_empty_leader = mapa_u32(smem_u32(&S.tmem_empty[0]), 0);
#ifdef GP_TRACE
    int tr_tile = 0;
#endif
    PairTile t;
    for (PairWalk w(S, P); w.next(&t);) {
      const DevProblem& d = probs[t.p];
      if (tile_tokm(d, t)) {
        // ---- token-major drain: TMEM lane = token (this CTA's half of the chunk), TMEM column = weight row of the tile.  A
        // thread owns ONE output row and walks its columns 32 at a time (64 contiguous bytes, 32 with the fused silu*mul, whose
        // (gate_i, up_i) pair sits in adjacent registers of the same thread): a third of the weight-major drain's instructions.
        // The warp then transposes the chunk through its private 2 KB of shared memory so that one store instruction writes
        // 8 rows x 64 contiguous bytes (16 x 32 with silu) instead of 16 bytes in each of 32 rows.
        const int N = d.N;
        const int epi = d.epi;
        const int shift = d.acc_shift;
        const int64_t ldc = d.ldc;
        const bool silu = (d.epi_flags & 1) != 0;
        const bool facc = d.mma_kind == MMA_F8;
        const int half_rows = ((t.nt + 15) & ~15) >> 1;   // token rows each CTA stages (producer: trow = ... + rank * half)
        const int row_w0 = int(rank) * half_rows + q * 32;  // token (inside the tile) of this warp's lane 0
        const int trel = row_w0 + lane;
        const bool t_ok = q * 32 + lane < half_rows && trel < t.nt;
        const int grow = S.row0[t.p] + t.t0 + trel;       // row of A (scale_a)
        // C may start at another row than A (DevProblem::c_row): results written straight into a peer GPU's buffer
        const int crow_w0 = (d.c_row != nullptr ? static_cast<int>(ld_nc_u32(d.c_row)) : S.row0[t.p]) + t.t0 + row_w0;
        __half* sb_buf = S.sa_stage + acc * GP_BT;        // the tile's 256 per-channel scales (the buffer of the other orientation)
        __half2 sa2 = __float2half2_rn(1.f);
        if (epi == EPI_INT_ROWCOL) {
          sb_buf[et] = d.scale_b[t.n0 + et];
          if (t_ok) sa2 = __half2half2(d.scale_a[static_cast<int64_t>(grow) * d.scale_a_ld]);
        }
        named_bar_sync(2, 256);
        const bool vec = (reinterpret_cast<uintptr_t>(d.C) & 15) == 0 && (ldc & 7) == 0 && ((silu ? (N >> 1) : N) & 7) == 0;
        __half* cw0 = d.C + static_cast<int64_t>(crow_w0) * ldc + (silu ? (t.n0 >> 1) : t.n0);  // row of lane 0, first column of the tile
        const uint32_t wb = smem_u32(S.c_stage) + (grp * 4 + q) * 2048;  // this warp's transpose buffer
        mbar_wait(&S.tmem_full[acc], acc_phase);
#ifdef GP_TRACE
        if (warp == 4) GP_TR(6, tr_tile);
#endif
        tc_fence_after();
#ifdef GP_X_NOEPI
        for (int c = GP_BN; c < GP_BN; c += 64) {
#else
        // a warp whose 32 lanes lie past the chunk's rows (half_rows <= 96: the last quadrant) has nothing to drain
        for (int c = q * 32 < half_rows ? grp * 32 : GP_BN; c < GP_BN; c += 64) {
#endif
          uint32_t v[32];
          tmem_ld_32x32(lane_addr + acc * 256 + c, v);
          tmem_ld_wait();
#ifdef GP_X_LDONLY  // timing experiment: the accumulator is read and dropped (the test below is never true)
          if (t.nt < 100000) continue;
#endif
          uint32_t h[16];
          if (epi == EPI_F32) {
#pragma unroll
            for (int j = 0; j < 16; ++j) h[j] = pack_half2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
          } else {
            if (!facc) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__int2float_rn(static_cast<int32_t>(v[j]) >> shift));
            }
            const uint4* sbp = reinterpret_cast<const uint4*>(sb_buf + c);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint4 s4 = sbp[g];
              const uint32_t sw[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const __half2 prod = __hmul2(sa2, *reinterpret_cast<const __half2*>(&sw[j]));
                const float2 pf = __half22float2(prod);
                h[g * 4 + j] = pack_half2(__uint_as_float(v[g * 8 + 2 * j]) * pf.x, __uint_as_float(v[g * 8 + 2 * j + 1]) * pf.y);
              }
            }
          }
          if (silu) {
            // fused act_fn(gate) * up (MXB_EPI_SILU_MUL): columns (2i, 2i+1) of the tile are (gate_i, up_i) = the two halves of
            // h[i]: 16 outputs per chunk, from the fp16-rounded gate / up the unfused GEMM would have stored
            uint32_t o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const __half2 a = *reinterpret_cast<const __half2*>(&h[2 * j]), b = *reinterpret_cast<const __half2*>(&h[2 * j + 1]);
              o[j] = pack_half2(silu_mul_h(__low2half(a), __high2half(a)), silu_mul_h(__low2half(b), __high2half(b)));
            }
            if (vec) {
              // 32 rows x 32 B: piece p (16 B) of row r at r * 32 + ((p ^ (r >> 2 & 1)) << 4): conflict-free both ways
              const uint32_t sx = static_cast<uint32_t>((lane >> 2) & 1) << 4;
              st_shared_v4(wb + lane * 32 + sx, make_uint4(o[0], o[1], o[2], o[3]));
              st_shared_v4(wb + lane * 32 + (sx ^ 16u), make_uint4(o[4], o[5], o[6], o[7]));
              __syncwarp();
#pragma unroll
              for (int it = 0; it < 2; ++it) {
                const int r = (lane >> 1) + 16 * it, pc = lane & 1;
                const uint4 val = ld_shared_v4(wb + r * 32 + ((static_cast<uint32_t>(pc) ^ static_cast<uint32_t>((r >> 2) & 1)) << 4));
                if (q * 32 + r < half_rows && row_w0 + r < t.nt GP_X_STG_COND)
                  st_global_v4(cw0 + static_cast<int64_t>(r) * ldc + (c >> 1) + pc * 8, val);
              }
              __syncwarp();
            } else if (t_ok) {
              __half* op = cw0 + static_cast<int64_t>(lane) * ldc + (c >> 1);
#pragma unroll
              for (int j = 0; j < 16; ++j) st_global_b16(op + j, static_cast<uint16_t>(j & 1 ? o[j >> 1] >> 16 : o[j >> 1] & 0xFFFFu));
            }
            continue;
          }
          if (vec) {
            // 32 rows x 64 B: piece p of row r at r * 64 + ((p ^ (r >> 1 & 3)) << 4)
            const uint32_t sx = static_cast<uint32_t>((lane >> 1) & 3);
#pragma unroll
            for (int pc = 0; pc < 4; ++pc)
              st_shared_v4(wb + lane * 64 + ((static_cast<uint32_t>(pc) ^ sx) << 4), make_uint4(h[4 * pc], h[4 * pc + 1], h[4 * pc + 2], h[4 * pc + 3]));
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 4; ++it) {
              const int r = (lane >> 2) + 8 * it, pc = lane & 3;
              const uint4 val = ld_shared_v4(wb + r * 64 + ((static_cast<uint32_t>(pc) ^ static_cast<uint32_t>((r >> 1) & 3)) << 4));
              if (q * 32 + r < half_rows && row_w0 + r < t.nt GP_X_STG_COND) st_global_v4(cw0 + static_cast<int64_t>(r) * ldc + c + pc * 8, val);
            }
            __syncwarp();
          } else if (t_ok) {
            __half* op = cw0 + static_cast<int64_t>(lane) * ldc + c;
#pragma unroll
            for (int j = 0; j < 32; ++j) st_global_b16(op + j, static_cast<uint16_t>(j & 1 ? h[j >> 1] >> 16 : h[j >> 1] & 0xFFFFu));
          }
        }
      } else {
      const int N = d.N;
      const int n = t.n0 + int(rank) * 128 + q * 32 + lane;  // this thread's output column (weight row)
      const bool n_ok = n < N;
      const int epi = d.epi;
      const int shift = d.acc_shift;
      const int64_t ldc = d.ldc;
      const bool silu = (d.epi_flags & 1) != 0;
      const bool facc = d.mma_kind == MMA_F8;
      const int grow0 = S.row0[t.p] + t.t0;  // global row of the tile's first token
      // C may start at another row than A (DevProblem::c_row): results written straight into a peer GPU's buffer
      const int crow0 = (d.c_row != nullptr ? static_cast<int>(ld_nc_u32(d.c_row)) : S.row0[t.p]) + t.t0;
      __half* sa_buf = S.sa_stage + acc * GP_BT;
      __half2 sb2 = __float2half2_rn(1.f);
      if (epi == EPI_INT_ROWCOL) {
        // stage the tile's per-token scales (each thread 2 of them), fetch this lane's per-channel scale
        const int64_t sald = d.scale_a_ld;  // 1, or the pitch of a scale buffer shared with group-wise experts
        const __half* sa = d.scale_a + grow0 * sald;
        sa_buf[et] = et < t.nt ? sa[et * sald] : __float2half(0.f);
        if (n_ok) sb2 = __half2half2(d.scale_b[n]);
      }
      // every tile, also the unscaled ones: keeps the eight warps within one tile of each other, which is what makes
      // the two-deep sa_stage buffer safe
      named_bar_sync(2, 256);
      __half* Cp = d.C + static_cast<int64_t>(crow0) * ldc + n;
      // Vector form of the store (the normal case): the 4 warps of the group transpose a 32-token chunk through shared
      // memory and every lane writes 16 bytes of a token's row (256 contiguous bytes per half warp, or 128 with the fused
      // silu*mul) instead of 2 bytes at a 2*ldc stride: 8x fewer store instructions and full 128-byte lines — what a store
      // into ANOTHER GPU's memory over NVLink needs (expert-parallel combine, MxbProblem.c_row).
      int32_t* ao = d.acc_out != nullptr ? d.acc_out + static_cast<int64_t>(t.t0) * N + n : nullptr;
      // Only where it pays: a store into a PEER's memory (c_row) wants 16-byte rows; into local HBM the 2-byte stores below
      // cost fewer instructions than the transpose (c2 GEMM pair 0.479 ms against 0.509 with the transpose everywhere), and
      // issue slots are what this kernel runs out of.  MXB_EPI_VEC=1 (epi_flags bit 8) forces the transpose for measurements.
      const bool vec = (d.c_row != nullptr || (d.epi_flags & 256) != 0) && (reinterpret_cast<uintptr_t>(d.C) & 15) == 0 && (ldc & 7) == 0 &&
                       ((silu ? (N >> 1) : N) & 7) == 0 && ao == nullptr;
      const uint32_t cs = smem_u32(S.c_stage) + grp * (32 * GP_CS_PITCH);
      const int ncol0 = t.n0 + int(rank) * 128;  // first weight row (output column) of this CTA's half of the tile
      mbar_wait(&S.tmem_full[acc], acc_phase);
#ifdef GP_TRACE
      if (warp == 4) GP_TR(6, tr_tile);
#endif
      tc_fence_after();
#ifdef GP_X_NOEPI  // timing experiment (wrong results): the accumulator is handed back undrained
      for (int c = t.nt; c < t.nt; c += 64) {
#else
      for (int c = grp * 32; c < t.nt; c += 64) {
#endif
        uint32_t v[32];
        tmem_ld_32x32(lane_addr + acc * 256 + c, v);
        tmem_ld_wait();
        const int valid = n_ok ? min(32, t.nt - c) : 0;
        uint32_t h[16];
        if (epi == EPI_F32) {
#pragma unroll
          for (int j = 0; j < 16; ++j) h[j] = pack_half2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
        } else {
          if (ao != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < valid) ao[static_cast<int64_t>(c + j) * N] = static_cast<int32_t>(v[j]) >> shift;
          }
          const uint4* sap = reinterpret_cast<const uint4*>(sa_buf + c);
          // integer accumulators (>> shift: the x16 of the in-smem w4 expansion) become floats in place; the fp32 ones of
          // kind::f8f6f4 already are (one branch per tile, not a select per element)
          if (!facc) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__int2float_rn(static_cast<int32_t>(v[j]) >> shift));
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint4 s4 = sap[g];
            const uint32_t sw[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const __half2 prod = __hmul2(*reinterpret_cast<const __half2*>(&sw[j]), sb2);
              const float2 pf = __half22float2(prod);
              h[g * 4 + j] = pack_half2(__uint_as_float(v[g * 8 + 2 * j]) * pf.x, __uint_as_float(v[g * 8 + 2 * j + 1]) * pf.y);
            }
          }
        }
        if (silu) {
          // Fused act_fn(gate) * up (MXB_EPI_SILU_MUL): even lanes hold gate_i, odd lanes up_i (interleaved weight rows).
          // The pair swaps one value per two tokens; the even lane finishes token c+2j, the odd lane token c+2j+1, each
          // from the fp16-rounded gate/up the unfused GEMM would have stored (silu_mul_h: quant_common.cuh).
          const int odd = lane & 1;
          if (vec) {
            const uint32_t colb = cs + odd * GP_CS_PITCH + ((q * 32 + lane) >> 1) * 2;  // this pair's output column, rows c+odd, c+odd+2, ..
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              const uint32_t send = odd ? ((h[j] & 0xFFFFu) | (h[j + 1] << 16)) : ((h[j] >> 16) | (h[j + 1] & 0xFFFF0000u));
              const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 1);
#pragma unroll
              for (int jj = 0; jj < 2; ++jj) {
                const uint16_t own = odd ? static_cast<uint16_t>(h[j + jj] >> 16) : static_cast<uint16_t>(h[j + jj] & 0xFFFFu);
                const uint16_t oth = static_cast<uint16_t>(jj ? (recv >> 16) : (recv & 0xFFFFu));
                const __half g = __ushort_as_half(odd ? oth : own), u = __ushort_as_half(odd ? own : oth);
                st_shared_b16(colb + 2 * (j + jj) * GP_CS_PITCH, __half_as_ushort(__float2half_rn(silu_mul_h(g, u))));
              }
            }
            named_bar_sync(3 + grp, 128);
            const int ntok = min(32, t.nt - c);
            const int hcol = (ncol0 >> 1) + (lane & 7) * 8;  // 64 outputs of this CTA per token: 8 lanes x 16 bytes
#pragma unroll
            for (int it = 0; it < 2; ++it) {
              const int row = q * 8 + it * 4 + (lane >> 3);
              const uint4 val = ld_shared_v4(cs + row * GP_CS_PITCH + (lane & 7) * 16);
              if (row < ntok && hcol < (N >> 1)) st_global_v4(d.C + static_cast<int64_t>(crow0 + c + row) * ldc + hcol, val);
            }
            named_bar_sync(3 + grp, 128);  // the chunk buffer is rewritten by the group's next chunk
            continue;
          }
          __half* hp = d.C + static_cast<int64_t>(crow0 + c + odd) * ldc + (n >> 1);
#pragma unroll
          for (int j = 0; j < 16; j += 2) {
            const uint32_t send = odd ? ((h[j] & 0xFFFFu) | (h[j + 1] << 16)) : ((h[j] >> 16) | (h[j + 1] & 0xFFFF0000u));
            const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 1);
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
              const uint16_t own = odd ? static_cast<uint16_t>(h[j + jj] >> 16) : static_cast<uint16_t>(h[j + jj] & 0xFFFFu);
              const uint16_t oth = static_cast<uint16_t>(jj ? (recv >> 16) : (recv & 0xFFFFu));
              const __half g = __ushort_as_half(odd ? oth : own), u = __ushort_as_half(odd ? own : oth);
              const __half r = __float2half_rn(silu_mul_h(g, u));
              if (2 * (j + jj) + odd < valid) st_global_b16(hp, __half_as_ushort(r));
              hp += 2 * ldc;
            }
          }
          continue;
        }
        if (vec) {
          const uint32_t colb = cs + (q * 32 + lane) * 2;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            st_shared_b16(colb + (2 * j) * GP_CS_PITCH, static_cast<uint16_t>(h[j] & 0xFFFFu));
            st_shared_b16(colb + (2 * j + 1) * GP_CS_PITCH, static_cast<uint16_t>(h[j] >> 16));
          }
          named_bar_sync(3 + grp, 128);
          const int ntok = min(32, t.nt - c);
          const int col = ncol0 + (lane & 15) * 8;  // 128 outputs of this CTA per token: 16 lanes x 16 bytes
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            const int row = q * 8 + it * 2 + (lane >> 4);
            const uint4 val = ld_shared_v4(cs + row * GP_CS_PITCH + (lane & 15) * 16);
            if (row < ntok && col < N) st_global_v4(d.C + static_cast<int64_t>(crow0 + c + row) * ldc + col, val);
          }
          named_bar_sync(3 + grp, 128);  // the chunk buffer is rewritten by the group's next chunk
          continue;
        }
        __half* cp = Cp + static_cast<int64_t>(c) * ldc;
        if (valid == 32) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            st_global_b16(cp, static_cast<uint16_t>(h[j] & 0xFFFFu));
            st_global_b16(cp + ldc, static_cast<uint16_t>(h[j] >> 16));
            cp += 2 * ldc;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (2 * j < valid) st_global_b16(cp, static_cast<uint16_t>(h[j] & 0xFFFFu));
            if (2 * j + 1 < valid) st_global_b16(cp + ldc, static_cast<uint16_t>(h[j] >> 16));
            cp += 2 * ldc;
          }
        }
      }
      }  // weight-major drain
      tc_fence_before();
      __syncwarp();
#ifdef GP_TRACE
      if (warp == 4) GP_TR(7, tr_tile++);
#endif
      if (lane == 0) mbar_arrive_cluster(tmem_empty_leader + acc * 8);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
#ifdef GP_CONV_W2
  } else if (warp == 8 || warp == 2 || warp == 10 || warp == 11) {
    // ============================================================ 4-bit weight converters (4 warps in EACH CTA): warp 2 (idle
    // after the TMEM allocation) instead of warp 9, which shares its scheduler with the MMA issuer
    const int ct = (warp == 2 ? 32 : (warp - 8) * 32) + lane;  // 0..127
#else
  } else if (warp >= 8 && warp < 12) {
    // ============================================================ 4-bit weight converters (4 warps in EACH CTA)
    const int ct = threadIdx.x - 256;  // 0..127
#endif
    int stage = 0;
    uint32_t cphase_bits = 0;
#ifdef GP_CONV_FULL
    int kbc = 0;  // k-blocks of every tile so far: full[stage] is in phase (kbc / GP_STAGES) & 1
#endif
    const uint32_t conv_leader = mapa_u32(smem_u32(&S.conv[0]), 0);
#ifdef GP_TRACE
    int tr_kb = 0;
#endif
    PairTile t;
    for (PairWalk w(S, P); w.next(&t);) {
      const DevProblem& d = probs[t.p];
      const int kblocks = d.kblocks;
      if (d.conv == CONV_NONE) {
        stage = (stage + kblocks) % GP_STAGES;
#ifdef GP_CONV_FULL
        kbc += kblocks;
#endif
        continue;
      }
      for (int kb = 0; kb < kblocks; ++kb) {
        mbar_wait(&S.cfull[stage], (cphase_bits >> stage) & 1);
        cphase_bits ^= 1u << stage;
#ifdef GP_TRACE
        if (warp == 8) GP_TR(4, tr_kb);
#endif
        uint8_t* wfin = S.stages + stage * GP_STAGE_BYTES;
        const uint8_t* wpk = wfin + GP_PACKED_OFF;
        // packed slab: 128 rows x 64 B; a 16-byte item holds 32 k (low nibbles k..k+15, high nibbles k+16..k+31).
        // Everything is read into registers before the first write: the expansion overwrites the packed bytes.
        uint4 w4[4];
#ifdef GP_X_NOCONV  // timing experiment (wrong results): the slab is handed on unexpanded
        (void)wpk; (void)wfin; (void)w4;
        for (int j = 4; j < 4; ++j) {
#else
#pragma unroll
        for (int j = 0; j < 4; ++j) w4[j] = *reinterpret_cast<const uint4*>(wpk + (ct + j * 128) * 16);
        named_bar_sync(1, 128);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#endif
          const int i = ct + j * 128;
          const int row = i >> 2, c = i & 3;
          uint4 lo, hi;
          lo.x = (w4[j].x << 4) & 0xF0F0F0F0u; hi.x = w4[j].x & 0xF0F0F0F0u;
          lo.y = (w4[j].y << 4) & 0xF0F0F0F0u; hi.y = w4[j].y & 0xF0F0F0F0u;
          lo.z = (w4[j].z << 4) & 0xF0F0F0F0u; hi.z = w4[j].z & 0xF0F0F0F0u;
          lo.w = (w4[j].w << 4) & 0xF0F0F0F0u; hi.w = w4[j].w & 0xF0F0F0F0u;
          const int sw = row & 7;
          *reinterpret_cast<uint4*>(wfin + row * 128 + (((2 * c) ^ sw) << 4)) = lo;
          *reinterpret_cast<uint4*>(wfin + row * 128 + (((2 * c + 1) ^ sw) << 4)) = hi;
        }
        fence_proxy_async_smem();
#ifdef GP_CONV_FULL
        // the leader's converters also wait for the tokens of both CTAs: `conv` complete then implies `full` complete and
        // the MMA issuer — whose time per k-block paces the tensor pipe — waits on ONE barrier
        if (leader) mbar_wait(&S.full[stage], (kbc / GP_STAGES) & 1);
        ++kbc;
#endif
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(conv_leader + stage * 8);
#ifdef GP_TRACE
        if (warp == 8) GP_TR(5, tr_kb++);
#endif
        if (++stage == GP_STAGES) stage = 0;
      }
    }
  }

  // ---- teardown: nobody leaves while the peer may still read its smem / arrive on its barriers
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

int gg_pair_configure() {
  return static_cast<int>(cudaFuncSetAttribute(gg_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GP_SMEM_BYTES));
}

int gg_pair_smem_bytes() { return GP_SMEM_BYTES; }
int gg_pair_stages() { return GP_STAGES; }

void gg_pair_launch(const DevProblem* probs, int P, int grid, cudaStream_t stream) {
  gg_pair_kernel<<<grid & ~1, GP_THREADS, GP_SMEM_BYTES, stream>>>(probs, P);
}

}  // namespace mxb

#ifdef GP_TRACE
extern "C" int mxb_debug_gp_trace(unsigned long long* out, int clear) {
  using namespace mxb;
  if (clear) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, gp_trace_buf) != cudaSuccess) return -1;
    return int(cudaMemset(p, 0, sizeof(gp_trace_buf)));
  }
  return int(cudaMemcpyFromSymbol(out, gp_trace_buf, sizeof(gp_trace_buf)));
}
#endif
