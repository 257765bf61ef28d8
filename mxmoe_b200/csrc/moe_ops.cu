// Pieces (1), (2) and (5) of the hot path plus the offline weight quantize/pack kernels.
// All of them are HBM-bound byte/short kernels: 16-byte vector accesses, one warp per token row,
// grids sized from the row count; no tensor cores except the K x K Hadamard factor (hadamard.cu).
//
// Reference semantics restated here (paths relative to the reference repo):
//   permute           stable sort of flattened expert ids, ref_bind.cu:47-64, :449-453
//   per-token quant   quant_minmax (mxmoe/quant/quant.py:40-84) in fp16 arithmetic, sym, with the
//                     scale==0 -> 1 guard of quantize.cuh:256; call sites ref_bind.cu:531-544, 701-716
//   silu*mul          HF act_fn(gate)*up in fp16 tensors (silu in fp32, rounded to fp16, product rounded)
//   combine           routing-weighted sum, weights as moe_tracer.py:79-86 (absent in ref_bind.cu:66)
//   quant_weight      quantize.cuh:218-279
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "moe_ops.cuh"
#include "quant_common.cuh"

namespace mxb {

// ============================================================================ permute
namespace {
constexpr int PERM_THREADS = 256;
constexpr int PERM_ITEMS = 8;
constexpr int PERM_CHUNK = PERM_THREADS * PERM_ITEMS;

// ws layout: [0] ticket, [1..1+nblk*E) block_hist -> turned into exclusive per-expert block bases
__global__ void __launch_bounds__(PERM_THREADS) permute_hist_kernel(const int32_t* __restrict__ ids, int n, int E,
                                                                     int32_t* __restrict__ ws,
                                                                     int32_t* __restrict__ expert_offsets) {
  extern __shared__ int32_t sh[];  // E ints (+1 flag)
  const int nblk = gridDim.x, b = blockIdx.x;
  for (int e = threadIdx.x; e < E; e += PERM_THREADS) sh[e] = 0;
  __syncthreads();
  const int base = b * PERM_CHUNK;
#pragma unroll
  for (int it = 0; it < PERM_ITEMS; ++it) {
    int i = base + it * PERM_THREADS + threadIdx.x;
    if (i < n) {
      int e = ids[i];
      if (e >= 0 && e < E) atomicAdd(&sh[e], 1);
    }
  }
  __syncthreads();
  int32_t* hist = ws + 1;
  for (int e = threadIdx.x; e < E; e += PERM_THREADS) hist[b * E + e] = sh[e];
  __threadfence();
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) last = (atomicAdd(&ws[0], 1) == nblk - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  // last block: per expert, exclusive scan over blocks; then scan totals over experts
  for (int e = threadIdx.x; e < E; e += PERM_THREADS) {
    int run = 0;
    for (int bb = 0; bb < nblk; ++bb) {
      int c = hist[bb * E + e];
      hist[bb * E + e] = run;
      run += c;
    }
    sh[e] = run;
  }
  __syncthreads();
  if (threadIdx.x < 32) {  // warp scan of totals, 32 at a time
    int carry = 0;
    const int lane = threadIdx.x;
    if (lane == 0) expert_offsets[0] = 0;
    for (int b0 = 0; b0 < E; b0 += 32) {
      int idx = b0 + lane;
      int v = idx < E ? sh[idx] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      if (idx < E) expert_offsets[idx + 1] = v + carry;
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
    if (lane == 0) ws[0] = 0;  // re-arm the ticket for the next call
  }
}

__global__ void __launch_bounds__(PERM_THREADS) permute_scatter_kernel(const int32_t* __restrict__ ids, int n, int k, int E,
                                                                        const int32_t* __restrict__ ws,
                                                                        const int32_t* __restrict__ expert_offsets,
                                                                        int32_t* __restrict__ src_row,
                                                                        int32_t* __restrict__ slot_of) {
  extern __shared__ int32_t sh[];
  int32_t* run = sh;            // [E] running count inside this block
  int32_t* wcnt = sh + E;       // [8][E] per-warp counts of the current round
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int32_t* base_tab = ws + 1 + b * E;
  for (int e = threadIdx.x; e < E; e += PERM_THREADS) run[e] = expert_offsets[e] + base_tab[e];
  const int base = b * PERM_CHUNK;
  for (int it = 0; it < PERM_ITEMS; ++it) {
    for (int j = threadIdx.x; j < 8 * E; j += PERM_THREADS) wcnt[j] = 0;
    __syncthreads();
    const int i = base + it * PERM_THREADS + threadIdx.x;
    int e = -1;
    if (i < n) {
      e = ids[i];
      if (e < 0 || e >= E) e = -1;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, e);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    if (e >= 0 && rank == 0) wcnt[warp * E + e] = __popc(peers);
    __syncthreads();
    if (e >= 0) {
      int off = run[e] + rank;
      for (int w2 = 0; w2 < warp; ++w2) off += wcnt[w2 * E + e];
      slot_of[i] = off;
      src_row[off] = i / k;
    }
    __syncthreads();
    for (int e2 = threadIdx.x; e2 < E; e2 += PERM_THREADS) {
      int s = 0;
#pragma unroll
      for (int w2 = 0; w2 < 8; ++w2) s += wcnt[w2 * E + e2];
      run[e2] += s;
    }
    __syncthreads();
  }
}

// n <= PERM_CHUNK (decode and small prefill batches): histogram, offsets and the stable scatter in ONE block and one launch
// (the two-kernel form costs a second launch latency on a step that lasts ~150 us).  Same order as the pair above: rounds
// of PERM_THREADS ids, inside a round by warp then lane.
__global__ void __launch_bounds__(PERM_THREADS) permute_small_kernel(const int32_t* __restrict__ ids, int n, int k, int E,
                                                                      int32_t* __restrict__ expert_offsets,
                                                                      int32_t* __restrict__ src_row, int32_t* __restrict__ slot_of) {
  extern __shared__ int32_t sh[];
  int32_t* run = sh;       // [E] histogram, then the running output position of every expert
  int32_t* wcnt = sh + E;  // [8][E]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int my_e[PERM_ITEMS];
  for (int e = threadIdx.x; e < E; e += PERM_THREADS) run[e] = 0;
  __syncthreads();
#pragma unroll
  for (int it = 0; it < PERM_ITEMS; ++it) {
    const int i = it * PERM_THREADS + threadIdx.x;
    int e = -1;
    if (i < n) {
      e = ids[i];
      if (e < 0 || e >= E) e = -1;
    }
    my_e[it] = e;
    if (e >= 0) atomicAdd(&run[e], 1);
  }
  __syncthreads();
  if (warp == 0) {  // exclusive scan of the counts, 32 experts at a time
    int carry = 0;
    if (lane == 0) expert_offsets[0] = 0;
    for (int b0 = 0; b0 < E; b0 += 32) {
      const int idx = b0 + lane;
      const int c = idx < E ? run[idx] : 0;
      int v = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      if (idx < E) {
        expert_offsets[idx + 1] = v + carry;
        run[idx] = v + carry - c;
      }
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
  }
  const int rounds = (n + PERM_THREADS - 1) / PERM_THREADS;
#pragma unroll
  for (int it = 0; it < PERM_ITEMS; ++it) {
    if (it >= rounds) break;
    for (int j = threadIdx.x; j < 8 * E; j += PERM_THREADS) wcnt[j] = 0;
    __syncthreads();
    const int i = it * PERM_THREADS + threadIdx.x;
    const int e = my_e[it];
    const unsigned peers = __match_any_sync(0xffffffffu, e);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    if (e >= 0 && rank == 0) wcnt[warp * E + e] = __popc(peers);
    __syncthreads();
    if (e >= 0) {
      int off = run[e] + rank;
      for (int w2 = 0; w2 < warp; ++w2) off += wcnt[w2 * E + e];
      slot_of[i] = off;
      src_row[off] = i / k;
    }
    __syncthreads();
    if (it + 1 < rounds) {
      for (int e2 = threadIdx.x; e2 < E; e2 += PERM_THREADS) {
        int s2 = 0;
#pragma unroll
        for (int w2 = 0; w2 < 8; ++w2) s2 += wcnt[w2 * E + e2];
        run[e2] += s2;
      }
      __syncthreads();
    }
  }
}
}  // namespace

size_t permute_ws_bytes(int n, int E) {
  int nblk = (n + PERM_CHUNK - 1) / PERM_CHUNK;
  return sizeof(int32_t) * (1 + static_cast<size_t>(nblk) * E);
}

cudaError_t permute_launch(const int32_t* ids, int T, int k, int E, int32_t* expert_offsets, int32_t* src_row,
                           int32_t* slot_of, int32_t* ws, cudaStream_t st) {
  const int n = T * k;
  const int nblk = (n + PERM_CHUNK - 1) / PERM_CHUNK;
  if (nblk == 0) {
    return cudaMemsetAsync(expert_offsets, 0, sizeof(int32_t) * (E + 1), st);
  }
  if (nblk == 1 && E <= 1024) {
    permute_small_kernel<<<1, PERM_THREADS, sizeof(int32_t) * 9 * E, st>>>(ids, n, k, E, expert_offsets, src_row, slot_of);
    return cudaGetLastError();
  }
  permute_hist_kernel<<<nblk, PERM_THREADS, sizeof(int32_t) * (E + 1), st>>>(ids, n, E, ws, expert_offsets);
  permute_scatter_kernel<<<nblk, PERM_THREADS, sizeof(int32_t) * 9 * E, st>>>(ids, n, k, E, ws, expert_offsets, src_row,
                                                                              slot_of);
  return cudaGetLastError();
}

// ---- permute without any allocation (mxb_permute_static; no counterpart in the reference) -------------------------------
// The same kernels, with the block-histogram scratch in a file-scope device array instead of the handle's cudaMalloc'ed
// workspace.  PERM_STATIC_INTS * 4 = 256 KiB per GPU that loads the library, held until exit.  In HOST code the array's
// name is only a stand-in for the symbol, not a device address: the address is looked up once per device, with that
// device current, and that pointer is what the kernels receive.
namespace {
constexpr int PERM_STATIC_INTS = 64 * 1024;
static __device__ __align__(256) int32_t g_permute_scratch[PERM_STATIC_INTS];
constexpr int PERM_STATIC_MAX_DEV = 64;
int32_t* g_permute_scratch_addr[PERM_STATIC_MAX_DEV] = {};
std::mutex g_permute_scratch_mu;
}  // namespace

size_t permute_static_capacity() { return sizeof(int32_t) * PERM_STATIC_INTS; }

// `device` must be the current device.  Every call on a device shares the one array: calls that overlap on the GPU
// (different streams, no ordering between them) race on the ticket and the block histograms.
cudaError_t permute_static_launch(int device, const int32_t* ids, int T, int k, int E, int32_t* expert_offsets, int32_t* src_row,
                                  int32_t* slot_of, cudaStream_t st) {
  if (device < 0 || device >= PERM_STATIC_MAX_DEV) return cudaErrorInvalidDevice;
  if (permute_ws_bytes(T * k, E) > permute_static_capacity()) return cudaErrorInvalidValue;
  int32_t* ws = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_permute_scratch_mu);
    ws = g_permute_scratch_addr[device];
    if (!ws) {
      const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ws), g_permute_scratch);
      if (e != cudaSuccess) return e;
      g_permute_scratch_addr[device] = ws;
    }
  }
  // the array keeps the previous call's contents: the histograms are fully rewritten by permute_hist_kernel, the
  // ticket ws[0] must start at zero (a call that was aborted mid-way would leave it armed)
  if (T * k > PERM_CHUNK) {  // the two-kernel form is the one that uses the scratch
    const cudaError_t e = cudaMemsetAsync(ws, 0, sizeof(int32_t), st);
    if (e != cudaSuccess) return e;
  }
  return permute_launch(ids, T, k, E, expert_offsets, src_row, slot_of, ws, st);
}

// ============================================================================ per-token quantization
namespace {
// One warp per output row (grid-stride).  SILU=false: x row gathered through src_row; SILU=true: row = [gate | up].
// dest != nullptr (expert-parallel dispatch over peer memory): row r of expert e is written to
// dest[e].{q8,f16,scales} at row dest[e].row0 + (r - expert_offsets[e]) — buffers of another GPU, reached over NVLink.
template <bool SILU>
__global__ void __launch_bounds__(256) quant_rows_kernel(const __half* __restrict__ x, int H, int64_t in_pitch,
                                                          const int32_t* __restrict__ src_row, int n_rows,
                                                          const int32_t* __restrict__ expert_offsets,
                                                          const int2* __restrict__ actq, int E, int8_t* __restrict__ out_q8,
                                                          __half* __restrict__ out_f16, __half* __restrict__ scales,
                                                          int scale_cols, const RowDest* __restrict__ dest) {
  const int lane = threadIdx.x & 31;
  if (E > 1 || dest != nullptr) n_rows = min(n_rows, expert_offsets[E]);  // rows past the last expert are not owned by anyone
  for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += gridDim.x * 8) {
    int a_bits = actq[0].x, gsize = actq[0].y;
    int64_t orow = row;
    int8_t* q8 = out_q8;
    __half* f16 = out_f16;
    __half* sc = scales;
    if (E > 1 || dest != nullptr) {
      const int e = E > 1 ? find_expert(expert_offsets, E, row) : 0;
      a_bits = actq[E > 1 ? e : 0].x;
      gsize = actq[E > 1 ? e : 0].y;
      if (dest != nullptr) {
        const RowDest d = dest[e];
        q8 = static_cast<int8_t*>(d.q8);
        f16 = static_cast<__half*>(d.f16);
        sc = static_cast<__half*>(d.scales);
        orow = d.row0 + (row - expert_offsets[e]);
      }
    }
    const int64_t srow = src_row ? src_row[row] : row;
    const __half* in = x + srow * in_pitch;
    auto ld = [&](int i, float (&v)[8]) { load8<SILU>(in, H, i, v); };
    quantize_row(ld, H, lane, a_bits, gsize, q8 ? q8 + orow * H : nullptr, f16 ? f16 + orow * H : nullptr,
                 sc ? sc + orow * scale_cols : nullptr);
  }
}
}  // namespace

cudaError_t quant_rows_launch(bool silu, const void* x, int H, int64_t in_pitch, const int32_t* src_row, int n_rows,
                              const int32_t* expert_offsets, const int32_t* actq_dev, int E, void* out_q8, void* out_f16,
                              void* scales, int scale_cols, const RowDest* dest, int max_ctas, cudaStream_t st) {
  if (n_rows == 0) return cudaSuccess;
  int grid = (n_rows + 7) / 8;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  if (silu)
    quant_rows_kernel<true><<<grid, 256, 0, st>>>(static_cast<const __half*>(x), H, in_pitch, src_row, n_rows, expert_offsets,
                                                  reinterpret_cast<const int2*>(actq_dev), E, static_cast<int8_t*>(out_q8),
                                                  static_cast<__half*>(out_f16), static_cast<__half*>(scales), scale_cols, dest);
  else
    quant_rows_kernel<false><<<grid, 256, 0, st>>>(static_cast<const __half*>(x), H, in_pitch, src_row, n_rows, expert_offsets,
                                                   reinterpret_cast<const int2*>(actq_dev), E, static_cast<int8_t*>(out_q8),
                                                   static_cast<__half*>(out_f16), static_cast<__half*>(scales), scale_cols, dest);
  return cudaGetLastError();
}

// ============================================================================ expert-parallel exchange plan
namespace {
// counts of this rank -> row `rank` of every peer's [W, E] count matrix (peer memory)
__global__ void ep_push_counts_kernel(const int32_t* __restrict__ expert_offsets, int E, int32_t* const* __restrict__ peer_counts, int W,
                                      int rank) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < W * E; i += gridDim.x * blockDim.x) {
    const int p = i / E, e = i - p * E;
    peer_counts[p][rank * E + e] = expert_offsets[e + 1] - expert_offsets[e];
  }
}

__device__ __forceinline__ int ep_bound(int r, int E, int W) { return static_cast<int>(static_cast<int64_t>(r) * E / W); }

// From the all-ranks count matrix: where my rows go (dest), how what I receive is laid out (seg_offs: source-major,
// local-expert-minor, mxmoe_b200/ep.py ep_plan) and where every received segment's results return to (back_row:
// the segment's first row in the SOURCE rank's expert-sorted order).
__global__ void ep_plan_kernel(const int32_t* __restrict__ counts, int W, int rank, int E, void* const* __restrict__ peer_q8,
                               void* const* __restrict__ peer_f16, void* const* __restrict__ peer_scales, int32_t* __restrict__ seg_offs,
                               int32_t* __restrict__ back_row, RowDest* __restrict__ dest) {
  const int lo = ep_bound(rank, E, W), hi = ep_bound(rank + 1, E, W), EL = hi - lo;
  const int tid = threadIdx.x;
  // dispatch destinations: expert e lives on rank p; my rows follow those of the ranks before me, experts in order
  for (int e = tid; e < E; e += blockDim.x) {
    int p = 0;
    while (ep_bound(p + 1, E, W) <= e) ++p;
    const int plo = ep_bound(p, E, W), phi = ep_bound(p + 1, E, W);
    int64_t r0 = 0;
    for (int s = 0; s < rank; ++s)
      for (int q = plo; q < phi; ++q) r0 += counts[s * E + q];
    for (int q = plo; q < e; ++q) r0 += counts[rank * E + q];
    RowDest d;
    d.q8 = peer_q8 ? peer_q8[p] : nullptr;
    d.f16 = peer_f16 ? peer_f16[p] : nullptr;
    d.scales = peer_scales ? peer_scales[p] : nullptr;
    d.row0 = r0;
    dest[e] = d;
  }
  // return rows: segment (s, i) = rows of source s for my expert lo+i; in s's sorted order they start after all of s's
  // rows for experts before lo+i
  for (int j = tid; j < W * EL; j += blockDim.x) {
    const int s = j / EL, i = j - s * EL;
    int r0 = 0;
    for (int q = 0; q < lo + i; ++q) r0 += counts[s * E + q];
    back_row[j] = r0;
  }
  __syncthreads();
  if (tid == 0) {  // receive layout (W*EL <= a few hundred)
    int acc = 0;
    seg_offs[0] = 0;
    for (int j = 0; j < W * EL; ++j) {
      const int s = j / EL, i = j - s * EL;
      acc += counts[s * E + lo + i];
      seg_offs[j + 1] = acc;
    }
  }
}
}  // namespace

cudaError_t ep_push_counts_launch(const int32_t* expert_offsets, int E, int32_t* const* peer_counts_dev, int W, int rank, cudaStream_t st) {
  ep_push_counts_kernel<<<1, 256, 0, st>>>(expert_offsets, E, peer_counts_dev, W, rank);
  return cudaGetLastError();
}
cudaError_t ep_plan_launch(const int32_t* counts, int W, int rank, int E, void* const* peer_q8, void* const* peer_f16,
                           void* const* peer_scales, int32_t* seg_offs, int32_t* back_row, RowDest* dest, cudaStream_t st) {
  ep_plan_kernel<<<1, 256, 0, st>>>(counts, W, rank, E, peer_q8, peer_f16, peer_scales, seg_offs, back_row, dest);
  return cudaGetLastError();
}

// ============================================================================ combine
namespace {
__global__ void __launch_bounds__(256) combine_kernel(const __half* __restrict__ down, int H, const int32_t* __restrict__ slot_of,
                                                       const float* __restrict__ topk_w, int T, int k,
                                                       const __half* __restrict__ shared_out,
                                                       const float* __restrict__ shared_scale, __half* __restrict__ out) {
  // one thread per (token, 8 output elements): a decode step (T of a few dozen) still spreads over the whole chip, and
  // the k gathered rows of a thread are independent loads in flight together
  const int n8 = H >> 3;
  const int64_t id = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x;
  if (id >= static_cast<int64_t>(T) * n8) return;
  const int t = static_cast<int>(id / n8), i = static_cast<int>(id - static_cast<int64_t>(t) * n8);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int j0 = 0; j0 < k; j0 += 4) {
    uint4 v[4];
    float w[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = j0 + jj;
      w[jj] = 0.f;
      v[jj] = make_uint4(0, 0, 0, 0);
      if (j < k) {
        w[jj] = topk_w[t * k + j];
        v[jj] = *reinterpret_cast<const uint4*>(down + static_cast<int64_t>(slot_of[t * k + j]) * H + i * 8);
      }
    }
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      if (j0 + jj < k) {  // fp32 accumulation in j order (oracle.combine)
        const __half* vh = reinterpret_cast<const __half*>(&v[jj]);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(w[jj], __half2float(vh[e]), acc[e]);
      }
    }
  }
  if (shared_out != nullptr) {
    uint4 v = *reinterpret_cast<const uint4*>(shared_out + static_cast<int64_t>(t) * H + i * 8);
    const __half* vh = reinterpret_cast<const __half*>(&v);
    // Qwen's shared-expert gate: sigmoid(shared_expert_gate(x)) per token, applied to the fp16 shared output
    const float ss = shared_scale != nullptr ? shared_scale[t] : 1.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = shared_scale != nullptr ? fmaf(ss, __half2float(vh[e]), acc[e]) : acc[e] + __half2float(vh[e]);
  }
  uint4 pk;
  __half2* ph = reinterpret_cast<__half2*>(&pk);
#pragma unroll
  for (int e = 0; e < 4; ++e) ph[e] = __floats2half2_rn(acc[2 * e], acc[2 * e + 1]);
  *reinterpret_cast<uint4*>(out + static_cast<int64_t>(t) * H + i * 8) = pk;
}
}  // namespace

cudaError_t combine_launch(const void* down, int H, const int32_t* slot_of, const float* topk_w, int T, int k,
                           const void* shared_out, const float* shared_scale, void* out, cudaStream_t st) {
  if (T == 0) return cudaSuccess;
  const int64_t units = static_cast<int64_t>(T) * (H >> 3);
  combine_kernel<<<static_cast<unsigned>((units + 255) / 256), 256, 0, st>>>(static_cast<const __half*>(down), H, slot_of, topk_w, T, k,
                                                                           static_cast<const __half*>(shared_out), shared_scale, static_cast<__half*>(out));
  return cudaGetLastError();
}

// ============================================================================ offline weight quantize + pack
namespace {
// One block per weight row.  Codes are staged in shared memory, then packed into the native layout.
//   layout 0: int8 two's complement [N,K]                         (w8a8)
//   layout 1: 4-bit two's complement, 32-k blocks, byte j = k j (low nibble) | k j+16 (high)   (w4a4)
//   layout 2: unsigned 4-bit codes, nibble order 0,2,4,6,1,3,5,7 inside each 32-bit word       (w4a16)
//   layout 3: unsigned 8-bit codes [N,K]                                                      (w8a16)
//   layout 4: E4M3 bytes [N,K], per-channel scale absmax / 448                                  (w8a8 E4M3)
__global__ void __launch_bounds__(256) quant_pack_weight_kernel(const __half* __restrict__ W, int N, int K, int bits, int gsize,
                                                                 int sym, int layout, uint8_t* __restrict__ packed,
                                                                 __half* __restrict__ scales) {
  extern __shared__ uint8_t codes[];  // K bytes
  const int n = blockIdx.x;
  const int g = gsize <= 0 ? K : gsize;
  const int G = K / g;
  const __half* w = W + static_cast<int64_t>(n) * K;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool e4m3 = layout == 4;  // 8-bit float codes: scale = absmax / 448, q = e4m3(w / scale), round to nearest even
  const float upper = e4m3 ? 448.f : (sym ? static_cast<float>((1 << (bits - 1)) - 1) : static_cast<float>((1 << bits) - 1));
  const float lower = sym ? -upper : 0.f;
  for (int gi = warp; gi < G; gi += 8) {
    float mx = -65504.f, mn = 65504.f, amx = 0.f;
    for (int i = lane; i < g; i += 32) {
      float v = __half2float(w[gi * g + i]);
      mx = fmaxf(mx, v);
      mn = fminf(mn, v);
      amx = fmaxf(amx, fabsf(v));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      amx = fmaxf(amx, __shfl_xor_sync(0xffffffffu, amx, o));
    }
    float zp = sym ? 0.f : mn;
    __half sh = sym ? __float2half_rn(__fdiv_rn(amx, upper)) : __float2half_rn(__fdiv_rn(h_round(mx - mn), upper));
    if (__half2float(sh) == 0.f) sh = __float2half(1.f);
    const float sf = __half2float(sh);
    if (lane == 0) {
      if (sym) scales[static_cast<int64_t>(gi) * N + n] = sh;
      else {
        scales[(static_cast<int64_t>(gi) * N + n) * 2] = sh;
        scales[(static_cast<int64_t>(gi) * N + n) * 2 + 1] = __float2half_rn(zp);
      }
    }
    const int bias = ((layout == 2 || layout == 3) && sym) ? (1 << (bits - 1)) - 1 : 0;
    for (int i = lane; i < g; i += 32) {
      float v = __half2float(w[gi * g + i]);
      if (e4m3) {
        codes[gi * g + i] = static_cast<uint8_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v, sf), __NV_SATFINITE, __NV_E4M3));
        continue;
      }
      float q = h_round(__fdiv_rn(h_round(v - zp), sf));
      q = rintf(fminf(fmaxf(q, lower), upper));
      codes[gi * g + i] = static_cast<uint8_t>(static_cast<int>(q) + bias);
    }
  }
  __syncthreads();
  if (layout == 0 || layout == 3 || layout == 4) {
    uint8_t* o = packed + static_cast<int64_t>(n) * K;
    for (int i = threadIdx.x; i < K; i += 256) o[i] = codes[i];
  } else if (layout == 1) {
    uint8_t* o = packed + static_cast<int64_t>(n) * (K / 2);
    for (int i = threadIdx.x; i < K / 2; i += 256) {
      const int blk = i >> 4, j = i & 15;
      o[i] = static_cast<uint8_t>((codes[blk * 32 + j] & 0xF) | ((codes[blk * 32 + 16 + j] & 0xF) << 4));
    }
  } else {
    uint32_t* o = reinterpret_cast<uint32_t*>(packed + static_cast<int64_t>(n) * (K / 2));
    for (int i = threadIdx.x; i < K / 8; i += 256) {
      const uint8_t* c = codes + i * 8;
      uint32_t wv = 0;
      const int order[8] = {0, 2, 4, 6, 1, 3, 5, 7};
#pragma unroll
      for (int p = 0; p < 8; ++p) wv |= static_cast<uint32_t>(c[order[p]] & 0xF) << (4 * p);
      o[i] = wv;
    }
  }
}

// Reference WxAx layout (pack_wxax, quantize.cuh:424-475): uint16 words along K, first element in the most
// significant bits -> native layouts 0 / 1.
__global__ void __launch_bounds__(256) repack_wxax_kernel(const uint16_t* __restrict__ ref, int N, int K, int bits,
                                                           uint8_t* __restrict__ packed) {
  extern __shared__ uint8_t codes[];
  const int n = blockIdx.x;
  const int pack = 16 / bits;
  const uint16_t* r = ref + static_cast<int64_t>(n) * (K / pack);
  for (int i = threadIdx.x; i < K / pack; i += 256) {
    uint16_t wv = r[i];
    for (int x = 0; x < pack; ++x) codes[i * pack + x] = static_cast<uint8_t>((wv >> (bits * (pack - 1 - x))) & ((1 << bits) - 1));
  }
  __syncthreads();
  if (bits == 8) {
    uint8_t* o = packed + static_cast<int64_t>(n) * K;
    for (int i = threadIdx.x; i < K; i += 256) o[i] = codes[i];
  } else {
    uint8_t* o = packed + static_cast<int64_t>(n) * (K / 2);
    for (int i = threadIdx.x; i < K / 2; i += 256) {
      const int blk = i >> 4, j = i & 15;
      o[i] = static_cast<uint8_t>((codes[blk * 32 + j] & 0xF) | ((codes[blk * 32 + 16 + j] & 0xF) << 4));
    }
  }
}

// Reference WxAx ACTIVATION layout (the A half of pack_wxax, quantize.cuh:424-475: uint16 words along K, first element in
// the most significant bits, two's complement) -> int8 containers [M, K] (4-bit values sign-extended).
__global__ void __launch_bounds__(256) repack_act_kernel(const uint16_t* __restrict__ ref, int64_t n_words, int bits,
                                                          int8_t* __restrict__ out) {
  const int pack = 16 / bits;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n_words; i += static_cast<int64_t>(gridDim.x) * 256) {
    const uint16_t wv = ref[i];
    if (bits == 8) {
      *reinterpret_cast<uint16_t*>(out + i * 2) = static_cast<uint16_t>((wv >> 8) | (wv << 8));
    } else {
      uint32_t o = 0;
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const int v = (wv >> (4 * (3 - x))) & 0xF;
        o |= static_cast<uint32_t>(static_cast<uint8_t>(static_cast<int8_t>((v ^ 8) - 8))) << (8 * x);
      }
      *reinterpret_cast<uint32_t*>(out + i * pack) = o;
    }
  }
}

// Reference W-only layout (permute_weight + pack_weightonly, quantize.cuh:318-421) -> native layouts 2 / 3.
// `inv` maps (row in a PACK*8-row block, k in a 16-k block) -> flat index of the value inside the
// permuted block as it was before packing; built on the host from the reference's permutation tables.
__global__ void __launch_bounds__(256) repack_wonly_kernel(const uint16_t* __restrict__ ref, int N, int K, int bits,
                                                            const int32_t* __restrict__ inv, uint8_t* __restrict__ packed) {
  extern __shared__ uint8_t codes[];
  const int n = blockIdx.x;
  const int pack = 16 / bits;
  const int rows_blk = pack * 8;
  const int nb = n / rows_blk, rin = n % rows_blk;
  for (int kk = threadIdx.x; kk < K; kk += 256) {
    const int kb = kk >> 4, kin = kk & 15;
    const int src = inv[rin * 16 + kin];      // index inside the permuted [rows_blk][16] block
    const int prow = src >> 4, pk = src & 15;  // permuted row / k
    const int frag = prow >> 3, jj = prow & 7;
    const uint16_t wv = ref[(static_cast<int64_t>(nb) * 8 + jj) * K + kb * 16 + pk];
    codes[kk] = static_cast<uint8_t>((wv >> (bits * (pack - 1 - frag))) & ((1 << bits) - 1));
  }
  __syncthreads();
  if (bits == 8) {
    uint8_t* o = packed + static_cast<int64_t>(n) * K;
    for (int i = threadIdx.x; i < K; i += 256) o[i] = codes[i];
  } else {
    uint32_t* o = reinterpret_cast<uint32_t*>(packed + static_cast<int64_t>(n) * (K / 2));
    for (int i = threadIdx.x; i < K / 8; i += 256) {
      const uint8_t* c = codes + i * 8;
      uint32_t wv = 0;
      const int order[8] = {0, 2, 4, 6, 1, 3, 5, 7};
#pragma unroll
      for (int p = 0; p < 8; ++p) wv |= static_cast<uint32_t>(c[order[p]] & 0xF) << (4 * p);
      o[i] = wv;
    }
  }
}
}  // namespace

cudaError_t quant_pack_weight_launch(const void* W, int N, int K, int bits, int gsize, int sym, int layout, void* packed,
                                     void* scales, cudaStream_t st) {
  quant_pack_weight_kernel<<<N, 256, K, st>>>(static_cast<const __half*>(W), N, K, bits, gsize, sym, layout,
                                              static_cast<uint8_t*>(packed), static_cast<__half*>(scales));
  return cudaGetLastError();
}

cudaError_t repack_wxax_launch(const void* ref, int N, int K, int bits, void* packed, cudaStream_t st) {
  repack_wxax_kernel<<<N, 256, K, st>>>(static_cast<const uint16_t*>(ref), N, K, bits, static_cast<uint8_t*>(packed));
  return cudaGetLastError();
}

cudaError_t repack_act_launch(const void* ref, int64_t M, int K, int bits, void* out, cudaStream_t st) {
  const int64_t n_words = M * K * bits / 16;
  if (n_words == 0) return cudaSuccess;
  int64_t grid = (n_words + 255) / 256;
  if (grid > 148 * 16) grid = 148 * 16;
  repack_act_kernel<<<static_cast<unsigned>(grid), 256, 0, st>>>(static_cast<const uint16_t*>(ref), n_words, bits, static_cast<int8_t*>(out));
  return cudaGetLastError();
}

cudaError_t repack_wonly_launch(const void* ref, int N, int K, int bits, const int32_t* inv_dev, void* packed, cudaStream_t st) {
  repack_wonly_kernel<<<N, 256, K, st>>>(static_cast<const uint16_t*>(ref), N, K, bits, inv_dev, static_cast<uint8_t*>(packed));
  return cudaGetLastError();
}

}  // namespace mxb
