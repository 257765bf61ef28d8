// Host side of the C-ABI (include/mxmoe_b200.h): context, plan cache, TMA tensor-map encoding and the
// launch of the persistent grouped-GEMM kernel and the MoE helper kernels.
//
// Replaces the reference's generated host launcher (mxmoe/kernels/kernel_sketch.py:82-145): no per-call
// cudaMalloc/cudaMemcpy/cudaFree of a tile prefix array (the prefix is computed on the device), no
// exceptions, no exit(); errors come back as codes + mxb_last_error().
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/mxmoe_b200.h"
#include "gg_kernel.cuh"
#include "moe_ops.cuh"

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CU_OK(expr)                                                                          \
  do {                                                                                       \
    cudaError_t e__ = (expr);                                                                \
    if (e__ != cudaSuccess) return fail(MXB_ECUDA, "%s: %s", #expr, cudaGetErrorString(e__)); \
  } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

uint64_t fnv1a(const void* p, size_t n, uint64_t h = 1469598103934665603ull) {
  const uint8_t* b = static_cast<const uint8_t*>(p);
  for (size_t i = 0; i < n; ++i) {
    h ^= b[i];
    h *= 1099511628211ull;
  }
  return h;
}

struct Plan {
  uint64_t key = 0;
  int P = 0;
  mxb::DevProblem* dev = nullptr;  // device array, capacity GG_MAXP per chunk list
  int stream_tasks = 0, tc_problems = 0, stream_classes = 0;
  bool paired = false;             // problems are encoded for gg_pair_kernel (CTA pairs, cta_group::2)
  size_t cap = 0;
  uint64_t last_use = 0;
};

struct DevTable {  // small constant tables mirrored on the device, keyed by content
  uint64_t key = 0;
  void* dev = nullptr;
  size_t bytes = 0;
  uint64_t last_use = 0;
};

}  // namespace

struct mxb_context {
  int device = 0;
  int sm_count = 0;  // CTAs the persistent kernels launch (mxb_set_sm_limit; default = every SM)
  int sm_phys = 0;
  int gemm_path = 0;  // 0 = choose per problem list, 1 = single-CTA kernel, 2 = CTA-pair kernel (error if not eligible)
  EncodeTiledFn encode = nullptr;
  std::vector<Plan> plans;
  std::vector<DevTable> tables;
  uint64_t tick = 0;
  int32_t* ws = nullptr;  // permute workspace
  size_t ws_bytes = 0;
  int32_t* inv_tab[3] = {nullptr, nullptr, nullptr};  // repack tables for 8/4/2-bit W-only reference layout
  std::vector<std::pair<int, void*>> hads;            // device fp16 copies of the K x K Hadamard factors
  cudaStream_t aux = nullptr;                         // fork/join side stream: tcgen05 kernel next to the streaming kernel
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // mxb_group_gemm_reflayout in reference-operand mode (mxb_set_ref_operands)
  int ref_flags = 0;
  struct RefWeight { const void* src; int N, K, w_bits, a_bits; void* native; };
  std::vector<RefWeight> ref_weights;                 // reference-packed B -> native copy, converted once per pointer
  void* ref_act = nullptr;                            // scratch for the converted A operands of one call
  size_t ref_act_bytes = 0;
  struct RefTables { const void* key[5]; int P; std::vector<void*> host[5]; };
  std::vector<RefTables> ref_tables;                  // fetched pointer tables (MXB_REF_TABLES_FIXED)
};

namespace {

constexpr int kMaxPlans = 512;  // a 28-layer stack keeps 4 lists per layer alive on one handle
constexpr int kMaxTables = 256;

// Encode a 2-D row-major tensor [rows, inner_elems] with a [box_rows, box_inner] box.
int encode_2d(mxb_context* c, CUtensorMap* tm, CUtensorMapDataType dt, int elsize, const void* base, uint64_t inner_elems,
              uint64_t rows, uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_rows, bool swizzle128) {
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return fail(MXB_EINVAL, "operand pointer %p is not 16-byte aligned", base);
  if (pitch_bytes % 16 != 0) return fail(MXB_EINVAL, "operand row pitch %llu B is not a multiple of 16", (unsigned long long)pitch_bytes);
  if (rows == 0) rows = 1;  // an empty problem never issues a load; keep the descriptor encodable
  cuuint64_t gdim[2] = {inner_elems, rows};
  cuuint64_t gstr[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  (void)elsize;
  CUresult r = c->encode(tm, dt, 2, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(MXB_ECUDA, "cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] box [%u x %u]", (int)r,
                                     (unsigned long long)rows, (unsigned long long)inner_elems, box_rows, box_inner);
  return MXB_OK;
}

struct KindInfo {
  int mma_kind, conv, epi, acc_shift;
  int a_elsize;        // bytes per A element in HBM (2 fp16, 1 int8)
  int k_per_block;     // K elements per 128-byte MMA slab
  int b_bytes_per_kb;  // bytes of B per row per k-block as stored in HBM
};

// (w_bits, a_bits, gsize, sym) -> pipeline, the run-time equivalent of CTA_KERNEL_MAP / BRANCH_CONDITION_MAP
// (mxmoe/kernels/compose_kernel.py:22-57).  Unsupported combinations are an error like compose_kernel.py:433.
int classify(const MxbProblem& p, KindInfo* k) {
  const int w = p.w_bits, a = p.a_bits, g = p.gsize;
  if (g != -1 && g != 128) return fail(MXB_EINVAL, "quant type not supported: gsize %d (only -1 or 128)", g);
  if ((a == MXB_BITS_E4M3) != (w == MXB_BITS_E4M3)) return fail(MXB_EINVAL, "quant type not supported: E4M3 needs both operands in E4M3");
  if (w == 16 && a == 16) {
    *k = {mxb::MMA_F16, mxb::CONV_NONE, mxb::EPI_F32, 0, 2, 64, 128};
    return MXB_OK;
  }
  if (a == 16 && (w == 4 || w == 8)) {
    *k = {mxb::MMA_F16, w == 4 ? mxb::CONV_W4_F16 : mxb::CONV_W8_F16, mxb::EPI_F32, 0, 2, 64, w == 4 ? 32 : 64};
    return MXB_OK;
  }
  if (a == 8 && w == 8 && p.sym) {
    *k = {mxb::MMA_I8, mxb::CONV_NONE, g == -1 ? mxb::EPI_INT_ROWCOL : mxb::EPI_INT_GROUP, 0, 1, 128, 128};
    return MXB_OK;
  }
  if (a == MXB_BITS_E4M3 && w == MXB_BITS_E4M3 && p.sym && g == -1) {  // w8a8_g-1_sym_E4M3 (tile_config.py:45,105): kind::f8f6f4
    *k = {mxb::MMA_F8, mxb::CONV_NONE, mxb::EPI_INT_ROWCOL, 0, 1, 128, 128};
    return MXB_OK;
  }
  if (a == 4 && w == 4 && p.sym) {
    *k = {mxb::MMA_I8, mxb::CONV_W4_I8, g == -1 ? mxb::EPI_INT_ROWCOL : mxb::EPI_INT_GROUP, 4, 1, 128, 64};
    return MXB_OK;
  }
  return fail(MXB_EINVAL, "quant type not supported: w%da%d_g%d_%s", w, a, g, p.sym ? "sym" : "asym");
}

int pick_tile_n(const MxbProblem& p, const KindInfo& k) {
  int cap = 256;
  if (k.conv >= mxb::CONV_W4_F16) cap = 128;  // weight-only converters own 2 rows per thread
  if (k.epi == mxb::EPI_INT_GROUP) cap = 64;
  int bn = p.tile_n > 0 ? p.tile_n : cap;
  if (bn > cap) bn = cap;
  int n32 = (p.N + 31) / 32 * 32;
  if (bn > n32) bn = n32;
  if (bn >= 256) bn = 256;
  else if (bn >= 128) bn = 128;
  else if (bn >= 64) bn = 64;
  else bn = 32;
  return bn;
}

// Problems the CTA-pair kernel covers: per-token / per-channel scaled integer kinds and plain fp16, on the
// tensor-core path, with no explicit tile_n request (tile_n addresses the single-CTA kernel's UMMA N).
bool pair_eligible(const MxbProblem& p) {
  KindInfo k = {};
  if (classify(p, &k) != MXB_OK) return false;
  if (p.tile_n != 0) return false;
  if (k.epi == mxb::EPI_INT_GROUP) return false;
  if ((p.flags & MXB_EPI_SILU_MUL) && p.N % 2 != 0) return false;
  return k.conv == mxb::CONV_NONE || k.conv == mxb::CONV_W4_I8;
}

int build_dev_problem(mxb_context* c, const MxbProblem& p, int32_t* acc_out, bool paired, mxb::DevProblem* d) {
  memset(d, 0, sizeof *d);
  KindInfo k = {};
  int rc = classify(p, &k);
  if (rc) return rc;
  if (p.M < 0 || p.N <= 0 || p.K <= 0) return fail(MXB_EINVAL, "bad problem size %d x %d x %d", p.M, p.N, p.K);
  if (p.N % 8 != 0) return fail(MXB_EINVAL, "N=%d must be a multiple of 8", p.N);
  if (p.K % 32 != 0) return fail(MXB_EINVAL, "K=%d must be a multiple of 32", p.K);
  if (p.gsize == 128 && p.K % 128 != 0) return fail(MXB_EINVAL, "K=%d must be a multiple of the group size 128", p.K);
  if (p.M == 0 && !p.m_range) {  // empty problem (the reference's callers skip these, ref_bind.cu:360): zero tiles
    d->n_tiles = 1;
    d->bn = 32;
    return MXB_OK;
  }
  if (!p.A || !p.B || !p.C) return fail(MXB_EINVAL, "null operand pointer");
  if (k.epi != mxb::EPI_F32 && (!p.scale_a || !p.scale_b)) return fail(MXB_EINVAL, "quantized problem without scales");
  if (k.conv >= mxb::CONV_W4_F16 && !p.scale_b) return fail(MXB_EINVAL, "weight-only problem without scale_b");
  const int bn = paired ? 256 : pick_tile_n(p, k);
  const int64_t lda = p.lda > 0 ? p.lda : ((p.flags & MXB_PRO_SILU_MUL) ? 2 * int64_t(p.K) : p.K);
  d->scale_a = static_cast<const __half*>(p.scale_a);
  d->scale_b = static_cast<const __half*>(p.scale_b);
  d->C = static_cast<__half*>(p.C);
  d->acc_out = acc_out;
  d->m_range = p.m_range;
  d->c_row = p.c_row;
  d->a_rows = p.a_rows;
  if (p.c_row && !p.m_range) return fail(MXB_EINVAL, "c_row needs m_range (device-resident rows)");
  if (p.flags & ~(MXB_EPI_SILU_MUL | MXB_PRO_SILU_MUL)) return fail(MXB_EINVAL, "unknown problem flags 0x%x", p.flags);
  d->epi_flags = p.flags;
  { static const int epi_vec = [] { const char* e = getenv("MXB_EPI_VEC"); return e ? atoi(e) : 0; }(); if (epi_vec) d->epi_flags |= 256; }  // measurement switch
  d->ldc = p.ldc > 0 ? p.ldc : ((p.flags & MXB_EPI_SILU_MUL) ? p.N / 2 : p.N);
  d->M = p.M;
  d->N = p.N;
  d->K = p.K;
  d->mma_kind = k.mma_kind;
  d->conv = k.conv;
  d->epi = k.epi;
  d->acc_shift = k.acc_shift;
  d->bn = bn;
  d->n_tiles = (p.N + bn - 1) / bn;
  d->kblocks = (p.K + k.k_per_block - 1) / k.k_per_block;
  d->a_kstep = k.k_per_block;  // A map dtype has one element per K element
  d->b_kstep = k.conv == mxb::CONV_NONE ? k.k_per_block : k.b_bytes_per_kb;
  d->tx_bytes = mxb::GG_A_BYTES + bn * k.b_bytes_per_kb;
  d->gsize = p.gsize;
  d->sym = p.sym;
  d->scale_a_ld = (k.epi == mxb::EPI_INT_GROUP) ? p.K / 128 : 1;
  if (p.lds_a > 0) {  // scale rows at the caller's pitch (a buffer shared by per-token and group-wise experts)
    if (p.lds_a < d->scale_a_ld || p.lds_a > (1 << 20)) return fail(MXB_EINVAL, "lds_a=%lld below the %d scales a row holds", (long long)p.lds_a, d->scale_a_ld);
    d->scale_a_ld = (int32_t)p.lds_a;
  }
  d->A_ptr = p.A;
  d->B_ptr = p.B;
  d->lda = lda;
  // Path choice (replaces the reference's per-qconfig tile lists for the memory-bound experts,
  // tile_config.py:330-610): 4-bit weight-only problems with at most 16 rows stream through w4a16_stream.cu.
  // With device-resident M the row capacity of the shared buffer decides (decode batches).
  // Up to 64 rows (four 16-token passes over L2-resident slabs) the streaming kernel still beats a 128-row UMMA tile
  // fed by the converter warps, and a decode step's shared expert then rides in the same launch.
  const int m_cap = p.m_range ? (p.M <= 512 ? 16 : p.M) : p.M;
  d->path = (!paired && !p.c_row && k.conv == mxb::CONV_W4_F16 && m_cap <= 64 && p.K % 128 == 0 && p.N % 16 == 0 && p.K <= mxb::w4a16_stream_max_k() &&
             p.tile_n == 0) ? 1 : 0;
  if ((p.flags & MXB_EPI_SILU_MUL) && !paired && d->path != 1)
    return fail(MXB_EINVAL, "MXB_EPI_SILU_MUL is served by the CTA-pair kernel and the streaming w4a16 kernel only (list not eligible, or gemm path 1)");
  if ((p.a_rows || (p.flags & MXB_PRO_SILU_MUL)) && d->path != 1)
    return fail(MXB_EINVAL, "a_rows / MXB_PRO_SILU_MUL are served by the streaming w4a16 kernel only (decode-sized weight-only problems)");
  if (k.epi == mxb::EPI_INT_GROUP) d->kb_per_group = 128 / k.k_per_block;            // int8 slabs hold 128 k
  else if (k.conv >= mxb::CONV_W4_F16) d->kb_per_group = p.gsize == 128 ? 2 : d->kblocks;  // fp16 slabs hold 64 k
  else d->kb_per_group = d->kblocks;
  if (d->kb_per_group < 1) d->kb_per_group = 1;
  const uint32_t cfmt = k.mma_kind == mxb::MMA_I8 ? 2u : 1u;
  const uint32_t abfmt = k.mma_kind == mxb::MMA_I8 ? 1u : 0u;
  // pair kernel: UMMA M = 256 (weights), the token extent (UMMA N) is patched in per tile on the device
  d->idesc = (cfmt << 4) | (abfmt << 7) | (abfmt << 10) |
             (paired ? ((256u >> 4) << 24) : (((uint32_t(bn) >> 3) << 17) | ((128u >> 4) << 24)));

  // A (tokens): [rows, K] of fp16 or int8, 128-byte swizzled box of 128 rows (pair kernel: 32-row boxes, so a
  // token tile of any multiple of 16 rows moves only the rows it needs)
  const CUtensorMapDataType adt = k.a_elsize == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  rc = encode_2d(c, &d->tmA, adt, k.a_elsize, p.A, p.K, p.M, lda * k.a_elsize, k.k_per_block, paired ? 32 : 128, true);
  if (rc) return rc;
  if (paired) {
    rc = encode_2d(c, &d->tmA2, adt, k.a_elsize, p.A, p.K, p.M, lda * k.a_elsize, k.k_per_block, 128, true);
    if (rc) return rc;
  }
  const int wbox = paired ? 128 : bn;  // weight rows per TMA box (pair kernel: each CTA stages 128 of the tile's 256)
  if (k.conv == mxb::CONV_NONE) {
    rc = encode_2d(c, &d->tmB, adt, k.a_elsize, p.B, p.K, p.N, uint64_t(p.K) * k.a_elsize, k.k_per_block, wbox, true);
  } else if (d->path == 1) {
    // streaming kernel (w4a16_stream.cu): 16-row slabs, 128 bytes (256 k) per row and stage, 128-byte swizzle
    const uint64_t row_bytes = uint64_t(p.K) / 2;
    rc = encode_2d(c, &d->tmB, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p.B, row_bytes, p.N, row_bytes, 128, 16, true);
  } else {
    const uint64_t row_bytes = uint64_t(p.K) * p.w_bits / 8;
    rc = encode_2d(c, &d->tmB, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p.B, row_bytes, p.N, row_bytes, k.b_bytes_per_kb, wbox, false);
  }
  return rc;
}

}  // namespace

// ============================================================================ exported C-ABI
namespace {

int device_table(mxb_context* c, const void* host, size_t bytes, cudaStream_t st, void** out) {
  const uint64_t key = fnv1a(host, bytes) ^ (uint64_t(bytes) << 48);
  ++c->tick;
  for (auto& t : c->tables)
    if (t.key == key && t.bytes == bytes) {
      t.last_use = c->tick;
      *out = t.dev;
      return MXB_OK;
    }
  DevTable* slot = nullptr;
  if ((int)c->tables.size() < kMaxTables) {
    c->tables.emplace_back();
    slot = &c->tables.back();
  } else {
    slot = &c->tables[0];
    for (auto& t : c->tables)
      if (t.last_use < slot->last_use) slot = &t;
    // the evicted table may still be read by work queued earlier on another stream
    CU_OK(cudaDeviceSynchronize());
    cudaFree(slot->dev);
    slot->dev = nullptr;
  }
  CU_OK(cudaMalloc(&slot->dev, bytes < 256 ? 256 : bytes));
  CU_OK(cudaMemcpyAsync(slot->dev, host, bytes, cudaMemcpyHostToDevice, st));
  slot->key = key;
  slot->bytes = bytes;
  slot->last_use = c->tick;
  *out = slot->dev;
  return MXB_OK;
}

int get_plan(mxb_context* c, const MxbProblem* problems, int P, int32_t* const* acc_out, cudaStream_t st, Plan** out) {
  uint64_t key = fnv1a(problems, sizeof(MxbProblem) * size_t(P));
  if (acc_out) key = fnv1a(acc_out, sizeof(int32_t*) * size_t(P), key ^ 0x9e3779b97f4a7c15ull);
  key ^= uint64_t(c->gemm_path) * 0x632be59bd9b4e019ull;
  ++c->tick;
  for (auto& pl : c->plans)
    if (pl.key == key && pl.P == P) {
      pl.last_use = c->tick;
      *out = &pl;
      return MXB_OK;
    }
  // Kernel choice for the whole list (one launch): the CTA-pair kernel when every problem is one it covers.
  bool paired = c->gemm_path != 1;
  for (int i = 0; i < P && paired; ++i)
    if (!(problems[i].M == 0 && !problems[i].m_range) && !pair_eligible(problems[i])) paired = false;
  if (c->gemm_path == 2 && !paired) return fail(MXB_EINVAL, "gemm path 2 (CTA pairs) requested but the list holds a problem it does not cover");
  std::vector<mxb::DevProblem> host(static_cast<size_t>(P));
  for (int i = 0; i < P; ++i) {
    int rc = build_dev_problem(c, problems[i], acc_out ? acc_out[i] : nullptr, paired, &host[i]);
    if (rc) {
      std::string m = g_err;
      return fail(rc, "problem %d: %s", i, m.c_str());
    }
  }
  int tc_problems = 0, stream_classes = 0, stream_units = 0;
  const int rows_per_group = mxb::w4a16_stream_rows_per_group();
  for (int i = 0; i < P; ++i) {
    const bool st_path = host[i].path != 0;
    if (st_path && host[i].K > stream_classes) stream_classes = host[i].K;  // largest K routed to the streaming kernel
    // grid estimate only: the kernel builds the exact unit prefix from the device-resident row counts
    if (st_path) stream_units += (host[i].N + rows_per_group - 1) / rows_per_group * (problems[i].m_range ? 1 : (problems[i].M + 15) / 16);
    if (!st_path && host[i].n_tiles > 0 && !(problems[i].M == 0 && !problems[i].m_range)) ++tc_problems;
  }
  Plan* slot = nullptr;
  if ((int)c->plans.size() < kMaxPlans) {
    c->plans.emplace_back();
    slot = &c->plans.back();
  } else {
    slot = &c->plans[0];
    for (auto& pl : c->plans)
      if (pl.last_use < slot->last_use) slot = &pl;
    CU_OK(cudaDeviceSynchronize());  // the evicted plan may still be in use by queued launches
  }
  if (slot->cap < size_t(P)) {
    if (slot->dev) cudaFree(slot->dev);
    slot->dev = nullptr;
    slot->cap = 0;
    CU_OK(cudaMalloc(reinterpret_cast<void**>(&slot->dev), sizeof(mxb::DevProblem) * size_t(P)));
    slot->cap = size_t(P);
  }
  slot->stream_tasks = stream_units;
  slot->tc_problems = tc_problems;
  slot->stream_classes = stream_classes;
  slot->paired = paired;
  // pageable source: the runtime stages the bytes before returning, so `host` may die right after
  CU_OK(cudaMemcpyAsync(slot->dev, host.data(), sizeof(mxb::DevProblem) * size_t(P), cudaMemcpyHostToDevice, st));
  slot->key = key;
  slot->P = P;
  slot->last_use = c->tick;
  *out = slot;
  return MXB_OK;
}

int group_gemm_impl(mxb_context* c, const MxbProblem* problems, int P, int32_t* const* acc_out, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (P < 0 || (P > 0 && !problems)) return fail(MXB_EINVAL, "bad problem list");
  if (P == 0) return MXB_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  Plan* pl = nullptr;
  int rc = get_plan(c, problems, P, acc_out, st, &pl);
  if (rc) return rc;
  // A list that holds both decode-shaped weight-only problems (streaming kernel) and tensor-core problems (e.g. the
  // shared expert of a decode step) runs the two kernels side by side: the persistent kernel's CTAs without tiles
  // exit at once and the streaming CTAs take those SMs.  Fork/join with events (CUDA-graph capturable).
  const bool both = pl->stream_tasks > 0 && pl->tc_problems > 0 && c->aux != nullptr;
  cudaStream_t tc_st = st;
  if (both) {
    CU_OK(cudaEventRecord(c->ev_fork, st));
    CU_OK(cudaStreamWaitEvent(c->aux, c->ev_fork, 0));
    tc_st = c->aux;
  }
  if (pl->tc_problems > 0 && both) {
    for (int off = 0; off < P; off += mxb::GG_MAXP) {
      const int n = P - off < mxb::GG_MAXP ? P - off : mxb::GG_MAXP;
      if (pl->paired) mxb::gg_pair_launch(pl->dev + off, n, c->sm_count, tc_st);
      else mxb::gg_launch(pl->dev + off, n, c->sm_count, tc_st);
    }
    CU_OK(cudaEventRecord(c->ev_join, tc_st));
  }
  if (pl->stream_tasks > 0) mxb::w4a16_stream_launch(pl->dev, P, pl->stream_tasks, pl->stream_classes, c->sm_count, st);
  if (both) CU_OK(cudaStreamWaitEvent(st, c->ev_join, 0));
  if (pl->tc_problems > 0 && !both) {
    for (int off = 0; off < P; off += mxb::GG_MAXP) {
      const int n = P - off < mxb::GG_MAXP ? P - off : mxb::GG_MAXP;
      if (pl->paired) mxb::gg_pair_launch(pl->dev + off, n, c->sm_count, st);
      else mxb::gg_launch(pl->dev + off, n, c->sm_count, st);
    }
  }
  CU_OK(cudaGetLastError());
  return MXB_OK;
}

}  // namespace

extern "C" {

const char* mxb_last_error(void) { return g_err.c_str(); }
const char* mxb_version(void) { return "mxmoe_b200 0.1 (sm_100a)"; }

int mxb_create(int device, mxb_handle_t* out) {
  if (!out) return fail(MXB_EINVAL, "null out");
  *out = nullptr;
  CU_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU_OK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(MXB_EUNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  mxb_context* c = new (std::nothrow) mxb_context();
  if (!c) return fail(MXB_ENOMEM, "out of host memory");
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->sm_phys = c->sm_count;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
    delete c;
    return fail(MXB_ECUDA, "cuTensorMapEncodeTiled entry point not available");
  }
  c->encode = reinterpret_cast<EncodeTiledFn>(fn);
  e = static_cast<cudaError_t>(mxb::gg_configure());
  if (e != cudaSuccess) {
    delete c;
    return fail(MXB_ECUDA, "cudaFuncSetAttribute(max dynamic smem %d): %s", mxb::GG_SMEM_BYTES, cudaGetErrorString(e));
  }
  e = static_cast<cudaError_t>(mxb::gg_pair_configure());
  if (e != cudaSuccess) {
    delete c;
    return fail(MXB_ECUDA, "cudaFuncSetAttribute(max dynamic smem %d): %s", mxb::gg_pair_smem_bytes(), cudaGetErrorString(e));
  }
  if (const char* env = getenv("MXB_GEMM_PATH")) c->gemm_path = atoi(env);
  if (cudaStreamCreateWithFlags(&c->aux, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    c->aux = nullptr;  // no side stream: the two kernels of a mixed list simply run back to back
    cudaGetLastError();
  }
  c->ws_bytes = 8u << 20;
  e = cudaMalloc(reinterpret_cast<void**>(&c->ws), c->ws_bytes);
  if (e != cudaSuccess) {
    delete c;
    return fail(MXB_ENOMEM, "workspace allocation failed: %s", cudaGetErrorString(e));
  }
  cudaMemset(c->ws, 0, c->ws_bytes);
  c->plans.reserve(kMaxPlans);
  c->tables.reserve(kMaxTables);
  *out = c;
  return MXB_OK;
}

int mxb_destroy(mxb_handle_t c) {
  if (!c) return MXB_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (auto& pl : c->plans) {
    if (pl.dev) cudaFree(pl.dev);
  }
  for (auto& t : c->tables)
    if (t.dev) cudaFree(t.dev);
  for (auto* p : c->inv_tab)
    if (p) cudaFree(p);
  for (auto& h : c->hads) cudaFree(h.second);
  for (auto& w : c->ref_weights) cudaFree(w.native);
  if (c->ref_act) cudaFree(c->ref_act);
  if (c->ws) cudaFree(c->ws);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->aux) cudaStreamDestroy(c->aux);
  delete c;
  return MXB_OK;
}

int mxb_sm_count(mxb_handle_t c) { return c ? c->sm_count : 0; }

int mxb_struct_size(int which) {
  switch (which) {
    case 0: return (int)sizeof(MxbProblem);
    case 1: return (int)sizeof(MxbActQ);
    case 2: return (int)sizeof(MxbRowDest);
    case 3: return (int)sizeof(MxbTileChoice);
    default: return -1;
  }
}

int mxb_set_sm_limit(mxb_handle_t c, int sms) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (sms <= 0 || sms > c->sm_phys) sms = c->sm_phys;
  sms &= ~1;  // the CTA-pair kernel launches clusters of two
  if (sms < 2) sms = 2;
  c->sm_count = sms;
  return MXB_OK;
}

int mxb_set_gemm_path(mxb_handle_t c, int path) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (path < 0 || path > 2) return fail(MXB_EINVAL, "gemm path %d (0 = auto, 1 = single-CTA kernel, 2 = CTA-pair kernel)", path);
  c->gemm_path = path;
  return MXB_OK;
}

int mxb_group_gemm(mxb_handle_t c, const MxbProblem* problems, int problem_count, void* stream) {
  return group_gemm_impl(c, problems, problem_count, nullptr, stream);
}

int mxb_group_gemm_debug_acc(mxb_handle_t c, const MxbProblem* problems, int problem_count, int32_t* const* acc_out,
                             void* stream) {
  return group_gemm_impl(c, problems, problem_count, acc_out, stream);
}

int mxb_set_ref_operands(mxb_handle_t c, int flags) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (flags & ~(MXB_REF_OPERANDS | MXB_REF_TABLES_FIXED)) return fail(MXB_EINVAL, "unknown flags 0x%x", flags);
  c->ref_flags = flags;
  return MXB_OK;
}

int mxb_repack_act_from_ref(mxb_handle_t c, const void* ref_packed, int64_t M, int K, int a_bits, void* out_q8, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (!(a_bits == 8 || a_bits == 4)) return fail(MXB_EINVAL, "a_bits %d: pack_wxax holds 8- or 4-bit values", a_bits);
  if (M < 0 || K <= 0 || K % (16 / a_bits) != 0) return fail(MXB_EINVAL, "bad shape %lld x %d", (long long)M, K);
  if (M > 0 && (!ref_packed || !out_q8)) return fail(MXB_EINVAL, "null pointer");
  CU_OK(cudaSetDevice(c->device));
  CU_OK(mxb::repack_act_launch(ref_packed, M, K, a_bits, out_q8, static_cast<cudaStream_t>(stream)));
  return MXB_OK;
}

int mxb_group_gemm_reflayout(mxb_handle_t c, void** ptr_As, void** ptr_Bs, void** ptr_scale_a, void** ptr_scale_b,
                             void** ptr_Cs, void** ptr_Ds, int64_t* ldas, int64_t* ldbs, int64_t* ldcs, int64_t* ldds,
                             const void* dev_problem_sizes, const void* host_problem_sizes, const void* dev_qparams,
                             const void* host_qparams, int P, void* stream) {
  (void)ptr_Ds; (void)ldas; (void)ldbs; (void)ldcs; (void)ldds; (void)dev_problem_sizes; (void)dev_qparams;  // unused by the reference too
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (P <= 0) return P == 0 ? MXB_OK : fail(MXB_EINVAL, "negative problem count");
  if (!host_problem_sizes || !host_qparams) return fail(MXB_EINVAL, "host copies of sizes/qparams are required");
  if (!ptr_As || !ptr_Bs || !ptr_Cs) return fail(MXB_EINVAL, "null pointer table");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  // The pointer tables live on the device in the reference's contract.  Default: fetch them on every call (synchronous,
  // like the reference launcher's own cudaMemcpy); MXB_REF_TABLES_FIXED: once per (table addresses, P).
  const void* key[5] = {ptr_As, ptr_Bs, ptr_Cs, ptr_scale_a, ptr_scale_b};
  mxb_context::RefTables* tab = nullptr;
  if (c->ref_flags & MXB_REF_TABLES_FIXED)
    for (auto& t : c->ref_tables)
      if (t.P == P && memcmp(t.key, key, sizeof key) == 0) tab = &t;
  mxb_context::RefTables local;
  if (!tab) {
    mxb_context::RefTables* dst = &local;
    if (c->ref_flags & MXB_REF_TABLES_FIXED) {
      if (c->ref_tables.size() >= 64) c->ref_tables.clear();
      c->ref_tables.emplace_back();
      dst = &c->ref_tables.back();
    }
    memcpy(dst->key, key, sizeof key);
    dst->P = P;
    for (int i = 0; i < 5; ++i) {
      dst->host[i].assign(size_t(P), nullptr);
      if (key[i]) CU_OK(cudaMemcpyAsync(dst->host[i].data(), key[i], sizeof(void*) * P, cudaMemcpyDeviceToHost, st));
    }
    CU_OK(cudaStreamSynchronize(st));
    tab = dst;
  }
  const std::vector<void*>&hA = tab->host[0], &hB = tab->host[1], &hC = tab->host[2], &hSa = tab->host[3], &hSb = tab->host[4];
  struct Dim3 { uint32_t x, y, z; };
  struct QP { int32_t a_bits, w_bits, gsize; uint8_t sym; uint8_t pad[3]; };
  const Dim3* sz = static_cast<const Dim3*>(host_problem_sizes);
  const QP* qp = static_cast<const QP*>(host_qparams);
  const bool ref_ops = (c->ref_flags & MXB_REF_OPERANDS) != 0;
  size_t act_need = 0;
  if (ref_ops)
    for (int i = 0; i < P; ++i)
      if (sz[i].x && qp[i].a_bits < 16) act_need += (size_t(sz[i].x) * sz[i].z + 255) & ~size_t(255);
  if (act_need > c->ref_act_bytes) {
    CU_OK(cudaStreamSynchronize(st));  // earlier launches may still read the old scratch
    if (c->ref_act) cudaFree(c->ref_act);
    c->ref_act = nullptr;
    c->ref_act_bytes = 0;
    CU_OK(cudaMalloc(&c->ref_act, act_need));
    c->ref_act_bytes = act_need;
  }
  size_t act_off = 0;
  std::vector<MxbProblem> pr;
  pr.reserve(P);
  for (int i = 0; i < P; ++i) {
    if (sz[i].x == 0) continue;  // the reference's callers skip empty problems (ref_bind.cu:360)
    MxbProblem p;
    memset(&p, 0, sizeof p);
    p.A = hA[i]; p.B = hB[i]; p.C = hC[i];
    p.scale_a = ptr_scale_a ? hSa[i] : nullptr;
    p.scale_b = ptr_scale_b ? hSb[i] : nullptr;
    p.M = (int)sz[i].x; p.N = (int)sz[i].y; p.K = (int)sz[i].z;
    p.a_bits = qp[i].a_bits; p.w_bits = qp[i].w_bits; p.gsize = qp[i].gsize; p.sym = qp[i].sym ? 1 : 0;
    if (p.a_bits == 16 && p.w_bits == 16) { p.gsize = -1; p.sym = 1; }
    if (ref_ops && !(p.a_bits == 16 && p.w_bits == 16)) {
      if (p.a_bits < 16) {  // A: the caller's packed halves -> int8 containers, every call (activations change)
        void* dstA = static_cast<char*>(c->ref_act) + act_off;
        act_off += (size_t(p.M) * p.K + 255) & ~size_t(255);
        int rc = mxb_repack_act_from_ref(c, p.A, p.M, p.K, p.a_bits, dstA, stream);
        if (rc) return rc;
        p.A = dstA;
      }
      void* nat = nullptr;  // B: converted once per weight pointer
      for (auto& w : c->ref_weights)
        if (w.src == p.B && w.N == p.N && w.K == p.K && w.w_bits == p.w_bits && w.a_bits == p.a_bits) nat = w.native;
      if (!nat) {
        CU_OK(cudaMalloc(&nat, size_t(mxb_packed_bytes(p.N, p.K, p.w_bits))));
        int rc = mxb_repack_from_ref(c, p.B, p.N, p.K, p.w_bits, p.a_bits, p.sym, nat, stream);
        if (rc) {
          cudaFree(nat);
          return rc;
        }
        c->ref_weights.push_back({p.B, p.N, p.K, p.w_bits, p.a_bits, nat});
      }
      p.B = nat;
    }
    pr.push_back(p);
  }
  return group_gemm_impl(c, pr.data(), (int)pr.size(), nullptr, stream);
}

}  // extern "C"

// ============================================================================ MoE helper ops + offline weight path
namespace {

int actq_table(mxb_context* c, const MxbActQ* actq, int E, cudaStream_t st, const int32_t** dev, int* any_g128, int* any_q,
               int* any_f16) {
  *any_g128 = *any_q = *any_f16 = 0;
  for (int e = 0; e < E; ++e) {
    const int a = actq[e].a_bits, g = actq[e].gsize;
    if (!(a == 16 || a == 8 || a == 4 || a == MXB_BITS_E4M3))
      return fail(MXB_EINVAL, "expert %d: a_bits %d not supported (16, 8, 4, MXB_BITS_E4M3)", e, a);
    if (a != 16 && g != -1 && g != 128) return fail(MXB_EINVAL, "expert %d: activation gsize %d not supported (-1, 128)", e, g);
    if (a == MXB_BITS_E4M3 && g != -1) return fail(MXB_EINVAL, "expert %d: E4M3 activations take per-token scales only", e);
    if (a == 16) *any_f16 = 1; else *any_q = 1;
    if (a != 16 && g == 128) *any_g128 = 1;
  }
  void* d = nullptr;
  int rc = device_table(c, actq, sizeof(MxbActQ) * size_t(E), st, &d);
  *dev = static_cast<const int32_t*>(d);
  return rc;
}

#include "hadk_data.inc"

// fp16 copy of the reference's K x K factor, zero padded to a multiple of 16, cached per K on the device
// A > 1: the Kronecker product hadK (x) H_A of order K * A (index (k, a) -> k * A + a, H_A the Sylvester matrix): a row of
// K blocks of m is then K * A blocks of m / A, and the top log2(A) butterfly stages of the length-m transform happen inside
// the tensor-core factor instead of on the CUDA cores (H_m = H_A (x) H_{m/A}).  Kp: padded order the kernels are built for.
int hadamard_matrix(mxb_context* c, int K, int A, const void** dev, int* Kp_out) {
  const int KA = K * A;
  int Kp = (KA + 15) / 16 * 16;
  if (A > 1 && Kp > 64) Kp = 176;  // the kernels exist for 16, 32, 48, 64 and 176
  *Kp_out = Kp;
  const int key = K + 1000 * A;
  for (auto& h : c->hads)
    if (h.first == key) {
      *dev = h.second;
      return MXB_OK;
    }
  const unsigned char* bits = nullptr;
  for (int i = 0; i < kHadTableSize; ++i)
    if (kHadTable[i].k == K) bits = kHadTable[i].bits;
  if (!bits) return fail(MXB_EINVAL, "no Hadamard matrix of order %d", K);
  std::vector<__half> h(size_t(Kp) * Kp, __float2half(0.f));
  for (int i = 0; i < KA; ++i)
    for (int j = 0; j < KA; ++j) {
      const int b = (i / A) * K + (j / A);
      const int bit = (bits[b >> 3] >> (7 - (b & 7))) & 1;
      const int sa = __builtin_popcount(unsigned((i % A) & (j % A))) & 1;  // H_A[a', a] = (-1)^popcount(a' & a)
      h[size_t(i) * Kp + j] = __float2half((bit != 0) != (sa != 0) ? 1.f : -1.f);
    }
  void* d = nullptr;
  CU_OK(cudaMalloc(&d, h.size() * sizeof(__half)));
  CU_OK(cudaMemcpy(d, h.data(), h.size() * sizeof(__half), cudaMemcpyHostToDevice));
  c->hads.emplace_back(key, d);
  *dev = d;
  return MXB_OK;
}

int quant_common(mxb_context* c, bool silu, const void* x, int H, const int32_t* src_row, int n_rows,
                 const int32_t* expert_offsets, const MxbActQ* actq, int E, int hadamard, void* out_q8, void* out_f16,
                 void* scales, int scale_cols, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (!x || !actq || E < 1 || n_rows < 0) return fail(MXB_EINVAL, "bad argument");
  if (H % 8 != 0) return fail(MXB_EINVAL, "row length %d must be a multiple of 8", H);
  if (E > 1 && !expert_offsets) return fail(MXB_EINVAL, "expert_offsets required when E > 1");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  const int32_t* tab = nullptr;
  int g128, anyq, anyf;
  int rc = actq_table(c, actq, E, st, &tab, &g128, &anyq, &anyf);
  if (rc) return rc;
  if (anyq && (!out_q8 || !scales)) return fail(MXB_EINVAL, "quantized experts need out_q8 and scales");
  if (anyf && !out_f16) return fail(MXB_EINVAL, "fp16 experts need out_f16");
  if (g128 && (H % 128 != 0 || scale_cols < H / 128)) return fail(MXB_EINVAL, "gsize 128 needs H %% 128 == 0 and scale_cols >= H/128");
  if (scale_cols < 1) scale_cols = 1;
  if (hadamard) {
    const int K = mxb::hadamard_factor(H);
    if (K == 0) return fail(MXB_EINVAL, "no Hadamard factorisation for row length %d (get_hadK would assert)", H);
    if (K > 1 && H / K < 8) return fail(MXB_EINVAL, "Hadamard row length %d = %d x %d: power-of-two part below 8 not supported", H, K, H / K);
    // group-wise and E4M3 quantization tails exist in the warp-per-row kernel only (quantize_row)
    bool generic = g128 != 0;
    for (int e = 0; e < E; ++e) generic = generic || actq[e].a_bits == MXB_BITS_E4M3;
    // The butterfly stages over m run on the CUDA cores and are bound by the shuffle unit.  MXB_HAD_KRON=<cap> (opt-in) moves
    // the top log2(A) of them into the tensor-core factor, hadK (x) H_A with K * A <= 176 and m / A >= 8 (2560 = 40 x 64 ->
    // 160 x 16: 381 -> 231 us for 65536 rows).  It is the same transform but NOT the reference's rounding point: the
    // reference rounds to fp16 after the whole length-m transform (fast_hadamard_transform returns fp16, then hadK @ in
    // fp16, hadamard_utils.py:96-105), here the rounding would come log2(A) stages earlier, which moves about half of the
    // outputs by one fp16 ulp.  Off by default: parity with the reference's staging comes first.
    int A = 1;
    const char* kron_env = getenv("MXB_HAD_KRON");
    if (K > 1 && !generic && kron_env && atoi(kron_env) > 1) A = mxb::hadamard_split(H, K, atoi(kron_env));
    const void* hmat = nullptr;
    int Kp = 16;
    if (K > 1) {
      rc = hadamard_matrix(c, K, A, &hmat, &Kp);
      if (rc) return rc;
    }
    cudaError_t e = mxb::hadamard_quant_launch(silu, x, H, src_row, n_rows, expert_offsets, tab, E, out_q8, out_f16, scales,
                                               scale_cols, hmat, K * A, Kp, c->sm_count, generic, st);
    if (e != cudaSuccess) return fail(MXB_ECUDA, "hadamard_quant_launch: %s", cudaGetErrorString(e));
    return MXB_OK;
  }
  const int64_t pitch = silu ? 2 * int64_t(H) : int64_t(H);
  CU_OK(mxb::quant_rows_launch(silu, x, H, pitch, src_row, n_rows, expert_offsets, tab, E, out_q8, out_f16, scales, scale_cols, nullptr,
                               c->sm_count * 32, st));
  return MXB_OK;
}

// (row in a PACK*8-row block, k in a 16-k block) -> position of that value inside the reference's permuted
// block; restates the index arithmetic of permute_weight (quantize.cuh:318-381, PermuteMode::Row).
void build_inv_table(int bits, std::vector<int32_t>* inv) {
  const int pack = 16 / bits;
  const int n = 4 * pack, plen = n / 2;
  static const int proj8[4] = {1, 0, 3, 2};
  static const int des8[8] = {0, 2, 4, 6, 1, 3, 5, 7};
  static const int proj4[8] = {3, 7, 2, 6, 1, 5, 0, 4};
  static const int des4[16] = {0, 4, 8, 12, 1, 5, 9, 13, 2, 6, 10, 14, 3, 7, 11, 15};
  static const int proj2[16] = {7, 15, 6, 14, 5, 13, 4, 12, 3, 11, 2, 10, 1, 9, 0, 8};
  static const int des2[32] = {0, 8, 16, 24, 1, 9, 17, 25, 2, 10, 18, 26, 3, 11, 19, 27,
                               4, 12, 20, 28, 5, 13, 21, 29, 6, 14, 22, 30, 7, 15, 23, 31};
  const int* proj = bits == 8 ? proj8 : bits == 4 ? proj4 : proj2;
  const int* des = bits == 8 ? des8 : bits == 4 ? des4 : des2;
  std::vector<int> perm(n);
  for (int i = 0; i < n; i += plen)
    for (int j = 0; j < plen; ++j) perm[proj[j] + i] = des[i + j];
  inv->assign(size_t(pack) * 8 * 16, 0);
  auto coord = [&](int x, int tj, int ti, int* row, int* k) {
    const int frag = x % pack, rest = x / pack;  // x = (ii*2 + trav_ii)*pack + frag
    const int tii = rest % 2, ii = rest / 2;
    *row = frag * 8 + tj;
    *k = ii * 8 + ti * 2 + tii;
  };
  for (int tj = 0; tj < 8; ++tj)
    for (int ti = 0; ti < 4; ++ti)
      for (int x = 0; x < n; ++x) {
        int prow, pk, r, kin;
        coord(x, tj, ti, &prow, &pk);        // where the value lands
        coord(perm[x], tj, ti, &r, &kin);    // where it came from
        (*inv)[size_t(r) * 16 + kin] = prow * 16 + pk;
      }
}

int native_layout(int w_bits, int a_bits) {
  if (w_bits == MXB_BITS_E4M3 || a_bits == MXB_BITS_E4M3) return (w_bits == a_bits) ? 4 : -1;
  if (a_bits < 16) return w_bits == 8 ? 0 : (w_bits == 4 ? 1 : -1);
  return w_bits == 4 ? 2 : (w_bits == 8 ? 3 : -1);
}

}  // namespace

extern "C" {

// ---- expert-parallel exchange over peer memory
int mxb_ep_push_counts(mxb_handle_t c, const int32_t* expert_offsets, int E, void* const* peer_counts, int W, int rank, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (!expert_offsets || !peer_counts || E < 1 || W < 1 || rank < 0 || rank >= W) return fail(MXB_EINVAL, "bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  void* dev = nullptr;
  int rc = device_table(c, peer_counts, sizeof(void*) * size_t(W), st, &dev);
  if (rc) return rc;
  CU_OK(mxb::ep_push_counts_launch(expert_offsets, E, static_cast<int32_t* const*>(dev), W, rank, st));
  return MXB_OK;
}

int mxb_ep_plan(mxb_handle_t c, const int32_t* counts_all, int W, int rank, int E, void* const* peer_q8, void* const* peer_f16,
                void* const* peer_scales, int32_t* seg_offs, int32_t* back_row, MxbRowDest* dest, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (!counts_all || !seg_offs || !back_row || !dest || E < 1 || W < 1 || rank < 0 || rank >= W) return fail(MXB_EINVAL, "bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  void* tabs[3] = {nullptr, nullptr, nullptr};
  void* const* src[3] = {peer_q8, peer_f16, peer_scales};
  for (int i = 0; i < 3; ++i)
    if (src[i]) {
      int rc = device_table(c, src[i], sizeof(void*) * size_t(W), st, &tabs[i]);
      if (rc) return rc;
    }
  static_assert(sizeof(MxbRowDest) == sizeof(mxb::RowDest), "MxbRowDest layout");
  CU_OK(mxb::ep_plan_launch(counts_all, W, rank, E, static_cast<void* const*>(tabs[0]), static_cast<void* const*>(tabs[1]),
                            static_cast<void* const*>(tabs[2]), seg_offs, back_row, reinterpret_cast<mxb::RowDest*>(dest), st));
  return MXB_OK;
}

int mxb_quant_act_push(mxb_handle_t c, const void* x, int H, const int32_t* src_row, int n_rows, const int32_t* expert_offsets,
                       const MxbActQ* actq, int E, const MxbRowDest* dest, int scale_cols, int max_ctas, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (!x || !actq || !dest || !expert_offsets || E < 1 || n_rows < 0) return fail(MXB_EINVAL, "bad argument");
  if (H % 8 != 0) return fail(MXB_EINVAL, "row length %d must be a multiple of 8", H);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  const int32_t* tab = nullptr;
  int g128, anyq, anyf;
  int rc = actq_table(c, actq, E, st, &tab, &g128, &anyq, &anyf);
  if (rc) return rc;
  if (g128 && (H % 128 != 0 || scale_cols < H / 128)) return fail(MXB_EINVAL, "gsize 128 needs H %% 128 == 0 and scale_cols >= H/128");
  if (scale_cols < 1) scale_cols = 1;
  // E == 1 still goes through the destination table: the kernel treats dest != null like a routed list
  CU_OK(mxb::quant_rows_launch(false, x, H, H, src_row, n_rows, expert_offsets, tab, E, nullptr, nullptr, nullptr, scale_cols,
                               reinterpret_cast<const mxb::RowDest*>(dest), max_ctas, st));
  return MXB_OK;
}

int mxb_permute(mxb_handle_t c, const int32_t* topk_ids, int T, int k, int E, int32_t* expert_offsets, int32_t* src_row,
                int32_t* slot_of, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (T < 0 || k < 1 || E < 1 || E > 1024) return fail(MXB_EINVAL, "bad permute shape T=%d k=%d E=%d (E <= 1024)", T, k, E);
  if (!expert_offsets || (T > 0 && (!topk_ids || !src_row || !slot_of))) return fail(MXB_EINVAL, "null pointer");
  if (mxb::permute_ws_bytes(T * k, E) > c->ws_bytes) return fail(MXB_ENOMEM, "permute workspace too small for T*k=%d, E=%d", T * k, E);
  CU_OK(cudaSetDevice(c->device));
  CU_OK(mxb::permute_launch(topk_ids, T, k, E, expert_offsets, src_row, slot_of, c->ws, static_cast<cudaStream_t>(stream)));
  return MXB_OK;
}

// No allocation, no handle: the scratch is moe_ops.cu's file-scope device array (see include/mxmoe_b200.h).
int mxb_permute_static(int device, const int32_t* topk_ids, int T, int k, int E, int32_t* expert_offsets, int32_t* src_row,
                       int32_t* slot_of, void* stream) {
  if (device < 0 || device >= 64) return fail(MXB_EINVAL, "mxb_permute_static: device %d outside [0, 64)", device);
  if (T < 0 || k < 1 || E < 1 || E > 1024) return fail(MXB_EINVAL, "bad permute shape T=%d k=%d E=%d (E <= 1024)", T, k, E);
  if (!expert_offsets || (T > 0 && (!topk_ids || !src_row || !slot_of))) return fail(MXB_EINVAL, "null pointer");
  if (int64_t(T) * k > (int64_t(1) << 30)) return fail(MXB_EINVAL, "mxb_permute_static: T*k = %lld too large", (long long)(int64_t(T) * k));
  static_assert(MXB_PERMUTE_STATIC_BYTES == 256 * 1024, "header and moe_ops.cu disagree");
  if (mxb::permute_static_capacity() != size_t(MXB_PERMUTE_STATIC_BYTES)) return fail(MXB_EINVAL, "static scratch size mismatch");
  const size_t need = mxb::permute_ws_bytes(T * k, E);
  if (need > mxb::permute_static_capacity())
    return fail(MXB_EINVAL, "mxb_permute_static: T*k=%d, E=%d needs %zu bytes of scratch, the static array holds %zu (use mxb_permute)",
                T * k, E, need, mxb::permute_static_capacity());
  CU_OK(cudaSetDevice(device));
  CU_OK(mxb::permute_static_launch(device, topk_ids, T, k, E, expert_offsets, src_row, slot_of, static_cast<cudaStream_t>(stream)));
  return MXB_OK;
}

int mxb_quant_act(mxb_handle_t c, const void* x, int H, const int32_t* src_row, int n_rows, const int32_t* expert_offsets,
                  const MxbActQ* actq, int E, int hadamard, void* out_q8, void* out_f16, void* scales, int scale_cols,
                  void* stream) {
  return quant_common(c, false, x, H, src_row, n_rows, expert_offsets, actq, E, hadamard, out_q8, out_f16, scales, scale_cols, stream);
}

int mxb_silu_mul_quant(mxb_handle_t c, const void* gate_up, int I, int n_rows, const int32_t* expert_offsets,
                       const MxbActQ* actq, int E, int hadamard, void* out_q8, void* out_f16, void* scales, int scale_cols,
                       void* stream) {
  return quant_common(c, true, gate_up, I, nullptr, n_rows, expert_offsets, actq, E, hadamard, out_q8, out_f16, scales, scale_cols, stream);
}

int mxb_combine(mxb_handle_t c, const void* down_out, int H, const int32_t* slot_of, const float* topk_w, int T, int k,
                const void* shared_out, void* out, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (T < 0 || k < 1 || H % 8 != 0) return fail(MXB_EINVAL, "bad combine shape T=%d k=%d H=%d", T, k, H);
  if (T > 0 && (!down_out || !slot_of || !topk_w || !out)) return fail(MXB_EINVAL, "null pointer");
  CU_OK(cudaSetDevice(c->device));
  CU_OK(mxb::combine_launch(down_out, H, slot_of, topk_w, T, k, shared_out, nullptr, out, static_cast<cudaStream_t>(stream)));
  return MXB_OK;
}

int mxb_combine_gated(mxb_handle_t c, const void* down_out, int H, const int32_t* slot_of, const float* topk_w, int T, int k,
                      const void* shared_out, const float* shared_scale, void* out, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (T < 0 || k < 1 || H % 8 != 0) return fail(MXB_EINVAL, "bad combine shape T=%d k=%d H=%d", T, k, H);
  if (T > 0 && (!down_out || !slot_of || !topk_w || !out)) return fail(MXB_EINVAL, "null pointer");
  if (shared_scale && !shared_out) return fail(MXB_EINVAL, "shared_scale without shared_out");
  CU_OK(cudaSetDevice(c->device));
  CU_OK(mxb::combine_launch(down_out, H, slot_of, topk_w, T, k, shared_out, shared_scale, out, static_cast<cudaStream_t>(stream)));
  return MXB_OK;
}

int mxb_interleave_gate_up(mxb_handle_t c, const void* src, int I, int64_t row_bytes, void* dst, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  if (!src || !dst || I < 1 || row_bytes < 1 || src == dst) return fail(MXB_EINVAL, "bad argument");
  CU_OK(cudaSetDevice(c->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const char* s8 = static_cast<const char*>(src);
  char* d8 = static_cast<char*>(dst);
  // two strided 2-D copies: gate rows to the even rows, up rows to the odd rows (offline weight path)
  CU_OK(cudaMemcpy2DAsync(d8, 2 * row_bytes, s8, row_bytes, row_bytes, I, cudaMemcpyDeviceToDevice, st));
  CU_OK(cudaMemcpy2DAsync(d8 + row_bytes, 2 * row_bytes, s8 + int64_t(I) * row_bytes, row_bytes, row_bytes, I,
                          cudaMemcpyDeviceToDevice, st));
  return MXB_OK;
}

int64_t mxb_packed_bytes(int N, int K, int w_bits) { return int64_t(N) * K * (w_bits < 0 ? -w_bits : w_bits) / 8; }

int64_t mxb_scale_elems(int N, int K, int w_bits, int a_bits, int gsize, int sym) {
  (void)w_bits; (void)a_bits;
  const int64_t G = gsize <= 0 ? 1 : K / gsize;
  return G * N * (sym ? 1 : 2);
}

int mxb_quant_pack_weight(mxb_handle_t c, const void* W, int N, int K, int w_bits, int a_bits, int gsize, int sym, void* packed,
                          void* scales, void* stream) {
  if (!c) return fail(MXB_EINVAL, "null handle");
  const int layout = native_layout(w_bits, a_bits);
  if (layout < 0) return fail(MXB_EINVAL, "quant type not supported: w%da%d", w_bits, a_bits);
  if (a_bits < 16 && !sym) return fail(MXB_EINVAL, "quant type not supported: asymmetric WxAx");
  if (layout == 4 && gsize > 0) return fail(MXB_EINVAL, "quant type not supported: E4M3 weights take per-channel scales only");
  if (K % 32 != 0 || (gsize > 0 && K % gsize != 0) || K > 96 * 1024) return fail(MXB_EINVAL, "bad K=%d for gsize %d", K, gsize);
  if (!W || !packed || !scales) return fail(MXB_EINVAL, "null pointer");
  CU_OK(cudaSetDevice(c->device));
  CU_OK(mxb::quant_pack_weight_launch(W, N, K, w_bits < 0 ? 8 : w_bits, gsize, sym, layout, packed, scales, static_cast<cudaStream_t>(stream)));
  return MXB_OK;
}

int mxb_repack_from_ref(mxb_handle_t c, const void* ref_packed, int N, int K, int w_bits, int a_bits, int sym, void* packed,
                        void* stream) {
  (void)sym;  // the bias of symmetric W-only codes is the same in both layouts
  if (!c) return fail(MXB_EINVAL, "null handle");
  const int layout = native_layout(w_bits, a_bits);
  if (layout < 0) return fail(MXB_EINVAL, "quant type not supported: w%da%d", w_bits, a_bits);
  if (!ref_packed || !packed) return fail(MXB_EINVAL, "null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU_OK(cudaSetDevice(c->device));
  if (a_bits < 16) {
    if (K % 32 != 0) return fail(MXB_EINVAL, "K=%d must be a multiple of 32", K);
    CU_OK(mxb::repack_wxax_launch(ref_packed, N, K, w_bits, packed, st));
    return MXB_OK;
  }
  const int pack = 16 / w_bits;
  if (N % (pack * 8) != 0 || K % 16 != 0) return fail(MXB_EINVAL, "reference W-only layout needs N %% %d == 0 and K %% 16 == 0", pack * 8);
  const int slot = w_bits == 8 ? 0 : 1;
  if (!c->inv_tab[slot]) {
    std::vector<int32_t> inv;
    build_inv_table(w_bits, &inv);
    CU_OK(cudaMalloc(reinterpret_cast<void**>(&c->inv_tab[slot]), inv.size() * 4));
    CU_OK(cudaMemcpy(c->inv_tab[slot], inv.data(), inv.size() * 4, cudaMemcpyHostToDevice));
  }
  CU_OK(mxb::repack_wonly_launch(ref_packed, N, K, w_bits, c->inv_tab[slot], packed, st));
  return MXB_OK;
}

}  // extern "C"

// ============================================================================ Blackwell tile selector / cost model
// Replaces the hand-curated sm80/sm89 tile lists (mxmoe/kernels/tile_config.py:330-610) and the hard-coded
// "best tiles" of bits_solver.py:425-515.  On sm_100 the MMA atom is fixed by the hardware (UMMA 128 x N x 32 B),
// so the choice reduces to tile N, the path (tensor-core GroupGEMM vs HBM-streaming weight-only) and a roofline
// estimate that the bit-width solver can use as T[e][n][s] (bits_solver.py:518-542).
extern "C" int mxb_select_tiles(mxb_handle_t c, const MxbProblem* problems, int P, MxbTileChoice* out) {
  if (P < 0 || (P > 0 && (!problems || !out))) return fail(MXB_EINVAL, "bad argument");
  const double sms = c ? c->sm_count : 148;
  const double clk = 1.9e9;     // SM clock of a kernel timed alone on this pool (MEASURED_PEAKS.json: 1965 MHz max)
  const double hbm = 6.5e12;    // measured copy bandwidth, B/s
  const double l2 = 10.0e12;    // measured L2 -> SM operand rate of the pair kernel (ncu l1tex__m_xbar2l1tex_read_bytes)
  const int gemm_path = c ? c->gemm_path : 0;
  for (int i = 0; i < P; ++i) {
    const MxbProblem& p = problems[i];
    KindInfo k = {};
    int rc = classify(p, &k);
    if (rc) return rc;
    MxbTileChoice t;
    const bool stream = p.a_bits == 16 && p.w_bits == 4 && p.M <= 64 && p.K % 128 == 0 && p.N % 16 == 0 &&
                        p.K <= mxb::w4a16_stream_max_k() && p.tile_n == 0;
    const bool pair = gemm_path != 1 && pair_eligible(p);
    const double w_bytes = double(p.N) * p.K * p.w_bits / 8.0;
    const double a_el = p.a_bits == 16 ? 2.0 : 1.0;
    double t_tensor, t_l2;
    if (pair) {
      // CTA pair: 256 weight rows x nt tokens per tile, nt = M / ceil(M/256) rounded up to 16 (gg_pair_kernel.cu PairWalk)
      const int tch = (p.M + 255) / 256 > 0 ? (p.M + 255) / 256 : 1;
      const int nt = (((p.M + tch - 1) / tch) + 15) & ~15;
      t.tile_m = 256;
      t.tile_n = nt;
      t.tile_k = k.k_per_block;
      t.stages = mxb::gg_pair_stages();
      t.path = 2;
      const double tiles = double(tch) * double((p.N + 255) / 256);
      const double kb = double((p.K + k.k_per_block - 1) / k.k_per_block);
      // UMMA 256 x nt x 32 B on two SMs: nt/2 cycles per instruction, 4 instructions per 128-byte slab
      t_tensor = tiles * kb * 4.0 * (nt / 2.0) / ((sms / 2.0) * clk);
      // per slab and pair: 256 weight rows (128 B, or 64 B packed) + nt token rows of 128 B cross the L2 -> SM fabric
      t_l2 = tiles * kb * (256.0 * (p.w_bits == 4 && p.a_bits == 4 ? 64.0 : 128.0) + nt * 128.0) / l2;
    } else {
      t.tile_m = 128;
      t.tile_n = pick_tile_n(p, k);
      t.tile_k = k.k_per_block;
      t.stages = mxb::GG_STAGES;
      t.path = stream ? 1 : 0;
      const double tiles = double((p.M + 127) / 128) * double((p.N + t.tile_n - 1) / t.tile_n);
      const double kb = double((p.K + k.k_per_block - 1) / k.k_per_block);
      double cyc = tiles * kb * 4.0 * (128.0 * t.tile_n / 256.0);
      t_tensor = cyc / (sms * clk);
      // Kinds paced by something other than the tensor pipe, at the rates measured on this pool
      // (profiles/cost_table_b200_r01.json, M = 8192): group-wise int scales drain TMEM once per 128-K group (274 TOPS);
      // weight-only kinds wait for the converter warps (4-bit 420 TOPS, 8-bit alike)
      const double ops = 2.0 * double((p.M + 127) / 128 * 128) * p.N * p.K;
      if (k.epi == mxb::EPI_INT_GROUP) t_tensor = ops / 274e12;
      else if (k.conv >= mxb::CONV_W4_F16) t_tensor = ops / 420e12;
      t_l2 = tiles * kb * (128.0 * 128.0 + t.tile_n * double(k.b_bytes_per_kb)) / l2;
      if (stream) t_tensor = t_l2 = 0.0;
    }
    const double bytes = w_bytes + double(p.M) * p.K * a_el + double(p.M) * p.N * 2.0;
    double t_mem = bytes / hbm;
    // the streaming kernel is instruction-bound at ~2.6 TB/s per 16-token pass (profiles/ncu_w4a16_stream_r01.txt)
    if (stream) t_mem = w_bytes * double((p.M + 15) / 16) / 2.6e12;
    double tt = t_tensor > t_mem ? t_tensor : t_mem;
    if (t_l2 > tt) tt = t_l2;
    // fixed cost of a problem inside a launch (tile-table entry, pipeline fill and drain of its first tile): ~1.3 us
    // measured on 8-expert lists of tiny problems
    t.est_us = float(1e6 * tt + 1.3);
    out[i] = t;
  }
  return MXB_OK;
}
