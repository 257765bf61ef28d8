/* mxmoe_b200 — C-ABI of the B200-native mixed-precision MoE GroupGEMM path.
 *
 * This header is the drop-in boundary for the hot path that the reference drives from
 * run_mxmoe_gg.py.  Every entry point is `extern "C"`, takes plain pointers / sizes / a
 * cudaStream_t passed as void*, returns 0 on success or a negative MXB_E* code, never throws and
 * never calls exit().  mxb_last_error() returns a thread-local message for the last failure.
 * All data pointers are DEVICE pointers unless a parameter is documented as host.
 *
 * Reference interfaces replaced (paths relative to the reference repository):
 *   mxb_group_gemm            mxmoe::FuncType, mxmoe/kernels/src/include/registry.cuh:28-39
 *                             (generated launcher: mxmoe/kernels/kernel_sketch.py:82-145)
 *   mxb_group_gemm_reflayout  the same FuncType argument list, verbatim (SoA of device arrays)
 *   mxb_permute               gg_permute_inp, mxmoe/kernels/src/ref_bind.cu:47-64
 *   mxb_quant_act             quant_inp_act / quant_act_kernel call site, ref_bind.cu:434-592
 *   mxb_silu_mul_quant        silu_mul_then_quant, ref_bind.cu:595-757 (+ online Hadamard,
 *                             mxmoe/quant/rotation.py:403-421, hadamard_utils.py:96-105)
 *   mxb_combine               gg_unpermute_out (empty in the reference, ref_bind.cu:66); routing
 *                             weights as mxmoe/quant/moe_tracer.py:79-86
 *   mxb_quant_pack_weight     quant_weight kernel + host packers, quantize.cuh:218-475
 *   mxb_repack_from_ref       reads the reference's packed layouts (quantize.cuh:318-475)
 *   mxb_select_tiles          get_possible_tile_list / get_best_tile_configs,
 *                             mxmoe/kernels/tile_config.py:330-610, mxmoe/quant/bits_solver.py:425-515
 */
#ifndef MXMOE_B200_H_
#define MXMOE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MXB_OK 0
#define MXB_EINVAL (-1)     /* bad argument / unsupported qconfig (the reference throws std::runtime_error) */
#define MXB_ECUDA (-2)      /* a CUDA runtime/driver call failed (the reference prints and exit()s) */
#define MXB_ENOMEM (-3)
#define MXB_EUNSUPPORTED (-4) /* device is not sm_100 */

typedef struct mxb_context* mxb_handle_t;

/* One GroupGEMM problem: C[M,N] = A[M,K] * B[N,K]^T in the precision named by (w_bits, a_bits, gsize, sym).
 * Mirrors one entry of the reference's (ptr_As[i], ptr_Bs[i], ptr_scale_a[i], ptr_scale_b[i], ptr_Cs[i],
 * problem_sizes[i], qbits_list[i]) SoA, registry.cuh:28-39 + quantize.cuh:14-25 (QParams).
 *
 * Operand formats ("native" layouts, produced by mxb_quant_pack_weight / mxb_repack_from_ref / mxb_quant_act):
 *   fp16      (w16 a16) : A fp16 [M,K] (row pitch lda elements), B fp16 [N,K]
 *   w8a8      sym g-1   : A int8 [M,K], B int8 [N,K], scale_a fp16 [M], scale_b fp16 [N]
 *   w8a8 E4M3 sym g-1   : a_bits = w_bits = MXB_BITS_E4M3: same shapes, bytes are E4M3 codes
 *   w4a4      sym g-1   : A int8 containers holding values in [-7,7] [M,K]; B 4-bit [N,K/2] bytes, inside
 *                         every 32-k block byte j holds k=j in the low and k=j+16 in the high nibble
 *                         (two's complement); scale_a fp16 [M], scale_b fp16 [N]
 *   w4a4      sym g128  : as above; scale_a fp16 [M, K/128], scale_b fp16 [K/128, N]
 *   wNa16     (N=4,8)   : A fp16 [M,K]; B unsigned codes (sym: q + 2^(N-1)-1), 4-bit: [N,K/2] bytes with
 *                         the nibble order k = 0,2,4,6,1,3,5,7 inside every 32-bit word; 8-bit: [N,K];
 *                         scale_b fp16 [K/g, N] (sym) or [K/g, N, 2] = (scale, zero-offset) (asym),
 *                         the group-major layout of permute_scale (quantize.cuh:299-315)
 * C is fp16 [M,N] with row pitch ldc elements.
 *
 * Device-resident M (no host sync after routing): when m_range is non-NULL it points at two consecutive
 * int32 on the DEVICE, {row_begin, row_end} (e.g. expert_offsets + e).  A, scale_a and C are then the bases
 * of buffers shared by all problems and this problem owns rows [row_begin,row_end); M is the row capacity
 * of those buffers.
 */
/* Bit-width code of 8-bit floating-point containers (E4M3: 4 exponent bits, 3 mantissa bits, max 448, no infinities):
 * the reference's `w8a8_g-1_sym_E4M3` (mxmoe/kernels/tile_config.py:45,105 MMA_E4M3_K32).  Wherever a_bits / w_bits
 * is taken, this value selects E4M3 instead of int8: per-token / per-channel scale = absmax / 448, codes rounded to
 * nearest even; the GEMM runs tcgen05.mma kind::f8f6f4 with fp32 accumulation and the int path's scale order
 * (acc * half(sa * sb)).  Both operands must be E4M3, scales per token and per channel (gsize -1). */
#define MXB_BITS_E4M3 (-8)

typedef struct MxbProblem {
  const void* A;
  const void* B;
  const void* scale_a;
  const void* scale_b;
  void* C;
  const int32_t* m_range;
  int64_t lda; /* elements of A's storage type per row; 0 = dense */
  int64_t ldc; /* fp16 elements per row of C; 0 = N */
  int32_t M, N, K;
  int32_t a_bits, w_bits, gsize, sym; /* gsize -1 = per-channel/per-token */
  int32_t tile_n;                     /* 0 = let the selector choose; else 64/128/256 */
  int32_t flags;                      /* MXB_EPI_* bits, 0 = plain GEMM */
  const int32_t* c_row;               /* optional, with m_range: DEVICE int32 = first row of this problem in C when it is not
                                         row_begin (C then is the base of a buffer laid out differently from A, e.g. the
                                         result buffer of ANOTHER GPU mapped over NVLink: mxmoe_b200/ep.py); NULL = row_begin */
  const int32_t* a_rows;              /* optional: DEVICE int32 per row of the problem's row range = the row of A to read (the
                                         token gather of gg_permute_inp, ref_bind.cu:47-64, folded into the operand load).  A is
                                         then the un-permuted [T, K] buffer.  Streaming (decode) path only. */
  int64_t lds_a;                      /* fp16 elements per row of scale_a; 0 = dense (1 for per-token scales, K/128 for gsize 128).
                                         A buffer shared by per-token and group-wise experts (the layer allocates [rows, K/128] as
                                         soon as one expert of the launch is g128) is read at this pitch by every problem. */
} MxbProblem;

/* Epilogue fusion of the step that follows the gate_up GEMM in the reference pipeline (silu_mul_then_quant,
 * ref_bind.cu:595-757; HF act_fn(gate) * up): B holds the gate and up rows INTERLEAVED (row 2i = gate_i,
 * row 2i+1 = up_i; scale_b interleaved the same way; mxb_interleave_gate_up converts the reference's
 * [gate; up] order), N still counts all 2I rows, and C is fp16 [M, N/2] = silu(gate) * up with default
 * ldc = N/2.  gate and up are first rounded to fp16 exactly as the unfused GEMM would store them, then combined
 * with the arithmetic of mxb_silu_mul_quant, so the result is bit-identical to the two-step path.
 * Served by the CTA-pair kernel (a list that is not eligible for it, see mxb_set_gemm_path, is MXB_EINVAL) and by the
 * streaming weight-only kernel (decode-sized w4a16 problems; group-wise scales are interleaved along N inside every group
 * row): there the 16-row weight slab holds 8 (gate, up) pairs and a lane gets its partner's value with one shuffle. */
#define MXB_EPI_SILU_MUL 1
/* Prologue fusion for the DOWN projection of a decode step: A holds the gate_up output [M, 2K] = [gate | up] (default
 * lda = 2K) and the kernel multiplies by act_fn(gate) * up, computed while the rows are staged, with the arithmetic of
 * mxb_silu_mul_quant in fp16 mode (bit-identical to running it first).  Served by the streaming weight-only kernel only
 * (w4a16, at most 64 rows, or device-resident rows with capacity <= 512); any other problem is MXB_EINVAL. */
#define MXB_PRO_SILU_MUL 2

/* Per-expert activation quantization parameters for mxb_quant_act / mxb_silu_mul_quant
 * (the reference's qparams_per_exp tuples, ref_bind.cu:321,365). */
typedef struct MxbActQ {
  int32_t a_bits; /* 16 = keep fp16, 8, 4 (stored in int8 containers) */
  int32_t gsize;  /* -1 or 128 */
} MxbActQ;

typedef struct MxbTileChoice {
  int32_t tile_m, tile_n, tile_k, stages;
  int32_t path;       /* 0 = single-CTA tcgen05 GroupGEMM, 1 = HBM-streaming w4a16, 2 = CTA-pair tcgen05 GroupGEMM
                         (tile_m = weight rows, tile_n = tokens per tile) */
  float est_us;       /* cost-model estimate for this problem alone */
} MxbTileChoice;

/* ---- lifetime ---- */
int mxb_create(int device, mxb_handle_t* out);
int mxb_destroy(mxb_handle_t h);
const char* mxb_last_error(void);
int mxb_sm_count(mxb_handle_t h);
const char* mxb_version(void);
/* sizeof of an ABI struct as this library was compiled (0 MxbProblem, 1 MxbActQ, 2 MxbRowDest, 3 MxbTileChoice; else -1):
 * lets a binding (ctypes, cgo, JNI) check its mirror of the layout before the first call. */
int mxb_struct_size(int which);
/* Which GroupGEMM kernel serves mxb_group_gemm*: 0 (default) = chosen per problem list — the CTA-pair kernel
 * (tcgen05 cta_group::2, 256 weight rows x <=256 tokens per tile) when every problem is fp16 / w8a8 / w4a4 with
 * g-1 scales and no explicit tile_n, else the single-CTA kernel; 1 = always the single-CTA kernel; 2 = the
 * CTA-pair kernel, MXB_EINVAL if the list holds a problem it does not cover.  The reference makes the
 * corresponding choice at code-generation time from tile_config.json (mxmoe/kernels/tile_config.py:330-610).
 * The environment variable MXB_GEMM_PATH sets the initial value. */
int mxb_set_gemm_path(mxb_handle_t h, int path);
/* Number of SMs the handle's persistent kernels (GroupGEMM, streaming w4a16, Hadamard) occupy: `sms` CTAs instead of
 * one per SM (rounded down to even; <= 0 or more than the device has restores the default).  A handle that runs
 * next to NCCL kernels or to another handle's launches on a second stream leaves the remaining SMs to them — a
 * persistent kernel that owns every SM would otherwise delay the collective until it exits (mxmoe_b200/ep.py runs
 * the replicated shared expert this way under the dispatch/combine all-to-all).  The reference has one stream and
 * no collectives (kernel_sketch.py:123). */
int mxb_set_sm_limit(mxb_handle_t h, int sms);

/* ---- piece (3)+(4): one launch over a mixed-precision problem list ---- */
/* `problems` is a HOST array.  Plans (tile tables, tensor maps) are cached in the handle keyed by its content. */
int mxb_group_gemm(mxb_handle_t h, const MxbProblem* problems, int problem_count, void* stream);
/* Same launch, but integer problems also dump their int32 accumulators (after removing the internal x16 of
 * the 4-bit path) to acc_out[i] (device int32 [M,N], may be NULL per problem).  For the bit-exact check. */
int mxb_group_gemm_debug_acc(mxb_handle_t h, const MxbProblem* problems, int problem_count,
                             int32_t* const* acc_out, void* stream);
/* Verbatim FuncType argument list (registry.cuh:28-39): device SoA arrays plus host copies of sizes/QParams.
 * sizes are 3x uint32 (dim3: M,N,K); qparams are {int a_bits; int w_bits; int gsize; bool sym} = 16 bytes
 * (quantize.cuh:14-25).  Problems with M = 0 are skipped like the reference's callers do (ref_bind.cu:360).  Operand
 * layouts: the native ones above, or — after mxb_set_ref_operands(h, MXB_REF_OPERANDS) — the REFERENCE's packed layouts
 * exactly as its callers hold them (A and B of WxAx problems from pack_wxax, B of weight-only problems from
 * permute_weight + pack_weightonly, scales in the permute_scale layout): the shim converts A on every call and B once
 * per weight pointer into handle-owned buffers, so a test.cu-style caller (test.cu:794-814) links unchanged.
 * stream == NULL is the default stream, as in the reference (kernel_sketch.py:123). */
int mxb_group_gemm_reflayout(mxb_handle_t h, void** ptr_As, void** ptr_Bs, void** ptr_scale_a, void** ptr_scale_b,
                             void** ptr_Cs, void** ptr_Ds, int64_t* ldas, int64_t* ldbs, int64_t* ldcs,
                             int64_t* ldds, const void* dev_problem_sizes, const void* host_problem_sizes,
                             const void* dev_qparams, const void* host_qparams, int problem_count, void* stream);

/* How mxb_group_gemm_reflayout reads its operands (bit mask, default 0):
 *   MXB_REF_OPERANDS      A / B are in the reference's packed layouts (see above); converted inside the call.
 *   MXB_REF_TABLES_FIXED  the five device pointer tables do not change between calls with the same table addresses and
 *                         problem count: they are fetched once instead of on every call (5 blocking D2H copies + a stream
 *                         synchronisation per call otherwise, which is what the reference's own launcher costs too). */
#define MXB_REF_OPERANDS 1
#define MXB_REF_TABLES_FIXED 2
int mxb_set_ref_operands(mxb_handle_t h, int flags);
/* The A half of pack_wxax (quantize.cuh:424-475; uint16 words along K, first element in the most significant bits)
 * -> int8 containers [M, K] as mxb_quant_act writes them.  a_bits 8 or 4. */
int mxb_repack_act_from_ref(mxb_handle_t h, const void* ref_packed, int64_t M, int K, int a_bits, void* out_q8, void* stream);

/* ---- piece (1): token -> expert permute ---- */
/* topk_ids int32 [T,k] with values in [0,E).  Outputs: expert_offsets int32 [E+1]; src_row int32 [T*k]
 * (source token of every sorted slot, stable in (token, j) order); slot_of int32 [T,k] (sorted slot of (t,j)). */
int mxb_permute(mxb_handle_t h, const int32_t* topk_ids, int T, int k, int E, int32_t* expert_offsets,
                int32_t* src_row, int32_t* slot_of, void* stream);
/* mxb_permute without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Needs no handle: `device` is made current, inputs and outputs are the caller's.  The block-histogram scratch that
 * mxb_permute keeps in the handle's cudaMalloc'ed workspace lives in a file-scope __device__ array of
 * MXB_PERMUTE_STATIC_BYTES (256 KiB, held by every GPU that loads the library until exit); its address is looked up with
 * cudaGetSymbolAddress on the first call per device (make that call — T = 0 is enough — before a capture starts).  The call
 * performs no cudaMalloc / cudaFree and calls no allocating library routine.
 * Limit: 4 * (1 + ceil(T*k / 2048) * E) <= MXB_PERMUTE_STATIC_BYTES (e.g. E = 64: T*k up to 2 095 104 rows), else
 * MXB_EINVAL with the numbers in mxb_last_error(); device < 64.
 * Every call on a device shares the one array, which keeps the previous call's contents (the call clears the ticket word
 * it needs at zero; the histograms are rewritten): calls that OVERLAP on the GPU — different streams without an ordering
 * between them — race.  Calls on one stream, or ordered by events, are safe.  Results are identical to mxb_permute's. */
#define MXB_PERMUTE_STATIC_BYTES (256 * 1024)
int mxb_permute_static(int device, const int32_t* topk_ids, int T, int k, int E, int32_t* expert_offsets,
                       int32_t* src_row, int32_t* slot_of, void* stream);

/* ---- piece (2): per-token activation quantization (+ optional Hadamard) ---- */
/* x fp16 [T,H] gathered through src_row (NULL = identity, rows = n_rows) into sorted order; expert e owns
 * rows [expert_offsets[e], expert_offsets[e+1]) and is quantized by actq[e] (HOST array, E entries).
 * out_q8 int8 [rows,H] receives 8/4-bit rows, out_f16 fp16 [rows,H] receives a_bits==16 rows (either may be
 * NULL if unused), scales fp16 [rows, H/128 if any gsize==128 else 1].  hadamard != 0 applies the online
 * Hadamard rotation of hadamard_utils.py:96-105 to every row before quantization. */
int mxb_quant_act(mxb_handle_t h, const void* x, int H, const int32_t* src_row, int n_rows,
                  const int32_t* expert_offsets, const MxbActQ* actq, int E, int hadamard, void* out_q8,
                  void* out_f16, void* scales, int scale_cols, void* stream);
/* gate_up fp16 [rows, 2I] (gate in the first I columns) -> h = silu(gate)*up -> optional Hadamard over I ->
 * per-token quantization as above into [rows, I] buffers. */
int mxb_silu_mul_quant(mxb_handle_t h, const void* gate_up, int I, int n_rows, const int32_t* expert_offsets,
                       const MxbActQ* actq, int E, int hadamard, void* out_q8, void* out_f16, void* scales,
                       int scale_cols, void* stream);

/* ---- expert-parallel exchange over peer memory (no counterpart in the reference, which is single-GPU; BASELINE north_star:
 * "sharding experts ... token dispatch and combine").  Rank r owns the experts [r*E/W, (r+1)*E/W).  Buffers reached through
 * the peer_* tables are device allocations of the OTHER ranks mapped into this process (CUDA IPC / symmetric memory);
 * entry [p] is rank p's buffer, entry [rank] the local one.  All three calls are stream-ordered and never synchronise. */
typedef struct MxbRowDest { /* where the rows of one expert go: written on the device by mxb_ep_plan */
  void* q8;      /* int8 containers [.., H] of the owning rank (or NULL) */
  void* f16;     /* fp16 rows [.., H] of the owning rank (or NULL) */
  void* scales;  /* fp16 [.., scale_cols] of the owning rank (or NULL) */
  int64_t row0;  /* first row of this (source rank, expert) segment in those buffers */
} MxbRowDest;
/* counts of this rank (expert_offsets[e+1]-expert_offsets[e]) -> row `rank` of every rank's [W,E] int32 count matrix.
 * peer_counts: HOST array of W device pointers. */
int mxb_ep_push_counts(mxb_handle_t h, const int32_t* expert_offsets, int E, void* const* peer_counts, int W, int rank, void* stream);
/* From the complete count matrix (after a cross-rank barrier): dest[E] = where each of my experts' rows go (segments on
 * the owner are ordered source-rank-major, expert-minor); seg_offs[W*EL+1] = prefix of the segments I receive;
 * back_row[W*EL] = first row of every received segment in its SOURCE rank's expert-sorted order (the c_row of its down
 * GEMM).  peer_q8 / peer_f16 / peer_scales: HOST arrays of W device pointers, any may be NULL. */
int mxb_ep_plan(mxb_handle_t h, const int32_t* counts_all, int W, int rank, int E, void* const* peer_q8, void* const* peer_f16,
                void* const* peer_scales, int32_t* seg_offs, int32_t* back_row, MxbRowDest* dest, void* stream);
/* mxb_quant_act whose output rows are scattered to their owners: row r of expert e -> dest[e] at row0 + (r - offsets[e]).
 * max_ctas > 0 caps the grid (the kernel is NVLink-bound; leave SMs to whatever runs beside it). */
int mxb_quant_act_push(mxb_handle_t h, const void* x, int H, const int32_t* src_row, int n_rows, const int32_t* expert_offsets,
                       const MxbActQ* actq, int E, const MxbRowDest* dest, int scale_cols, int max_ctas, void* stream);

/* ---- piece (5): gather + top-k weighted combine ---- */
/* out[t] = fp16( sum_j topk_w[t,j] * down_out[slot_of[t,j]] (+ shared_out[t]) ), fp32 accumulation in j order. */
int mxb_combine(mxb_handle_t h, const void* down_out, int H, const int32_t* slot_of, const float* topk_w, int T,
                int k, const void* shared_out, void* out, void* stream);
/* As mxb_combine with the shared expert's output scaled per token: out[t] = sum_j w[t,j] * down[slot(t,j)] +
 * shared_scale[t] * shared_out[t] (fp32 accumulation).  shared_scale is sigmoid(shared_expert_gate(x)) of
 * Qwen2MoeSparseMoeBlock; NULL = 1 (the behaviour of mxb_combine). */
int mxb_combine_gated(mxb_handle_t h, const void* down_out, int H, const int32_t* slot_of, const float* topk_w, int T, int k,
                      const void* shared_out, const float* shared_scale, void* out, void* stream);

/* ---- offline weight path ---- */
/* W fp16 [N,K] -> native packed weights + scales for (w_bits, a_bits, gsize, sym); numerics of
 * quant_weight (quantize.cuh:218-279).  packed/scales sizes: see mxb_packed_bytes / mxb_scale_elems. */
int mxb_quant_pack_weight(mxb_handle_t h, const void* W, int N, int K, int w_bits, int a_bits, int gsize, int sym,
                          void* packed, void* scales, void* stream);
/* Reference-layout packed weights (pack_wxax when a_bits<16, permute_weight+pack_weightonly when a_bits==16)
 * -> native layout.  Scales already share the reference's permute_scale layout and need no conversion. */
int mxb_repack_from_ref(mxb_handle_t h, const void* ref_packed, int N, int K, int w_bits, int a_bits, int sym,
                        void* packed, void* stream);
/* Row interleave for MXB_EPI_SILU_MUL: src holds 2I rows of row_bytes bytes ([gate rows; up rows], the reference's
 * fused gate_up order, gen_workload.py:84-93); dst row 2i = src row i, dst row 2i+1 = src row I+i.  Works on packed
 * weight rows of any bit width and, with row_bytes = 2 (4 for asym pairs), on per-channel scale vectors. */
int mxb_interleave_gate_up(mxb_handle_t h, const void* src, int I, int64_t row_bytes, void* dst, void* stream);
int64_t mxb_packed_bytes(int N, int K, int w_bits);
int64_t mxb_scale_elems(int N, int K, int w_bits, int a_bits, int gsize, int sym);

/* ---- Blackwell tile selector / cost model ---- */
int mxb_select_tiles(mxb_handle_t h, const MxbProblem* problems, int problem_count, MxbTileChoice* out);

#ifdef __cplusplus
}
#endif
#endif /* MXMOE_B200_H_ */
