// Device-side problem descriptor shared between the host planner (api.cu) and the persistent
// grouped-GEMM kernel (gg_kernel.cu).
//
// Replaces the reference's SoA of device arrays + QParams + host-computed tile prefix
// (mxmoe/kernels/src/include/registry.cuh:28-39, quantize.cuh:14-25, kernel_sketch.py:87-104):
// one AoS record per problem that also carries the two TMA tensor maps; the tile prefix sum is
// computed on the device by every CTA so that M may live in device memory (m_range).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace mxb {

enum MmaKind : int32_t { MMA_F16 = 0, MMA_I8 = 1, MMA_F8 = 2 };  // tcgen05.mma kind::f16 / kind::i8 / kind::f8f6f4 (E4M3 x E4M3)
enum ConvMode : int32_t {
  CONV_NONE = 0,
  CONV_W4_I8 = 1,   // 4-bit two's complement weights -> int8 x16 (nibble split), w4a4
  CONV_W4_F16 = 2,  // 4-bit unsigned codes -> fp16 (LOP3 magic) * scale + zp, w4a16
  CONV_W8_F16 = 3,  // 8-bit unsigned codes -> fp16, w8a16
};
enum EpiMode : int32_t {
  EPI_F32 = 0,        // fp32 accumulator -> fp16
  EPI_INT_ROWCOL = 1, // int32 accumulator * half(sa[m]*sb[n]) (mm_tile.cuh:472-496)
  EPI_INT_GROUP = 2,  // per-128-K-group: out += float(acc_g) * half(sa[m,g]*sb[g,n]) (cta_gemm.cuh:580-602)
};

struct alignas(128) DevProblem {
  CUtensorMap tmA;
  CUtensorMap tmB;
  const __half* scale_a;
  const __half* scale_b;
  __half* C;
  int32_t* acc_out;
  const int32_t* m_range;
  int64_t ldc;
  int32_t M, N, K;
  int32_t mma_kind, conv, epi;
  int32_t acc_shift;
  int32_t bn;       // tile N (UMMA N): 32/64/128/256 (<= 128 for the weight-only converters)
  int32_t n_tiles;  // ceil(N / bn)
  int32_t kblocks;  // number of 128-byte K slabs of the MMA operand
  int32_t a_kstep;  // tensor-map coordinate step per k-block (elements of the map's dtype)
  int32_t b_kstep;
  uint32_t tx_bytes;  // bytes TMA delivers per stage (A box + B box)
  uint32_t idesc;
  int32_t gsize, sym;
  int32_t scale_a_ld;  // fp16 elements per row of scale_a (1 or K/gsize)
  int32_t kb_per_group; // k-blocks per quantization group (EPI_INT_GROUP / grouped W-only)
  int32_t path;         // 0 = tcgen05 persistent GroupGEMM, 1 = HBM-streaming w4a16 kernel (w4a16_stream.cu)
  const void* A_ptr;    // raw operand pointers for the streaming path
  const void* B_ptr;
  int64_t lda;          // elements per row of A
  int32_t epi_flags;    // MXB_EPI_* (pair kernel): 1 = silu(gate)*up over interleaved weight rows, C is [M, N/2]
  const int32_t* c_row; // device int32: first row of this problem in C when it differs from m_range[0] (else null)
  const int32_t* a_rows; // streaming path: row of A behind every row of the problem's range (token gather), or null
  CUtensorMap tmA2;     // pair kernel only: tokens with a 128-row box (tmA has the 32-row box for ragged tiles)
};
static_assert(sizeof(DevProblem) == 640, "DevProblem layout");

constexpr int GG_MAXP = 256;       // problems per launch (host splits longer lists)
constexpr int GG_THREADS = 384;    // 12 warps: TMA, MMA, TMEM-alloc, spare, 4 epilogue, 4 converter
constexpr int GG_BM = 128;
constexpr int GG_STAGES = 3;
constexpr int GG_A_BYTES = 128 * 128;   // 128 rows x 128 B
constexpr int GG_B_BYTES = 256 * 128;   // up to 256 rows x 128 B
constexpr int GG_PACKED_BYTES = 256 * 64;  // raw 4/8-bit weight slab staged for the converter warps
constexpr int GG_STAGE_BYTES = GG_A_BYTES + GG_B_BYTES + GG_PACKED_BYTES;
constexpr int GG_PACKED_OFF = GG_B_BYTES;  // offset of the packed staging from the start of the B region
constexpr int GG_EPI_STAGE_BYTES = 4 * 32 * 64;
constexpr int GG_SMEM_BYTES = 1024 /*align slack*/ + GG_STAGES * GG_STAGE_BYTES + GG_EPI_STAGE_BYTES + 256 /*barriers*/ +
                              (3 * GG_MAXP + 4) * 4;

void gg_launch(const DevProblem* probs, int P, int grid, cudaStream_t stream);
void w4a16_stream_launch(const DevProblem* probs, int P, int est_units, int max_k, int sm_count, cudaStream_t st);
int w4a16_stream_rows_per_group();
int w4a16_stream_max_k();
int gg_configure();  // cudaFuncSetAttribute(max dynamic smem); returns cudaError_t as int

// CTA-pair kernel (gg_pair_kernel.cu): tcgen05 cta_group::2, weights as the UMMA M operand, tokens as UMMA N.
// DevProblem fields are read with the pair meaning: tmA = tokens with a 32-row box, tmB = weights with a 128-row
// box, n_tiles = ceil(N / 256), idesc carries M = 256 and no N (the token extent is patched in per tile).
void gg_pair_launch(const DevProblem* probs, int P, int grid, cudaStream_t stream);
int gg_pair_configure();
int gg_pair_smem_bytes();
int gg_pair_stages();

}  // namespace mxb
