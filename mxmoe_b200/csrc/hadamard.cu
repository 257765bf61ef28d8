// Online Hadamard rotation fused into the activation-quantization prologue (piece (2) of the hot path).
//
// Semantics (reference): matmul_hadU_cuda, mxmoe/quant/hadamard_utils.py:96-105, used by the down_proj
// pre-hook at mxmoe/quant/rotation.py:403-421: view the row as (K, m) with n = K*m, m a power of two and K
// the first factor get_hadK accepts (hadamard_utils.py:5-59); Sylvester FWHT over m scaled by 1/sqrt(n),
// result rounded to fp16; then hadK[K,K] @ over the K axis in fp16 with fp32 accumulation, rounded to fp16.
// The reference needs a separate FHT launch (3rdparty/fast-hadamard-transform) plus a torch matmul; here the
// rotation, silu*mul and the per-token quantization are one kernel and the rotated row never touches HBM.
//
// Mapping: persistent CTAs, ONE WARP PER ROW (only warp-level syncs).  The row is staged in shared memory
// as fp32 for the butterflies (strides < 8 are done in registers), converted in place to fp16, multiplied by
// the +-1 factor with mma.sync.m16n8k16 (A = hadK resident in shared memory for the CTA's lifetime, B = the
// row read with ldmatrix.trans), and quantized straight from shared memory.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "moe_ops.cuh"
#include "ptx.cuh"
#include "quant_common.cuh"

namespace mxb {
namespace {

__device__ __forceinline__ uint32_t s_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2_trans(uint32_t addr, uint32_t (&r)[2]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__device__ __forceinline__ uint4 ptx_ld_nc(const __half* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

struct HadParams {
  const __half* x;
  int n;              // row length after silu*mul (I) or H
  int64_t in_pitch;
  const int32_t* src_row;
  int n_rows;
  const int32_t* expert_offsets;
  const int2* actq;
  int E;
  int8_t* out_q8;
  __half* out_f16;
  __half* scales;
  int scale_cols;
  const __half* hmat;  // [Kp][Kp] fp16 (+-1, zero padded) or nullptr when K == 1
  int K, Kp, m;
  float scale;         // 1/sqrt(n)
  int warp_bytes;      // per-warp staging bytes
  int h_pitch;         // bytes per row of the shared copy of hmat
  unsigned long long* trace;  // MXB_HAD_TRACE=1: globaltimer stamps of block 0's pipeline events (hadamard_pipe.inc), else null
  int dbg;             // MXB_HAD_DBG bit mask (profiling experiments only): 1 skip load-phase math, 2 skip MMA, 4 skip quant
};

template <bool SILU>
__global__ void __launch_bounds__(256) hadamard_quant_kernel(HadParams p) {
  extern __shared__ __align__(16) uint8_t hsm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int hbytes = p.K > 1 ? p.Kp * p.h_pitch : 0;
  // ---- CTA-resident copy of hadK
  if (p.K > 1) {
    const int chunks_per_row = p.Kp / 8;  // 16-byte chunks
    for (int i = threadIdx.x; i < p.Kp * chunks_per_row; i += blockDim.x) {
      const int r = i / chunks_per_row, c = i % chunks_per_row;
      *reinterpret_cast<uint4*>(hsm + r * p.h_pitch + c * 16) = *reinterpret_cast<const uint4*>(p.hmat + r * p.Kp + c * 8);
    }
  }
  __syncthreads();
  uint8_t* wbuf = hsm + hbytes + warp * p.warp_bytes;
  float* F = reinterpret_cast<float*>(wbuf);
  __half* Xh = reinterpret_cast<__half*>(wbuf);
  const int n = p.n, n8 = n >> 3, m = p.m;
  __half* Yh = p.K > 1 ? Xh + p.Kp * m : Xh;

  for (int row = blockIdx.x * nwarps + warp; row < p.n_rows; row += gridDim.x * nwarps) {
    int a_bits = p.actq[0].x, gsize = p.actq[0].y;
    if (p.E > 1) {
      if (row >= p.expert_offsets[p.E]) break;
      const int e = find_expert(p.expert_offsets, p.E, row);
      a_bits = p.actq[e].x;
      gsize = p.actq[e].y;
    }
    const int64_t srow = p.src_row ? p.src_row[row] : row;
    const __half* in = p.x + srow * p.in_pitch;
    // ---- load (+ silu*mul), butterflies of stride 1,2,4 in registers, stage as fp32
    for (int i = lane; i < n8; i += 32) {
      float v[8];
      load8<SILU>(in, n, i, v);
      if (m >= 8) {
#pragma unroll
        for (int h = 1; h < 8; h <<= 1) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if ((j & h) == 0) {
              const float a = v[j], b = v[j + h];
              v[j] = a + b;
              v[j + h] = a - b;
            }
          }
        }
      }
      *reinterpret_cast<float4*>(F + i * 8) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(F + i * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    __syncwarp();
    // ---- remaining butterflies in shared memory
    for (int h = (m >= 8 ? 8 : 1); h < m; h <<= 1) {
      for (int t = lane; t < n / 2; t += 32) {
        const int i = ((t / h) * 2 * h) + (t % h);
        const float a = F[i], b = F[i + h];
        F[i] = a + b;
        F[i + h] = a - b;
      }
      __syncwarp();
    }
    // ---- scale by 1/sqrt(n), round to fp16, convert in place (fp32 bytes [32i,32i+32) -> fp16 bytes [16i,16i+16))
    for (int it = 0; it * 32 < n8; ++it) {
      const int i = it * 32 + lane;
      uint4 pk = make_uint4(0, 0, 0, 0);
      if (i < n8) {
        const float4 a = *reinterpret_cast<const float4*>(F + i * 8);
        const float4 b = *reinterpret_cast<const float4*>(F + i * 8 + 4);
        __half2* ph = reinterpret_cast<__half2*>(&pk);
        ph[0] = __floats2half2_rn(a.x * p.scale, a.y * p.scale);
        ph[1] = __floats2half2_rn(a.z * p.scale, a.w * p.scale);
        ph[2] = __floats2half2_rn(b.x * p.scale, b.y * p.scale);
        ph[3] = __floats2half2_rn(b.z * p.scale, b.w * p.scale);
      }
      __syncwarp();
      if (i < n8) *reinterpret_cast<uint4*>(Xh + i * 8) = pk;
    }
    __syncwarp();
    const __half* R = Xh;
    if (p.K > 1) {
      // zero the padding rows K..Kp-1 of the B operand
      for (int i = n + lane * 8; i < p.Kp * m; i += 32 * 8) *reinterpret_cast<uint4*>(Xh + i) = make_uint4(0, 0, 0, 0);
      __syncwarp();
      // ---- Y[K, m] = hadK[K, K] @ X[K, m]
      const int mt_n = p.Kp / 16, nt_n = m / 8, kt_n = p.Kp / 16;
      const uint32_t hs = s_u32(hsm), xs = s_u32(Xh);
      for (int mt = 0; mt < mt_n; ++mt) {
        for (int nc = 0; nc < nt_n; nc += 4) {
          float acc[4][4];
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
          for (int kt = 0; kt < kt_n; ++kt) {
            uint32_t af[4];
            ldsm_x4(hs + (mt * 16 + (lane & 15)) * p.h_pitch + (kt * 16 + (lane >> 4) * 8) * 2, af);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (nc + q < nt_n) {
                uint32_t bf[2];
                ldsm_x2_trans(xs + ((kt * 16 + (lane & 15)) * m + (nc + q) * 8) * 2, bf);
                mma_16816(acc[q], af, bf);
              }
            }
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (nc + q < nt_n) {
              const int r0 = mt * 16 + (lane >> 2), c0 = (nc + q) * 8 + (lane & 3) * 2;
              if (r0 < p.K) *reinterpret_cast<__half2*>(Yh + r0 * m + c0) = __floats2half2_rn(acc[q][0], acc[q][1]);
              if (r0 + 8 < p.K) *reinterpret_cast<__half2*>(Yh + (r0 + 8) * m + c0) = __floats2half2_rn(acc[q][2], acc[q][3]);
            }
          }
        }
      }
      __syncwarp();
      R = Yh;
    }
    // ---- per-token quantization straight from shared memory
    auto ld = [&](int i, float (&v)[8]) {
      uint4 u = *reinterpret_cast<const uint4*>(R + i * 8);
      const __half* uh = reinterpret_cast<const __half*>(&u);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = __half2float(uh[j]);
    };
    quantize_row(ld, n, lane, a_bits, gsize, p.out_q8 ? p.out_q8 + static_cast<int64_t>(row) * n : nullptr,
                 p.out_f16 ? p.out_f16 + static_cast<int64_t>(row) * n : nullptr,
                 p.scales ? p.scales + static_cast<int64_t>(row) * p.scale_cols : nullptr);
    __syncwarp();
  }
}


// ---------------------------------------------------------------------------------------------------------
// Long rows (power-of-two part m > 128, e.g. the fused shared expert of Qwen2-57B: n = 20480 = 40 x 512): one CTA of
// 16 warps per row instead of one warp.  hadamard_quant_kernel above keeps a whole row per WARP in shared memory
// (4n bytes), which leaves two warps per SM at this length (measured 5.06 ms for 8192 x 20480, 58 % of the c5 step).
// Same arithmetic, step for step (butterfly network and order, fp16 staging, the same m16n8k16 tiles in the same k
// order, quant8), so the two kernels agree bit for bit; only the work distribution differs.
template <bool SILU>
__global__ void __launch_bounds__(512, 2) hadamard_wide_kernel(HadParams p) {
  extern __shared__ __align__(16) uint8_t hsm[];
  __shared__ float red[16];
  if (p.E > 1 && p.expert_offsets != nullptr) p.n_rows = min(p.n_rows, p.expert_offsets[p.E]);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int hbytes = p.Kp * p.h_pitch;
  {
    const int chunks_per_row = p.Kp / 8;
    for (int i = tid; i < p.Kp * chunks_per_row; i += 512) {
      const int r = i / chunks_per_row, c = i % chunks_per_row;
      *reinterpret_cast<uint4*>(hsm + r * p.h_pitch + c * 16) = *reinterpret_cast<const uint4*>(p.hmat + r * p.Kp + c * 8);
    }
  }
  uint8_t* wbuf = hsm + hbytes;
  float* F = reinterpret_cast<float*>(wbuf);
  __half* Xh = reinterpret_cast<__half*>(wbuf);
  const int n = p.n, n8 = n >> 3, m = p.m;
  __half* Yh = Xh + p.Kp * m;
  __syncthreads();

  for (int row = blockIdx.x; row < p.n_rows; row += gridDim.x) {
    int a_bits = p.actq[0].x;
    if (p.E > 1) a_bits = p.actq[find_expert(p.expert_offsets, p.E, row)].x;
    const int64_t srow = p.src_row ? p.src_row[row] : row;
    const __half* in = p.x + srow * p.in_pitch;
    // ---- load (+ silu*mul), butterflies of stride 1,2,4 in registers, stage as fp32
    for (int i = tid; i < n8; i += 512) {
      float v[8];
      load8<SILU>(in, n, i, v);
#pragma unroll
      for (int h = 1; h < 8; h <<= 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if ((j & h) == 0) {
            const float a = v[j], b = v[j + h];
            v[j] = a + b;
            v[j + h] = a - b;
          }
        }
      }
      *reinterpret_cast<float4*>(F + i * 8) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(F + i * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    __syncthreads();
    // ---- remaining butterflies in shared memory
    for (int h = 8; h < m; h <<= 1) {
      for (int t = tid; t < n / 2; t += 512) {
        const int i = ((t / h) * 2 * h) + (t % h);
        const float a = F[i], b = F[i + h];
        F[i] = a + b;
        F[i + h] = a - b;
      }
      __syncthreads();
    }
    // ---- scale by 1/sqrt(n), round to fp16, convert in place: a round's fp16 bytes land on fp32 bytes of this or
    // earlier rounds only, and every round reads before it writes
    for (int i0 = 0; i0 < n8; i0 += 512) {
      const int i = i0 + tid;
      uint4 pk = make_uint4(0, 0, 0, 0);
      if (i < n8) {
        const float4 a = *reinterpret_cast<const float4*>(F + i * 8);
        const float4 b = *reinterpret_cast<const float4*>(F + i * 8 + 4);
        __half2* ph = reinterpret_cast<__half2*>(&pk);
        ph[0] = __floats2half2_rn(a.x * p.scale, a.y * p.scale);
        ph[1] = __floats2half2_rn(a.z * p.scale, a.w * p.scale);
        ph[2] = __floats2half2_rn(b.x * p.scale, b.y * p.scale);
        ph[3] = __floats2half2_rn(b.z * p.scale, b.w * p.scale);
      }
      __syncthreads();
      if (i < n8) *reinterpret_cast<uint4*>(Xh + i * 8) = pk;
    }
    // zero the padding rows K..Kp-1 of the B operand (past the converted data; the fp32 copy is dead)
    for (int i = n + tid * 8; i < p.Kp * m; i += 512 * 8) *reinterpret_cast<uint4*>(Xh + i) = make_uint4(0, 0, 0, 0);
    __syncthreads();
    // ---- Y[K, m] = hadK[K, K] @ X[K, m]: (16-row, 32-column) tile groups dealt to the 16 warps
    {
      const int mt_n = p.Kp / 16, nt_n = m / 8, kt_n = p.Kp / 16, ngrp = (nt_n + 3) / 4;
      const uint32_t hs = s_u32(hsm), xs = s_u32(Xh);
      for (int g = warp; g < mt_n * ngrp; g += 16) {
        const int mt = g / ngrp, nc = (g - mt * ngrp) * 4;
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
        for (int kt = 0; kt < kt_n; ++kt) {
          uint32_t af[4];
          ldsm_x4(hs + (mt * 16 + (lane & 15)) * p.h_pitch + (kt * 16 + (lane >> 4) * 8) * 2, af);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (nc + q < nt_n) {
              uint32_t bf[2];
              ldsm_x2_trans(xs + ((kt * 16 + (lane & 15)) * m + (nc + q) * 8) * 2, bf);
              mma_16816(acc[q], af, bf);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (nc + q < nt_n) {
            const int r0 = mt * 16 + (lane >> 2), c0 = (nc + q) * 8 + (lane & 3) * 2;
            if (r0 < p.K) *reinterpret_cast<__half2*>(Yh + r0 * m + c0) = __floats2half2_rn(acc[q][0], acc[q][1]);
            if (r0 + 8 < p.K) *reinterpret_cast<__half2*>(Yh + (r0 + 8) * m + c0) = __floats2half2_rn(acc[q][2], acc[q][3]);
          }
        }
      }
    }
    __syncthreads();
    // ---- per-token quantization straight from shared memory (per-token scale: gsize -1)
    if (a_bits >= 16) {
      __half* o = p.out_f16 + static_cast<int64_t>(row) * n;
      for (int i = tid; i < n8; i += 512) *reinterpret_cast<uint4*>(o + i * 8) = *reinterpret_cast<const uint4*>(Yh + i * 8);
    } else {
      const int iupper = (1 << (a_bits - 1)) - 1;
      float amax = 0.f;
      for (int i = tid; i < n8; i += 512) {
        const uint4 u = *reinterpret_cast<const uint4*>(Yh + i * 8);
        const __half* uh = reinterpret_cast<const __half*>(&u);
#pragma unroll
        for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(__half2float(uh[j])));
      }
      amax = warp_max(amax);
      if (lane == 0) red[warp] = amax;
      __syncthreads();
      amax = red[0];
#pragma unroll
      for (int w = 1; w < 16; ++w) amax = fmaxf(amax, red[w]);
      const __half sh = make_scale(amax, static_cast<float>(iupper));
      const float sf = __half2float(sh), rcp = __frcp_rn(sf);
      if (tid == 0) p.scales[static_cast<int64_t>(row) * p.scale_cols] = sh;
      int8_t* o = p.out_q8 + static_cast<int64_t>(row) * n;
      for (int i = tid; i < n8; i += 512) {
        const uint4 u = *reinterpret_cast<const uint4*>(Yh + i * 8);
        const __half* uh = reinterpret_cast<const __half*>(&u);
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __half2float(uh[j]);
        *reinterpret_cast<uint2*>(o + i * 8) = quant8(v, sf, rcp, iupper);
      }
    }
    __syncthreads();  // the staging is reused by the next row
  }
}


#include "hadamard_mma.inc"
#include "hadamard_umma.inc"
#include "hadamard_pipe.inc"

}  // namespace

// get_hadK's factor for n (hadamard_utils.py:5-59): first K in the reference's test order with n % K == 0
// and n / K a power of two; K = 1 for powers of two; 0 if the reference would assert.
int hadamard_factor(int n) {
  static const int order[12] = {176, 172, 156, 140, 108, 60, 52, 36, 28, 40, 20, 12};
  for (int k : order)
    if (n % k == 0) {
      const int m = n / k;
      return (m & (m - 1)) == 0 ? k : 0;
    }
  return (n & (n - 1)) == 0 ? 1 : 0;
}

// Largest power of two A <= cap such that the row, read as K * A blocks of m / A, is served by the pipelined kernel with at
// least two staging slots (hadK (x) H_A as the tensor-core factor, api.cu hadamard_matrix); 1 keeps the reference's split.
int hadamard_split(int n, int K, int cap) {
  int best = 1;
  for (int A = 2; A <= cap && K * A <= 176 && n / (K * A) >= 8; A *= 2) {
    const int Kp = (K * A + 15) / 16 * 16;
    HadParams p{};
    p.n = n;
    p.K = K * A;
    p.m = n / (K * A);
    int smem = 0, slots = 0;
    if (Kp <= 16) slots = pipe_slots<1>(p, &smem);
    else if (Kp <= 32) slots = pipe_slots<2>(p, &smem);
    else if (Kp <= 48) slots = pipe_slots<3>(p, &smem);
    else if (Kp <= 64) slots = pipe_slots<4>(p, &smem);
    else slots = pipe_slots<11>(p, &smem);
    if (slots >= 2 && p.m <= 128) best = A;
  }
  return best;
}

cudaError_t hadamard_quant_launch(bool silu, const void* x, int n, const int32_t* src_row, int n_rows,
                                  const int32_t* expert_offsets, const int32_t* actq_dev, int E, void* out_q8, void* out_f16,
                                  void* scales, int scale_cols, const void* hmat_dev, int K, int Kp, int sm_count,
                                  bool any_g128, cudaStream_t st) {
  if (n_rows == 0) return cudaSuccess;
  HadParams p;
  p.x = static_cast<const __half*>(x);
  p.n = n;
  p.in_pitch = silu ? 2 * int64_t(n) : int64_t(n);
  p.src_row = src_row;
  p.n_rows = n_rows;
  p.expert_offsets = expert_offsets;
  p.actq = reinterpret_cast<const int2*>(actq_dev);
  p.E = E;
  p.out_q8 = static_cast<int8_t*>(out_q8);
  p.out_f16 = static_cast<__half*>(out_f16);
  p.scales = static_cast<__half*>(scales);
  p.scale_cols = scale_cols;
  p.hmat = static_cast<const __half*>(hmat_dev);
  p.K = K;
  p.Kp = Kp;
  p.m = n / K;
  p.scale = static_cast<float>(1.0 / sqrt(static_cast<double>(n)));
  p.h_pitch = Kp * 2 + 16;
  { const char* e = getenv("MXB_HAD_DBG"); p.dbg = e ? atoi(e) : 0; }
  p.trace = nullptr;
  static unsigned long long* trace_buf = nullptr;
  const char* tr_env = getenv("MXB_HAD_TRACE");
  const bool tracing = tr_env && atoi(tr_env) != 0;
  if (tracing) {  // profiling aid: [role 0..3][event slot 0..255] nanosecond stamps of block 0, printed after the launch
    if (!trace_buf) cudaMalloc(reinterpret_cast<void**>(&trace_buf), 4 * 256 * sizeof(unsigned long long));
    cudaMemsetAsync(trace_buf, 0, 4 * 256 * sizeof(unsigned long long), st);
    p.trace = trace_buf;
  }
  struct TraceDump {
    bool on; cudaStream_t st; unsigned long long* buf;
    ~TraceDump() {
      if (!on) return;
      static unsigned long long h[4 * 256];
      cudaStreamSynchronize(st);
      cudaMemcpy(h, buf, sizeof h, cudaMemcpyDeviceToHost);
      unsigned long long t0 = ~0ull;
      for (int i = 0; i < 4 * 256; ++i) if (h[i] && h[i] < t0) t0 = h[i];
      const char* names[4] = {"producer(slot issued)", "loader0(slot full seen, slot done)", "mma(issued)", "drain0(t_full seen, barrier passed, stored)"};
      for (int r = 0; r < 4; ++r) {
        fprintf(stderr, "[had trace] %s:", names[r]);
        for (int i = 0; i < 256 && h[r * 256 + i]; ++i) fprintf(stderr, " %.2f", (h[r * 256 + i] - t0) * 1e-3);
        fprintf(stderr, "\n");
      }
    }
  } trace_dump{tracing, st, trace_buf};
  static const int use_umma = [] { const char* e = getenv("MXB_HAD_UMMA"); return e ? atoi(e) : 1; }();
  // warp-specialised tcgen05 pipeline (hadamard_pipe.inc); MXB_HAD_PIPE=0 keeps the phase-by-phase kernel below (tests
  // compare the two bit for bit)
  const char* pipe_env = getenv("MXB_HAD_PIPE");
  const int use_pipe = pipe_env ? atoi(pipe_env) : 1;
  // m = 256 / 512 (long rows: 14336 = 28 x 512, 20480 = 40 x 512, 7168 = 28 x 256) are served when K' <= 64
  // (MXB_HAD_PIPE_WIDE=0 keeps hadamard_wide_kernel for them)
  const char* pw_env = getenv("MXB_HAD_PIPE_WIDE");
  const int m_max = (pw_env && atoi(pw_env) == 0) ? 128 : 512;
  if (use_pipe && use_umma && !silu && K > 1 && p.m >= 8 && p.m <= m_max && !any_g128) {
    cudaError_t perr = cudaSuccess;  // (a (K, m) the pipeline does not serve falls through to the phase kernel)
#define MXB_TRY_PIPE(KT_) \
    if (Kp == 16 * KT_ && launch_pipe<KT_>(p, sm_count, st, &perr)) return perr;
    MXB_TRY_PIPE(11) MXB_TRY_PIPE(4) MXB_TRY_PIPE(3) MXB_TRY_PIPE(2) MXB_TRY_PIPE(1)
#undef MXB_TRY_PIPE
  }
  if (use_umma && K > 1 && p.m >= 8 && p.m <= 128 && !any_g128) {  // tcgen05 form (hadamard_umma.inc)
    if (Kp == 176) return launch_umma<11>(silu, p, sm_count, st);
    if (Kp == 48) return launch_umma<3>(silu, p, sm_count, st);
    if (Kp == 32) return launch_umma<2>(silu, p, sm_count, st);
    if (Kp == 16) return launch_umma<1>(silu, p, sm_count, st);
    if (Kp == 64) return launch_umma<4>(silu, p, sm_count, st);
  }
  if (K > 1 && p.m >= 8 && p.m <= 128 && !any_g128) {
    if (Kp == 176) return launch_mma<11>(silu, p, sm_count, st);
    if (Kp == 48) return launch_mma<3>(silu, p, sm_count, st);
    if (Kp == 32) return launch_mma<2>(silu, p, sm_count, st);
    if (Kp == 16) return launch_mma<1>(silu, p, sm_count, st);
    if (Kp == 64) return launch_mma<4>(silu, p, sm_count, st);
  }
  int wb = 4 * n;
  if (K > 1 && 2 * Kp * p.m + 2 * n > wb) wb = 2 * Kp * p.m + 2 * n;
  wb = (wb + 15) & ~15;
  p.warp_bytes = wb;
  const int hbytes = K > 1 ? Kp * p.h_pitch : 0;
  // long rows: one CTA per row (hadamard_wide_kernel); MXB_HAD_WIDE=0 keeps the warp-per-row kernel (tests compare the two)
  const char* wide_env = getenv("MXB_HAD_WIDE");
  if ((!wide_env || atoi(wide_env) != 0) && K > 1 && p.m > 128 && !any_g128 && hbytes + wb <= 226 * 1024) {
    const int smem = hbytes + wb;
    int grid = sm_count * (2 * (smem + 1024) <= 227 * 1024 ? 2 : 1);
    if (grid > n_rows) grid = n_rows;
    cudaError_t e;
    if (silu) {
      e = cudaFuncSetAttribute(hadamard_wide_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);  // + 64 B static
      if (e != cudaSuccess) return e;
      hadamard_wide_kernel<true><<<grid, 512, smem, st>>>(p);
    } else {
      e = cudaFuncSetAttribute(hadamard_wide_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
      if (e != cudaSuccess) return e;
      hadamard_wide_kernel<false><<<grid, 512, smem, st>>>(p);
    }
    return cudaGetLastError();
  }
  const int budget = 227 * 1024 - hbytes;
  int warps = budget / wb;
  if (warps < 1) return cudaErrorInvalidValue;
  if (warps > 8) warps = 8;
  const int smem = hbytes + warps * wb;
  int ctas_per_sm = (227 * 1024) / (smem + 1024);
  if (ctas_per_sm < 1) ctas_per_sm = 1;
  if (ctas_per_sm > 4) ctas_per_sm = 4;
  int grid = sm_count * ctas_per_sm;
  const int need = (n_rows + warps - 1) / warps;
  if (grid > need) grid = need;
  cudaError_t e;
  if (silu) {
    e = cudaFuncSetAttribute(hadamard_quant_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    hadamard_quant_kernel<true><<<grid, warps * 32, smem, st>>>(p);
  } else {
    e = cudaFuncSetAttribute(hadamard_quant_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    hadamard_quant_kernel<false><<<grid, warps * 32, smem, st>>>(p);
  }
  return cudaGetLastError();
}

}  // namespace mxb
