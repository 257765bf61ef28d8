// Per-token activation quantization shared by the plain (moe_ops.cu) and Hadamard (hadamard.cu) kernels.
// Numerics: quant_minmax (mxmoe/quant/quant.py:40-84) on an fp16 tensor, symmetric, clip=(1,1), every
// operation rounded to fp16 the way torch does for half tensors, plus the scale==0 -> 1 guard of the
// reference kernel (mxmoe/kernels/src/include/quantize.cuh:256).
#pragma once
#include <cuda_fp8.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mxb {

constexpr int kBitsE4M3 = -8;  // MXB_BITS_E4M3: bit-width code of 8-bit E4M3 containers (include/mxmoe_b200.h)


__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float half_warp_max(float v) {  // reduce inside each group of 16 lanes
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float h_round(float x) { return __half2float(__float2half_rn(x)); }

// act_fn(gate) * up on fp16 tensors: silu evaluated in fp32 and rounded to fp16, product rounded to fp16.
// ex2.approx / rcp.approx (a few fp32 ulp) instead of expf + IEEE division: after the fp16 rounding the result
// equals the exactly-rounded one except in ~5e-4 of the elements (one fp16 ulp) — the kernel is issue-bound.
__device__ __forceinline__ float silu_mul_h(__half g, __half u) {
  const float gf = __half2float(g);
  const float s = h_round(__fdividef(gf, 1.f + __expf(-gf)));
  return h_round(s * __half2float(u));
}

template <bool SILU>
__device__ __forceinline__ void load8(const __half* __restrict__ row, int H, int idx8, float (&x)[8]) {
  if (SILU) {
    uint4 g = *reinterpret_cast<const uint4*>(row + idx8 * 8);
    uint4 u = *reinterpret_cast<const uint4*>(row + H + idx8 * 8);
    const __half* gh = reinterpret_cast<const __half*>(&g);
    const __half* uh = reinterpret_cast<const __half*>(&u);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = silu_mul_h(gh[j], uh[j]);
  } else {
    uint4 v = *reinterpret_cast<const uint4*>(row + idx8 * 8);
    const __half* vh = reinterpret_cast<const __half*>(&v);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = __half2float(vh[j]);
  }
}

__device__ __forceinline__ int find_expert(const int32_t* __restrict__ offs, int E, int row) {
  int lo = 0, hi = E;  // largest e with offs[e] <= row
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (offs[mid] <= row) lo = mid; else hi = mid;
  }
  return lo;
}

// q = rint(clamp(half(x / s))) with torch's fp16 semantics (fp32 divide, round to fp16).  The IEEE division is
// replaced by a per-row correctly-rounded reciprocal plus one Newton correction per element:
//   q0 = x*r;  q1 = q0 + (x - q0*s)*r   (|q1 - x/s| <= 0.5 ulp + 2^-40)
// x and s are fp16 values, so x/s is either exactly representable in fp32 or at least one fp32 ulp away from every
// fp16 rounding boundary: half(q1) == half(x/s) always (and the GPU tests compare bit-for-bit with the oracle).
__device__ __forceinline__ uint2 quant8(const float (&v)[8], float sf, float rcp, int upper) {
  // The tail runs on packed halves and never touches the conversion (XU) pipe: the two quotients are rounded to fp16
  // together (one F2FP), clamped with HMNMX2, and rounded to integers (ties to even, torch.round) by adding 1536 in
  // fp16 — in [1024, 2048) the fp16 spacing is 1, so the low byte of the sum's bit pattern is the two's-complement code.
  const __half2 hi2 = __float2half2_rn(static_cast<float>(upper)), lo2 = __float2half2_rn(static_cast<float>(-upper));
  const __half2 magic = __float2half2_rn(1536.f);
  uint32_t w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float a0 = v[2 * j] * rcp, b0 = v[2 * j + 1] * rcp;
    const float a1 = fmaf(fmaf(-a0, sf, v[2 * j]), rcp, a0), b1 = fmaf(fmaf(-b0, sf, v[2 * j + 1]), rcp, b0);
    __half2 q = __floats2half2_rn(a1, b1);
    q = __hadd2(__hmax2(__hmin2(q, hi2), lo2), magic);
    w[j] = *reinterpret_cast<uint32_t*>(&q);
  }
  uint2 pk;
  pk.x = __byte_perm(w[0], w[1], 0x6420);
  pk.y = __byte_perm(w[2], w[3], 0x6420);
  return pk;
}
// ---- the same 8 codes from PACKED fp16 values, entirely on the half2 pipe (no fp32 unpack, no Newton step in fp32).
// q = RN16(x / s) is produced by Markstein's division in fp16 FMAs on operands normalised by a power of two:
//   s' = s * 2^k in [1, 2),  x' = x * 2^k (exact),  r = RN16(1 / s'),
//   q0 = RN16(x' r),  e = RN16(x' - q0 s')  (one FMA),  q1 = RN16(q0 + e r)  (one FMA)
// Without the normalisation the residual e underflows for small scales and codes flip; with it, EVERY (x, s) pair with s a
// normal fp16 number and |x / s| <= 160 gives the code of the reference arithmetic, rint(clamp(RN16(RN32(x / s)))), for both
// the 8-bit and the 4-bit clamp — verified exhaustively (1.4e9 pairs, tools/verify_h2div.c).  Rows whose scale is an fp16
// denormal (< 6.1e-5) take quant8.
struct H2Div {
  __half2 fac, r, sp;  // 2^k, RN16(1/s'), s' — each duplicated in both halves
  bool fast;           // false: the scale is denormal, use quant8
};
__device__ __forceinline__ H2Div h2div_prepare(__half s) {
  H2Div d;
  const unsigned short sb = __half_as_ushort(s);
  const int es = (sb >> 10) & 31;
  d.fast = es != 0;
  const unsigned short fac = es == 30 ? 0x0200u : static_cast<unsigned short>((30 - es) << 10);  // 2^(15 - es); 2^-15 is denormal
  const unsigned short sp = static_cast<unsigned short>(0x3C00u | (sb & 0x3FFu));
  const __half spv = __ushort_as_half(sp);
  d.fac = __half2half2(__ushort_as_half(fac));
  d.sp = __half2half2(spv);
  d.r = __half2half2(__float2half_rn(__frcp_rn(__half2float(spv))));
  return d;
}
__device__ __forceinline__ uint2 quant8_h2(const uint32_t (&hv)[4], const __half2 fac, const __half2 r, const __half2 sp, int upper) {
  const __half2 hi2 = __float2half2_rn(static_cast<float>(upper)), lo2 = __float2half2_rn(static_cast<float>(-upper));
  const __half2 magic = __float2half2_rn(1536.f);
  uint32_t w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const __half2 xp = __hmul2(*reinterpret_cast<const __half2*>(&hv[j]), fac);
    const __half2 q0 = __hmul2(xp, r);
    const __half2 e = __hfma2(__hneg2(q0), sp, xp);
    __half2 q = __hfma2(e, r, q0);
    q = __hadd2(__hmax2(__hmin2(q, hi2), lo2), magic);
    w[j] = *reinterpret_cast<uint32_t*>(&q);
  }
  uint2 pk;
  pk.x = __byte_perm(w[0], w[1], 0x6420);
  pk.y = __byte_perm(w[2], w[3], 0x6420);
  return pk;
}
// 8 quotients -> 8 E4M3 bytes (round to nearest even, saturating at +-448): the fp8 counterpart of quant8.  The E4M3
// rounding boundaries are fp16-representable, so the argument above carries over: e4m3(q1) == e4m3(x / s) always.
__device__ __forceinline__ uint2 quant8_e4m3(const float (&v)[8], float sf, float rcp) {
  uint32_t w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float a0 = v[2 * j] * rcp, b0 = v[2 * j + 1] * rcp;
    const float a1 = fmaf(fmaf(-a0, sf, v[2 * j]), rcp, a0), b1 = fmaf(fmaf(-b0, sf, v[2 * j + 1]), rcp, b0);
    w[j] = __nv_cvt_float2_to_fp8x2(make_float2(a1, b1), __NV_SATFINITE, __NV_E4M3);
  }
  uint2 pk;
  pk.x = (w[0] & 0xFFFFu) | (w[1] << 16);
  pk.y = (w[2] & 0xFFFFu) | (w[3] << 16);
  return pk;
}
__device__ __forceinline__ __half make_scale(float amax, float upper) {
  __half sh = __float2half_rn(__fdiv_rn(amax, upper));
  if (__half2float(sh) == 0.f) sh = __float2half(1.f);
  return sh;
}

// One warp quantizes one row of H values delivered 8 at a time by `ld(idx8, v)`.
// a_bits 16: fp16 copy to o_f16; 8 / 4: int8 containers to o_q8 with scale(s) to so.
template <typename Loader>
__device__ __forceinline__ void quantize_row(Loader ld, int H, int lane, int a_bits, int gsize, int8_t* __restrict__ o_q8,
                                             __half* __restrict__ o_f16, __half* __restrict__ so) {
  const int n8 = H >> 3;
  if (a_bits >= 16) {
    for (int i = lane; i < n8; i += 32) {
      float v[8];
      ld(i, v);
      uint4 pk;
      __half2* ph = reinterpret_cast<__half2*>(&pk);
#pragma unroll
      for (int j = 0; j < 4; ++j) ph[j] = __floats2half2_rn(v[2 * j], v[2 * j + 1]);
      *reinterpret_cast<uint4*>(o_f16 + i * 8) = pk;
    }
    return;
  }
  const bool e4m3 = a_bits == kBitsE4M3;  // 8-bit float containers (per-token scale only)
  const int iupper = e4m3 ? 448 : (1 << (a_bits - 1)) - 1;
  const float upper = static_cast<float>(iupper);
  if (gsize <= 0 || e4m3) {
    float amax = 0.f;
    for (int i = lane; i < n8; i += 32) {
      float v[8];
      ld(i, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[j]));
    }
    amax = warp_max(amax);
    const __half sh = make_scale(amax, upper);
    const float sf = __half2float(sh), rcp = __frcp_rn(sf);
    if (lane == 0) so[0] = sh;
    for (int i = lane; i < n8; i += 32) {
      float v[8];
      ld(i, v);
      *reinterpret_cast<uint2*>(o_q8 + i * 8) = e4m3 ? quant8_e4m3(v, sf, rcp) : quant8(v, sf, rcp, iupper);
    }
  } else {  // gsize == 128: 16 lanes x 8 elements per group
    const int iters = (n8 + 31) / 32;
    for (int it = 0; it < iters; ++it) {
      const int i = it * 32 + lane;
      float v[8];
      float amax = 0.f;
      if (i < n8) {
        ld(i, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[j]));
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = 0.f;
      }
      amax = half_warp_max(amax);
      const __half sh = make_scale(amax, upper);
      if (i < n8) {
        if ((lane & 15) == 0) so[i >> 4] = sh;
        const float sf = __half2float(sh);
        *reinterpret_cast<uint2*>(o_q8 + i * 8) = quant8(v, sf, __frcp_rn(sf), iupper);
      }
    }
  }
}

}  // namespace mxb
