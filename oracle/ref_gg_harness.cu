// TEST INFRASTRUCTURE (oracle/): the REFERENCE's own GroupGEMM device code, compiled where it lies, as a GPU-side
// checker and same-box speed anchor.  Nothing of the reference is copied: this file only instantiates its templates
//   cta_gemm_multistage_v2      mxmoe/kernels/src/include/cta_gemm.cuh:8-107    (fp16)
//   cta_gemm_multistage_qab_v2  mxmoe/kernels/src/include/cta_gemm.cuh:424-608  (w8a8 / w4a4, per-token x per-channel scales)
//   GemmTile* / QuantTile* / GlobalTileC  mm_tile.cuh,  TileScheduler  tile_scheduler.cuh:5-50,  QParams  quantize.cuh:14-25
// with three tile configurations of the reference's sm_89 lists (tile_config.py:375-395; equal warp counts, so they
// may share one fused launch, compose_kernel.py:69-71) behind the calling convention its generator emits
// (compose_kernel.py:153-224 kernel body, kernel_sketch.py:82-145 host launcher).  The reference ships generated
// sources for fp16 only (src/generated/hz_fused_*.cu); the quantized instantiations below are what its generator would
// produce for qcfg_list = [fp16, w4a4_g-1_sym, w8a8_g-1_sym].
//
// Built by oracle/Makefile into oracle/_ref/libref_gg.so (git-ignored; travels to the GPU box).  Its mma.sync kernels
// assemble for sm_100a (the s4 MMA is emulated with two s8 IMMAs there).  Used by
// tests/test_gpu_ref_kernel.py (epilogue scale order pinned by EXECUTION of the reference's code) and
// tools/ref_kernel_compare.py (speed of the reference's kernel on the same B200).  Never loaded by the product.
#include "cta_gemm.cuh"
#include "registry.cuh"
#include "tile_scheduler.cuh"

namespace mxmoe {
using namespace tiled_gemm;
using MI = RCR_FP16FP16FP16;

using QA8 = QConfig<half, true, 8, -1, PackDim::K, false, half>;
using QA4 = QConfig<half, true, 4, -1, PackDim::K, false, half>;
using TileF16 = TileConfig<128, 128, 64, 2, 2, 1, 3, -1, MMA_FP16_FP32, NO_QUANT, NO_QUANT>;
using TileW8A8 = TileConfig<128, 128, 128, 2, 2, 1, 3, -1, MMA_S8_K32, QA8, QA8>;
using TileW4A4 = TileConfig<128, 128, 128, 2, 2, 1, 4, -1, MMA_S4_K64, QA4, QA4>;  // BK 128: K = 1408 (c2 down) has no 256 multiple
constexpr int kTileSmem = 98304;  // STAGE x (128 x BK x bits + 128 x BK x bits) / 8: 96 KB (fp16, w8a8), 64 KB (w4a4)
constexpr int kScaleSmem = 2048;

enum Kind { K_F16 = 0, K_W8A8 = 1, K_W4A4 = 2, K_BAD = -1 };
__host__ __device__ inline int kind_of(const QParams& q) {
  if (q.qbits.x == 16 && q.qbits.y == 16) return K_F16;
  if (q.qbits.x == 8 && q.qbits.y == 8 && q.gsize == -1 && q.sym) return K_W8A8;
  if (q.qbits.x == 4 && q.qbits.y == 4 && q.gsize == -1 && q.sym) return K_W4A4;
  return K_BAD;
}

template <typename Cfg>
__device__ __forceinline__ void quant_tile(half* A, half* B, half* sa, half* sb, half* C, int M, int N, int K, dim3 tc, uint8_t* smem,
                                           half* smem_scale, int lane, int tidx) {
  using TA_ = QuantTileA<half, MI::LAYOUT_A, Cfg, typename Cfg::QConfigA>;
  using TB_ = QuantTileB<half, MI::LAYOUT_B, Cfg, typename Cfg::QConfigB>;
  using TC_ = GlobalTileC<MI::TC, MI::LAYOUT_C, Cfg>;
  auto wx = Cfg::warp_x();
  auto wy = Cfg::warp_y();
  auto wz = Cfg::warp_z();
  auto ta = TA_(A, M, K, tc.y, tc.x, wy, wz, tidx);
  auto tb = TB_(B, N, K, tc.y, tc.x, wx, wz, tidx);
  auto tcc = TC_(C, M, N, tc.y, tc.x, tidx);
  cta_gemm_multistage_qab_v2<Cfg>(ta, tb, tcc, smem, sa + tc.x * TA_::CTA_SCALE_SIZE, sb + tc.y * TB_::CTA_SCALE_SIZE, smem_scale,
                                  smem_scale + TA_::CTA_SCALE_SIZE, M, N, K, lane, wx, wy, wz, tidx);
}

__global__ void ref_gg_kernel(half** As, half** Bs, half** sa, half** sb, half** Cs, dim3* sizes, const QParams* qp, int P, int* prefix) {
  const int lane = threadIdx.x, tidx = lane + threadIdx.y * WarpSize;
  extern __shared__ uint8_t smem[];
  half* smem_scale = reinterpret_cast<half*>(smem + kTileSmem);
  TileScheduler sched(sizes, P, prefix);
  for (int tile = blockIdx.x;; tile += gridDim.x) {
    const int p = sched.get_problem_idx(tile);
    if (p == -1) break;
    const int kind = kind_of(qp[p]);
    const dim3 tc = sched.get_tile_coord<128, 128>(p, tile);  // every configuration above is 128 x 128
    const int M = sizes[p].x, N = sizes[p].y, K = sizes[p].z;
    if (kind == K_F16) {
      using TA_ = GemmTileA<MI::TA, MI::LAYOUT_A, TileF16>;
      using TB_ = GemmTileB<MI::TB, MI::LAYOUT_B, TileF16>;
      using TC_ = GlobalTileC<MI::TC, MI::LAYOUT_C, TileF16>;
      auto wx = TileF16::warp_x();
      auto wy = TileF16::warp_y();
      auto wz = TileF16::warp_z();
      auto ta = TA_(As[p], M, K, tc.y, tc.x, wy, wz, tidx);
      auto tb = TB_(Bs[p], N, K, tc.y, tc.x, wx, wz, tidx);
      auto tcc = TC_(Cs[p], M, N, tc.y, tc.x, tidx);
      cta_gemm_multistage_v2<TileF16>(ta, tb, tcc, smem, K, lane, wx, wy, wz);
    } else if (kind == K_W8A8) {
      quant_tile<TileW8A8>(As[p], Bs[p], sa[p], sb[p], Cs[p], M, N, K, tc, smem, smem_scale, lane, tidx);
    } else if (kind == K_W4A4) {
      quant_tile<TileW4A4>(As[p], Bs[p], sa[p], sb[p], Cs[p], M, N, K, tc, smem, smem_scale, lane, tidx);
    }
  }
}
}  // namespace mxmoe

// The reference's FuncType argument list (registry.cuh:28-39) without the unused ld* / Ds; `d_prefix` is caller-owned
// scratch of P ints so that repeated timed launches do not allocate.  Returns 0, -1 for an unsupported QParams entry
// (the reference throws std::runtime_error), or the cudaError_t.
extern "C" int ref_gg_launch(void** As, void** Bs, void** sa, void** sb, void** Cs, void* d_sizes, const void* h_sizes, void* d_q,
                             const void* h_q, int P, int* d_prefix, int n_ctas, void* stream) {
  using namespace mxmoe;
  const dim3* hs = static_cast<const dim3*>(h_sizes);
  const QParams* hq = static_cast<const QParams*>(h_q);
  std::vector<int> prefix(P);
  int total = 0;
  for (int i = 0; i < P; ++i) {
    if (kind_of(hq[i]) == K_BAD) return -1;
    total += cu_cdiv(int(hs[i].x), 128) * cu_cdiv(int(hs[i].y), 128);
    prefix[i] = total;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemcpyAsync(d_prefix, prefix.data(), sizeof(int) * P, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return int(e);
  static bool attr = false;
  if (!attr) {
    e = cudaFuncSetAttribute(ref_gg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTileSmem + kScaleSmem);
    if (e != cudaSuccess) return int(e);
    attr = true;
  }
  if (n_ctas <= 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_ctas, cudaDevAttrMultiProcessorCount, dev);  // grid = #SM like kernel_sketch.py:111-118
  }
  ref_gg_kernel<<<n_ctas, dim3(32, 4), kTileSmem + kScaleSmem, st>>>(reinterpret_cast<half**>(As), reinterpret_cast<half**>(Bs),
                                                                      reinterpret_cast<half**>(sa), reinterpret_cast<half**>(sb),
                                                                      reinterpret_cast<half**>(Cs), static_cast<dim3*>(d_sizes),
                                                                      static_cast<const QParams*>(d_q), P, d_prefix);
  return int(cudaGetLastError());
}

extern "C" int ref_gg_qparams_size() { return int(sizeof(mxmoe::QParams)); }
