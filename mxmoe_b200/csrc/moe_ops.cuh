// Launchers of the non-GEMM kernels (moe_ops.cu, hadamard.cu, w4a16_stream.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mxb {

size_t permute_ws_bytes(int n, int E);
cudaError_t permute_launch(const int32_t* ids, int T, int k, int E, int32_t* expert_offsets, int32_t* src_row,
                           int32_t* slot_of, int32_t* ws, cudaStream_t st);
// the same launch with the scratch in a file-scope device array (no allocation; `device` must be current)
size_t permute_static_capacity();
cudaError_t permute_static_launch(int device, const int32_t* ids, int T, int k, int E, int32_t* expert_offsets, int32_t* src_row,
                                  int32_t* slot_of, cudaStream_t st);

// destination of one expert's rows in an expert-parallel dispatch (same layout as MxbRowDest, include/mxmoe_b200.h)
struct RowDest {
  void* q8;
  void* f16;
  void* scales;
  int64_t row0;
};

// actq_dev: device int32 pairs (a_bits, gsize) per expert; dest (device [E]) redirects every expert's rows (may be null);
// max_ctas > 0 caps the grid (rows are walked grid-stride)
cudaError_t quant_rows_launch(bool silu, const void* x, int H, int64_t in_pitch, const int32_t* src_row, int n_rows,
                              const int32_t* expert_offsets, const int32_t* actq_dev, int E, void* out_q8, void* out_f16,
                              void* scales, int scale_cols, const RowDest* dest, int max_ctas, cudaStream_t st);
cudaError_t ep_push_counts_launch(const int32_t* expert_offsets, int E, int32_t* const* peer_counts_dev, int W, int rank, cudaStream_t st);
cudaError_t ep_plan_launch(const int32_t* counts, int W, int rank, int E, void* const* peer_q8, void* const* peer_f16,
                           void* const* peer_scales, int32_t* seg_offs, int32_t* back_row, RowDest* dest, cudaStream_t st);

int hadamard_factor(int n);
int hadamard_split(int n, int K, int cap);  // Kronecker widening of the tensor-core factor (hadamard.cu)
cudaError_t hadamard_quant_launch(bool silu, const void* x, int n, const int32_t* src_row, int n_rows,
                                  const int32_t* expert_offsets, const int32_t* actq_dev, int E, void* out_q8, void* out_f16,
                                  void* scales, int scale_cols, const void* hmat_dev, int K, int Kp, int sm_count,
                                  bool any_g128, cudaStream_t st);

cudaError_t combine_launch(const void* down, int H, const int32_t* slot_of, const float* topk_w, int T, int k,
                           const void* shared_out, const float* shared_scale, void* out, cudaStream_t st);

cudaError_t quant_pack_weight_launch(const void* W, int N, int K, int bits, int gsize, int sym, int layout, void* packed,
                                     void* scales, cudaStream_t st);
cudaError_t repack_wxax_launch(const void* ref, int N, int K, int bits, void* packed, cudaStream_t st);
cudaError_t repack_act_launch(const void* ref, int64_t M, int K, int bits, void* out, cudaStream_t st);
cudaError_t repack_wonly_launch(const void* ref, int N, int K, int bits, const int32_t* inv_dev, void* packed,
                                cudaStream_t st);

}  // namespace mxb
