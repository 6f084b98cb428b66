// Internal interface between nets.cu (FFMA path, dispatch) and gemm_tc.cu (tcgen05 path).
#pragma once
#include <algorithm>

#include "common.cuh"

namespace l2hmc {

struct TcDims {
  int kbK, kbH;        // 32-wide K blocks of the input features / hidden width
  int h_rows;          // packed rows of W1^T, W2^T (H padded to 256)
  int d_tiles;         // 64-output head tiles
  int rows_p;          // chains padded to 128
  size_t w1_floats, w2_floats, w3_floats;   // packed weights (hi + lo)
  size_t acat_floats, h_floats;             // activations in TS32 form
};

TcDims tc_dims(int K, int H, int D, int N);
size_t tc_workspace_floats(int K, int H, int D, int N);
int tc_min_rows(int K, int H, int D);
bool tc_eligible(const l2hmc_net_weights* w, const l2hmc_net_call* c);
bool tc_supported(const l2hmc_net_weights* w, const l2hmc_net_call* c);
int net_run_tc(const l2hmc_net_weights* w, const l2hmc_net_call* c, cudaStream_t st);
int net_run_simt(const l2hmc_net_weights* w, const l2hmc_net_call* c, cudaStream_t st);
int heads_col_tiles(int D, int N);
// chained tcgen05 products (gemm_tc.cu): see tc_gemm_chain
bool tc_gemm_chainable(const l2hmc_gemm_args* g);
size_t tc_gemm_ts_floats(int M, int K);
int tc_gemm_chain(const l2hmc_gemm_args* g, const float* a_ts, float* out_ts, void* workspace, size_t workspace_bytes,
                  cudaStream_t st);
bool dense_grad_nt_ok(const l2hmc_gemm_args* g);
int dense_grad_nt(const l2hmc_gemm_args* g, cudaStream_t st);

}  // namespace l2hmc
