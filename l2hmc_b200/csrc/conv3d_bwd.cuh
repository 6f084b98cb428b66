// Internal interface between conv3d_bwd.cu (dispatch, recompute kernel) and conv3d_bwd_tape.cu (tape kernel).
#pragma once
#include "common.cuh"

namespace l2hmc {

// -1: the shape does not fit shared memory (the caller falls back to the recompute kernel)
int conv3d_backward_tape(const l2hmc_conv3d_bwd_args* a, int rows, int psize, cudaStream_t stream);

}  // namespace l2hmc
