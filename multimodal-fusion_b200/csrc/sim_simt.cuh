// FP32 SIMT tile kernels (sim_simt.cu): dense similarity and plain-FP32 fused top-k.
#pragma once
#include "common.cuh"

namespace hg {

struct SimArgs {
    const float* a; const float* b;          // [m, d], [n, d]
    int64_t m, n, d;
    const float* pa; const float* pb;        // positions or null
    int pos_dim;
    float lambda_h, lambda_g;
    int metric;                              // top-k mode only
    float* out; int64_t ld_out;              // dense mode
    int k; int64_t row_begin, row_end; int exclude_self;   // top-k mode
    int32_t* out_idx; float* out_val;
};

int launch_dense(const SimArgs& p, int form, cudaStream_t st);
int launch_topk_fp32(const SimArgs& p, cudaStream_t st);

}  // namespace hg
