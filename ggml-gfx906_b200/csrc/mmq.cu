#include "common.cuh"
#include "mmq_tc.cuh"
using namespace b200;
extern "C" size_t b200_mul_mat_q_workspace_size(int64_t m, int64_t k, int64_t n) { return 256; }
extern "C" int b200_mul_mat_q(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                               const float * x_dev, size_t x_row_stride_bytes, int64_t n,
                               float * dst_dev, size_t dst_col_stride_bytes,
                               void * workspace_dev, size_t workspace_bytes, void * stream) { return B200_ERR_TYPE; }
