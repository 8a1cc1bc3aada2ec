// tcgen05 policy path (TFODE_POLICY_TC) -- placeholder until the tensor-core kernel lands:
// selecting it fails loudly instead of silently running the fp32 kernel.
#pragma once
#include "tfode_device.cuh"
namespace tfode {
struct PolicyTC { void* w = nullptr; };
inline int policy_tc_upload(PolicyTC&, const float*, cudaStream_t) { return 0; }
inline void policy_tc_free(PolicyTC&) {}
inline int policy_tc_launch(PolicyTC&, const float*, const int*, const int*, int, int, const float*, const float*, const float*, const float*, EnvRec*, float*, cudaStream_t) { return -1; }
}  // namespace tfode
