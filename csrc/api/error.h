// error.h -- error plumbing shared by the C ABI translation units.
//
// Replaces the reference's checkCuda (S4:20-27), which prints and calls exit(): here a failing CUDA call
// throws pb::CudaError, the C ABI turns it into a negative return code plus pb_last_error(), and the
// Python layer raises RuntimeError.
#pragma once
#include <stdexcept>
#include <string>

namespace pb {

void set_last_error(const std::string& msg);

struct CudaError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

}  // namespace pb

#if defined(__CUDACC__) || defined(PB_WITH_CUDA_RUNTIME)
#include <cuda_runtime.h>
#define PB_CUDA_CHECK(expr)                                                                              \
  do {                                                                                                   \
    cudaError_t pb_err_ = (expr);                                                                        \
    if (pb_err_ != cudaSuccess) {                                                                        \
      throw pb::CudaError(std::string(#expr) + " failed at " + __FILE__ + ":" + std::to_string(__LINE__) + \
                          ": " + cudaGetErrorString(pb_err_));                                           \
    }                                                                                                    \
  } while (0)
#endif
