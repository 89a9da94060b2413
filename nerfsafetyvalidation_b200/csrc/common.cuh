// common.cuh -- error plumbing shared by every translation unit of libngp_b200.so
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "ngp_b200.h"

namespace ngp {

// thread-local message returned by ngp_last_error()
void set_error(const char *fmt, ...);
// bumps the library-wide kernel launch counter (ngp_launch_count) and checks the launch
int after_launch(const char *what, uint64_t n = 1);

#define NGP_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            ngp::set_error(__VA_ARGS__);  \
            return 1;                     \
        }                                 \
    } while (0)

#define NGP_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            ngp::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return 2;                                                                    \
        }                                                                                \
    } while (0)

static inline uint32_t div_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
static inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// true iff p is a device (or managed) pointer; the library has no CPU fallback
bool is_device_ptr(const void *p);

#define NGP_DEVPTR(p) NGP_REQUIRE((p) != nullptr && ngp::is_device_ptr(p), #p " must be a CUDA device pointer (no CPU fallback)")

}  // namespace ngp
