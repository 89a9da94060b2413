// common.cu -- error string, launch counter, pointer checks of libngp_b200.so
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace ngp {

static thread_local char g_err[1024] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int after_launch(const char *what, uint64_t n) {
    g_launches.fetch_add(n, std::memory_order_relaxed);
    const cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        set_error("%s: kernel launch failed: %s", what, cudaGetErrorString(e));
        return 3;
    }
    return 0;
}

bool is_device_ptr(const void *p) {
    cudaPointerAttributes attr;
    const cudaError_t e = cudaPointerGetAttributes(&attr, p);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

}  // namespace ngp

extern "C" {
int ngp_abi_version(void) { return NGP_ABI_VERSION; }
const char *ngp_last_error(void) { return ngp::g_err; }
uint64_t ngp_launch_count(void) { return ngp::g_launches.load(std::memory_order_relaxed); }
}
