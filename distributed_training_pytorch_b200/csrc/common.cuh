// Shared helpers for libddp_b200.so: thread-local error text and CUDA error mapping.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include "../../include/ddp_b200.h"

#define DDPB_API extern "C" __attribute__((visibility("default")))

namespace ddpb {

char *error_buffer();                                  // 512 bytes, thread-local
int fail(int code, const char *fmt, ...);              // records text, returns code
// Poll a pinned host word until it no longer holds `sentinel_bits` (util.cu, ddpb_wait_f32).
int wait_slot(const float *host_slot, unsigned sentinel_bits, long long spin_ns, cudaStream_t stream, float *out);

inline int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    return fail(1000 + (int)e, "%s failed: %s (%s) at %s:%d", what, cudaGetErrorName(e),
                cudaGetErrorString(e), file, line);
}

// Every kernel launch of the library is counted here (ddpb_launch_count): the number bench.py
// reports as gpu_launches is read from this counter, not kept by the caller.
void count_launch();

}  // namespace ddpb

#define DDPB_CUDA(call)                                                              \
    do {                                                                             \
        cudaError_t e__ = (call);                                                    \
        if (e__ != cudaSuccess) return ::ddpb::cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)

#define DDPB_REQUIRE(cond, ...)                                   \
    do {                                                          \
        if (!(cond)) return ::ddpb::fail(2, __VA_ARGS__);         \
    } while (0)
