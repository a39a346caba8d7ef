// Introspection, error text and thin copy/sync helpers of the C ABI.
#include "common.cuh"
#include <cstring>

namespace ddpb {

char *error_buffer() {
    static thread_local char buf[512] = "";
    return buf;
}

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace ddpb

DDPB_API int ddpb_version(void) { return DDPB_VERSION; }

DDPB_API const char *ddpb_last_error(void) { return ddpb::error_buffer(); }

DDPB_API int ddpb_device_info(int device, int *cc_major, int *cc_minor, int *sm_count) {
    cudaDeviceProp prop;
    DDPB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (sm_count) *sm_count = prop.multiProcessorCount;
    return 0;
}

DDPB_API int ddpb_stream_sync(void *stream) {
    DDPB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

DDPB_API int ddpb_memcpy_h2d_async(void *dst_dev, const void *src_host, int64_t bytes, void *stream) {
    DDPB_REQUIRE(bytes >= 0, "ddpb_memcpy_h2d_async: negative size");
    DDPB_CUDA(cudaMemcpyAsync(dst_dev, src_host, (size_t)bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
}

DDPB_API int ddpb_memcpy_d2h_async(void *dst_host, const void *src_dev, int64_t bytes, void *stream) {
    DDPB_REQUIRE(bytes >= 0, "ddpb_memcpy_d2h_async: negative size");
    DDPB_CUDA(cudaMemcpyAsync(dst_host, src_dev, (size_t)bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return 0;
}
