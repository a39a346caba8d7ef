// Device-resident batch assembly: the dataset (512 x [2] + 512 x [1] floats for ToyData,
// toy_model_and_data.py:27-36) lives in HBM and a batch is a row gather by the sampler's
// indices -- replaces the per-step pinned H2D copies of demo.py:102-103.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float *__restrict__ X, const float *__restrict__ Y, const int32_t *__restrict__ idx,
                   long long B, int dx, int dy, float *__restrict__ outX, float *__restrict__ outY) {
    const long long per_row = dx + dy;
    const long long total = B * per_row;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const long long b = t / per_row;
        const int c = (int)(t - b * per_row);
        const long long src = idx[b];
        if (c < dx) outX[b * dx + c] = X[src * dx + c];
        else outY[b * dy + (c - dx)] = Y[src * dy + (c - dx)];
    }
}

}  // namespace

DDPB_API int ddpb_gather_batch(const float *X, const float *Y, const int32_t *idx, int64_t B, int dx, int dy,
                               float *outX, float *outY, void *stream) {
    DDPB_REQUIRE(X && Y && idx && outX && outY && B >= 0 && dx > 0 && dy > 0, "ddpb_gather_batch: bad arguments");
    if (B == 0) return 0;
    long long total = (long long)B * (dx + dy);
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    gather_rows_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(X, Y, idx, B, dx, dy, outX, outY);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
