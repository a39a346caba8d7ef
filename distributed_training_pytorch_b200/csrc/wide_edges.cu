// The thin ends of the widened toy model: layer 0 (2 -> H), layer 4 (H -> 1), the MSE loss and
// their backward.  These are not GEMM-shaped (K = 2, N = 1): they are streaming kernels over
// [B,H] activations (B <= 256), coalesced on the H axis, bounded by HBM/L2 bandwidth on a few MB.
// Reference lines: toy_model_and_data.py:13-14 (first Linear + LeakyReLU), :21 (last Linear),
// demo.py:106-109 (MSELoss + backward).  Arithmetic order per element follows
// oracle/ddp_oracle.c (fmaf from the bias, inputs ascending); batch sums run b ascending.
#include "common.cuh"

namespace {

constexpr float kSlope = 0.01f;
constexpr int kMaxB = 256;

__device__ __forceinline__ float lrelu(float z) { return z > 0.f ? z : z * kSlope; }

// act[b][j] and act_t[j][b] for the first layer; x rows optionally gathered through idx.
__global__ void __launch_bounds__(256)
first_fwd_kernel(const float *__restrict__ x, const int32_t *__restrict__ idx, const float *__restrict__ W0,
                 const float *__restrict__ b0, float *__restrict__ act, float *__restrict__ act_t, int B, int H, int ld) {
    const long long total = (long long)B * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        {   // j fastest: coalesced row-major store
            const int b = (int)(t / H), j = (int)(t - (long long)b * H);
            const long long s = idx ? idx[b] : b;
            float z = b0[j];
            z = fmaf(W0[2 * j], x[2 * s], z);
            z = fmaf(W0[2 * j + 1], x[2 * s + 1], z);
            act[t] = lrelu(z);
        }
        {   // b fastest: coalesced store of the transposed copy
            const int j = (int)(t / B), b = (int)(t - (long long)j * B);
            const long long s = idx ? idx[b] : b;
            float z = b0[j];
            z = fmaf(W0[2 * j], x[2 * s], z);
            z = fmaf(W0[2 * j + 1], x[2 * s + 1], z);
            act_t[(long long)j * ld + b] = lrelu(z);
        }
    }
}

// One CTA per sample: out[b] = b4 + W4 . act3[b]; dz4[b] = (2/B)(out - y); sq[b] = (out - y)^2.
__global__ void __launch_bounds__(256)
last_fwd_kernel(const float *__restrict__ act3, const float *__restrict__ W4, const float *__restrict__ b4,
                const float *__restrict__ y, const int32_t *__restrict__ idx, float *__restrict__ out,
                float *__restrict__ dz4, float *__restrict__ sq, int B, int H) {
    __shared__ float red[8];
    const int b = blockIdx.x;
    const float *row = act3 + (long long)b * H;
    float s = 0.f;
    for (int k = threadIdx.x; k < H; k += blockDim.x) s = fmaf(W4[k], row[k], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float total = b4[0];
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) total += red[w];
        out[b] = total;
        if (y) {
            const float diff = total - y[idx ? idx[b] : b];
            dz4[b] = __fdiv_rn(2.0f, (float)B) * diff;
            sq[b] = diff * diff;
        }
    }
}

// Backward of the last layer for a 32-wide slab of hidden units k:
//   dW4[k] = sum_b dz4[b] act3[b][k];  dz3[b][k] = LeakyReLU'(act3[b][k]) dz4[b] W4[k] (+ transposed copy);
//   block 0 also reduces the loss and db4.
__global__ void __launch_bounds__(256)
last_bwd_kernel(const float *__restrict__ act3, const float *__restrict__ W4, const float *__restrict__ dz4,
                const float *__restrict__ sq, float *__restrict__ dz3, float *__restrict__ dz3_t, float *__restrict__ gW4,
                float *__restrict__ gb4, float *__restrict__ sq_sum_out, float *__restrict__ loss_slot, int B, int H, int ld) {
    __shared__ float tile[32][kMaxB + 1];
    __shared__ float part[8][32];
    const int k0 = blockIdx.x * 32, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k = k0 + lane;
    float acc = 0.f;
    if (k < H) {
        const float w = W4[k];
        for (int b = warp; b < B; b += 8) {                       // 8 warps stride the batch; lanes = consecutive k
            const float a = act3[(long long)b * H + k], d = dz4[b];
            acc = fmaf(d, a, acc);
            const float g = d * w;
            const float v = a > 0.f ? g : g * kSlope;
            dz3[(long long)b * H + k] = v;
            tile[lane][b] = v;
        }
    }
    part[warp][lane] = acc;
    __syncthreads();
    if (warp == 0 && k < H) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += part[w][lane];
        gW4[k] = s;
    }
    for (int r = warp; r < 32; r += 8)                            // transposed copy: lanes = consecutive b
        if (k0 + r < H)
            for (int b = lane; b < B; b += 32) dz3_t[(long long)(k0 + r) * ld + b] = tile[r][b];
    if (blockIdx.x == 0 && warp == 0) {                           // loss and bias gradient
        float s = 0.f, d = 0.f;
        for (int b = lane; b < B; b += 32) { s += sq[b]; d += dz4[b]; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); d += __shfl_xor_sync(0xffffffffu, d, o); }
        if (lane == 0) {
            gb4[0] = d;
            sq_sum_out[0] = s;
            if (loss_slot) *loss_slot = __fdiv_rn(s, (float)B);
        }
    }
}

// Backward of the first layer: one thread per hidden unit j, batch ascending.
__global__ void __launch_bounds__(256)
first_bwd_kernel(const float *__restrict__ dz0, const float *__restrict__ x, const int32_t *__restrict__ idx,
                 float *__restrict__ gW0, float *__restrict__ gb0, int B, int H) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= H) return;
    float g0 = 0.f, g1 = 0.f, gb = 0.f;
    for (int b = 0; b < B; ++b) {
        const float d = dz0[(long long)b * H + j];
        const long long s = idx ? idx[b] : b;
        g0 = fmaf(d, x[2 * s], g0);
        g1 = fmaf(d, x[2 * s + 1], g1);
        gb += d;
    }
    gW0[2 * j] = g0;
    gW0[2 * j + 1] = g1;
    gb0[j] = gb;
}

int grid_for(long long n) {
    long long blocks = (n + 255) / 256;
    return (int)(blocks < 148 * 8 ? (blocks < 1 ? 1 : blocks) : 148 * 8);
}

}  // namespace

DDPB_API int ddpb_wide_first_fwd(const float *x, const int32_t *idx, const float *W0, const float *b0, float *act,
                                 float *act_t, int B, int H, int ld_t, void *stream) {
    DDPB_REQUIRE(x && W0 && b0 && act && act_t && B >= 1 && B <= kMaxB && H >= 1 && ld_t >= B, "ddpb_wide_first_fwd: bad arguments");
    first_fwd_kernel<<<grid_for((long long)B * H), 256, 0, (cudaStream_t)stream>>>(x, idx, W0, b0, act, act_t, B, H, ld_t);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_wide_last_fwd(const float *act3, const float *W4, const float *b4, const float *y, const int32_t *idx,
                                float *out, float *dz4, float *sq, int B, int H, void *stream) {
    DDPB_REQUIRE(act3 && W4 && b4 && out && B >= 1 && B <= kMaxB && H >= 1 && (!y || (dz4 && sq)), "ddpb_wide_last_fwd: bad arguments");
    last_fwd_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(act3, W4, b4, y, idx, out, dz4, sq, B, H);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_wide_last_bwd(const float *act3, const float *W4, const float *dz4, const float *sq, float *dz3,
                                float *dz3_t, float *gW4, float *gb4, float *sq_sum_out, float *loss_slot, int B, int H,
                                int ld_t, void *stream) {
    DDPB_REQUIRE(act3 && W4 && dz4 && sq && dz3 && dz3_t && gW4 && gb4 && sq_sum_out && B >= 1 && B <= kMaxB && ld_t >= B,
                 "ddpb_wide_last_bwd: bad arguments");
    last_bwd_kernel<<<(H + 31) / 32, 256, 0, (cudaStream_t)stream>>>(act3, W4, dz4, sq, dz3, dz3_t, gW4, gb4, sq_sum_out,
                                                                      loss_slot, B, H, ld_t);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_wide_first_bwd(const float *dz0, const float *x, const int32_t *idx, float *gW0, float *gb0, int B, int H,
                                 void *stream) {
    DDPB_REQUIRE(dz0 && x && gW0 && gb0 && B >= 1 && B <= kMaxB && H >= 1, "ddpb_wide_first_bwd: bad arguments");
    first_bwd_kernel<<<(H + 255) / 256, 256, 0, (cudaStream_t)stream>>>(dz0, x, idx, gW0, gb0, B, H);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
