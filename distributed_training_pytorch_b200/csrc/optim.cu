// Fused "average + optimizer" pass over a flat gradient bucket.
//
// Reference path (per bucket, per step): the DDP reducer scales every gradient by 1/W
// before the all-reduce and copies the bucket back into .grad
// (torch:include/torch/csrc/distributed/c10d/reducer.hpp:276-286), then optim.Adam runs
// ~10 multi-tensor kernels over g, p, m, v (torch:optim/adam.py:554+; demo.py:110-111).
// Here the reduced bucket is read ONCE: g = bucket * (1/W), then m, v, p are updated in
// the same pass -- 28 B/param of HBM traffic (read g,p,m,v; write p,m,v), which is the
// floor for Adam with fp32 state.  The kernel is a grid-stride float4 stream sized to
// a multiple of the SM count; it is HBM-bound for large buckets and latency-bound for
// the 371-parameter toy bucket.
//
// Op order = torch's (_single_tensor_adam, torch:optim/adam.py:457,476,531-547) with
// explicit round-to-nearest intrinsics so nvcc cannot contract or reassociate: given the
// same g the update is bit-identical to oracle/ddp_oracle.c.
#include "common.cuh"
#include <cmath>

namespace {

struct AdamScalars {
    float w1, b2, w2, bc2_sqrt, eps, neg_step, grad_scale;
};

__device__ __forceinline__ void adam_one(float g_raw, float &p, float &m, float &v, const AdamScalars &k) {
    const float g = __fmul_rn(g_raw, k.grad_scale);
    m = __fmaf_rn(k.w1, __fsub_rn(g, m), m);
    v = __fadd_rn(__fmul_rn(v, k.b2), __fmul_rn(__fmul_rn(k.w2, g), g));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), k.bc2_sqrt), k.eps);
    p = __fadd_rn(p, __fmul_rn(k.neg_step, __fdiv_rn(m, denom)));
}

__device__ __forceinline__ AdamScalars make_scalars(double lr, double beta1, double beta2, double eps,
                                                    double b1pow, double b2pow, float grad_scale) {
    AdamScalars k;
    k.w1 = (float)(1.0 - beta1);
    k.b2 = (float)beta2;
    k.w2 = (float)(1.0 - beta2);
    k.bc2_sqrt = (float)sqrt(1.0 - b2pow);
    k.eps = (float)eps;
    k.neg_step = (float)(-(lr / (1.0 - b1pow)));
    k.grad_scale = grad_scale;
    return k;
}

// b1pow/b2pow are beta^step for the step being applied (host doubles).  If `state` is given the
// bias-correction scalars come from the host-built table indexed by the device step counter
// instead (graph replay).
__global__ void __launch_bounds__(256)
avg_adam_kernel(const float *__restrict__ bucket, float *__restrict__ p, float *__restrict__ m,
                float *__restrict__ v, long long n, float grad_scale, double lr, double beta1, double beta2,
                double eps, double b1pow, double b2pow, const ddpb_opt_state *state,
                const float2 *__restrict__ bias_table, int bias_table_n) {
    AdamScalars k;
    if (state) {
        const long long step = state->step;
        const float2 bias = bias_table[step < bias_table_n ? step : bias_table_n - 1];
        k.w1 = (float)(1.0 - beta1); k.b2 = (float)beta2; k.w2 = (float)(1.0 - beta2);
        k.bc2_sqrt = bias.y; k.eps = (float)eps; k.neg_step = bias.x; k.grad_scale = grad_scale;
    } else {
        k = make_scalars(lr, beta1, beta2, eps, b1pow, b2pow, grad_scale);
    }
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    const bool aligned = ((((uintptr_t)bucket) | ((uintptr_t)p) | ((uintptr_t)m) | ((uintptr_t)v)) & 15u) == 0;
    long long done = 0;
    if (aligned) {
        const long long n4 = n >> 2;
        const float4 *g4 = reinterpret_cast<const float4 *>(bucket);
        float4 *p4 = reinterpret_cast<float4 *>(p);
        float4 *m4 = reinterpret_cast<float4 *>(m);
        float4 *v4 = reinterpret_cast<float4 *>(v);
        for (long long i = tid; i < n4; i += nthreads) {
            float4 g = __ldcs(g4 + i);            // the bucket is dead after this pass: stream it
            float4 pp = p4[i], mm = m4[i], vv = v4[i];
            adam_one(g.x, pp.x, mm.x, vv.x, k);
            adam_one(g.y, pp.y, mm.y, vv.y, k);
            adam_one(g.z, pp.z, mm.z, vv.z, k);
            adam_one(g.w, pp.w, mm.w, vv.w, k);
            p4[i] = pp; m4[i] = mm; v4[i] = vv;
        }
        done = n4 << 2;
    }
    for (long long i = done + tid; i < n; i += nthreads) {
        float pp = p[i], mm = m[i], vv = v[i];
        adam_one(bucket[i], pp, mm, vv, k);
        p[i] = pp; m[i] = mm; v[i] = vv;
    }
}

// The counters are advanced by a separate single-thread kernel enqueued right after the
// update: same stream, so it is ordered after every block of the pass above.
__global__ void advance_state_kernel(ddpb_opt_state *state) { state->step += 1; }

__global__ void __launch_bounds__(256)
avg_sgd_kernel(const float *__restrict__ bucket, float *__restrict__ p, float *__restrict__ buf, long long n,
               float grad_scale, float lr, float momentum, int first_step) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    for (long long i = tid; i < n; i += nthreads) {
        float d = __fmul_rn(bucket[i], grad_scale);
        if (momentum != 0.f) {
            d = first_step ? d : __fadd_rn(__fmul_rn(buf[i], momentum), d);
            buf[i] = d;
        }
        p[i] = __fsub_rn(p[i], __fmul_rn(lr, d));
    }
}

int stream_grid(long long n, int per_thread) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long blocks = (n / per_thread + 255) / 256;
    if (blocks < 1) blocks = 1;
    long long cap = (long long)sms * 8;          // 8 resident 256-thread CTAs per SM
    return (int)(blocks < cap ? blocks : cap);
}

}  // namespace

DDPB_API int ddpb_fused_avg_adam(const float *bucket, float *p, float *m, float *v, int64_t n, float grad_scale,
                                 double lr, double beta1, double beta2, double eps, int64_t step, void *stream) {
    DDPB_REQUIRE(bucket && p && m && v && n >= 0 && step >= 1, "ddpb_fused_avg_adam: bad arguments");
    if (n == 0) return 0;
    avg_adam_kernel<<<stream_grid(n, 4), 256, 0, (cudaStream_t)stream>>>(
        bucket, p, m, v, n, grad_scale, lr, beta1, beta2, eps, std::pow(beta1, (double)step),
        std::pow(beta2, (double)step), nullptr, nullptr, 0);
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_fused_avg_adam_dev(const float *bucket, float *p, float *m, float *v, int64_t n, float grad_scale,
                                     double beta1, double beta2, double eps, const float *bias_table,
                                     int bias_table_n, ddpb_opt_state *state, int advance, void *stream) {
    DDPB_REQUIRE(bucket && p && m && v && n >= 0 && state && bias_table && bias_table_n >= 1,
                 "ddpb_fused_avg_adam_dev: bad arguments");
    if (n > 0)
        avg_adam_kernel<<<stream_grid(n, 4), 256, 0, (cudaStream_t)stream>>>(
            bucket, p, m, v, n, grad_scale, 0.0, beta1, beta2, eps, 0.0, 0.0, state,
            reinterpret_cast<const float2 *>(bias_table), bias_table_n);
    if (advance) advance_state_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state);
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_fused_avg_sgd(const float *bucket, float *p, float *buf, int64_t n, float grad_scale, double lr,
                                double momentum, int64_t step, void *stream) {
    DDPB_REQUIRE(bucket && p && n >= 0 && step >= 1 && (momentum == 0.0 || buf), "ddpb_fused_avg_sgd: bad arguments");
    if (n == 0) return 0;
    avg_sgd_kernel<<<stream_grid(n, 1), 256, 0, (cudaStream_t)stream>>>(bucket, p, buf, n, grad_scale, (float)lr,
                                                                         (float)momentum, step == 1);
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
