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
               float grad_scale, float lr, float momentum, int first_step, const ddpb_opt_state *state) {
    if (state) first_step = state->step == 0;
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

// ---- bucket all-reduce fused with Adam over peer windows (ddpb_bucket_reduce_adam) -----------
__device__ __forceinline__ float4 ld_sys_f4(const float4 *p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys_f4(float4 *p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float ld_sys_f1(const float *p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys_f1(float *p, float v) {
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys_i32(int *p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys_i32(const int *p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Spin until *flag >= ticket; after ~5 s record the failure in state->reserved[0] and give up.
// Returns false on a time-out.
__device__ __forceinline__ bool wait_flag(const int *flag, int ticket, ddpb_opt_state *state) {
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys_i32(flag) < ticket) {
        if (global_ns() - t0 > 5000000000ull) {
            state->reserved[0] = 1;
            return false;
        }
        __nanosleep(100);
    }
    return true;
}

// Flags of one window: ready[bucket][source rank], done[bucket][source rank], then a CTA counter
// per bucket (used only by the window's owner).
struct BucketFlags {
    int ready[DDPB_BUCKET_MAX][DDPB_MAX_PEERS];
    int done[DDPB_BUCKET_MAX][DDPB_MAX_PEERS];
    unsigned counter[DDPB_BUCKET_MAX];
};
static_assert(sizeof(BucketFlags) == DDPB_BUCKET_FLAGS_BYTES, "flag block layout");

template <int W>
__global__ void __launch_bounds__(256)
bucket_reduce_adam_kernel(const __grid_constant__ ddpb_bucket_job job) {
    const int r = job.rank, b = job.bucket_index, t = (int)job.ticket;
    BucketFlags *mine = reinterpret_cast<BucketFlags *>(job.window[r]);
    __shared__ int s_last, s_timeout;
    if (threadIdx.x == 0) s_timeout = 0;
    __syncthreads();

    // 1. every rank has produced this bucket (and, having got that far, no longer reads the layer's
    //    old parameters): tell the peers that we have, wait until they all have.  A peer that does
    //    not show up within ~5 s: every CTA of this kernel sees the same missing flag and leaves
    //    WITHOUT touching parameters or optimizer state (state->reserved[0] tells the host).
    if (job.mode) {
#pragma unroll
        for (int q = 0; q < W; ++q)
            if ((int)threadIdx.x == q && q != r) {
                if (blockIdx.x == 0) st_release_sys_i32(&reinterpret_cast<BucketFlags *>(job.window[q])->ready[b][r], t);
                if (!wait_flag(&mine->ready[b][q], t, job.state)) s_timeout = 1;
            }
        __syncthreads();
        if (s_timeout) return;
    }

    AdamScalars k;
    {
        const long long step = job.state->step;
        const float2 bias = reinterpret_cast<const float2 *>(job.bias_table)[step < job.bias_table_n ? step : job.bias_table_n - 1];
        k.w1 = (float)(1.0 - job.beta1); k.b2 = (float)job.beta2; k.w2 = (float)(1.0 - job.beta2);
        k.bc2_sqrt = bias.y; k.eps = (float)job.eps; k.neg_step = bias.x; k.grad_scale = job.grad_scale;
    }

    // 2-4. this rank's shard, 16 bytes per thread and iteration
    const long long n4 = job.n >> 2, per = (n4 + W - 1) / W;
    const long long g0 = per * r, g1 = (g0 + per < n4) ? g0 + per : n4;
    const long long first4 = job.start >> 2;
    const float4 *grads[W];
    float4 *params[W];
#pragma unroll
    for (int q = 0; q < W; ++q) {
        grads[q] = reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(job.window[q]) + job.grads_off) + first4;
        params[q] = reinterpret_cast<float4 *>(reinterpret_cast<char *>(job.window[q]) + job.params_off) + first4;
    }
    const float4 *lg = reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(mine) + job.grads_off) + first4;
    float4 *lp = reinterpret_cast<float4 *>(reinterpret_cast<char *>(mine) + job.params_off) + first4;
    float4 *m4 = reinterpret_cast<float4 *>(job.m) + first4, *v4 = reinterpret_cast<float4 *>(job.v) + first4;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    for (long long i = g0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < g1; i += nthreads) {
        float4 part[W];
        if (job.mode) {
#pragma unroll
            for (int q = 0; q < W; ++q)                      // all peers' loads in flight together; fixed order of the sum
                part[q] = (q == r || (job.reserved & 1)) ? __ldcs(lg + i) : ld_sys_f4(grads[q] + i);
        } else {
            part[0] = __ldcs(lg + i);
        }
        float4 pp = lp[i], mm = m4[i], vv = v4[i];
        float4 g = part[0];
        if (job.mode) {
#pragma unroll
            for (int q = 1; q < W; ++q) {
                g.x = __fadd_rn(g.x, part[q].x); g.y = __fadd_rn(g.y, part[q].y);
                g.z = __fadd_rn(g.z, part[q].z); g.w = __fadd_rn(g.w, part[q].w);
            }
        }
        adam_one(g.x, pp.x, mm.x, vv.x, k);
        adam_one(g.y, pp.y, mm.y, vv.y, k);
        adam_one(g.z, pp.z, mm.z, vv.z, k);
        adam_one(g.w, pp.w, mm.w, vv.w, k);
        m4[i] = mm; v4[i] = vv;
#pragma unroll
        for (int q = 0; q < W; ++q) {
            if (q == r) params[q][i] = pp;
            else if (!(job.reserved & 2)) st_sys_f4(params[q] + i, pp);
        }
    }
    // the 1-3 elements after the last full group of four (only the last layer's bucket has them)
    if (r == W - 1 && blockIdx.x == 0 && threadIdx.x < (job.n & 3)) {
        const long long e = job.start + (n4 << 2) + threadIdx.x;
        float g = 0.f;
        for (int q = 0; q < W; ++q) {
            if (!job.mode && q != r) continue;
            const float *gq = reinterpret_cast<const float *>(reinterpret_cast<const char *>(job.window[q]) + job.grads_off);
            g = __fadd_rn(g, q == r ? gq[e] : ld_sys_f1(gq + e));
        }
        float *pr = reinterpret_cast<float *>(reinterpret_cast<char *>(job.window[r]) + job.params_off);
        float pp = pr[e], mm = job.m[e], vv = job.v[e];
        adam_one(g, pp, mm, vv, k);
        job.m[e] = mm; job.v[e] = vv;
        for (int q = 0; q < W; ++q) {
            float *pq = reinterpret_cast<float *>(reinterpret_cast<char *>(job.window[q]) + job.params_off);
            if (q == r) pq[e] = pp; else st_sys_f1(pq + e, pp);
        }
    }

    // the last CTA to finish tells the peers that our shard has been pushed everywhere and waits for
    // theirs; the other CTAs leave (only one CTA per GPU ever spins here, so grids larger than the
    // machine cannot deadlock against each other)
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        s_last = atomicAdd(&mine->counter[b], 1u) == gridDim.x - 1;
        __threadfence();
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) mine->counter[b] = 0;
#pragma unroll
        for (int q = 0; q < W; ++q)
            if ((int)threadIdx.x == q && q != r) {
                st_release_sys_i32(&reinterpret_cast<BucketFlags *>(job.window[q])->done[b][r], t);
                wait_flag(&mine->done[b][q], t, job.state);
            }
    }
}

int stream_grid(long long n, int per_thread, int ctas_per_sm = 8) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long blocks = (n / per_thread + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (ctas_per_sm < 1 || ctas_per_sm > 8) ctas_per_sm = 8;
    long long cap = (long long)sms * ctas_per_sm;          // at most 8 resident 256-thread CTAs per SM
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
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_fused_avg_adam_dev(const float *bucket, float *p, float *m, float *v, int64_t n, float grad_scale,
                                     double beta1, double beta2, double eps, const float *bias_table,
                                     int bias_table_n, ddpb_opt_state *state, int advance, void *stream) {
    return ddpb_fused_avg_adam_dev_grid(bucket, p, m, v, n, grad_scale, beta1, beta2, eps, bias_table, bias_table_n, state,
                                        advance, 8, stream);
}

DDPB_API int ddpb_fused_avg_adam_dev_grid(const float *bucket, float *p, float *m, float *v, int64_t n, float grad_scale,
                                          double beta1, double beta2, double eps, const float *bias_table,
                                          int bias_table_n, ddpb_opt_state *state, int advance, int ctas_per_sm,
                                          void *stream) {
    DDPB_REQUIRE(bucket && p && m && v && n >= 0 && state && bias_table && bias_table_n >= 1,
                 "ddpb_fused_avg_adam_dev: bad arguments");
    if (n > 0) {
        avg_adam_kernel<<<stream_grid(n, 4, ctas_per_sm), 256, 0, (cudaStream_t)stream>>>(
            bucket, p, m, v, n, grad_scale, 0.0, beta1, beta2, eps, 0.0, 0.0, state,
            reinterpret_cast<const float2 *>(bias_table), bias_table_n);
        ddpb::count_launch();
    }
    if (advance) {
        advance_state_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state);
        ddpb::count_launch();
    }
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_fused_avg_sgd(const float *bucket, float *p, float *buf, int64_t n, float grad_scale, double lr,
                                double momentum, int64_t step, void *stream) {
    DDPB_REQUIRE(bucket && p && n >= 0 && step >= 1 && (momentum == 0.0 || buf), "ddpb_fused_avg_sgd: bad arguments");
    if (n == 0) return 0;
    avg_sgd_kernel<<<stream_grid(n, 1), 256, 0, (cudaStream_t)stream>>>(bucket, p, buf, n, grad_scale, (float)lr,
                                                                         (float)momentum, step == 1, nullptr);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_fused_avg_sgd_dev(const float *bucket, float *p, float *buf, int64_t n, float grad_scale, double lr,
                                    double momentum, ddpb_opt_state *state, int advance, void *stream) {
    DDPB_REQUIRE(bucket && p && n >= 0 && state && (momentum == 0.0 || buf), "ddpb_fused_avg_sgd_dev: bad arguments");
    if (n > 0) {
        avg_sgd_kernel<<<stream_grid(n, 1), 256, 0, (cudaStream_t)stream>>>(bucket, p, buf, n, grad_scale, (float)lr,
                                                                             (float)momentum, 0, state);
        ddpb::count_launch();
    }
    if (advance) {
        advance_state_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state);
        ddpb::count_launch();
    }
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

namespace {
template <int W>
int launch_bucket(const ddpb_bucket_job &job, cudaStream_t stream) {
    const long long n4 = job.n >> 2, per = (n4 + W - 1) / W;
    int grid = stream_grid(per * 4, 4);
    const int cap = (job.reserved >> 8) & 0xffff;            // diagnostic: cap the grid
    if (cap && grid > cap) grid = cap;
    bucket_reduce_adam_kernel<W><<<grid, 256, 0, stream>>>(job);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
}  // namespace

DDPB_API int ddpb_bucket_reduce_adam(const ddpb_bucket_job *job, void *stream) {
    DDPB_REQUIRE(job && job->m && job->v && job->bias_table && job->state && job->bias_table_n >= 1,
                 "ddpb_bucket_reduce_adam: bad arguments");
    DDPB_REQUIRE(job->world >= 2 && job->world <= DDPB_MAX_PEERS && job->rank >= 0 && job->rank < job->world,
                 "ddpb_bucket_reduce_adam: world %d / rank %d out of range (2..%d ranks)", job->world, job->rank, DDPB_MAX_PEERS);
    DDPB_REQUIRE(job->bucket_index >= 0 && job->bucket_index < DDPB_BUCKET_MAX, "ddpb_bucket_reduce_adam: bucket index %d >= %d",
                 job->bucket_index, DDPB_BUCKET_MAX);
    DDPB_REQUIRE(job->n >= 0 && job->start >= 0 && job->start % 4 == 0 && job->grads_off % 16 == 0 && job->params_off % 16 == 0 &&
                     (((uintptr_t)job->m | (uintptr_t)job->v) & 15u) == 0,
                 "ddpb_bucket_reduce_adam: the bucket must start on a 16-byte boundary of 16-byte aligned buffers");
    DDPB_REQUIRE(job->ticket >= 1 && job->ticket < 0x7fffffffLL, "ddpb_bucket_reduce_adam: ticket out of range");
    for (int q = 0; q < job->world; ++q)
        DDPB_REQUIRE(job->window[q] != nullptr, "ddpb_bucket_reduce_adam: window of rank %d is not mapped", q);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = 0;
    if (job->n > 0) {
        switch (job->world) {
            case 2: rc = launch_bucket<2>(*job, s); break;
            case 3: rc = launch_bucket<3>(*job, s); break;
            case 4: rc = launch_bucket<4>(*job, s); break;
            case 5: rc = launch_bucket<5>(*job, s); break;
            case 6: rc = launch_bucket<6>(*job, s); break;
            case 7: rc = launch_bucket<7>(*job, s); break;
            default: rc = launch_bucket<8>(*job, s); break;
        }
        if (rc) return rc;
    }
    if (job->advance) {
        advance_state_kernel<<<1, 1, 0, s>>>(job->state);
        ddpb::count_launch();
        DDPB_CUDA(cudaGetLastError());
    }
    return 0;
}
