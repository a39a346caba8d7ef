// A pipeline SEGMENT of the toy MLP: up to three Linear(+LeakyReLU) layers of width <= 10, run
// forward only, or forward + backward (+ fused Adam/SGD) given either targets (the segment ends
// in the MSE loss) or the upstream gradient of its output.
//
// Reference: demo_one_model_multi_gpu.py:17-42 splits the model in two stages (layers0: 2->10->10
// on dev0, layers1: 10->10->10->1 on dev1) and moves the [B,10] activation with .to(dev1)
// (cudaMemcpyPeer) and its gradient back through autograd.  Here a stage is one kernel and the
// transfer is part of it: `out` / `d_in` may point into the OTHER GPU's memory (peer access
// enabled with ddpb_enable_peer), so the activation / gradient is stored straight over NVLink by
// the kernel that produces it, ordered against the consuming kernel by an event.  Stage 0 does
// not keep its activations between its forward and backward launch: it recomputes them from the
// 2-wide input (cheaper than a round trip through HBM).
//
// This is the secondary configuration (BASELINE.json configs[3]), so the kernel is deliberately
// simple: one CTA, one sample per thread, register-resident chain, gradients via a shared
// [rows][256 samples] table with one parameter per thread, same arithmetic order as
// oracle/ddp_oracle.c (forward values and the per-sample backward are bit-identical).
#include "common.cuh"

namespace {

constexpr int kMaxW = DDPB_SEG_MAX_WIDTH;      // 10
constexpr int kMaxL = DDPB_SEG_MAX_LAYERS;     // 3
constexpr int kThreads = 256;
constexpr int kRowStride = kThreads + 4;       // floats; 16-byte aligned rows, shifted banks
constexpr int kRows = 2 * kMaxL * kMaxW + 1;   // layer inputs + dz rows + ones
constexpr int kWRow = 12;
constexpr float kSlope = 0.01f;

struct SegSmem {
    float w[kMaxL][kMaxW][kWRow];
    float b[kMaxL][kWRow];
    float red[8];
    float table[kRows][kRowStride];
};

__device__ __forceinline__ float lrelu(float z) { return z > 0.f ? z : z * kSlope; }

__global__ void __launch_bounds__(kThreads, 1) segment_kernel(const __grid_constant__ ddpb_seg_job job) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SegSmem &S = *reinterpret_cast<SegSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nl = job.nl, B = job.B, mode = job.mode;
    int off[kMaxL + 1];
    off[0] = 0;
#pragma unroll
    for (int l = 0; l < kMaxL; ++l) off[l + 1] = off[l] + (l < nl ? job.dims[l + 1] * job.dims[l] + job.dims[l + 1] : 0);
    const int P = off[kMaxL];

    // optimizer operands of the parameter this thread owns (one parameter per thread)
    float p_old = 0.f, m_old = 0.f, v_old = 0.f;
    float2 bias = make_float2(0.f, 1.f);
    long long step = 0, seq = 0;
    if (mode != DDPB_MODE_FWD) seq = job.state->seq;
    if (mode == DDPB_MODE_STEP) {
        step = job.state->step;
        if (tid < P) {
            p_old = job.params[tid];
            m_old = job.m[tid];
            if (job.opt == DDPB_OPT_ADAM) {
                v_old = job.v[tid];
                bias = reinterpret_cast<const float2 *>(job.bias_table)[step < job.bias_table_n ? step : job.bias_table_n - 1];
            }
        }
    }
    // stage the weights
    for (int i = tid; i < kMaxL * kMaxW * kWRow + kMaxL * kWRow; i += kThreads) (&S.w[0][0][0])[i] = 0.f;
    for (int i = tid; i < kRowStride; i += kThreads) S.table[kRows - 1][i] = 1.f;
    __syncthreads();
    for (int i = tid; i < P; i += kThreads) {
        int l = 0;
        while (l + 1 < nl && i >= off[l + 1]) ++l;
        const int r = i - off[l], di = job.dims[l], dout = job.dims[l + 1];
        if (r < dout * di) S.w[l][r / di][r % di] = job.params[i];
        else S.b[l][r - dout * di] = job.params[i];
    }
    __syncthreads();

    // which table rows make up this thread's parameter gradient
    int row_dz = kRows - 1, row_a = kRows - 1;
    if (tid < P) {
        int l = 0;
        while (l + 1 < nl && tid >= off[l + 1]) ++l;
        const int r = tid - off[l], di = job.dims[l], dout = job.dims[l + 1];
        if (r < dout * di) { row_dz = (kMaxL + l) * kMaxW + r / di; row_a = l * kMaxW + r % di; }
        else row_dz = (kMaxL + l) * kMaxW + (r - dout * di);
    }
    float gacc = 0.f, sq_acc = 0.f;
    const int d0 = job.dims[0], dlast = job.dims[nl];
    const float norm = __fdiv_rn(2.0f, (float)((long long)B * dlast));     // MSE over all B*dout elements

    for (int chunk0 = 0; chunk0 < B; chunk0 += kThreads) {
        const int b = chunk0 + tid;
        const bool valid = b < B;
        float a[kMaxL + 1][kMaxW];
        unsigned mask[kMaxL];
#pragma unroll
        for (int k = 0; k < kMaxW; ++k) a[0][k] = (valid && k < d0) ? job.x[(long long)b * d0 + k] : 0.f;
#pragma unroll
        for (int l = 0; l < kMaxL; ++l) {
            mask[l] = 0u;
#pragma unroll
            for (int j = 0; j < kMaxW; ++j) a[l + 1][j] = 0.f;
            if (l < nl) {
                const bool act = (l + 1 < nl) || job.act_last;
#pragma unroll
                for (int j = 0; j < kMaxW; ++j) {
                    if (j < job.dims[l + 1]) {
                        float acc = S.b[l][j];
#pragma unroll
                        for (int k = 0; k < kMaxW; ++k)
                            if (k < job.dims[l]) acc = fmaf(S.w[l][j][k], a[l][k], acc);
                        mask[l] |= (acc > 0.f ? 1u : 0u) << j;
                        a[l + 1][j] = act ? lrelu(acc) : acc;
                    }
                }
            }
        }
        // the segment's output is a[nl] (layers beyond nl do not exist)
        float outv[kMaxW];
#pragma unroll
        for (int j = 0; j < kMaxW; ++j) outv[j] = nl == 1 ? a[1][j] : (nl == 2 ? a[2][j] : a[3][j]);
        if (mode == DDPB_MODE_FWD) {
            if (valid)
#pragma unroll
                for (int j = 0; j < kMaxW; ++j)
                    if (j < dlast) job.out[(long long)b * dlast + j] = outv[j];
            continue;
        }
        // ---- gradient of the output
        float dz[kMaxW];
#pragma unroll
        for (int j = 0; j < kMaxW; ++j) {
            dz[j] = 0.f;
            if (valid && j < dlast) {
                if (job.has_loss) {
                    const float diff = outv[j] - job.y[(long long)b * dlast + j];
                    sq_acc = fmaf(diff, diff, sq_acc);
                    dz[j] = norm * diff;
                } else {
                    dz[j] = job.d_out[(long long)b * dlast + j];
                }
            }
        }
        // ---- backward through the layers, last first; table rows are written on the way
#pragma unroll
        for (int l = kMaxL - 1; l >= 0; --l) {
            if (l < nl) {
                const bool act = (l + 1 < nl) || job.act_last;
#pragma unroll
                for (int j = 0; j < kMaxW; ++j) {
                    if (act) dz[j] = ((mask[l] >> j) & 1u) ? dz[j] : dz[j] * kSlope;
                    S.table[(kMaxL + l) * kMaxW + j][tid] = dz[j];
                    S.table[l * kMaxW + j][tid] = a[l][j];
                }
                if (l > 0 || job.d_in) {
                    float din[kMaxW];
#pragma unroll
                    for (int k = 0; k < kMaxW; ++k) {
                        float s = 0.f;
#pragma unroll
                        for (int j = 0; j < kMaxW; ++j)
                            if (j < job.dims[l + 1]) s = fmaf(dz[j], S.w[l][j][k], s);
                        din[k] = (k < job.dims[l]) ? s : 0.f;
                    }
                    if (l == 0) {
                        if (valid)
#pragma unroll
                            for (int k = 0; k < kMaxW; ++k)
                                if (k < d0) job.d_in[(long long)b * d0 + k] = din[k];
                    } else {
#pragma unroll
                        for (int k = 0; k < kMaxW; ++k) dz[k] = din[k];
                    }
                }
            }
        }
        __syncthreads();
        if (tid < P) {
            const float4 *pd = reinterpret_cast<const float4 *>(S.table[row_dz]);
            const float4 *pa = reinterpret_cast<const float4 *>(S.table[row_a]);
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 8
            for (int c = 0; c < kThreads / 4; ++c) {
                const float4 d = pd[c], q = pa[c];
                s0 = fmaf(d.x, q.x, s0); s1 = fmaf(d.y, q.y, s1);
                s2 = fmaf(d.z, q.z, s2); s3 = fmaf(d.w, q.w, s3);
            }
            gacc += (s0 + s1) + (s2 + s3);
        }
        __syncthreads();
    }
    if (mode == DDPB_MODE_FWD) return;

    // ---- loss (segments that end in it)
    if (job.has_loss) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq_acc += __shfl_xor_sync(0xffffffffu, sq_acc, o);
        if (lane == 0) S.red[warp] = sq_acc;
        __syncthreads();
        if (tid == 0) {
            float total = 0.f;
            for (int w = 0; w < kThreads / 32; ++w) total += S.red[w];
            job.grads[P] = total;
            if (job.loss_ring) job.loss_ring[(unsigned)seq % (unsigned)job.loss_ring_n] = __fdiv_rn(total, (float)((long long)B * dlast));
        }
    }
    if (tid < P) {
        const float g = gacc;
        job.grads[tid] = g;
        if (mode == DDPB_MODE_STEP) {
            if (job.opt == DDPB_OPT_ADAM) {
                const float w1 = (float)(1.0 - job.beta1), b2 = (float)job.beta2;
                const float w2 = (float)(1.0 - job.beta2), eps = (float)job.eps;
                const float m_new = __fmaf_rn(w1, __fsub_rn(g, m_old), m_old);
                const float v_new = __fadd_rn(__fmul_rn(v_old, b2), __fmul_rn(__fmul_rn(w2, g), g));
                const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v_new), bias.y), eps);
                job.params[tid] = __fadd_rn(p_old, __fmul_rn(bias.x, __fdiv_rn(m_new, denom)));
                job.m[tid] = m_new;
                job.v[tid] = v_new;
            } else {
                float d = g;
                if (job.momentum != 0.0) {
                    d = (step == 0) ? g : __fadd_rn(__fmul_rn(m_old, (float)job.momentum), g);
                    job.m[tid] = d;
                }
                job.params[tid] = __fsub_rn(p_old, __fmul_rn((float)job.lr, d));
            }
        }
    }
    if (tid == 0) {
        job.state->seq = seq + 1;
        if (mode == DDPB_MODE_STEP) job.state->step = step + 1;
    }
}

}  // namespace

DDPB_API int ddpb_segment_step(const ddpb_seg_job *job, void *stream) {
    DDPB_REQUIRE(job != nullptr, "ddpb_segment_step: null job");
    const ddpb_seg_job &q = *job;
    DDPB_REQUIRE(q.nl >= 1 && q.nl <= DDPB_SEG_MAX_LAYERS, "ddpb_segment_step: %d layers, supported 1..%d", q.nl, DDPB_SEG_MAX_LAYERS);
    int P = 0;
    for (int l = 0; l <= q.nl; ++l)
        DDPB_REQUIRE(q.dims[l] >= 1 && q.dims[l] <= DDPB_SEG_MAX_WIDTH, "ddpb_segment_step: width %d outside 1..%d", q.dims[l], DDPB_SEG_MAX_WIDTH);
    for (int l = 0; l < q.nl; ++l) P += q.dims[l + 1] * q.dims[l] + q.dims[l + 1];
    DDPB_REQUIRE(P <= 255, "ddpb_segment_step: %d parameters, one CTA handles at most 255", P);
    DDPB_REQUIRE(q.params && q.x && q.B >= 1 && q.B <= (1 << 20), "ddpb_segment_step: params, x and 1 <= B <= 2^20 are required");
    DDPB_REQUIRE(q.mode == DDPB_MODE_FWD || q.mode == DDPB_MODE_GRAD || q.mode == DDPB_MODE_STEP, "ddpb_segment_step: unknown mode %d", q.mode);
    if (q.mode == DDPB_MODE_FWD) DDPB_REQUIRE(q.out, "ddpb_segment_step: FWD needs out");
    else {
        DDPB_REQUIRE(q.grads && q.state, "ddpb_segment_step: grads and state are required");
        DDPB_REQUIRE(q.has_loss ? q.y != nullptr : q.d_out != nullptr, "ddpb_segment_step: needs targets (has_loss) or the upstream gradient");
        if (q.loss_ring) DDPB_REQUIRE(q.loss_ring_n >= 1, "ddpb_segment_step: empty loss ring");
    }
    if (q.mode == DDPB_MODE_STEP)
        DDPB_REQUIRE(q.m && (q.opt == DDPB_OPT_SGD || (q.opt == DDPB_OPT_ADAM && q.v && q.bias_table && q.bias_table_n >= 1)),
                     "ddpb_segment_step: optimizer state missing");
    static bool configured[64] = {};
    int dev = 0;
    DDPB_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        DDPB_CUDA(cudaFuncSetAttribute(segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem)));
        configured[dev] = true;
    }
    segment_kernel<<<1, kThreads, sizeof(SegSmem), (cudaStream_t)stream>>>(q);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
