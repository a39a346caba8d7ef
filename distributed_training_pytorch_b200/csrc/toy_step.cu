// The fused training step of the reference's toy model (H = 10, 371 parameters).
//
// What the reference runs per model and per iteration (demo.py:104-111):
//   model(data)            ToyModel.forward, toy_model_and_data.py:24-25   (9 library kernels)
//   loss(out, target)      nn.MSELoss, demo.py:86,106                      (2-3)
//   loss.backward()        autograd + DDP reducer hooks, demo.py:107       (~45)
//   optimizer.step()       optim.Adam foreach path, demo.py:110            (~10)
// Here it is ONE launch for all models of the process: a thread-block cluster of 8 CTAs
// per model (blockIdx.y = model).  At this size (1 MFLOP, 10 KB) nothing but latency
// matters, so the design minimises the dependent-instruction chain:
//
//   * CTA c owns samples [32c, 32c+32) of the batch, one sample per lane.  Its four warps
//     run the same per-sample forward/backward chain redundantly (one warp per SM
//     sub-partition, so redundancy costs no time) and each writes a quarter of the
//     per-sample table (activations + pre-activation gradients, [84 rows][32 samples]).
//   * The weight gradient dW = dZ^T A is then a [371 x 32] x [32] contraction per CTA:
//     thread t computes parameters t, t+128, t+256 with conflict-free LDS.128 rows.
//   * The eight per-CTA partials meet through distributed shared memory: CTA c sums
//     slice [47c, 47c+47) over the cluster in a fixed order (deterministic), and - in
//     STEP mode - applies Adam/SGD to that slice with operands prefetched at kernel
//     start.  Element 371 of the slice table carries sum (o-t)^2, i.e. the loss.
//
// Arithmetic order of the per-sample chain (fmaf from the bias, k ascending) and of the
// optimizer (torch's op order, explicit round-to-nearest intrinsics, no contraction) is
// the one oracle/ddp_oracle.c uses, so forward values and the update are bit-identical
// given identical gradients; only the batch-sum order differs.
#include "common.cuh"
#include <cooperative_groups.h>
#include <cstddef>

namespace cg = cooperative_groups;

namespace {

constexpr int H = DDPB_TOY_HIDDEN;
constexpr int P = DDPB_TOY_PARAMS;
constexpr int kCluster = 8;
constexpr int kThreads = 128;
constexpr int kLanes = 32;                       // samples per CTA and pass
constexpr int kChunk = kCluster * kLanes;        // samples per cluster pass
constexpr int kWRow = 12;                        // padded weight row, 3 x float4
constexpr int kRowStride = 36;                   // floats per table row: 144 B keeps LDS.128 conflict-free
constexpr float kSlope = 0.01f;                  // nn.LeakyReLU() default

// offsets inside the flat parameter vector (module.parameters() order)
constexpr int OW0 = 0, OB0 = 2 * H;
constexpr int OW1 = OB0 + H;
constexpr int kHidden = H * H + H;               // one hidden layer: W then b
constexpr int OW4 = OW1 + 3 * kHidden, OB4 = OW4 + H;
static_assert(OB4 + 1 == P, "flat layout must add up to the parameter count");

// rows of the per-sample table
constexpr int R_IN = 0;                          // x0, x1
constexpr int R_ACT = 2;                         // act_0 .. act_3 (H rows each)
constexpr int R_ONE = R_ACT + 4 * H;             // ones (bias gradients)
constexpr int R_DZ = R_ONE + 1;                  // dz_0 .. dz_3 (H rows each), then dz_4
constexpr int kRows = R_DZ + 4 * H + 1;
constexpr int kSlice = (P + 1 + kCluster - 1) / kCluster;   // 47 values per CTA (371 grads + loss sum)

struct JobList {
    ddpb_toy_job jobs[DDPB_MAX_JOBS];
};

struct __align__(16) Smem {
    float w0[H][4];            // layer 0: [j][k<2]
    float wh[3][H][kWRow];     // layers 1..3
    float w4[kWRow];           // layer 4
    float b[4][kWRow];         // biases of layers 0..3
    float b4[4];
    float table[kRows][kRowStride];
    float partial[kSlice * kCluster];
};

__device__ __forceinline__ float lrelu(float z) { return z > 0.f ? z : z * kSlope; }

// Store per-sample values; the four redundant warps split the rows between them.  The
// split is resolved at compile time inside a warp-uniform switch: selecting rows with a
// run-time predicate makes nvcc index the register arrays dynamically (local memory).
template <int ROW, int W>
__device__ __forceinline__ void put_part(Smem &S, int lane, float val) {
    if ((ROW & 3) == W) S.table[ROW][lane] = val;
}

template <int ROW>
__device__ __forceinline__ void put(Smem &S, int warp, int lane, float val) {
    switch (warp) {
        case 0: put_part<ROW, 0>(S, lane, val); break;
        case 1: put_part<ROW, 1>(S, lane, val); break;
        case 2: put_part<ROW, 2>(S, lane, val); break;
        default: put_part<ROW, 3>(S, lane, val); break;
    }
}

template <int ROW0, int W>
__device__ __forceinline__ void put_vec_part(Smem &S, int lane, const float (&val)[H]) {
#pragma unroll
    for (int j = 0; j < H; ++j)
        if (((ROW0 + j) & 3) == W) S.table[ROW0 + j][lane] = val[j];
}

template <int ROW0>
__device__ __forceinline__ void put_vec(Smem &S, int warp, int lane, const float (&val)[H]) {
    switch (warp) {
        case 0: put_vec_part<ROW0, 0>(S, lane, val); break;
        case 1: put_vec_part<ROW0, 1>(S, lane, val); break;
        case 2: put_vec_part<ROW0, 2>(S, lane, val); break;
        default: put_vec_part<ROW0, 3>(S, lane, val); break;
    }
}

__device__ __forceinline__ void load_row(const float *row, float (&w)[kWRow]) {
    const float4 *q = reinterpret_cast<const float4 *>(row);
    float4 a = q[0], b = q[1], c = q[2];
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w;
}

// Hidden layer forward: nxt = lrelu(W cur + b); mask bit j = (pre-activation j > 0).
__device__ __forceinline__ void hidden_fwd(const Smem &S, int l, const float (&cur)[H], float (&nxt)[H], unsigned &mask) {
    mask = 0;
#pragma unroll
    for (int j = 0; j < H; ++j) {
        float w[kWRow];
        load_row(S.wh[l][j], w);
        float acc = S.b[l + 1][j];
#pragma unroll
        for (int k = 0; k < H; ++k) acc = fmaf(w[k], cur[k], acc);
        mask |= (acc > 0.f ? 1u : 0u) << j;
        nxt[j] = lrelu(acc);
    }
}

// Hidden layer backward: dprev = leaky'(mask_prev) * (W^T dz).
__device__ __forceinline__ void hidden_bwd(const Smem &S, int l, const float (&dz)[H], unsigned mask_prev, float (&dprev)[H]) {
#pragma unroll
    for (int k = 0; k < H; ++k) dprev[k] = 0.f;
#pragma unroll
    for (int j = 0; j < H; ++j) {
        float w[kWRow];
        load_row(S.wh[l][j], w);
#pragma unroll
        for (int k = 0; k < H; ++k) dprev[k] = fmaf(dz[j], w[k], dprev[k]);
    }
#pragma unroll
    for (int k = 0; k < H; ++k) dprev[k] = ((mask_prev >> k) & 1u) ? dprev[k] : dprev[k] * kSlope;
}

// Which table rows produce the gradient of flat parameter i:  grad_i = sum_b dz_row[b] * a_row[b].
__device__ __forceinline__ void rows_of_param(int i, int &row_dz, int &row_a) {
    if (i < OW1) {                      // layer 0
        if (i < OB0) { row_dz = R_DZ + i / 2; row_a = R_IN + (i & 1); }
        else         { row_dz = R_DZ + (i - OB0); row_a = R_ONE; }
    } else if (i < OW4) {               // layers 1..3
        int q = i - OW1, l = q / kHidden, r = q - l * kHidden;
        if (r < H * H) { row_dz = R_DZ + (l + 1) * H + r / H; row_a = R_ACT + l * H + r % H; }
        else           { row_dz = R_DZ + (l + 1) * H + (r - H * H); row_a = R_ONE; }
    } else {                            // layer 4
        row_dz = R_DZ + 4 * H;
        row_a = (i < OB4) ? R_ACT + 3 * H + (i - OW4) : R_ONE;
    }
}

__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kThreads, 1)
toy_step_kernel(const __grid_constant__ JobList jl) {
    __shared__ Smem S;
    const ddpb_toy_job &job = jl.jobs[blockIdx.y];
    cg::cluster_group cluster = cg::this_cluster();
    const int cta = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = job.B, mode = job.mode;

    // ---- 0. early loads: counters, and the optimizer operands of the slice this thread owns
    const long long seq = job.state ? job.state->seq : 0;
    const int own = cta * kSlice + tid;                 // index into the 372-entry slice table
    const bool owner = tid < kSlice && own <= P;
    float p_old = 0.f, m_old = 0.f, v_old = 0.f;
    long long step = 0;
    double b1pow = 1.0, b2pow = 1.0;
    if (mode == DDPB_MODE_STEP) {
        step = job.state->step; b1pow = job.state->beta1_pow; b2pow = job.state->beta2_pow;
        if (owner && own < P) {
            p_old = job.params[own];
            m_old = job.m[own];
            if (job.opt == DDPB_OPT_ADAM) v_old = job.v[own];
        }
    }

    // ---- 1. stage the weights (16-byte aligned, zero padded rows)
    {
        float *flat = reinterpret_cast<float *>(&S);
        constexpr int kWeightFloats = (int)(offsetof(Smem, table) / sizeof(float));
        for (int i = tid; i < kWeightFloats; i += kThreads) flat[i] = 0.f;
        if (tid < kRowStride) S.table[R_ONE][tid] = 1.f;
        for (int i = tid; i < kSlice * kCluster; i += kThreads) S.partial[i] = 0.f;
        __syncthreads();
        const float *prm = job.params;
        for (int i = tid; i < P; i += kThreads) {
            float val = prm[i];
            if (i < OB0) S.w0[i / 2][i & 1] = val;
            else if (i < OW1) S.b[0][i - OB0] = val;
            else if (i < OW4) {
                int q = i - OW1, l = q / kHidden, r = q - l * kHidden;
                if (r < H * H) S.wh[l][r / H][r % H] = val;
                else S.b[l + 1][r - H * H] = val;
            } else if (i < OB4) S.w4[i - OW4] = val;
            else S.b4[0] = val;
        }
        __syncthreads();
    }

    // gradient rows of the (up to three) parameters this thread accumulates
    int row_dz[3], row_a[3];
    float gacc[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        int i = tid + r * kThreads;
        row_dz[r] = R_ONE; row_a[r] = R_ONE;
        if (i < P) rows_of_param(i, row_dz[r], row_a[r]);
    }
    float sq_acc = 0.f;
    const float norm = (float)(2.0 / (double)B);        // mse_loss_backward: 2/numel
    const int32_t *idx_row = job.idx ? job.idx + (seq % job.idx_rows) * (long long)job.idx_stride : nullptr;

    for (int chunk0 = 0; chunk0 < B; chunk0 += kChunk) {
        const int first = chunk0 + cta * kLanes;
        const bool cta_has_samples = first < B;          // uniform per CTA
        if (cta_has_samples) {
            const int b = first + lane;
            const bool valid = b < B;
            float x0 = 0.f, x1 = 0.f, y = 0.f;
            if (valid) {
                const long long s = idx_row ? (long long)idx_row[b] : (long long)b;
                const float2 xv = *reinterpret_cast<const float2 *>(job.x + 2 * s);
                x0 = xv.x; x1 = xv.y;
                if (mode != DDPB_MODE_FWD) y = job.y[s];
            }
            // ---- forward, one sample per lane
            float a0[H], a1[H], a2[H], a3[H];
            unsigned k0 = 0, k1, k2, k3;
#pragma unroll
            for (int j = 0; j < H; ++j) {
                float acc = S.b[0][j];
                acc = fmaf(S.w0[j][0], x0, acc);
                acc = fmaf(S.w0[j][1], x1, acc);
                k0 |= (acc > 0.f ? 1u : 0u) << j;
                a0[j] = lrelu(acc);
            }
            hidden_fwd(S, 0, a0, a1, k1);
            asm volatile("" ::: "memory");
            hidden_fwd(S, 1, a1, a2, k2);
            asm volatile("" ::: "memory");
            hidden_fwd(S, 2, a2, a3, k3);
            float out = S.b4[0];
            {
                float w[kWRow];
                load_row(S.w4, w);
#pragma unroll
                for (int k = 0; k < H; ++k) out = fmaf(w[k], a3[k], out);
            }
            if (mode == DDPB_MODE_FWD) {
                if (valid && warp == 0) job.out[b] = out;
                continue;
            }
            put<R_IN + 0>(S, warp, lane, x0);
            put<R_IN + 1>(S, warp, lane, x1);
            put_vec<R_ACT + 0 * H>(S, warp, lane, a0);
            put_vec<R_ACT + 1 * H>(S, warp, lane, a1);
            put_vec<R_ACT + 2 * H>(S, warp, lane, a2);
            put_vec<R_ACT + 3 * H>(S, warp, lane, a3);

            // ---- loss and backward.  The barrier stops nvcc from keeping all 340 forward
            // weights live in registers for reuse here (it spills otherwise); reloading a
            // 48-byte row from shared memory is cheaper than a local-memory round trip.
            asm volatile("" ::: "memory");
            const float diff = valid ? out - y : 0.f;
            if (warp == 0) {
                float sq = diff * diff;
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, off);
                sq_acc += sq;
            }
            const float dz4 = norm * diff;
            put<R_DZ + 4 * H>(S, warp, lane, dz4);
            float d3[H], d2[H], d1[H], d0[H];
            {
                float w[kWRow];
                load_row(S.w4, w);
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    float g = fmaf(dz4, w[k], 0.f);
                    d3[k] = ((k3 >> k) & 1u) ? g : g * kSlope;
                }
            }
            put_vec<R_DZ + 3 * H>(S, warp, lane, d3);
            hidden_bwd(S, 2, d3, k2, d2);
            put_vec<R_DZ + 2 * H>(S, warp, lane, d2);
            asm volatile("" ::: "memory");
            hidden_bwd(S, 1, d2, k1, d1);
            put_vec<R_DZ + 1 * H>(S, warp, lane, d1);
            asm volatile("" ::: "memory");
            hidden_bwd(S, 0, d1, k0, d0);
            put_vec<R_DZ + 0 * H>(S, warp, lane, d0);
        }
        if (mode == DDPB_MODE_FWD) continue;
        __syncthreads();
        // ---- weight gradients of this CTA's 32 samples
        if (cta_has_samples) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                if (tid + r * kThreads < P) {
                    const float4 *pd = reinterpret_cast<const float4 *>(S.table[row_dz[r]]);
                    const float4 *pa = reinterpret_cast<const float4 *>(S.table[row_a[r]]);
                    float s = 0.f;
#pragma unroll
                    for (int c = 0; c < kLanes / 4; ++c) {
                        float4 d = pd[c], a = pa[c];
                        s = fmaf(d.x, a.x, s); s = fmaf(d.y, a.y, s);
                        s = fmaf(d.z, a.z, s); s = fmaf(d.w, a.w, s);
                    }
                    gacc[r] += s;
                }
            }
        }
        __syncthreads();
    }
    if (mode == DDPB_MODE_FWD) return;

    // ---- 2. publish the partials, reduce my slice across the cluster
#pragma unroll
    for (int r = 0; r < 3; ++r)
        if (tid + r * kThreads < P) S.partial[tid + r * kThreads] = gacc[r];
    if (tid == 0) S.partial[P] = sq_acc;
    cluster.sync();

    if (owner) {
        float total = 0.f;
#pragma unroll
        for (int c = 0; c < kCluster; ++c) total += cluster.map_shared_rank(S.partial, c)[own];
        if (own == P) {
            job.grads[P] = total;                                        // sum (o-t)^2 over the local batch
            if (job.loss_ring) job.loss_ring[seq % job.loss_ring_n] = total / (float)B;
        } else {
            job.grads[own] = total;
            if (mode == DDPB_MODE_STEP) {
                const float g = total;
                if (job.opt == DDPB_OPT_ADAM) {
                    // torch:optim/adam.py:457 (lerp), :476 (mul, addcmul), :531-536 (bias corrections
                    // in double), :545 (denom), :547 (addcdiv)
                    const double bc1 = 1.0 - b1pow * job.beta1;
                    const double bc2 = 1.0 - b2pow * job.beta2;
                    const float neg_step = (float)(-(job.lr / bc1));
                    const float bc2_sqrt = (float)sqrt(bc2);
                    const float w1 = (float)(1.0 - job.beta1), b2 = (float)job.beta2;
                    const float w2 = (float)(1.0 - job.beta2), eps = (float)job.eps;
                    const float m_new = __fmaf_rn(w1, __fsub_rn(g, m_old), m_old);
                    const float v_new = __fadd_rn(__fmul_rn(v_old, b2), __fmul_rn(__fmul_rn(w2, g), g));
                    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v_new), bc2_sqrt), eps);
                    job.params[own] = __fadd_rn(p_old, __fmul_rn(neg_step, __fdiv_rn(m_new, denom)));
                    job.m[own] = m_new;
                    job.v[own] = v_new;
                } else {
                    // torch:optim/sgd.py _single_tensor_sgd: buf = g (first step) | momentum*buf + g
                    float d = g;
                    if (job.momentum != 0.0) {
                        const float mo = (float)job.momentum;
                        d = (step == 0) ? g : __fadd_rn(__fmul_rn(m_old, mo), g);
                        job.m[own] = d;
                    }
                    job.params[own] = __fsub_rn(p_old, __fmul_rn((float)job.lr, d));
                }
            }
        }
    }
    cluster.sync();   // nobody leaves while a peer may still read its partials

    // ---- 3. advance the device-resident counters
    if (cta == 0 && tid == 0 && job.state) {
        job.state->seq = seq + 1;
        if (mode == DDPB_MODE_STEP) {
            job.state->step = step + 1;
            job.state->beta1_pow = b1pow * job.beta1;
            job.state->beta2_pow = b2pow * job.beta2;
        }
    }
}

}  // namespace

DDPB_API int ddpb_toy_step(const ddpb_toy_job *jobs, int njobs, void *stream) {
    DDPB_REQUIRE(jobs != nullptr && njobs >= 1 && njobs <= DDPB_MAX_JOBS,
                 "ddpb_toy_step: njobs must be 1..%d, got %d", DDPB_MAX_JOBS, njobs);
    JobList jl;
    for (int j = 0; j < njobs; ++j) {
        const ddpb_toy_job &q = jobs[j];
        DDPB_REQUIRE(q.params && q.x, "ddpb_toy_step: job %d: params and x are required", j);
        DDPB_REQUIRE(q.B >= 1 && q.B <= 65536, "ddpb_toy_step: job %d: batch %d outside 1..65536", j, q.B);
        DDPB_REQUIRE(q.mode == DDPB_MODE_FWD || q.mode == DDPB_MODE_GRAD || q.mode == DDPB_MODE_STEP,
                     "ddpb_toy_step: job %d: unknown mode %d", j, q.mode);
        if (q.mode == DDPB_MODE_FWD) DDPB_REQUIRE(q.out, "ddpb_toy_step: job %d: FWD needs out", j);
        else DDPB_REQUIRE(q.y && q.grads && q.state, "ddpb_toy_step: job %d: y, grads and state are required", j);
        if (q.mode == DDPB_MODE_STEP) {
            DDPB_REQUIRE(q.opt == DDPB_OPT_ADAM || q.opt == DDPB_OPT_SGD, "ddpb_toy_step: job %d: unknown optimizer", j);
            DDPB_REQUIRE(q.m && (q.opt == DDPB_OPT_SGD || q.v), "ddpb_toy_step: job %d: optimizer state missing", j);
        }
        if (q.loss_ring) DDPB_REQUIRE(q.loss_ring_n >= 1, "ddpb_toy_step: job %d: empty loss ring", j);
        if (q.idx) DDPB_REQUIRE(q.idx_rows >= 1 && q.idx_stride >= q.B, "ddpb_toy_step: job %d: index rows too short", j);
        jl.jobs[j] = q;
    }
    toy_step_kernel<<<dim3(kCluster, njobs, 1), kThreads, 0, (cudaStream_t)stream>>>(jl);
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
