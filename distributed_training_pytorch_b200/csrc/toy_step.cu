// The fused training step of the reference's toy model (H = 10, 371 parameters).
//
// What the reference runs per model and per iteration (demo.py:104-111):
//   model(data)            ToyModel.forward, toy_model_and_data.py:24-25   (9 library kernels)
//   loss(out, target)      nn.MSELoss, demo.py:86,106                      (2-3)
//   loss.backward()        autograd + DDP reducer hooks, demo.py:107       (~45)
//   optimizer.step()       optim.Adam foreach path, demo.py:110            (~10)
// Here it is ONE launch for all models of the process: a thread-block cluster of 8 CTAs
// per model (blockIdx.y = model).  At this size (1 MFLOP, 10 KB) only latency matters:
// ncu of the first version (profiles/r01_v0_*) showed one warp per scheduler crawling
// through a 1900-instruction dependent chain (IPC 0.14).  This version shortens the chain:
//
//   * CTA c owns samples [32c, 32c+32) of the batch, one sample per lane.  Its four compute
//     warps split each layer's OUTPUT neurons (warp w owns neurons w, w+4, w+8): a layer is
//     <= 30 FMAs per warp, the activations meet in a shared-memory table
//     ([84 rows][32 samples]: inputs, activations, pre-activation gradients), one named
//     barrier per layer.  The same table is the operand of the weight-gradient contraction
//     dW = dZ^T A ([371 x 32] x [32] per CTA), thread t computing parameters t, t+128, t+256
//     with conflict-free LDS.128 rows and four independent partial sums.
//   * Two more warps per CTA are the optimizer: they prefetch p, m, v and the bias-correction
//     scalars of their 47-parameter slice at kernel start, sleep on the cluster barrier, then
//     sum the eight per-CTA partials through distributed shared memory in a fixed order
//     (deterministic) and apply Adam/SGD.  Element 371 of the slice table carries
//     sum (o-t)^2, i.e. the loss.
//
// Arithmetic order of the per-sample chain (fmaf from the bias, k ascending) and of the
// optimizer (torch's op order, explicit round-to-nearest intrinsics, no contraction) is
// the one oracle/ddp_oracle.c uses: forward values are bit-identical, and so is the update
// given identical gradients; only the batch-sum order of the gradients differs.
#include "common.cuh"
#include <cooperative_groups.h>
#include <cstddef>

namespace cg = cooperative_groups;

namespace {

constexpr int H = DDPB_TOY_HIDDEN;
constexpr int P = DDPB_TOY_PARAMS;
constexpr int kCluster = 8;
constexpr int kComputeThreads = 128;             // 4 warps: the per-sample chain and the weight gradients
constexpr int kThreads = 192;                    // + 2 optimizer warps
constexpr int kLanes = 32;                       // samples per CTA and pass
constexpr int kChunk = kCluster * kLanes;        // samples per cluster pass
constexpr int kWRow = 12;                        // padded weight row, 3 x float4
constexpr int kRowStride = 36;                   // floats per table row: 144 B keeps LDS.128 conflict-free
constexpr int kOwned = 3;                        // neurons per warp and layer (ceil(10/4))
constexpr float kSlope = 0.01f;                  // nn.LeakyReLU() default

// offsets inside the flat parameter vector (module.parameters() order)
constexpr int OB0 = 2 * H;
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
static_assert(kSlice <= kThreads - kComputeThreads, "the optimizer warps must cover a slice");

struct JobList {
    ddpb_toy_job jobs[DDPB_MAX_JOBS];
};

struct __align__(16) Smem {
    float w0[H][4];            // layer 0: [j][k<2]
    float wh[3][H][kWRow];     // layers 1..3, [j][k]   (forward: row j)
    float wt[3][H][kWRow];     // layers 1..3, [k][j]   (backward: column k as a row)
    float w4[kWRow];           // layer 4
    float b[4][kWRow];         // biases of layers 0..3
    float b4[4];
    float table[kRows][kRowStride];
    float partial[kSlice * kCluster];
};

#ifdef DDPB_TIMELINE
// Development aid (tools/toy_timeline.py): per-phase SM clock stamps written through job.out.
#define DDPB_STAMP(slot)                                                                              \
    do {                                                                                              \
        if (lane == 0 && job.out) reinterpret_cast<long long *>(job.out)[(blockIdx.y * kCluster + cta) * 64 + warp * 8 + (slot)] = clock64(); \
    } while (0)
#else
#define DDPB_STAMP(slot) do { } while (0)
#endif

__device__ __forceinline__ float lrelu(float z) { return z > 0.f ? z : z * kSlope; }

// seq % n without a 64-bit division: a mask when n is a power of two (exact for ever), else a
// 32-bit modulo of the low word (exact for the first 2^32 evaluations).
__device__ __forceinline__ unsigned wrap_index(long long seq, int n) {
    const unsigned u = (unsigned)n, lo = (unsigned)seq;
    return (u & (u - 1)) == 0 ? (lo & (u - 1)) : lo % u;
}

__device__ __forceinline__ void compute_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(kComputeThreads) : "memory"); }

// Second cluster-barrier phase: pure execution ordering (keeps every CTA's shared memory
// alive until all remote reads are done), so no release/acquire fence is paid for.
__device__ __forceinline__ void cluster_rendezvous_relaxed() {
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
}

__device__ __forceinline__ void load_row(const float *row, float (&w)[kWRow]) {
    const float4 *q = reinterpret_cast<const float4 *>(row);
    float4 a = q[0], b = q[1], c = q[2];
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w;
}

// Read H consecutive table rows of this lane's sample.
__device__ __forceinline__ void read_rows(const Smem &S, int row0, int lane, float (&v)[H]) {
#pragma unroll
    for (int k = 0; k < H; ++k) v[k] = S.table[row0 + k][lane];
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

__device__ __forceinline__ void stage_weights(Smem &S, const float *__restrict__ prm, int tid) {
    for (int i = tid; i < P; i += kThreads) {
        const float val = prm[i];
        if (i < OB0) S.w0[i / 2][i & 1] = val;
        else if (i < OW1) S.b[0][i - OB0] = val;
        else if (i < OW4) {
            int q = i - OW1, l = q / kHidden, r = q - l * kHidden;
            if (r < H * H) { S.wh[l][r / H][r % H] = val; S.wt[l][r % H][r / H] = val; }
            else S.b[l + 1][r - H * H] = val;
        } else if (i < OB4) S.w4[i - OW4] = val;
        else S.b4[0] = val;
    }
    if (tid < kRowStride) S.table[R_ONE][tid] = 1.f;
}

__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kThreads, 1)
toy_step_kernel(const __grid_constant__ JobList jl) {
    __shared__ Smem S;
    const ddpb_toy_job &job = jl.jobs[blockIdx.y];
    cg::cluster_group cluster = cg::this_cluster();
    const int cta = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = job.B, mode = job.mode;
    const bool compute = tid < kComputeThreads;
    DDPB_STAMP(0);
    const long long seq = job.state ? job.state->seq : 0;

    // ---- optimizer warps: everything they need from HBM is requested before anything else
    const int own = cta * kSlice + (tid - kComputeThreads);     // index into the 372-entry slice table
    const bool owner = !compute && (tid - kComputeThreads) < kSlice && own <= P;
    float p_old = 0.f, m_old = 0.f, v_old = 0.f;
    float2 bias = make_float2(0.f, 1.f);
    long long step = 0;
    if (mode == DDPB_MODE_STEP && !compute) {
        step = job.state->step;
        if (owner && own < P) {
            p_old = job.params[own];
            m_old = job.m[own];
            if (job.opt == DDPB_OPT_ADAM) {
                v_old = job.v[own];
                const long long row = step < job.bias_table_n ? step : job.bias_table_n - 1;
                bias = reinterpret_cast<const float2 *>(job.bias_table)[row];
            }
        }
    }

    // ---- compute warps: the first batch rows are requested before the weights are staged
    const int32_t *idx_row = job.idx ? job.idx + (long long)wrap_index(seq, job.idx_rows) * job.idx_stride : nullptr;
    float x0 = 0.f, x1 = 0.f, y = 0.f;
    auto fetch_sample = [&](int chunk0) {
        const int b = chunk0 + cta * kLanes + lane;
        x0 = 0.f; x1 = 0.f; y = 0.f;
        if (compute && b < B) {
            const long long s = idx_row ? (long long)idx_row[b] : (long long)b;
            const float2 xv = *reinterpret_cast<const float2 *>(job.x + 2 * s);
            x0 = xv.x; x1 = xv.y;
            if (mode != DDPB_MODE_FWD) y = job.y[s];
        }
    };
    fetch_sample(0);
    stage_weights(S, job.params, tid);
    __syncthreads();
    DDPB_STAMP(1);

    if (!compute) {
        if (mode == DDPB_MODE_FWD) return;
        // ================= optimizer warps =================
        cluster.sync();                                   // partials of all eight CTAs are published
        DDPB_STAMP(2);
        if (owner) {
            float total = 0.f;
#pragma unroll
            for (int c = 0; c < kCluster; ++c) total += cluster.map_shared_rank(S.partial, c)[own];
            if (own == P) {
                job.grads[P] = total;                                        // sum (o-t)^2 over the local batch
                if (job.loss_ring) job.loss_ring[wrap_index(seq, job.loss_ring_n)] = total / (float)B;
            } else {
                job.grads[own] = total;
                if (mode == DDPB_MODE_STEP) {
                    const float g = total;
                    if (job.opt == DDPB_OPT_ADAM) {
                        // torch:optim/adam.py:457 (lerp), :476 (mul, addcmul), :545 (denom), :547 (addcdiv);
                        // bias.x = -(lr / (1 - beta1^t)), bias.y = sqrt(1 - beta2^t): host table (:531-536)
                        const float w1 = (float)(1.0 - job.beta1), b2 = (float)job.beta2;
                        const float w2 = (float)(1.0 - job.beta2), eps = (float)job.eps;
                        const float m_new = __fmaf_rn(w1, __fsub_rn(g, m_old), m_old);
                        const float v_new = __fadd_rn(__fmul_rn(v_old, b2), __fmul_rn(__fmul_rn(w2, g), g));
                        const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v_new), bias.y), eps);
                        job.params[own] = __fadd_rn(p_old, __fmul_rn(bias.x, __fdiv_rn(m_new, denom)));
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
        if (cta == 0 && tid == kComputeThreads) {          // advance the device-resident counters
            job.state->seq = seq + 1;
            if (mode == DDPB_MODE_STEP) job.state->step = step + 1;
        }
        DDPB_STAMP(3);
        cluster_rendezvous_relaxed();
        DDPB_STAMP(4);
        return;
    }

    // ================= compute warps =================
    int row_dz[3], row_a[3];
    float gacc[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const int i = tid + r * kComputeThreads;
        row_dz[r] = R_ONE; row_a[r] = R_ONE;
        if (i < P) rows_of_param(i, row_dz[r], row_a[r]);
    }
    float sq_acc = 0.f;
    const float norm = (float)(2.0 / (double)B);        // mse_loss_backward: 2/numel

    for (int chunk0 = 0; chunk0 < B; chunk0 += kChunk) {
        const bool cta_has_samples = chunk0 + cta * kLanes < B;          // uniform per CTA
        const bool valid = chunk0 + cta * kLanes + lane < B;
        if (cta_has_samples) {
            // neurons owned by this warp in every layer: warp, warp+4, warp+8 (the last: warps 0,1 only)
            unsigned mask[4] = {0u, 0u, 0u, 0u};          // bit i: pre-activation of owned neuron i > 0
            float cur[H];
            // ---- layer 0 (2 -> H)
            if (warp == 0) { S.table[R_IN][lane] = x0; S.table[R_IN + 1][lane] = x1; }
#pragma unroll
            for (int i = 0; i < kOwned; ++i) {
                const int j = warp + 4 * i;
                if (j < H) {
                    float acc = S.b[0][j];
                    acc = fmaf(S.w0[j][0], x0, acc);
                    acc = fmaf(S.w0[j][1], x1, acc);
                    mask[0] |= (acc > 0.f ? 1u : 0u) << i;
                    S.table[R_ACT + j][lane] = lrelu(acc);
                }
            }
            compute_barrier();
            // ---- layers 1..3 (H -> H)
#pragma unroll
            for (int l = 0; l < 3; ++l) {
                read_rows(S, R_ACT + l * H, lane, cur);
#pragma unroll
                for (int i = 0; i < kOwned; ++i) {
                    const int j = warp + 4 * i;
                    if (j < H) {
                        float w[kWRow];
                        load_row(S.wh[l][j], w);
                        float acc = S.b[l + 1][j];
#pragma unroll
                        for (int k = 0; k < H; ++k) acc = fmaf(w[k], cur[k], acc);
                        mask[l + 1] |= (acc > 0.f ? 1u : 0u) << i;
                        S.table[R_ACT + (l + 1) * H + j][lane] = lrelu(acc);
                    }
                }
                compute_barrier();
            }
            // ---- layer 4 (H -> 1): every warp computes it, all of them need the loss gradient
            DDPB_STAMP(2);
            read_rows(S, R_ACT + 3 * H, lane, cur);
            float out = S.b4[0];
            float w4[kWRow];
            load_row(S.w4, w4);
#pragma unroll
            for (int k = 0; k < H; ++k) out = fmaf(w4[k], cur[k], out);
            if (mode == DDPB_MODE_FWD) {
                if (valid && warp == 0) job.out[chunk0 + cta * kLanes + lane] = out;
            } else {
                // ---- loss, then backward: warp w produces dz for the neurons it owns
                const float diff = valid ? out - y : 0.f;
                const float dz4 = norm * diff;
                if (warp == 0) {
                    float sq = diff * diff;
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, off);
                    sq_acc += sq;
                    S.table[R_DZ + 4 * H][lane] = dz4;
                }
#pragma unroll
                for (int i = 0; i < kOwned; ++i) {
                    const int k = warp + 4 * i;
                    if (k < H) {
                        const float g = fmaf(dz4, w4[k], 0.f);
                        S.table[R_DZ + 3 * H + k][lane] = ((mask[3] >> i) & 1u) ? g : g * kSlope;
                    }
                }
                compute_barrier();
#pragma unroll
                for (int l = 2; l >= 0; --l) {            // through W_{l+1}: dz_{l+1} -> dz_l
                    read_rows(S, R_DZ + (l + 1) * H, lane, cur);
#pragma unroll
                    for (int i = 0; i < kOwned; ++i) {
                        const int k = warp + 4 * i;
                        if (k < H) {
                            float w[kWRow];
                            load_row(S.wt[l][k], w);
                            float s = 0.f;
#pragma unroll
                            for (int j = 0; j < H; ++j) s = fmaf(cur[j], w[j], s);
                            S.table[R_DZ + l * H + k][lane] = ((mask[l] >> i) & 1u) ? s : s * kSlope;
                        }
                    }
                    compute_barrier();
                }
            }
        }
        if (mode == DDPB_MODE_FWD) {
            if (chunk0 + kChunk < B) { compute_barrier(); fetch_sample(chunk0 + kChunk); }
            continue;
        }
        // ---- weight gradients of this CTA's 32 samples (the table is complete: last barrier above)
        DDPB_STAMP(3);
        if (cta_has_samples) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                if (tid + r * kComputeThreads < P) {
                    const float4 *pd = reinterpret_cast<const float4 *>(S.table[row_dz[r]]);
                    const float4 *pa = reinterpret_cast<const float4 *>(S.table[row_a[r]]);
                    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
                    for (int c = 0; c < kLanes / 4; ++c) {
                        const float4 d = pd[c], a = pa[c];
                        s0 = fmaf(d.x, a.x, s0); s1 = fmaf(d.y, a.y, s1);
                        s2 = fmaf(d.z, a.z, s2); s3 = fmaf(d.w, a.w, s3);
                    }
                    gacc[r] += (s0 + s1) + (s2 + s3);
                }
            }
        }
        if (chunk0 + kChunk < B) {                         // another pass: the table is about to be overwritten
            compute_barrier();
            fetch_sample(chunk0 + kChunk);
        }
    }
    if (mode == DDPB_MODE_FWD) return;

    // ---- publish the partials; the optimizer warps of the cluster take it from here
#pragma unroll
    for (int r = 0; r < 3; ++r)
        if (tid + r * kComputeThreads < P) S.partial[tid + r * kComputeThreads] = gacc[r];
    if (tid == 0) S.partial[P] = sq_acc;
    DDPB_STAMP(4);
    cluster.sync();
    DDPB_STAMP(5);
    cluster_rendezvous_relaxed();
    DDPB_STAMP(6);
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
            DDPB_REQUIRE(q.m && (q.opt == DDPB_OPT_SGD || (q.v && q.bias_table && q.bias_table_n >= 1)),
                         "ddpb_toy_step: job %d: optimizer state (m, v, bias table) missing", j);
        }
        if (q.loss_ring) DDPB_REQUIRE(q.loss_ring_n >= 1, "ddpb_toy_step: job %d: empty loss ring", j);
        if (q.idx) DDPB_REQUIRE(q.idx_rows >= 1 && q.idx_stride >= q.B, "ddpb_toy_step: job %d: index rows too short", j);
        jl.jobs[j] = q;
    }
    toy_step_kernel<<<dim3(kCluster, njobs, 1), kThreads, 0, (cudaStream_t)stream>>>(jl);
    DDPB_CUDA(cudaGetLastError());
    return 0;
}
