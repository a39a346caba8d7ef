// The fused training step of the reference's toy model (H = 10, 371 parameters).
//
// What the reference runs per model and per iteration (demo.py:104-111):
//   model(data)            ToyModel.forward, toy_model_and_data.py:24-25   (9 library kernels)
//   loss(out, target)      nn.MSELoss, demo.py:86,106                      (2-3)
//   loss.backward()        autograd + DDP reducer hooks, demo.py:107       (~45)
//   optimizer.step()       optim.Adam foreach path, demo.py:110            (~10)
// Here it is ONE launch for all models of the process: a thread-block cluster of 8 CTAs
// per model (blockIdx.y = model).  At this size (1 MFLOP, 10 KB) only latency matters; the
// structure below is the result of two measured iterations (profiles/r01_v0_*, r01_v2_*:
// ncu stall samples and an in-kernel clock64 timeline, tools/toy_timeline.py):
//
//   * CTA c owns samples [32c, 32c+32) of the batch, one sample per lane.  Its four compute
//     warps split each layer's output neurons (warp w owns neurons 3w..3w+2 of a layer padded
//     to 12, so every warp runs the same branch-free code and its three FMA chains
//     interleave).  Activations and pre-activation gradients meet in a shared-memory table
//     ([100 rows][32 samples]), one named barrier per layer.
//   * The same table is the operand of the weight-gradient contraction dW = dZ^T A
//     ([371 x 32] x [32] per CTA): 106 threads each own a 2x2 tile of (dz row, activation
//     row) pairs, which halves the shared-memory traffic of one-output-per-thread.
//   * Reduction over the 8 CTAs is push-style: every tile owner stores its partial sums
//     straight into the inbox of the CTA that owns that parameter slice (distributed shared
//     memory) with st.async, whose completion is counted by the destination's mbarrier
//     (complete_tx), so data and signal travel together: no fence, no cluster-wide barrier on
//     the critical path (the one that makes the mbarriers visible is arrived at, relaxed, at
//     kernel start and waited for much later).
//   * Two more warps per CTA are the optimizer: they prefetch p, m, v and the bias-correction
//     scalars of their 47-parameter slice at kernel start, sleep on the mbarrier, sum the
//     eight inbox rows in a fixed order (deterministic) and apply Adam/SGD.  Element 371 of
//     the slice table carries sum (o-t)^2, i.e. the loss.
//
// Arithmetic order of the per-sample chain (fmaf from the bias, k ascending) and of the
// optimizer (torch's op order, explicit round-to-nearest intrinsics, no contraction) is
// the one oracle/ddp_oracle.c uses: forward values are bit-identical, and so is the update
// given identical gradients; only the batch-sum order of the gradients differs.
#include "common.cuh"
#include <cooperative_groups.h>
#include <atomic>
#include <cstddef>

namespace cg = cooperative_groups;

namespace {

constexpr int H = DDPB_TOY_HIDDEN;
constexpr int P = DDPB_TOY_PARAMS;
constexpr int HP = 12;                           // neurons per layer after padding (3 per compute warp)
constexpr int kCluster = 8;
constexpr int kComputeThreads = 128;             // 4 warps: the per-sample chain and the weight gradients
constexpr int kThreads = 192;                    // + 2 optimizer warps
constexpr int kLanes = 32;                       // samples per CTA and pass
constexpr int kChunk = kCluster * kLanes;        // samples per cluster pass
constexpr int kRowStride = 36;                   // floats per table row: 144 B keeps LDS.128 conflict-free
constexpr int kOwned = HP / 4;                   // neurons per warp and layer
constexpr float kSlope = 0.01f;                  // nn.LeakyReLU() default

// offsets inside the flat parameter vector (module.parameters() order)
constexpr int OB0 = 2 * H;
constexpr int OW1 = OB0 + H;
constexpr int kHidden = H * H + H;               // one hidden layer: W then b
constexpr int OW4 = OW1 + 3 * kHidden, OB4 = OW4 + H;
static_assert(OB4 + 1 == P, "flat layout must add up to the parameter count");

// rows of the per-sample table (HP rows per layer; rows of padding neurons hold zeros)
constexpr int R_IN = 0;                          // x0, x1
constexpr int R_ACT = 2;                         // act_0 .. act_3
constexpr int R_ONE = R_ACT + 4 * HP;            // ones (bias gradients)
constexpr int R_ZERO = R_ONE + 1;                // zeros (padding operand of 2x2 tiles)
constexpr int R_DZ = R_ZERO + 1;                 // dz_0 .. dz_3, then dz_4
constexpr int kRows = R_DZ + 4 * HP + 1;
constexpr int kSlice = (P + 1 + kCluster - 1) / kCluster;   // 47 values per CTA (371 grads + loss sum)
constexpr int kInboxRow = 48;
constexpr int kXchgRow = DDPB_XCHG_ROW;          // floats per (parity, source rank) row of a peer window
constexpr int kMaxStagedRows = 1024;             // datasets up to this many rows are staged in shared memory
constexpr int kStageX = (2 * kMaxStagedRows / 4 + kThreads - 1) / kThreads;   // float4 loads per thread
constexpr int kStageY = (kMaxStagedRows / 4 + kThreads - 1) / kThreads;
static_assert(kSlice <= kThreads - kComputeThreads && kSlice <= kInboxRow, "the optimizer warps must cover a slice");

// weight-gradient tiles: 3 hidden layers x (5 neuron pairs x 6 input pairs) + layer 0 (5 x 2) + layer 4 (1 x 6)
constexpr int kTilesHidden = 30, kTilesL0 = 10, kTilesL4 = 6;
constexpr int kTiles = 3 * kTilesHidden + kTilesL0 + kTilesL4;
static_assert(kTiles <= kComputeThreads, "one tile per compute thread");

struct JobList {
    ddpb_toy_job jobs[DDPB_MAX_JOBS];
};

struct __align__(16) Smem {
    float w0[HP][4];           // layer 0: [j][k<2]
    float wh[3][HP][HP];       // layers 1..3, [j][k]   (forward: row j)
    float wt[3][HP][HP];       // layers 1..3, [k][j]   (backward: column k as a row)
    float w4[HP];              // layer 4
    float b[4][HP];            // biases of layers 0..3
    float b4[4];
    float table[kRows][kRowStride];
    float data_x[2 * kMaxStagedRows];   // the whole dataset when it is small (gather without a dependent HBM trip)
    float data_y[kMaxStagedRows];
    float inbox[kCluster][kInboxRow];   // partial sums pushed by the cluster's CTAs for MY slice
    unsigned long long mbar;            // counts the 8 "pushed" arrivals
};
constexpr int kWeightFloats = (int)(offsetof(Smem, table) / sizeof(float));

#ifdef DDPB_TIMELINE
// Development aid (tools/toy_timeline.py): per-phase SM clock stamps written through job.out.
#define DDPB_STAMP(slot)                                                                              \
    do {                                                                                              \
        if (lane == 0 && job.out) reinterpret_cast<long long *>(job.out)[(blockIdx.y * kCluster + cta) * 64 + warp * 8 + (slot)] = clock64(); \
    } while (0)
#else
#define DDPB_STAMP(slot) do { } while (0)
#endif

#ifdef DDPB_XCHG_TIMELINE
// Development aid (tools/toy_exchange_timeline.py): %globaltimer stamps (ns) of the thread that owns bucket
// element 0, written through job.out: 0 kernel entry, 1 local gradients reduced, 2 exchange stores issued,
// 3 every rank's ticket seen, 4 update stored, 5 = polling rounds.
#define DDPB_XSTAMP(slot)                                                                                       \
    do {                                                                                                        \
        if (cta == 0 && tid == kComputeThreads && job.out) reinterpret_cast<unsigned long long *>(job.out)[slot] = global_ns(); \
    } while (0)
#else
#define DDPB_XSTAMP(slot) do { } while (0)
#endif

__device__ __forceinline__ float lrelu(float z) { return z > 0.f ? z : z * kSlope; }

// seq % n without a 64-bit division: a mask when n is a power of two (exact for ever), else a
// 32-bit modulo of the low word (exact for the first 2^32 evaluations).
__device__ __forceinline__ unsigned wrap_index(long long seq, int n) {
    const unsigned u = (unsigned)n, lo = (unsigned)seq;
    return (u & (u - 1)) == 0 ? (lo & (u - 1)) : lo % u;
}

__device__ __forceinline__ void compute_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(kComputeThreads) : "memory"); }

__device__ __forceinline__ void optimizer_barrier() {
    asm volatile("bar.sync 2, %0;" ::"n"(kThreads - kComputeThreads) : "memory");
}

// System-scope 8-byte {value, ticket} slots in peer-mapped windows (other GPUs' HBM over NVLink,
// or our own): one aligned v2 access = one transaction, so value and ticket arrive together.
__device__ __forceinline__ void st_slot_sys(uint2 *p, uint2 v) {
    asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ uint2 ld_slot_sys(const uint2 *p) {
    uint2 v;
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
constexpr unsigned long long kPeerTimeoutNs = 2000000000ull;   // a peer that has not delivered after 2 s is gone
constexpr unsigned kTimeoutSentinel = DDPB_LOSS_TIMEOUT_BITS;  // what the loss slot holds after a time-out

__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// The receiving side of the push: one arrival that also announces how many bytes the other
// CTAs of the cluster will deliver into this CTA's inbox.
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}

// The sending side: a 4-byte store into the shared memory of CTA `rank` of this cluster that
// reports its own completion to that CTA's mbarrier (st.async ... complete_tx).  Data and
// signal travel together, so no fence and no CTA barrier precede the hand-over.
__device__ __forceinline__ void push_f32(float *local_slot, unsigned long long *local_bar, unsigned rank, float val) {
    unsigned dst, bar;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(dst) : "r"(smem_addr(local_slot)), "r"(rank));
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(bar) : "r"(smem_addr(local_bar)), "r"(rank));
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
                 ::"r"(dst), "r"(__float_as_uint(val)), "r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    unsigned done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    } while (!done);
}

// Split cluster barrier.  The arrive is relaxed on purpose: the release form compiles to
// MEMBAR.ALL.GPU, which would wait for every global load of the prologue (measured: +1000
// cycles); the only thing it has to publish, the mbarrier initialisation, is covered by
// fence.mbarrier_init.release.cluster.
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

__device__ __forceinline__ void load_row(const float *row, float (&w)[HP]) {
    const float4 *q = reinterpret_cast<const float4 *>(row);
    float4 a = q[0], b = q[1], c = q[2];
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w;
}

// Read the H real rows of a layer for this lane's sample.
__device__ __forceinline__ void read_rows(const Smem &S, int row0, int lane, float (&v)[H]) {
#pragma unroll
    for (int k = 0; k < H; ++k) v[k] = S.table[row0 + k][lane];
}

// Scatter flat parameter i into the staged (padded, partly transposed) copies.
__device__ __forceinline__ void stage_param(Smem &S, int i, float val) {
    if (i < OB0) S.w0[i / 2][i & 1] = val;
    else if (i < OW1) S.b[0][i - OB0] = val;
    else if (i < OW4) {
        int q = i - OW1, l = q / kHidden, r = q - l * kHidden;
        if (r < H * H) { S.wh[l][r / H][r % H] = val; S.wt[l][r % H][r / H] = val; }
        else S.b[l + 1][r - H * H] = val;
    } else if (i < OB4) S.w4[i - OW4] = val;
    else S.b4[0] = val;
}

// The 2x2 weight-gradient tile of compute thread t: two dz rows x two operand rows, and the
// flat parameter index of each of the four sums (-1: padding, nothing to store).
struct Tile {
    int rd[2], ra[2], out[2][2];
};

__device__ __forceinline__ Tile tile_of_thread(int t) {
    Tile T;
    T.rd[0] = T.rd[1] = R_ZERO; T.ra[0] = T.ra[1] = R_ZERO;
    T.out[0][0] = T.out[0][1] = T.out[1][0] = T.out[1][1] = -1;
    if (t < 3 * kTilesHidden) {                       // hidden layer L+1: weights [j][k], bias [j]
        const int L = t / kTilesHidden, r = t - L * kTilesHidden, jp = r / 6, kp = r - jp * 6;
        const int base = OW1 + L * kHidden;
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * jp + jj;
            T.rd[jj] = R_DZ + (L + 1) * HP + j;
            if (kp < 5) {                             // inputs k and k+5: consecutive threads read
                T.out[jj][0] = base + j * H + kp;     // consecutive table rows (no bank conflicts)
                T.out[jj][1] = base + j * H + kp + 5;
            } else {
                T.out[jj][0] = base + H * H + j;      // bias: operand row of ones
            }
        }
        if (kp < 5) { T.ra[0] = R_ACT + L * HP + kp; T.ra[1] = T.ra[0] + 5; }
        else T.ra[0] = R_ONE;
    } else if (t < 3 * kTilesHidden + kTilesL0) {     // layer 0: inputs x0, x1 and the bias
        const int r = t - 3 * kTilesHidden, jp = r >> 1, kp = r & 1;
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * jp + jj;
            T.rd[jj] = R_DZ + j;
            if (kp == 0) { T.out[jj][0] = 2 * j; T.out[jj][1] = 2 * j + 1; }
            else T.out[jj][0] = OB0 + j;
        }
        if (kp == 0) { T.ra[0] = R_IN; T.ra[1] = R_IN + 1; }
        else T.ra[0] = R_ONE;
    } else if (t < kTiles) {                          // layer 4: one output neuron
        const int kp = t - 3 * kTilesHidden - kTilesL0;
        T.rd[0] = R_DZ + 4 * HP;
        if (kp < 5) {
            T.ra[0] = R_ACT + 3 * HP + kp; T.ra[1] = T.ra[0] + 5;
            T.out[0][0] = OW4 + kp; T.out[0][1] = OW4 + kp + 5;
        } else {
            T.ra[0] = R_ONE;
            T.out[0][0] = OB4;
        }
    }
    return T;
}

__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kThreads, 1)
toy_step_kernel(const __grid_constant__ JobList jl) {
    __shared__ Smem S;
    // By-value copy: ~20 independent constant-bank loads issued back to back.  Reading the fields
    // on demand through a dynamically indexed reference chained 17 LDC -> compare -> branch
    // round trips (~1000 cycles before the first useful instruction, seen in the timeline).
    const ddpb_toy_job job = jl.jobs[blockIdx.y];
    const ddpb_toy_job &jref = jl.jobs[blockIdx.y];     // the peer pointer arrays stay in the constant bank
    cg::cluster_group cluster = cg::this_cluster();
    const int cta = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = job.B, mode = job.mode;
    const bool compute = tid < kComputeThreads;
    DDPB_STAMP(0);
    DDPB_XSTAMP(0);

    // ---- 0. every global load this kernel needs is issued before anything waits on one
    const float *__restrict__ prm = job.params;
    const float pv0 = prm[tid];                                   // 371 parameters over 192 threads
    const float pv1 = (tid + kThreads < P) ? prm[tid + kThreads] : 0.f;
    const bool need_seq = mode != DDPB_MODE_FWD && (job.loss_ring_n > 1 || (job.idx && job.idx_rows > 1) || job.world > 1);
    const long long seq = (job.state && need_seq) ? job.state->seq : 0;

    const int own = cta * kSlice + (tid - kComputeThreads);     // index into the 372-entry slice table
    const bool owner = !compute && (tid - kComputeThreads) < kSlice && own <= P;
    float p_old = 0.f, m_old = 0.f, v_old = 0.f;
    float2 bias = make_float2(0.f, 1.f);
    long long step = 0;
    if (mode == DDPB_MODE_STEP && !compute) {
        step = job.state->step;
        if (owner && own < P) {
            p_old = prm[own];
            m_old = job.m[own];
            if (job.opt == DDPB_OPT_ADAM) v_old = job.v[own];
        }
    }
    // The batch: rows idx_row[b] of the dataset (or row b of a ready-made batch).  A dependent
    // index -> row load costs a second ~800-cycle L2 round trip, so a small dataset is pulled
    // into shared memory wholesale, in parallel with the index load, and gathered from there.
    const int32_t *idx_row = nullptr;
    if (job.idx) idx_row = job.idx + (job.idx_rows > 1 ? (long long)wrap_index(seq, job.idx_rows) * job.idx_stride : 0);
    const int n_staged = (idx_row && job.n_data > 0 && job.n_data <= kMaxStagedRows && B <= kChunk) ? job.n_data : 0;
    float4 stage_x[kStageX], stage_y[kStageY];
    int my_row = -1;
    if (n_staged) {
        const int b = cta * kLanes + lane;
        if (compute && b < B) my_row = idx_row[b];
#pragma unroll
        for (int r = 0; r < kStageX; ++r) {
            const int i = tid + r * kThreads;
            stage_x[r] = (4 * i < 2 * n_staged) ? reinterpret_cast<const float4 *>(job.x)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int r = 0; r < kStageY; ++r) {
            const int i = tid + r * kThreads;
            stage_y[r] = (4 * i < n_staged && mode != DDPB_MODE_FWD) ? reinterpret_cast<const float4 *>(job.y)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
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
    if (!n_staged) fetch_sample(0);
    const Tile T = tile_of_thread(tid);                // integer work, done while the loads fly

    // ---- 1. stage the weights: zero the padded arrays while the loads fly, then scatter
    {
        float *flat = reinterpret_cast<float *>(&S);
        for (int i = tid; i < kWeightFloats; i += kThreads) flat[i] = 0.f;
        if (tid < kRowStride) { S.table[R_ONE][tid] = 1.f; S.table[R_ZERO][tid] = 0.f; }
        if (tid == kComputeThreads && mode != DDPB_MODE_FWD) {
            const int entries = min(kSlice, P + 1 - cta * kSlice);          // slice entries this CTA owns
            mbar_init(&S.mbar, 1);
            mbar_expect_tx(&S.mbar, (unsigned)(entries * kCluster * sizeof(float)));
        }
        __syncthreads();
        DDPB_STAMP(1);
        if (mode != DDPB_MODE_FWD) cluster_arrive();             // waited for just before the first remote access
        stage_param(S, tid, pv0);
        if (tid + kThreads < P) stage_param(S, tid + kThreads, pv1);
        if (n_staged) {
#pragma unroll
            for (int r = 0; r < kStageX; ++r)
                if (4 * (tid + r * kThreads) < 2 * n_staged) reinterpret_cast<float4 *>(S.data_x)[tid + r * kThreads] = stage_x[r];
#pragma unroll
            for (int r = 0; r < kStageY; ++r)
                if (4 * (tid + r * kThreads) < n_staged) reinterpret_cast<float4 *>(S.data_y)[tid + r * kThreads] = stage_y[r];
        }
        __syncthreads();
        if (n_staged && my_row >= 0) { x0 = S.data_x[2 * my_row]; x1 = S.data_x[2 * my_row + 1]; y = S.data_y[my_row]; }
    }
    DDPB_STAMP(2);

    if (!compute) {
        if (mode == DDPB_MODE_FWD) return;
        // ================= optimizer warps =================
        cluster_wait();
        // the bias-correction row depends on the step counter loaded above: a second round trip,
        // started only now so that it never delays the staging barriers, and hidden by the wait
        if (mode == DDPB_MODE_STEP && job.opt == DDPB_OPT_ADAM && owner && own < P) {
            const long long row = step < job.bias_table_n ? step : job.bias_table_n - 1;
            bias = reinterpret_cast<const float2 *>(job.bias_table)[row];
        }
        // everything that does not depend on the gradients is in registers before the wait
        const float w1 = (float)(1.0 - job.beta1), b2 = (float)job.beta2;
        const float w2 = (float)(1.0 - job.beta2), eps = (float)job.eps;
        const float lr_f = (float)job.lr, mo = (float)job.momentum;
        const int opt = job.opt;
        float *const g_out = job.grads, *const p_out = job.params, *const m_out = job.m, *const v_out = job.v;
        float *const loss_slot = job.loss_ring ? job.loss_ring + wrap_index(seq, job.loss_ring_n) : nullptr;
        float *const gloss_slot = job.global_loss_ring ? job.global_loss_ring + wrap_index(seq, job.loss_ring_n) : nullptr;
        const float inv_B = __fdiv_rn(1.0f, (float)B);
        mbar_wait(&S.mbar, 0);                            // all eight CTAs have pushed their partials
        DDPB_STAMP(3);
        DDPB_XSTAMP(1);
        const int slot = tid - kComputeThreads;
        float total = 0.f;
        if (owner) {
#pragma unroll
            for (int c = 0; c < kCluster; ++c) total += S.inbox[c][slot];
            if (own == P && loss_slot)                                       // local mean loss (what loss.cpu() returns)
                *loss_slot = (B & (B - 1)) == 0 ? total * inv_B : __fdiv_rn(total, (float)B);
        }
        bool timed_out = false;
        if (job.world > 1 && mode == DDPB_MODE_STEP) {
            // ---- one-shot all-reduce over NVLink, fused into the step (no NCCL launch, no second
            // kernel): slice c of every rank's bucket meets in every rank's receive window.
            // Every element travels as one 8-byte {value, ticket} store (the low-latency protocol:
            // data and "it has arrived" in a single NVLink transaction, so no fence, no flag round
            // trip); the receiver polls its own window until the ticket of this step shows up.
            // DDP's arithmetic: scale by 1/W first, then sum (default_hooks.py:24-29); ranks are
            // summed in ascending order on every GPU, so replicas stay bit-identical.
            const unsigned xseq = (unsigned)seq + 1u;                        // this exchange's ticket (never 0)
            const int par = (int)(seq & 1);                                  // double-buffered: a peer is at most one step ahead
            const int W = job.world, me = job.rank;
            if (owner) {
                const float contrib = (own < P) ? __fmul_rn(total, job.inv_world) : total;
                const size_t at = (size_t)(par * DDPB_MAX_PEERS + me) * kXchgRow + own;
                for (int dst = 0; dst < W; ++dst)
                    st_slot_sys(reinterpret_cast<uint2 *>(jref.peer_data[dst]) + at, make_uint2(__float_as_uint(contrib), xseq));
                DDPB_XSTAMP(2);
                const uint2 *mine = reinterpret_cast<const uint2 *>(jref.peer_data[me]) + (size_t)(par * DDPB_MAX_PEERS) * kXchgRow + own;
                // All W slots are polled together: the loads of one round are in flight at the same
                // time (one NVLink/L2 round trip per round, not W).  The wait is bounded by the global
                // timer (kPeerTimeoutNs): a peer that never delivers makes this rank give up LOUDLY --
                // state->reserved[0] is set, the loss slot gets the time-out sentinel, and neither the
                // parameters, the moments nor the counters are touched, so the state stays the one of
                // the last completed step.
                uint2 got[DDPB_MAX_PEERS];
                unsigned long long t0 = 0;
                unsigned rounds = 0;
                for (;;) {
                    bool all = true;
#pragma unroll
                    for (int src = 0; src < DDPB_MAX_PEERS; ++src)
                        if (src < W) got[src] = ld_slot_sys(mine + (size_t)src * kXchgRow);
#pragma unroll
                    for (int src = 0; src < DDPB_MAX_PEERS; ++src)
                        if (src < W) all = all && got[src].y == xseq;
                    if (all) break;
                    if ((++rounds & 255u) == 0) {
                        const unsigned long long now = global_ns();
                        if (t0 == 0) t0 = now;
                        else if (now - t0 > kPeerTimeoutNs) { timed_out = true; break; }
                    }
                }
                DDPB_XSTAMP(3);
#ifdef DDPB_XCHG_TIMELINE
                if (cta == 0 && tid == kComputeThreads && job.out) reinterpret_cast<unsigned long long *>(job.out)[5] = rounds;
#endif
                total = 0.f;
#pragma unroll
                for (int src = 0; src < DDPB_MAX_PEERS; ++src)
                    if (src < W) total += __uint_as_float(got[src].x);
                if (timed_out) {
                    job.state->reserved[0] = 1;
                    if (own == P && loss_slot) *loss_slot = __uint_as_float(kTimeoutSentinel);
                    if (own == P && gloss_slot) *gloss_slot = __uint_as_float(kTimeoutSentinel);
                } else if (own == P && gloss_slot) {
                    // mean loss over all ranks: what demo.py:112-121 obtains with two gloo all-reduces
                    *gloss_slot = __fdiv_rn(total, __fmul_rn((float)B, (float)W));
                }
            }
        }
        if (owner && !timed_out) {
            if (own == P) {
                g_out[P] = total;                                            // sum (o-t)^2 (over all ranks when W > 1)
            } else {
                g_out[own] = total;
                if (mode == DDPB_MODE_STEP) {
                    const float g = total;
                    if (opt == DDPB_OPT_ADAM) {
                        // torch:optim/adam.py:457 (lerp), :476 (mul, addcmul), :545 (denom), :547 (addcdiv);
                        // bias.x = -(lr / (1 - beta1^t)), bias.y = sqrt(1 - beta2^t): host table (:531-536)
                        const float m_new = __fmaf_rn(w1, __fsub_rn(g, m_old), m_old);
                        const float v_new = __fadd_rn(__fmul_rn(v_old, b2), __fmul_rn(__fmul_rn(w2, g), g));
                        const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v_new), bias.y), eps);
                        p_out[own] = __fadd_rn(p_old, __fmul_rn(bias.x, __fdiv_rn(m_new, denom)));
                        m_out[own] = m_new;
                        v_out[own] = v_new;
                    } else {
                        // torch:optim/sgd.py _single_tensor_sgd: buf = g (first step) | momentum*buf + g
                        float d = g;
                        if (mo != 0.f) {
                            d = (step == 0) ? g : __fadd_rn(__fmul_rn(m_old, mo), g);
                            m_out[own] = d;
                        }
                        p_out[own] = __fsub_rn(p_old, __fmul_rn(lr_f, d));
                    }
                }
            }
        }
        if (cta == 0 && tid == kComputeThreads && !timed_out) {          // advance the device-resident counters
            const long long seq_now = need_seq ? seq : job.state->seq;
            job.state->seq = seq_now + 1;
            if (mode == DDPB_MODE_STEP) job.state->step = step + 1;
        }
        DDPB_STAMP(4);
        DDPB_XSTAMP(4);
        return;
    }

    // ================= compute warps =================
    float gacc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
    float sq_acc = 0.f;
    const float norm = __fdiv_rn(2.0f, (float)B);       // mse_loss_backward: 2/numel (== (float)(2.0/B) for B < 2^24)
    const int j0 = warp * kOwned;                       // first neuron owned by this warp in every layer

    for (int chunk0 = 0; chunk0 < B; chunk0 += kChunk) {
        const bool cta_has_samples = chunk0 + cta * kLanes < B;          // uniform per CTA
        const bool valid = chunk0 + cta * kLanes + lane < B;
        if (cta_has_samples) {
            unsigned mask[4] = {0u, 0u, 0u, 0u};          // bit i: pre-activation of owned neuron i > 0
            float cur[H];
            // ---- layer 0 (2 -> H)
            if (warp == 0) { S.table[R_IN][lane] = x0; S.table[R_IN + 1][lane] = x1; }
#pragma unroll
            for (int i = 0; i < kOwned; ++i) {
                const int j = j0 + i;
                float acc = S.b[0][j];
                acc = fmaf(S.w0[j][0], x0, acc);
                acc = fmaf(S.w0[j][1], x1, acc);
                mask[0] |= (acc > 0.f ? 1u : 0u) << i;
                S.table[R_ACT + j][lane] = lrelu(acc);
            }
            compute_barrier();
            DDPB_STAMP(3);
            // ---- layers 1..3 (H -> H): three interleaved 10-FMA chains per warp
#pragma unroll
            for (int l = 0; l < 3; ++l) {
                read_rows(S, R_ACT + l * HP, lane, cur);
                float w[kOwned][HP], acc[kOwned];
#pragma unroll
                for (int i = 0; i < kOwned; ++i) { load_row(S.wh[l][j0 + i], w[i]); acc[i] = S.b[l + 1][j0 + i]; }
#pragma unroll
                for (int k = 0; k < H; ++k)
#pragma unroll
                    for (int i = 0; i < kOwned; ++i) acc[i] = fmaf(w[i][k], cur[k], acc[i]);
#pragma unroll
                for (int i = 0; i < kOwned; ++i) {
                    mask[l + 1] |= (acc[i] > 0.f ? 1u : 0u) << i;
                    S.table[R_ACT + (l + 1) * HP + j0 + i][lane] = lrelu(acc[i]);
                }
                compute_barrier();
            }
            // ---- layer 4 (H -> 1): every warp computes it, all of them need the loss gradient
            DDPB_STAMP(4);
            read_rows(S, R_ACT + 3 * HP, lane, cur);
            float out = S.b4[0];
            float w4[HP];
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
                    sq_acc = fmaf(diff, diff, sq_acc);                       // per lane; reduced once, after the loop
                    S.table[R_DZ + 4 * HP][lane] = dz4;
                }
#pragma unroll
                for (int i = 0; i < kOwned; ++i) {
                    const float g = fmaf(dz4, S.w4[j0 + i], 0.f);            // padding neurons: weight 0
                    S.table[R_DZ + 3 * HP + j0 + i][lane] = ((mask[3] >> i) & 1u) ? g : g * kSlope;
                }
                compute_barrier();
#pragma unroll
                for (int l = 2; l >= 0; --l) {            // through W_{l+1}: dz_{l+1} -> dz_l
                    read_rows(S, R_DZ + (l + 1) * HP, lane, cur);
                    float w[kOwned][HP], s[kOwned];
#pragma unroll
                    for (int i = 0; i < kOwned; ++i) { load_row(S.wt[l][j0 + i], w[i]); s[i] = 0.f; }
#pragma unroll
                    for (int j = 0; j < H; ++j)
#pragma unroll
                        for (int i = 0; i < kOwned; ++i) s[i] = fmaf(cur[j], w[i][j], s[i]);
#pragma unroll
                    for (int i = 0; i < kOwned; ++i)
                        S.table[R_DZ + l * HP + j0 + i][lane] = ((mask[l] >> i) & 1u) ? s[i] : s[i] * kSlope;
                    compute_barrier();
                }
            }
        }
        if (mode == DDPB_MODE_FWD) {
            if (chunk0 + kChunk < B) { compute_barrier(); fetch_sample(chunk0 + kChunk); }
            continue;
        }
        // ---- weight gradients of this CTA's 32 samples: one 2x2 tile per thread
        DDPB_STAMP(5);
        if (cta_has_samples && tid < kTiles) {
            const float4 *d0 = reinterpret_cast<const float4 *>(S.table[T.rd[0]]);
            const float4 *d1 = reinterpret_cast<const float4 *>(S.table[T.rd[1]]);
            const float4 *a0 = reinterpret_cast<const float4 *>(S.table[T.ra[0]]);
            const float4 *a1 = reinterpret_cast<const float4 *>(S.table[T.ra[1]]);
            float s00[4] = {0.f, 0.f, 0.f, 0.f}, s01[4] = {0.f, 0.f, 0.f, 0.f};
            float s10[4] = {0.f, 0.f, 0.f, 0.f}, s11[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < kLanes / 4; ++c) {
                const float4 p = d0[c], q = d1[c], u = a0[c], v = a1[c];
                s00[0] = fmaf(p.x, u.x, s00[0]); s00[1] = fmaf(p.y, u.y, s00[1]); s00[2] = fmaf(p.z, u.z, s00[2]); s00[3] = fmaf(p.w, u.w, s00[3]);
                s01[0] = fmaf(p.x, v.x, s01[0]); s01[1] = fmaf(p.y, v.y, s01[1]); s01[2] = fmaf(p.z, v.z, s01[2]); s01[3] = fmaf(p.w, v.w, s01[3]);
                s10[0] = fmaf(q.x, u.x, s10[0]); s10[1] = fmaf(q.y, u.y, s10[1]); s10[2] = fmaf(q.z, u.z, s10[2]); s10[3] = fmaf(q.w, u.w, s10[3]);
                s11[0] = fmaf(q.x, v.x, s11[0]); s11[1] = fmaf(q.y, v.y, s11[1]); s11[2] = fmaf(q.z, v.z, s11[2]); s11[3] = fmaf(q.w, v.w, s11[3]);
            }
            gacc[0][0] += (s00[0] + s00[1]) + (s00[2] + s00[3]);
            gacc[0][1] += (s01[0] + s01[1]) + (s01[2] + s01[3]);
            gacc[1][0] += (s10[0] + s10[1]) + (s10[2] + s10[3]);
            gacc[1][1] += (s11[0] + s11[1]) + (s11[2] + s11[3]);
        }
        if (chunk0 + kChunk < B) {                         // another pass: the table is about to be overwritten
            compute_barrier();
            fetch_sample(chunk0 + kChunk);
        }
    }
    if (mode == DDPB_MODE_FWD) return;

    // ---- push the partials into the inbox of the CTA that owns each parameter slice
    DDPB_STAMP(6);
    cluster_wait();                                       // every CTA runs and has initialised its mbarrier
#pragma unroll
    for (int jj = 0; jj < 2; ++jj)
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
            const int i = T.out[jj][kk];
            if (i >= 0) {
                const int dst = i / kSlice;
                push_f32(&S.inbox[cta][i - dst * kSlice], &S.mbar, (unsigned)dst, gacc[jj][kk]);
            }
        }
    if (warp == 0) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) sq_acc += __shfl_xor_sync(0xffffffffu, sq_acc, off);
        if (lane == 0) push_f32(&S.inbox[cta][P % kSlice], &S.mbar, (unsigned)(P / kSlice), sq_acc);
    }
    DDPB_STAMP(7);
}

}  // namespace

static int launch_toy_step(const ddpb_toy_job *jobs, int njobs, cudaStream_t stream) {
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
        if (q.world > 1 && q.mode == DDPB_MODE_STEP) {
            DDPB_REQUIRE(q.world <= DDPB_MAX_PEERS && q.rank >= 0 && q.rank < q.world,
                         "ddpb_toy_step: job %d: world %d / rank %d outside the %d-peer exchange", j, q.world, q.rank, DDPB_MAX_PEERS);
            for (int r = 0; r < q.world; ++r)
                DDPB_REQUIRE(q.peer_data[r] && ((uintptr_t)q.peer_data[r] & 7) == 0, "ddpb_toy_step: job %d: peer window %d not mapped", j, r);
        }
        jl.jobs[j] = q;
        // the shared-memory staging of a small dataset needs whole, aligned float4 rows
        if (!q.idx || q.n_data < 0 || (q.n_data & 3) || ((uintptr_t)q.x & 15) || ((uintptr_t)q.y & 15)) jl.jobs[j].n_data = 0;
    }
    toy_step_kernel<<<dim3(kCluster, njobs, 1), kThreads, 0, stream>>>(jl);
    ddpb::count_launch();
    DDPB_CUDA(cudaGetLastError());
    return 0;
}

DDPB_API int ddpb_toy_step(const ddpb_toy_job *jobs, int njobs, void *stream) {
    return launch_toy_step(jobs, njobs, (cudaStream_t)stream);
}

DDPB_API int ddpb_toy_step_h2d(const ddpb_toy_job *jobs, int njobs, void *dst_dev, const void *src_host,
                               int64_t bytes, void *stream) {
    if (bytes > 0) {
        DDPB_REQUIRE(dst_dev && src_host, "ddpb_toy_step_h2d: null copy operand");
        DDPB_CUDA(cudaMemcpyAsync(dst_dev, src_host, (size_t)bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    }
    return launch_toy_step(jobs, njobs, (cudaStream_t)stream);
}

DDPB_API int ddpb_toy_iteration(ddpb_toy_iter *it) {
    DDPB_REQUIRE(it != nullptr, "ddpb_toy_iteration: null descriptor");
    DDPB_REQUIRE(it->nwait >= 0 && it->nwait <= 2 * DDPB_MAX_JOBS, "ddpb_toy_iteration: nwait %d outside 0..%d", it->nwait,
                 2 * DDPB_MAX_JOBS);
    if (it->idx) {
        if (int rc = ddpb_host_gather(it->X, it->Y, it->idx, it->B, it->dx, it->dy, it->out_x, it->out_y)) return rc;
    }
    if (it->njobs > 0)
        if (int rc = launch_toy_step(it->jobs, it->njobs, (cudaStream_t)it->stream)) return rc;
    for (int w = 0; w < it->nwait; ++w) {
        DDPB_REQUIRE(it->wait_slot[w] != nullptr, "ddpb_toy_iteration: wait slot %d is null", w);
        if (int rc = ddpb::wait_slot(it->wait_slot[w], it->sentinel_bits, it->spin_ns, (cudaStream_t)it->stream, &it->wait_value[w]))
            return rc;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// Several replicas of one model on ONE GPU, exchanging gradients without any allocation.
//
// The fused multi-rank step above needs an exchange window per rank, and somebody has to
// allocate those.  Some callers cannot: a service that handed the whole device to another
// memory pool at start-up (a raw cudaMalloc would fail or fragment that pool), a
// latency-critical loop that must never take the allocator's device-wide synchronisation, a
// test rig that wants the multi-rank arithmetic on a single-GPU box.  For them the windows
// of up to DDPB_MAX_VIRTUAL_RANKS replicas live in the file-scope array below (188 KiB of
// every GPU that loads this library, held until the process exits), and all replicas run in
// one launch (blockIdx.y = replica), so they are co-resident by construction.
//
// Rules that follow from sharing one array:
//   * It keeps whatever the previous call left, including arrival tickets that could match a
//     later caller's step number by accident, so the windows in use are cleared on the stream at
//     the start of every call.
//   * Two calls whose kernels overlap on the GPU (different streams, no dependency) race on
//     the array; serialise them.
//   * The device address of the array is looked up once per device, on the first call made
//     with that device current; make that first call before any stream capture begins.
namespace {

static_assert(DDPB_MAX_VIRTUAL_RANKS <= DDPB_MAX_JOBS && DDPB_MAX_VIRTUAL_RANKS <= DDPB_MAX_PEERS,
              "every virtual rank is one job of the launch and one peer of the exchange");
constexpr size_t kWindowBytes = DDPB_XCHG_WINDOW_BYTES;
constexpr size_t kVirtualBytes = DDPB_MAX_VIRTUAL_RANKS * kWindowBytes;

// layout: [rank 0 window][rank 1 window]...
static __device__ __align__(256) unsigned char g_virtual_windows[kVirtualBytes];

// In host code `g_virtual_windows` is only a handle for the runtime, never an address a kernel
// may use: the real device address comes from cudaGetSymbolAddress, per device.
// Host threads may race to fill an entry; they would all store the same address (the lookup is
// idempotent), so a relaxed atomic is enough.  cudaDeviceReset invalidates the entries: a process that
// resets a device must not call this entry point on it afterwards.
std::atomic<unsigned char *> g_virtual_base[64] = {};

int virtual_window_base(unsigned char **base_out) {
    int dev = -1;
    DDPB_CUDA(cudaGetDevice(&dev));
    DDPB_REQUIRE(dev >= 0 && dev < 64, "ddpb_toy_step_virtual_ranks: device index %d not supported", dev);
    unsigned char *base = g_virtual_base[dev].load(std::memory_order_relaxed);
    if (!base) {
        void *ptr = nullptr;
        DDPB_CUDA(cudaGetSymbolAddress(&ptr, g_virtual_windows));
        base = static_cast<unsigned char *>(ptr);
        g_virtual_base[dev].store(base, std::memory_order_relaxed);
    }
    *base_out = base;
    return 0;
}

}  // namespace

DDPB_API int ddpb_toy_step_virtual_ranks(const ddpb_toy_job *jobs, int nranks, void *stream) {
    DDPB_REQUIRE(jobs != nullptr, "ddpb_toy_step_virtual_ranks: null jobs");
    DDPB_REQUIRE(nranks >= 2 && nranks <= DDPB_MAX_VIRTUAL_RANKS,
                 "ddpb_toy_step_virtual_ranks: %d replicas requested, the built-in exchange windows hold 2..%d "
                 "(use ddpb_toy_step with caller-provided windows beyond that)", nranks, DDPB_MAX_VIRTUAL_RANKS);
    unsigned char *base = nullptr;
    if (int rc = virtual_window_base(&base)) return rc;
    ddpb_toy_job local[DDPB_MAX_VIRTUAL_RANKS];
    for (int r = 0; r < nranks; ++r) {
        DDPB_REQUIRE(jobs[r].mode == DDPB_MODE_STEP, "ddpb_toy_step_virtual_ranks: replica %d is not in STEP mode", r);
        local[r] = jobs[r];
        local[r].world = nranks;
        local[r].rank = r;
        local[r].inv_world = 1.0f / (float)nranks;
        for (int p = 0; p < DDPB_MAX_PEERS; ++p) {
            local[r].peer_data[p] = p < nranks ? reinterpret_cast<float *>(base + p * kWindowBytes) : nullptr;
            local[r].peer_flags[p] = nullptr;
        }
    }
    DDPB_CUDA(cudaMemsetAsync(base, 0, nranks * kWindowBytes, (cudaStream_t)stream));     // stale tickets must not match
    return launch_toy_step(local, nranks, (cudaStream_t)stream);
}
