/*
 * ddp_b200.h -- C ABI of libddp_b200.so: the B200 (sm_100a) implementation of the
 * data-parallel training step of ammunk/distributed-training-pytorch's toy model.
 *
 * Conventions
 *   - every entry point returns 0 on success; non-zero = failure, text in
 *     ddpb_last_error() (thread-local).  CUDA failures return 1000+cudaError_t,
 *     NCCL failures 2000+ncclResult_t, argument errors 1..99.
 *   - the caller (PyTorch) owns every tensor buffer: pointers are raw device (or,
 *     where stated, host) addresses, sizes are element counts, streams are
 *     cudaStream_t passed as void*.  The library owns only the opaque handles it
 *     creates (NCCL communicator, peer workspaces, TMA descriptors).
 *   - fp32 storage everywhere, exactly like the reference's models.
 *
 * The reference has no FFI of its own (it is 6 Python files that call PyTorch);
 * each declaration cites the reference line whose work it replaces.  Paths
 * without a prefix are relative to the reference repository root; "torch:" paths
 * are relative to site-packages/torch 2.11.0.
 */
#ifndef DDP_B200_H
#define DDP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDPB_VERSION 100           /* 0.1.0 */
#define DDPB_TOY_HIDDEN 10         /* toy_model_and_data.py:13-21 */
#define DDPB_TOY_PARAMS 371        /* 3H^2 + 7H + 1 */
#define DDPB_MAX_JOBS 4            /* models per ddpb_toy_step launch */
#define DDPB_MAX_PEERS 8           /* GPUs of one NVSwitch box taking part in the fused exchange */
#define DDPB_XCHG_ROW 376          /* 8-byte {value, ticket} slots per (parity, source rank) row (372 used) */
#define DDPB_XCHG_WINDOW_BYTES (2 * DDPB_MAX_PEERS * DDPB_XCHG_ROW * 8)   /* one model's window per rank: 48 128 B */
#define DDPB_LOSS_TIMEOUT_BITS 0x7FC0DEAEu       /* loss-slot value (NaN payload) after a peer time-out */

/* ---- introspection ------------------------------------------------------- */
int ddpb_version(void);
const char *ddpb_last_error(void);
/* device properties the host layer sizes its grids with (148 SMs on B200). */
int ddpb_device_info(int device, int *cc_major, int *cc_minor, int *sm_count);
/* kernels this library has launched in this process so far (every launch site counts itself);
 * bench.py reports the difference over its timed region as gpu_launches. */
long long ddpb_launch_count(void);

/* ---- sampler: DistributedSampler(dataset, shuffle=True) ------------------- */
/* demo.py:144 builds it, demo.py:96-98 reseeds it per epoch.  Host code, bit-exact
 * restatement of torch.randperm on the CPU generator (mt19937, forward
 * Fisher-Yates) and of torch:utils/data/distributed.py:84-141. */
int ddpb_randperm(uint64_t seed, int64_t n, int64_t *out /* host [n] */);
/* returns the number of indices written (ceil(n/world) without drop_last) or -1 */
int64_t ddpb_sampler_indices(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                             uint64_t epoch, int drop_last, int64_t *out /* host */, int64_t cap);
/* The same stream for epochs [epoch0, epoch0+nepochs) as int32 rows of `stride`
 * entries (row e holds the indices of epoch0+e), ready to be uploaded once for the
 * device-resident data path. */
int ddpb_sampler_epochs_i32(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                            uint64_t epoch0, int nepochs, int drop_last,
                            int32_t *out /* host [nepochs*stride] */, int64_t stride);

/* int64 variant for the host loader's epoch cache. */
int ddpb_sampler_epochs_i64(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                            uint64_t epoch0, int nepochs, int drop_last,
                            int64_t *out /* host [nepochs*stride] */, int64_t stride);

/* ---- batch assembly ------------------------------------------------------- */
/* Host: replaces ToyData.__getitem__ x B + default_collate (toy_model_and_data.py:32-33,
 * demo.py:99) by one gather from the dense dataset into (pinned) staging rows. */
int ddpb_host_gather(const float *X, const float *Y, const int64_t *idx, int64_t B, int dx, int dy,
                     float *outX, float *outY);
/* Device: dataset resident in HBM, batch = rows idx[0..B) (replaces the per-step
 * H2D copies of demo.py:102-103). */
int ddpb_gather_batch(const float *X, const float *Y, const int32_t *idx, int64_t B, int dx, int dy,
                      float *outX, float *outY, void *stream);

/* ---- the fused toy-model step (H = 10) ------------------------------------ */
/* One launch runs, for up to DDPB_MAX_JOBS independent models, what demo.py:104-111
 * spends ~70 library kernels on: ToyModel.forward (toy_model_and_data.py:24-25),
 * nn.MSELoss (demo.py:106,108), loss.backward() (demo.py:107,109) and, in STEP mode,
 * optim.Adam.step (demo.py:110-111) -- one 8-CTA thread-block cluster per model,
 * gradients reduced through distributed shared memory. */
enum {
    DDPB_MODE_FWD = 0,    /* out[B,1] = model(x) */
    DDPB_MODE_GRAD = 1,   /* grads[P] = d(mean loss)/dparams (local batch), grads[P] = sum (o-t)^2 */
    DDPB_MODE_STEP = 2    /* GRAD + in-place Adam/SGD update of params (single-rank fused path) */
};
enum { DDPB_OPT_ADAM = 0, DDPB_OPT_SGD = 1 };

/* Device-resident optimizer counters (32 bytes, zero-initialised by the host layer).
 * Kept on the device so a captured CUDA graph can replay steps without new arguments. */
typedef struct ddpb_opt_state {
    long long step;    /* optimizer steps applied so far */
    long long seq;     /* gradient evaluations so far (selects loss slot / index row) */
    long long reserved[2];
} ddpb_opt_state;

/* Adam's bias corrections are Python doubles in torch (torch:optim/adam.py:531-536).  The host
 * layer tabulates them once per optimizer: entry t-1 = { (float)-(lr / (1 - beta1^t)),
 * (float)sqrt(1 - beta2^t) } for step t = 1..n; the table must run until both have converged
 * (t > n reuses entry n-1).  Kernels index it with the device-resident step counter. */

typedef struct ddpb_toy_job {
    float *params;            /* [371] flat, module.parameters() order: W0,b0,...,W4,b4 */
    float *grads;             /* [372] flat gradient bucket; element 371 carries sum (o-t)^2 */
    float *m, *v;             /* Adam moments [371] (m = SGD momentum buffer); STEP only */
    ddpb_opt_state *state;    /* device counters */
    const float *bias_table;  /* Adam: [bias_table_n][2] bias-correction scalars (see above), 8-byte aligned */
    float *loss_ring;         /* [loss_ring_n] mean-loss slots, slot = seq % loss_ring_n */
    const float *x;           /* [B,2] batch, or the whole dataset when idx != NULL */
    const float *y;           /* [B,1] targets, or the whole dataset */
    const int32_t *idx;       /* NULL, or [idx_rows][idx_stride] sample indices; row = seq % idx_rows */
    float *out;               /* FWD: [B,1] */
    float *global_loss_ring;  /* optional, world > 1 STEP: [loss_ring_n] slots for the mean loss over ALL ranks,
                               * sum_r sum_b (o-t)^2 / (B * world) -- the value demo.py:112-121 logs after two gloo
                               * all-reduces, here a by-product of the fused exchange (bucket element 371) */
    int32_t B;                /* local batch (1..65536) */
    int32_t mode;             /* DDPB_MODE_* */
    int32_t opt;              /* DDPB_OPT_* */
    int32_t loss_ring_n;
    int32_t idx_rows, idx_stride;
    int32_t bias_table_n;
    int32_t n_data;           /* rows of the dataset behind idx (0 = unknown): <= 1024 rows are staged on chip */
    double lr, beta1, beta2, eps, momentum;
    /* world > 1 with DDPB_MODE_STEP: the gradient all-reduce (demo.py:107,109 via the c10d reducer)
     * is fused into the step as a one-shot exchange over peer-mapped windows (NVLink P2P stores):
     * peer_data[r] is rank r's window as mapped into THIS process (ddpb_peer_alloc /
     * ddpb_peer_open), zero-initialised, DDPB_XCHG_WINDOW_BYTES bytes, 8-byte aligned; entry [rank]
     * is this rank's own window.  state->seq must advance in lock step on all ranks;
     * A peer that has not delivered its slices within ~2 s (global timer) makes the kernel give up
     * loudly: state->reserved[0] = 1, the step's loss slot holds DDPB_LOSS_TIMEOUT_BITS (a NaN), and
     * params / m / v / step / seq are left as they were (the step can be repeated).  Callers must look
     * at one of the two; the Python layer raises.  peer_flags is unused (reserved). */
    float *peer_data[DDPB_MAX_PEERS];
    uint32_t *peer_flags[DDPB_MAX_PEERS];
    int32_t world, rank;
    float inv_world;          /* 1/world (DDP's div factor) */
    int32_t reserved2;
} ddpb_toy_job;

int ddpb_toy_step(const ddpb_toy_job *jobs /* host */, int njobs, void *stream);
/* The same launch preceded, on the same stream, by one pinned-host -> device copy: the
 * `data.cuda(); target.cuda()` of demo.py:102-103 and the step in one host call. */
int ddpb_toy_step_h2d(const ddpb_toy_job *jobs, int njobs, void *dst_dev, const void *src_host,
                      int64_t bytes, void *stream);

/* One host call per iteration of demo.py:99-118 (batch assembly, the step, the blocking loss read):
 *   1. idx != NULL: gather rows idx[0..B) of the dense host dataset (X [n,dx], Y [n,dy]) into the pinned
 *      staging rows out_x / out_y (ddpb_host_gather; replaces ToyData.__getitem__ x B + default_collate);
 *   2. launch ddpb_toy_step(jobs, njobs) on `stream` -- a job whose x / y point at that pinned staging
 *      reads its batch over PCIe itself (mapped pinned memory), which replaces the two `.cuda()` copies
 *      of demo.py:102-103: no copy-engine operation, no second host call;
 *   3. nwait > 0: poll the pinned loss slots wait_slot[0..nwait) (ddpb_wait_f32 semantics: sentinel,
 *      spin_ns, then a stream synchronisation) and return their values in wait_value -- the
 *      loss.detach().cpu() of demo.py:115,117.
 * The struct lives in host memory owned by the caller and may be reused from call to call. */
typedef struct ddpb_toy_iter {
    const ddpb_toy_job *jobs;
    void *stream;
    const float *X, *Y;             /* dense host dataset (gather source) */
    const int64_t *idx;             /* NULL: no gather */
    float *out_x, *out_y;           /* pinned staging rows */
    const float *wait_slot[2 * DDPB_MAX_JOBS];
    int64_t B, spin_ns;
    int32_t njobs, dx, dy, nwait;
    uint32_t sentinel_bits;
    float wait_value[2 * DDPB_MAX_JOBS];
    int32_t reserved;
} ddpb_toy_iter;
int ddpb_toy_iteration(ddpb_toy_iter *it /* host */);

/* No counterpart in the reference.  `nranks` replicas (2..DDPB_MAX_VIRTUAL_RANKS) of one model
 * on the CURRENT GPU take one fused data-parallel step in a single launch: jobs[r] is replica
 * r's job (DDPB_MODE_STEP, its own params/m/v/grads/state and its own batch; all replicas'
 * state->seq must be equal).  Unlike ddpb_toy_step with world > 1, the caller provides no
 * exchange windows and nothing is allocated, by this call or ever: the windows are a
 * 188 KiB array compiled into the library (every GPU that loads it holds one until exit).
 * It is meant for callers that cannot allocate device memory themselves -- the device was
 * handed to another pool, or the loop must not take the allocator's device-wide sync.
 * world/rank/inv_world/peer_* of the jobs are ignored and filled in here.
 * The array is shared by every call and keeps the previous call's contents: the windows in use
 * are cleared on `stream` first, and calls whose kernels overlap on the GPU race -- serialise
 * them.  The first call on each device looks the array's device address up; make that call
 * before starting any stream capture.  (Host threads that make that first call at the same time all
 * look up and record the same address, which is harmless; after cudaDeviceReset the recorded address
 * is stale, so do not use this entry point on a device the process has reset.)  A replica that does
 * not see the others' slices within ~2 s gives up as described for ddpb_toy_step: check
 * state->reserved[0] of the replicas.  More than DDPB_MAX_VIRTUAL_RANKS replicas is an error. */
#define DDPB_MAX_VIRTUAL_RANKS 4
int ddpb_toy_step_virtual_ranks(const ddpb_toy_job *jobs /* host */, int nranks, void *stream);

/* ---- a pipeline stage of the split model ------------------------------------------- */
/* demo_one_model_multi_gpu.py:17-42: MultiGPUModel.layers0 (2->10->10, both activated) on dev0,
 * layers1 (10->10->10->1) on dev1, activation moved with .to(dev1) and its gradient back through
 * autograd.  One launch runs a stage: FWD (out = stage(x)), or GRAD/STEP (forward recomputed,
 * backward from targets `y` when has_loss, else from the upstream gradient `d_out`; parameter
 * gradients to `grads`, optional input gradient to `d_in`, STEP also applies Adam/SGD).  `out`
 * and `d_in` may lie in the other GPU's memory (ddpb_enable_peer): the transfer is the kernel's
 * own stores over NVLink.  Launch with the stage's device current. */
#define DDPB_SEG_MAX_LAYERS 3
#define DDPB_SEG_MAX_WIDTH 10
typedef struct ddpb_seg_job {
    float *params, *grads;    /* flat [P], [P+1] (grads[P] = sum (o-t)^2 when has_loss) */
    float *m, *v;
    ddpb_opt_state *state;
    const float *bias_table;
    float *loss_ring;
    const float *x;           /* [B, dims[0]] */
    const float *y;           /* [B, dims[nl]] targets (has_loss) */
    const float *d_out;       /* [B, dims[nl]] upstream gradient (!has_loss) */
    float *out;               /* FWD: [B, dims[nl]] */
    float *d_in;              /* optional [B, dims[0]] */
    int32_t B, mode, opt, nl;
    int32_t dims[DDPB_SEG_MAX_LAYERS + 1];
    int32_t act_last;         /* LeakyReLU after the last layer too (stage 0) */
    int32_t has_loss;         /* the stage ends in nn.MSELoss (demo_one_model_multi_gpu.py:107,126) */
    int32_t loss_ring_n, bias_table_n;
    double lr, beta1, beta2, eps, momentum;
} ddpb_seg_job;
int ddpb_segment_step(const ddpb_seg_job *job /* host */, void *stream);

/* ---- widened model: tensor-core linear layers (hidden width a multiple of 4, >= 32) ---- */
/* The H x H nn.Linear layers of the widened toy model (toy_model_and_data.py:13-21 with the
 * widths scaled up, BASELINE.json configs[4]) as tcgen05 kind::tf32 GEMMs fed by TMA; PyTorch
 * runs them as cuBLAS sgemm in forward and in autograd (demo.py:104-109).  fp32 storage, tf32
 * multiplication, fp32 accumulation.  B = local batch.  Transposed copies ([H, ld_t], ld_t >= B,
 * ld_t % 4 == 0) are written by the producing epilogues so that every wgrad operand is K-major.
 *   fwd:   act_out[B,Hout] = act(act_in[B,Hin] W^T + bias); act_out_t (optional) = its transpose
 *   dgrad: dz_prev[B,Hin] = (dz[B,Hout] W) * LeakyReLU'(act_prev); dz_prev_t = its transpose
 *   wgrad: dW[Hout,Hin] = dz_t[Hout,:B] act_t_ext[:Hin,:B]^T; db[Hout] = dz_t . ones, where
 *          act_t_ext has Hin+1 rows and row Hin is all ones */
int ddpb_linear_fwd(const float *act_in, const float *W, const float *bias, float *act_out, float *act_out_t,
                    int B, int Hin, int Hout, int ld_t, int apply_act, void *stream);
int ddpb_linear_dgrad(const float *dz, const float *W, const float *act_prev, float *dz_prev, float *dz_prev_t,
                      int B, int Hin, int Hout, int ld_t, void *stream);
int ddpb_linear_wgrad(const float *dz_t, const float *act_t_ext, float *dW, float *db, int B, int Hin, int Hout,
                      int ld_t, void *stream);

/* wgrad of one layer with Adam applied in its epilogue (one rank, no gradient exchange): replaces, for that layer,
 * loss.backward()'s weight-gradient GEMM (demo.py:107,109) AND optim.Adam.step (demo.py:110-111).  The gradient tile goes
 * from the tensor cores' accumulator through registers into the update; p / m / v ([Hout][Hin] views of the flat buffers,
 * 16-byte aligned; pb / mb / vb = the layer's bias slices) are read and written once, dW and db are still stored (28
 * instead of 32 B/param of HBM traffic and no optimizer launch for the layer).  Same arithmetic, bit for bit, as
 * ddpb_linear_wgrad followed by ddpb_fused_avg_adam_dev with grad_scale 1.  The step counter is read from `state`
 * (bias-correction row) and NOT advanced: advance it once per step, after every layer.  The caller orders this call
 * after the dgrad that still reads the layer's old weights. */
int ddpb_linear_wgrad_adam(const float *dz_t, const float *act_t_ext, float *dW, float *db, float *p, float *pb,
                           float *m, float *mb, float *v, float *vb, int B, int Hin, int Hout, int ld_t,
                           double beta1, double beta2, double eps, const float *bias_table, int bias_table_n,
                           const ddpb_opt_state *state, void *stream);

/* The thin ends of the widened model (first Linear 2->H + LeakyReLU, toy_model_and_data.py:13-14;
 * last Linear H->1, :21; nn.MSELoss and its backward, demo.py:106-109): streaming kernels, B <= 256.
 * idx (optional) gathers rows of x / y from the HBM-resident dataset. */
int ddpb_wide_first_fwd(const float *x, const int32_t *idx, const float *W0, const float *b0, float *act, float *act_t,
                        int B, int H, int ld_t, void *stream);
int ddpb_wide_last_fwd(const float *act3, const float *W4, const float *b4, const float *y, const int32_t *idx,
                       float *out, float *dz4, float *sq, int B, int H, void *stream);
int ddpb_wide_last_bwd(const float *act3, const float *W4, const float *dz4, const float *sq, float *dz3, float *dz3_t,
                       float *gW4, float *gb4, float *sq_sum_out, float *loss_slot, int B, int H, int ld_t, void *stream);
int ddpb_wide_first_bwd(const float *dz0, const float *x, const int32_t *idx, float *gW0, float *gb0, int B, int H,
                        void *stream);

/* ---- optimizer over a flat bucket ----------------------------------------- */
/* One pass over a (reduced) gradient bucket: g = bucket * grad_scale (the 1/W of
 * DDP's averaging, torch:distributed/algorithms/ddp_comm_hooks/default_hooks.py:24-29),
 * then torch's Adam op order (torch:optim/adam.py:457-547; demo.py:80-81,110-111).
 * 28 B/param of HBM traffic.  `step` is the 1-based step being applied; the bias
 * corrections are computed by the host in double, as torch does. */
int ddpb_fused_avg_adam(const float *bucket, float *p, float *m, float *v, int64_t n,
                        float grad_scale, double lr, double beta1, double beta2, double eps,
                        int64_t step, void *stream);
/* Same with the device-resident step counter and the bias table (graph-replayable);
 * advance != 0 increments state->step afterwards. */
int ddpb_fused_avg_adam_dev(const float *bucket, float *p, float *m, float *v, int64_t n,
                            float grad_scale, double beta1, double beta2, double eps,
                            const float *bias_table, int bias_table_n,
                            ddpb_opt_state *state, int advance, void *stream);
/* Same pass with the grid capped at `ctas_per_sm` (1..8) 256-thread CTAs per SM: a pass that runs BESIDE the
 * backward GEMMs of other layers (its own stream) must leave thread slots for their CTAs, or the GEMM
 * launches queue behind it; 2 per SM still keeps enough loads in flight to use the HBM bandwidth that is left. */
int ddpb_fused_avg_adam_dev_grid(const float *bucket, float *p, float *m, float *v, int64_t n,
                                 float grad_scale, double beta1, double beta2, double eps,
                                 const float *bias_table, int bias_table_n,
                                 ddpb_opt_state *state, int advance, int ctas_per_sm, void *stream);
/* SGD with optional momentum (north_star lists it beside Adam; torch:optim/sgd.py). */
int ddpb_fused_avg_sgd(const float *bucket, float *p, float *buf, int64_t n, float grad_scale,
                       double lr, double momentum, int64_t step, void *stream);
/* Same with the device-resident counter: state->step == 0 means "first step" (the momentum buffer is
 * initialised with the gradient), exactly as the fused toy step reads it, so eager and fused steps of
 * one model can be mixed; advance != 0 increments state->step afterwards. */
int ddpb_fused_avg_sgd_dev(const float *bucket, float *p, float *buf, int64_t n, float grad_scale,
                           double lr, double momentum, ddpb_opt_state *state, int advance, void *stream);

/* ---- bucket all-reduce fused with the optimizer, over NVLink peer windows -------------------
 * Replaces, for one gradient bucket of a large model, the c10d Reducer's ncclAllReduce
 * (demo.py:107,109 via torch:include/torch/csrc/distributed/c10d/reducer.hpp:276-286) AND
 * optim.Adam.step on that bucket (demo.py:110-111) by ONE kernel per rank:
 *   1. wait until every rank has produced the bucket (a flag per rank in the peer windows),
 *   2. reduce-scatter by pull: rank r sums shard r of the bucket from every rank's gradient buffer
 *      (16-byte system-scope loads over NVLink), scales by grad_scale = 1/W,
 *   3. Adam on shard r only (this rank's m and v hold state for its shards; the other entries are
 *      never touched and stay zero),
 *   4. all-gather by push: the new parameters of shard r are stored into every rank's parameter
 *      buffer, then a "done" flag; the kernel ends when every rank's "done" has arrived, so stream
 *      order makes the next forward see complete parameters.
 * Every rank's window has the same layout: [flags DDPB_BUCKET_FLAGS_BYTES | gradients | parameters]
 * at the byte offsets given here; windows come from ddpb_peer_alloc / ddpb_peer_open and start
 * zeroed.  `ticket` must grow from call to call for a given bucket (the step sequence number).
 * mode 1 = reduce over the peers; mode 0 = the local gradient buffer already holds the sum
 * (ncclAllReduce ran, eager backward) and only steps 3-4 are done.  A peer that does not arrive
 * at step 1 within ~5 s sets state->reserved[0] = 1 and the kernel leaves without touching
 * parameters or optimizer state; callers must look at that flag (the Python layer raises).  The optimizer counters are
 * read from `state`; with advance != 0 they are incremented after the pass (last bucket of a step). */
#define DDPB_BUCKET_MAX 1024                                         /* buckets per model */
#define DDPB_BUCKET_FLAGS_BYTES (2 * DDPB_BUCKET_MAX * DDPB_MAX_PEERS * 4 + DDPB_BUCKET_MAX * 4)
typedef struct ddpb_bucket_job {
    void *window[DDPB_MAX_PEERS];       /* every rank's window as mapped in this process; [rank] is the own one */
    int64_t grads_off, params_off;      /* byte offsets of the flat gradient / parameter buffers in a window (16-byte aligned) */
    float *m, *v;                       /* local Adam state, flat, indexed like the parameters */
    const float *bias_table;            /* [bias_table_n][2], as for ddpb_fused_avg_adam_dev */
    ddpb_opt_state *state;
    int64_t start, n;                   /* the bucket: elements [start, start + n) of the flat buffers, start % 4 == 0 */
    int64_t ticket;
    int32_t bias_table_n, world, rank, bucket_index, mode, advance;
    double beta1, beta2, eps;
    float grad_scale;                   /* 1 / world */
    int32_t reserved;                   /* 0.  Diagnostic bits for tools/bench_wide.py (results are then WRONG on
                                         * purpose): 1 = read the local buffer instead of the peers', 2 = do not
                                         * push parameters, bits 8..23 = cap on the grid */
} ddpb_bucket_job;
int ddpb_bucket_reduce_adam(const ddpb_bucket_job *job, void *stream);



/* ---- communicator: NCCL over NVLink --------------------------------------- */
/* Replaces the c10d Reducer's ncclAllReduce per bucket (demo.py:107,109 via
 * torch:include/torch/csrc/distributed/c10d/reducer.hpp) and DDP's construction-time
 * parameter broadcast (demo.py:70,72; torch:nn/parallel/distributed.py:879-890). */
int ddpb_nccl_load(const char *path_or_null);           /* dlopen the NCCL torch already loaded */
int ddpb_get_unique_id(char out[128]);                  /* rank 0, then shared through the c10d store */
int ddpb_comm_create(const char id[128], int rank, int world, int device, void **comm_out);
int ddpb_comm_destroy(void *comm);
int ddpb_broadcast_f32(void *comm, float *buf, int64_t n, int root, void *stream);
int ddpb_allreduce_sum_f32(void *comm, float *buf, int64_t n, void *stream);

/* ---- peer windows (same box, one process per GPU or one process, several GPUs) ---- */
/* cudaMalloc'ed + zeroed window of `bytes` on the current device and its 64-byte IPC handle;
 * other processes map it with ddpb_peer_open (peer access is enabled lazily by the driver),
 * the same process on another GPU with ddpb_enable_peer + the raw pointer.  The library owns
 * these allocations (they cannot come from PyTorch's caching allocator: IPC handles cover
 * whole cudaMalloc blocks). */
int ddpb_peer_alloc(int64_t bytes, void **ptr_out, char handle_out[64]);
int ddpb_peer_open(const char handle[64], void **ptr_out);
int ddpb_peer_close(void *mapped_ptr);
int ddpb_peer_free(void *ptr);
/* cudaDeviceEnablePeerAccess in both directions (model-parallel demo: demo_one_model_multi_gpu.py:41). */
int ddpb_enable_peer(int dev_a, int dev_b);

/* ---- streams / events helpers the host layer needs without torch calls ----- */
int ddpb_stream_sync(void *stream);
/* Blocks until *host_slot (pinned, device-mapped memory a kernel writes, e.g. a loss_ring slot)
 * no longer holds `sentinel_bits`; after spin_ns of polling it synchronises the stream
 * instead.  Replaces the blocking loss.detach().cpu() of demo.py:115,117. */
int ddpb_wait_f32(const float *host_slot, uint32_t sentinel_bits, int64_t spin_ns, void *stream, float *out);
int ddpb_memcpy_h2d_async(void *dst_dev, const void *src_host, int64_t bytes, void *stream);
int ddpb_memcpy_d2h_async(void *dst_host, const void *src_dev, int64_t bytes, void *stream);

/* ---- CUDA graphs --------------------------------------------------------------- */
/* The reference drives 1000 iterations from Python (demo.py:88-129).  Steps whose inputs are
 * already in HBM can instead be captured once (begin ... launches on `stream` ... end) and
 * replayed: every kernel above takes its step / row counters from ddpb_opt_state. */
int ddpb_graph_begin(void *stream);
int ddpb_graph_end(void *stream, void **graph_exec_out);
int ddpb_graph_launch(void *graph_exec, void *stream);
int ddpb_graph_destroy(void *graph_exec);

#ifdef __cplusplus
}
#endif
#endif /* DDP_B200_H */
