/*
 * ddp_oracle.c -- CPU restatement of the DDP toy-model training step.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import, link or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference leg use it, and only as the checker.
 *
 * What it restates.  The reference (ammunk/distributed-training-pytorch) contains
 * no arithmetic of its own: every FLOP is executed by PyTorch, a third-party
 * dependency that is not under /root/reference (pinned torch==1.10.1+cu113 in
 * Pipfile:18; the box and the golden fixtures use torch 2.11.0+cu128).  Each
 * function below restates the published algorithm of the torch routine that the
 * cited reference line calls, in plain scalar C (fp32 storage, fp32 arithmetic,
 * bias corrections in double exactly like torch's Python code).
 *
 * Parity pin: tests/test_oracle_golden.py checks this file against the fixtures in
 * tests/golden/, which were produced by running the reference's own demo.py under
 * the CPU/gloo shim in tests/golden/make_golden.py (sampler indices and
 * permutations bit-exact; loss curves / final parameters within the fp32
 * tolerance written in that test).
 *
 * "torch:" paths are relative to site-packages/torch (2.11.0).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- */
/* MT19937 + randperm: torch.randperm(n, generator=g) on CPU, as used by       */
/* DistributedSampler.__iter__ (torch:utils/data/distributed.py:107-112),       */
/* which the reference builds at demo.py:144 and reseeds at demo.py:98.         */
/* at::mt19937 is the textbook 32-bit Mersenne Twister seeded with              */
/* init_genrand(seed & 0xffffffff); randperm_cpu for n < 2^32/20 is the         */
/* forward Fisher-Yates  swap(r[i], r[i + random() % (n - i)]), i = 0..n-2.     */
/* ------------------------------------------------------------------------- */
typedef struct { uint32_t s[624]; int pos; } mt_t;

static void mt_seed(mt_t *g, uint64_t seed)
{
    g->s[0] = (uint32_t)(seed & 0xffffffffu);
    for (int j = 1; j < 624; ++j)
        g->s[j] = 1812433253u * (g->s[j - 1] ^ (g->s[j - 1] >> 30)) + (uint32_t)j;
    g->pos = 624;
}

static uint32_t mt_next(mt_t *g)
{
    if (g->pos >= 624) {
        for (int k = 0; k < 624; ++k) {
            uint32_t y = (g->s[k] & 0x80000000u) | (g->s[(k + 1) % 624] & 0x7fffffffu);
            g->s[k] = g->s[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        g->pos = 0;
    }
    uint32_t y = g->s[g->pos++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

ORACLE_API int oracle_randperm(uint64_t seed, int64_t n, int64_t *out)
{
    if (n < 0 || n >= (int64_t)(0xffffffffu / 20)) return -1; /* only the small-n branch is restated */
    mt_t g;
    mt_seed(&g, seed);
    for (int64_t i = 0; i < n; ++i) out[i] = i;
    for (int64_t i = 0; i + 1 < n; ++i) {
        int64_t z = (int64_t)(mt_next(&g) % (uint64_t)(n - i));
        int64_t t = out[i]; out[i] = out[z + i]; out[z + i] = t;
    }
    return 0;
}

/* DistributedSampler index stream for one rank and one epoch.
 * torch:utils/data/distributed.py:84-141 (num_samples :90-100, shuffle :107-112,
 * padding :116-125, drop_last tail :127, subsample rank::world :134).
 * Reference call sites: demo.py:144 (shuffle=True, seed 0, drop_last False),
 * demo.py:96-98 (set_epoch).  Returns the number of indices written, or -1. */
ORACLE_API int64_t oracle_sampler_indices(int64_t n, int world, int rank, int shuffle,
                                          uint64_t seed, uint64_t epoch, int drop_last,
                                          int64_t *out, int64_t out_cap)
{
    if (n <= 0 || world <= 0 || rank < 0 || rank >= world) return -1;
    int64_t num_samples;
    if (drop_last && n % world != 0) num_samples = (n - world + world - 1) / world; /* ceil((n-world)/world) */
    else num_samples = (n + world - 1) / world;
    int64_t total = num_samples * world;
    if (num_samples > out_cap) return -1;
    int64_t *perm = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    if (!perm) return -1;
    if (shuffle) { if (oracle_randperm(seed + epoch, n, perm)) { free(perm); return -1; } }
    else for (int64_t i = 0; i < n; ++i) perm[i] = i;
    /* padding wraps around the (shuffled) list; with drop_last the list is cut at total <= n */
    for (int64_t k = 0; k < num_samples; ++k) {
        int64_t pos = rank + k * world;          /* position in the padded list */
        if (pos >= total) { free(perm); return -1; }
        out[k] = perm[pos % n];
    }
    free(perm);
    return num_samples;
}

/* ------------------------------------------------------------------------- */
/* MLP: nn.Linear / nn.LeakyReLU chain, toy_model_and_data.py:10-25.            */
/* Linear: y = x W^T + b with W row-major [out,in] (torch:nn/modules/linear.py  */
/* :117-128 + F.linear); LeakyReLU slope 0.01 (torch:nn/modules/activation.py   */
/* LeakyReLU default negative_slope=1e-2): x > 0 ? x : 0.01 x.                  */
/* Parameters are one flat fp32 vector in module.parameters() order:            */
/* W0,b0,W1,b1,...  dims[] has nl+1 entries.  Every layer but the last is        */
/* followed by LeakyReLU; act_last!=0 also activates the last one (stage 0 of   */
/* MultiGPUModel, demo_one_model_multi_gpu.py:24-29).                            */
/* ------------------------------------------------------------------------- */
#define LRELU_SLOPE 0.01f

static int64_t param_count(const int *dims, int nl)
{
    int64_t p = 0;
    for (int l = 0; l < nl; ++l) p += (int64_t)dims[l + 1] * dims[l] + dims[l + 1];
    return p;
}

ORACLE_API int64_t oracle_param_count(const int *dims, int nl) { return param_count(dims, nl); }

/* pre[l] holds the pre-activation of layer l ([B,dims[l+1]]), post[l] the layer's
 * output after the (optional) activation.  in0 is [B,dims[0]]. */
static void forward(const float *params, const int *dims, int nl, int act_last,
                    const float *in0, int B, float **pre, float **post)
{
    const float *w = params;
    const float *in = in0;
    for (int l = 0; l < nl; ++l) {
        int di = dims[l], dout = dims[l + 1];
        const float *bias = w + (int64_t)dout * di;
        for (int b = 0; b < B; ++b)
            for (int j = 0; j < dout; ++j) {
                float acc = bias[j];
                for (int k = 0; k < di; ++k) acc = fmaf(w[(int64_t)j * di + k], in[(int64_t)b * di + k], acc);
                pre[l][(int64_t)b * dout + j] = acc;
                int act = (l + 1 < nl) || act_last;
                post[l][(int64_t)b * dout + j] = act ? (acc > 0.f ? acc : acc * LRELU_SLOPE) : acc;
            }
        in = post[l];
        w = bias + dout;
    }
}

/* Backward of the chain given dOut = dL/d(post[nl-1]).  Writes parameter
 * gradients (same flat layout) and, when dIn != NULL, dL/d(in0).
 * Linear backward: dW = dY^T X, db = sum_B dY, dX = dY W (autograd of F.linear);
 * leaky_relu_backward: g * (x > 0 ? 1 : 0.01) on the pre-activation. */
static void backward(const float *params, const int *dims, int nl, int act_last,
                     const float *in0, int B, float **pre, float **post,
                     const float *dOut, float *grads, float *dIn)
{
    int maxw = 0;
    for (int l = 0; l <= nl; ++l) if (dims[l] > maxw) maxw = dims[l];
    float *dcur = (float *)malloc(sizeof(float) * (size_t)B * maxw);
    float *dnext = (float *)malloc(sizeof(float) * (size_t)B * maxw);
    memcpy(dcur, dOut, sizeof(float) * (size_t)B * dims[nl]);
    /* offsets of each layer inside the flat vector */
    int64_t *off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nl + 1));
    off[0] = 0;
    for (int l = 0; l < nl; ++l) off[l + 1] = off[l] + (int64_t)dims[l + 1] * dims[l] + dims[l + 1];
    for (int l = nl - 1; l >= 0; --l) {
        int di = dims[l], dout = dims[l + 1];
        const float *w = params + off[l];
        float *gw = grads + off[l];
        float *gb = gw + (int64_t)dout * di;
        const float *in = (l == 0) ? in0 : post[l - 1];
        int act = (l + 1 < nl) || act_last;
        if (act)
            for (int64_t i = 0; i < (int64_t)B * dout; ++i)
                dcur[i] = pre[l][i] > 0.f ? dcur[i] : dcur[i] * LRELU_SLOPE;
        for (int j = 0; j < dout; ++j) {
            float sb = 0.f;
            for (int b = 0; b < B; ++b) sb += dcur[(int64_t)b * dout + j];
            gb[j] = sb;
            for (int k = 0; k < di; ++k) {
                float s = 0.f;
                for (int b = 0; b < B; ++b) s = fmaf(dcur[(int64_t)b * dout + j], in[(int64_t)b * di + k], s);
                gw[(int64_t)j * di + k] = s;
            }
        }
        if (l > 0 || dIn) {
            float *dst = (l == 0) ? dIn : dnext;
            for (int b = 0; b < B; ++b)
                for (int k = 0; k < di; ++k) {
                    float s = 0.f;
                    for (int j = 0; j < dout; ++j) s = fmaf(dcur[(int64_t)b * dout + j], w[(int64_t)j * di + k], s);
                    dst[(int64_t)b * di + k] = s;
                }
            if (l > 0) { float *t = dcur; dcur = dnext; dnext = t; }
        }
    }
    free(off); free(dcur); free(dnext);
}

static float **alloc_acts(const int *dims, int nl, int B)
{
    float **a = (float **)malloc(sizeof(float *) * (size_t)nl);
    for (int l = 0; l < nl; ++l) a[l] = (float *)malloc(sizeof(float) * (size_t)B * dims[l + 1]);
    return a;
}
static void free_acts(float **a, int nl) { for (int l = 0; l < nl; ++l) free(a[l]); free(a); }

/* Forward only: out is [B, dims[nl]]. */
ORACLE_API void oracle_mlp_forward(const float *params, const int *dims, int nl, int act_last,
                                   const float *x, int B, float *out)
{
    float **pre = alloc_acts(dims, nl, B), **post = alloc_acts(dims, nl, B);
    forward(params, dims, nl, act_last, x, B, pre, post);
    memcpy(out, post[nl - 1], sizeof(float) * (size_t)B * dims[nl]);
    free_acts(pre, nl); free_acts(post, nl);
}

/* nn.MSELoss() (reduction='mean'), demo.py:86,106,108:  mean((o-t)^2) over all
 * B*dout elements; backward 2 (o-t) / numel (ATen mse_loss_backward). */
ORACLE_API float oracle_mse(const float *o, const float *t, int64_t n, float *dO)
{
    float s = 0.f;
    for (int64_t i = 0; i < n; ++i) { float d = o[i] - t[i]; s += d * d; }
    if (dO) {
        float norm = (float)(2.0 / (double)n);
        for (int64_t i = 0; i < n; ++i) dO[i] = norm * (o[i] - t[i]);
    }
    return s / (float)n;
}

/* forward + MSE + backward of the whole chain for one local batch
 * (model(data); loss(out,target); loss.backward(): demo.py:104-109).
 * Returns the loss; grads gets dLoss/dparams of the LOCAL mean loss. */
ORACLE_API float oracle_mlp_loss_grad(const float *params, const int *dims, int nl,
                                      const float *x, const float *y, int B, float *grads)
{
    float **pre = alloc_acts(dims, nl, B), **post = alloc_acts(dims, nl, B);
    forward(params, dims, nl, 0, x, B, pre, post);
    int64_t n = (int64_t)B * dims[nl];
    float *dO = (float *)malloc(sizeof(float) * (size_t)n);
    float loss = oracle_mse(post[nl - 1], y, n, dO);
    backward(params, dims, nl, 0, x, B, pre, post, dO, grads, NULL);
    free(dO); free_acts(pre, nl); free_acts(post, nl);
    return loss;
}

/* A pipeline segment (demo_one_model_multi_gpu.py:39-42): forward of layers
 * [0,nl) of `params`, backward given dOut; used for the 2-stage split. */
ORACLE_API void oracle_segment_backward(const float *params, const int *dims, int nl, int act_last,
                                        const float *x, int B, const float *dOut,
                                        float *grads, float *dIn)
{
    float **pre = alloc_acts(dims, nl, B), **post = alloc_acts(dims, nl, B);
    forward(params, dims, nl, act_last, x, B, pre, post);
    backward(params, dims, nl, act_last, x, B, pre, post, dOut, grads, dIn);
    free_acts(pre, nl); free_acts(post, nl);
}

/* ------------------------------------------------------------------------- */
/* DDP gradient averaging: the Reducer copies grad * (1/W) into the bucket and  */
/* all-reduces with SUM (torch:include/torch/csrc/distributed/c10d/reducer.hpp  */
/* :276 mark_variable_ready_dense; the arithmetic is restated in                 */
/* torch:distributed/algorithms/ddp_comm_hooks/default_hooks.py:24-29).          */
/* Reference call sites: demo.py:70,72 (wrap), demo.py:107,109 (backward).       */
/* rank_grads is [W][n]; summation in ascending rank order.                      */
/* ------------------------------------------------------------------------- */
ORACLE_API void oracle_ddp_average(const float *rank_grads, int world, int64_t n, float *bucket)
{
    float inv = 1.0f / (float)world; /* DDP multiplies by 1/W (div_factor), exact for W = 2^k */
    for (int64_t i = 0; i < n; ++i) {
        float s = 0.f;
        for (int r = 0; r < world; ++r) s += rank_grads[(int64_t)r * n + i] * inv;
        bucket[i] = s;
    }
}

/* ------------------------------------------------------------------------- */
/* optim.Adam(params, lr=1e-3).step()  (demo.py:80-81,110-111): betas            */
/* (0.9,0.999), eps 1e-8, no weight decay, no amsgrad                            */
/* (torch:optim/adam.py:35-49).  Op order of _single_tensor_adam                 */
/* (torch:optim/adam.py:347+): lerp :457, mul/addcmul :476, bias corrections     */
/* :531-536 (Python doubles), denom :545, addcdiv :547.  `step` is the 1-based   */
/* step count after the increment at :413.                                       */
/* ------------------------------------------------------------------------- */
ORACLE_API void oracle_adam_step(float *p, const float *g, float *m, float *v, int64_t n,
                                 double lr, double beta1, double beta2, double eps, int64_t step)
{
    double bc1 = 1.0 - pow(beta1, (double)step);
    double bc2 = 1.0 - pow(beta2, (double)step);
    double step_size = lr / bc1;
    double bc2_sqrt = sqrt(bc2);
    float w1 = (float)(1.0 - beta1);      /* lerp weight, scalar cast to the tensor dtype */
    float b2 = (float)beta2;
    float w2 = (float)(1.0 - beta2);
    float inv_bc2 = (float)bc2_sqrt;
    float eps_f = (float)eps;
    float neg_step = (float)(-step_size);
    for (int64_t i = 0; i < n; ++i) {
        float gi = g[i];
        float mi = fmaf(w1, gi - m[i], m[i]);                 /* exp_avg.lerp_(grad, 1-beta1), weight < 0.5 branch */
        float vi = v[i] * b2;                                 /* exp_avg_sq.mul_(beta2) */
        vi = vi + (w2 * gi) * gi;                             /* .addcmul_(grad, grad, value=1-beta2) */
        float denom = sqrtf(vi) / inv_bc2 + eps_f;            /* (sqrt / bc2_sqrt).add_(eps) */
        p[i] = p[i] + neg_step * (mi / denom);                /* addcdiv_(exp_avg, denom, value=-step_size) */
        m[i] = mi; v[i] = vi;
    }
}

/* optim.SGD(params, lr, momentum).step(): torch:optim/sgd.py _single_tensor_sgd:
 * buf = g on the first step, else buf = momentum*buf + g; p -= lr * buf.
 * (north_star lists SGD beside Adam; the reference demos themselves use Adam.) */
ORACLE_API void oracle_sgd_step(float *p, const float *g, float *buf, int64_t n,
                                double lr, double momentum, int64_t step)
{
    float lr_f = (float)lr, mo = (float)momentum;
    for (int64_t i = 0; i < n; ++i) {
        float d = g[i];
        if (momentum != 0.0) {
            buf[i] = (step == 1) ? d : buf[i] * mo + d;
            d = buf[i];
        }
        p[i] = p[i] - lr_f * d;
    }
}

/* ------------------------------------------------------------------------- */
/* The whole loop of demo.py:87-129 for `world` simulated ranks and `nmodels`    */
/* independent models (demo.py trains two, X and Y), `iters` iterations:        */
/*   per epoch: sampler.set_epoch(epoch) (:98); per batch of `batch` indices:   */
/*   forward/backward per rank (:104-109) -> DDP average -> Adam (:110-111);    */
/*   logged value = sum_r(loss_r * B) / (B * W)  (:114-121).                     */
/* dataloader_mode 0 = 'distributed' (sampler), 1 = 'standard' (every rank sees */
/* the whole dataset in order, demo.py:149-154).                                 */
/* params: [nmodels][P] in/out.  losses: [iters][nmodels] out.                   */
/* ------------------------------------------------------------------------- */
ORACLE_API int oracle_train_loop(float *params, int nmodels, const int *dims, int nl,
                                 const float *data_x, const float *data_y, int64_t ndata,
                                 int world, int batch, int dataloader_mode, int iters,
                                 double lr, float *losses)
{
    int64_t P = param_count(dims, nl);
    int di = dims[0], dout = dims[nl];
    int64_t cap = (ndata + world - 1) / world + 1;
    if (dataloader_mode == 1) cap = ndata;
    int64_t *idx = (int64_t *)malloc(sizeof(int64_t) * (size_t)cap * world);
    float *m = (float *)calloc((size_t)(P * nmodels), sizeof(float));
    float *v = (float *)calloc((size_t)(P * nmodels), sizeof(float));
    float *rg = (float *)malloc(sizeof(float) * (size_t)P * world);
    float *bucket = (float *)malloc(sizeof(float) * (size_t)P);
    float *bx = (float *)malloc(sizeof(float) * (size_t)batch * di);
    float *by = (float *)malloc(sizeof(float) * (size_t)batch * dout);
    int it = 0;
    for (uint64_t epoch = 0; it < iters; ++epoch) {
        int64_t per_rank = 0;
        for (int r = 0; r < world; ++r) {
            if (dataloader_mode == 0)
                per_rank = oracle_sampler_indices(ndata, world, r, 1, 0, epoch, 0, idx + r * cap, cap);
            else {
                per_rank = ndata;
                for (int64_t i = 0; i < ndata; ++i) idx[r * cap + i] = i;
            }
            if (per_rank < 0) return -1;
        }
        for (int64_t start = 0; start < per_rank && it < iters; start += batch, ++it) {
            int B = (int)((per_rank - start) < batch ? (per_rank - start) : batch);
            for (int mdl = 0; mdl < nmodels; ++mdl) {
                float *p = params + (int64_t)mdl * P;
                float logged = 0.f;   /* fp32 CPU tensors summed by the gloo all_reduce (:114-118) */
                for (int r = 0; r < world; ++r) {
                    for (int b = 0; b < B; ++b) {
                        int64_t s = idx[r * cap + start + b];
                        memcpy(bx + (int64_t)b * di, data_x + s * di, sizeof(float) * (size_t)di);
                        memcpy(by + (int64_t)b * dout, data_y + s * dout, sizeof(float) * (size_t)dout);
                    }
                    float l = oracle_mlp_loss_grad(p, dims, nl, bx, by, B, rg + (int64_t)r * P);
                    logged += l * (float)B;
                }
                oracle_ddp_average(rg, world, P, bucket);
                oracle_adam_step(p, bucket, m + (int64_t)mdl * P, v + (int64_t)mdl * P, P,
                                 lr, 0.9, 0.999, 1e-8, (int64_t)it + 1);
                losses[(int64_t)it * nmodels + mdl] = logged / (float)(B * world);
            }
        }
    }
    free(idx); free(m); free(v); free(rg); free(bucket); free(bx); free(by);
    return 0;
}
