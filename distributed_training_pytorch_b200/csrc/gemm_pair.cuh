// Internal interface of gemm_pair.cu (the CTA-pair, weight-stationary tf32 GEMMs) for gemm_tf32.cu.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace ddpb {
namespace pairgemm {

enum { EPI_FWD = 0, EPI_DGRAD = 1, EPI_WGRAD = 2, EPI_WGRAD_ADAM = 3 };

struct EpiParams {
    const float *bias;        // FWD: [M]
    float *bias_grad;         // WGRAD: [M], column n_real of the product
    const float *act_prev;    // DGRAD: [N samples][ld_act] activation whose sign is the LeakyReLU mask
    int ld_act;
    int has_t;                // FWD / DGRAD: also store the neuron-major (transposed) copy
    int apply_act;            // FWD: LeakyReLU after the bias
    int n_real;               // WGRAD: Hin (column Hin of the product is the bias gradient)
    int debug;                // tools/gemm_probe.py only (results are then WRONG on purpose): 1 = load the ring once and
                              // re-multiply it (tensor + shared-memory read rate), 2 = load everything, multiply nothing
};

// Tensor maps are cached per (pointer, shape, pitch, box, swizzle).
int cached_map(CUtensorMap *out, const float *base, long long rows, long long cols, long long ld, int box_rows, int box_cols,
               int swizzle);

int linear_fwd(const float *act_in, const float *W, const float *bias, float *act_out, float *act_out_t, int B, int Hin,
               int Hout, int ld_t, int apply_act, cudaStream_t stream);
int linear_dgrad(const float *dz, const float *W, const float *act_prev, float *dz_prev, float *dz_prev_t, int B, int Hin,
                 int Hout, int ld_t, cudaStream_t stream);
int linear_wgrad(const float *dz_t, const float *act_t_ext, float *dW, float *db, int B, int Hin, int Hout, int ld_t,
                 cudaStream_t stream);

// wgrad with Adam applied to the layer in the epilogue (one rank): the weight-gradient tile never makes the round trip
// through HBM to a separate optimizer pass.  p / m / v are the layer's [Hout][Hin] weight views of the flat buffers,
// pb / mb / vb its bias slices; the step counter is read from `state` and NOT advanced.
int linear_wgrad_adam(const float *dz_t, const float *act_t_ext, float *dW, float *db, float *p, float *pb, float *m,
                      float *mb, float *v, float *vb, int B, int Hin, int Hout, int ld_t, double beta1, double beta2,
                      double eps, const float *bias_table, int bias_table_n, const void *state, cudaStream_t stream);
int max_active_clusters(int split, int *out);

}  // namespace pairgemm
}  // namespace ddpb
