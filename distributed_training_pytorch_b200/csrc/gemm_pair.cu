// Weight-stationary CTA-PAIR GEMMs for the widened toy model (tcgen05 cta_group::2, kind::tf32).
//
// Reference: the same nn.Linear layers as gemm_tf32.cu (toy_model_and_data.py:13-21 widened; cuBLAS
// sgemm in torch's forward and autograd, demo.py:104-109).
//
// Why a second GEMM kernel.  ncu on the 128 x 128 single-CTA kernel (profiles/r01_wide_h4096_v2_raw.csv)
// shows what bounds it: 277 MB arrive in shared memory per launch for 67 MB of weights
// (l1tex__m_xbar2l1tex_read_bytes), i.e. ~2.2 MB per SM at the ~64 B/clk an SM can take from L2 --
// every weight tile was fetched by both batch halves, every activation tile by every tile column.
// At B <= 256 a layer is a weight STREAM (4*H*H bytes, 128 FLOP per byte), so the fix is to touch each
// weight byte once and to halve the activation bytes per SM:
//
//   D[m][n] = sum_k A(m,k) * B(n,k)        m = NEURON (256 per CTA pair), n = SAMPLE (up to 256)
//
//   * weights are the M operand: a CTA pair owns 256 neurons; CTA r of the pair loads rows
//     [128r, 128r+128) of the weight tile and samples [128r, 128r+128) of the activation tile, and ONE
//     tcgen05.mma.cta_group::2 (M = 256, N = 256, issued by the pair's leader) multiplies them: the
//     tensor cores of both SMs read both CTAs' shared memory, so each SM receives only half of the
//     activation tile.  Shared-memory fill per launch: 4*H*H (weights, once) + 4*H*H (activations:
//     H/256 pairs x 256 x H x 4 B... x 2 halves) = 2x the weight bytes instead of 4x.
//   * the accumulator (128 neurons x 256 samples per CTA) lives in 256 TMEM columns.
//   * too few pairs for 148 SMs (16 at H = 4096), so K is split over the z dimension of a cluster
//     (2 x 1 x S CTAs, S = 4 at H = 4096): every CTA parks its partial tile in its own shared memory and
//     then REDUCES AND FINISHES only its share of the 32-sample column chunks (chunk c belongs to K-rank
//     c mod S): partial sums are pulled over distributed shared memory in ascending K-rank order
//     (deterministic), so the reduction and the epilogue are spread over all CTAs instead of serialised
//     in one.
//   * epilogue per (warp, chunk): a thread owns one neuron and 32 samples.  The tile leaves through
//     shared memory as TMA bulk stores in both layouts every consumer needs: neuron-major ([H][ld],
//     what wgrad reads as a K-major operand) and sample-major ([B][H], what the next layer reads).
//       forward: + bias, LeakyReLU
//       dgrad:   x LeakyReLU'(act_prev), the mask bits fetched while the main loop runs
//       wgrad:   dW straight into the flat gradient bucket; the operand row of ones makes column Hin
//                of the product the bias gradient
//
// Roles in a CTA (256 threads): warp 0 lane 0 = TMA producer (both CTAs of a pair load; all loads
// signal the LEADER's full barrier, cp.async.bulk.tensor...cta_group::2), warp 1 lane 0 of the leader =
// MMA issuer (tcgen05.commit...multicast::cluster frees the stage in both CTAs), warp 2 = TMEM
// allocator, warps 4..7 = epilogue.
#include "common.cuh"
#include "gemm_pair.cuh"
#include <cuda.h>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace ddpb {
namespace pairgemm {

namespace {

constexpr int BK = 32;                               // tf32 per k-block: 128 bytes = one swizzle row
constexpr int kRows = 128;                           // operand rows a CTA loads per tile (half of the pair's 256)
constexpr int kTileBytes = kRows * BK * 4;           // 16 KB
constexpr int kStageBytes = 2 * kTileBytes;          // A half + B half
constexpr int kThreads = 256;
constexpr int kTmemCols = 256;                       // accumulator: 128 lanes x 256 fp32 columns
constexpr int kChunks = 8;                           // 32-column chunks of the accumulator
constexpr int kParkRow = 144;                        // bytes per parked row of 32 floats (+16: conflict-free v4)
constexpr int kParkChunk = 128 * kParkRow;
constexpr int kParkBytes = kChunks * kParkChunk;     // 147456
constexpr int kWarpStage = 8192;                     // per epilogue warp: neuron-major chunk + sample-major chunk
constexpr float kSlope = 0.01f;
constexpr int kFuseWarpStage = 2 * 3 * 4096;         // fused Adam: two buffers of {p, m, v} chunks per epilogue warp

template <int EPI> struct Cfg {
    static constexpr int kStages = EPI >= EPI_WGRAD ? 3 : 6;
    static constexpr int kCtasPerSm = EPI >= EPI_WGRAD ? 2 : 1;
    static constexpr int kRing = kStages * kStageBytes;
    static constexpr int kSmemBytes = kRing + 1024 /*alignment slack*/ + 256 /*barriers*/;
    static_assert(EPI >= EPI_WGRAD || kParkBytes + 4 * kWarpStage <= kRing, "parking + staging must fit in the dead ring");
    static_assert(EPI != EPI_WGRAD_ADAM || 4 * kFuseWarpStage <= kRing, "fused-Adam staging must fit in the dead ring");
    static_assert(4 * kWarpStage <= kRing, "staging must fit in the dead ring");
};

// ---- PTX wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned saddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ unsigned map_to_rank(unsigned local_saddr, unsigned rank) {
    unsigned remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_saddr), "r"(rank));
    return remote;
}
// Load a box into THIS CTA's shared memory; the bytes are reported to `bar`, a shared::cluster address
// that may name the pair leader's barrier (the .cta_group::2 form allows exactly that).
__device__ __forceinline__ void tma_load_pair(unsigned dst, const CUtensorMap *map, unsigned bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *map, unsigned src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ unsigned cluster_ctarank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(unsigned slot_smem, unsigned cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(unsigned taddr, unsigned cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_pair_tf32(unsigned tmem_d, unsigned long long desc_a, unsigned long long desc_b,
                                               unsigned idesc, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// Arrive (once every MMA issued so far has retired) on the barrier at this offset in every CTA of `mask`.
__device__ __forceinline__ void umma_commit_pair(unsigned bar, unsigned short mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_ld32(unsigned taddr, float (&v)[32]) {
    unsigned r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// r[0..32) = the 128 contiguous bytes at `addr` (a shared::cluster address: any CTA of the cluster).
__device__ __forceinline__ void ld_cluster_row(unsigned addr, float (&r)[32]) {
    asm volatile(
        "ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%32];\n\t"
        "ld.shared::cluster.v4.f32 {%4, %5, %6, %7}, [%32+16];\n\t"
        "ld.shared::cluster.v4.f32 {%8, %9, %10, %11}, [%32+32];\n\t"
        "ld.shared::cluster.v4.f32 {%12, %13, %14, %15}, [%32+48];\n\t"
        "ld.shared::cluster.v4.f32 {%16, %17, %18, %19}, [%32+64];\n\t"
        "ld.shared::cluster.v4.f32 {%20, %21, %22, %23}, [%32+80];\n\t"
        "ld.shared::cluster.v4.f32 {%24, %25, %26, %27}, [%32+96];\n\t"
        "ld.shared::cluster.v4.f32 {%28, %29, %30, %31}, [%32+112];"
        : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7]),
          "=f"(r[8]), "=f"(r[9]), "=f"(r[10]), "=f"(r[11]), "=f"(r[12]), "=f"(r[13]), "=f"(r[14]), "=f"(r[15]),
          "=f"(r[16]), "=f"(r[17]), "=f"(r[18]), "=f"(r[19]), "=f"(r[20]), "=f"(r[21]), "=f"(r[22]), "=f"(r[23]),
          "=f"(r[24]), "=f"(r[25]), "=f"(r[26]), "=f"(r[27]), "=f"(r[28]), "=f"(r[29]), "=f"(r[30]), "=f"(r[31])
        : "r"(addr) : "memory");
}

// Shared-memory matrix descriptors, as in gemm_tf32.cu: K-major tile [rows][32 tf32] with 128-byte
// swizzle (type 2, SBO 1024); MN-major tile of 32-bit elements = boxes [32 k][32 mn] in the "128B
// swizzle, 32B atoms" layout (type 1, LBO 4096 between 32-wide mn atoms, SBO 512 between 4-row k groups).
__device__ __forceinline__ unsigned long long make_desc(unsigned smem_byte_addr, unsigned lbo_bytes, unsigned sbo_bytes,
                                                        unsigned layout_type) {
    unsigned long long d = 0;
    d |= (unsigned long long)((smem_byte_addr & 0x3FFFFu) >> 4);
    d |= (unsigned long long)(lbo_bytes >> 4) << 16;
    d |= (unsigned long long)(sbo_bytes >> 4) << 32;
    d |= 1ull << 46;
    d |= (unsigned long long)layout_type << 61;
    return d;
}
// Instruction descriptor: fp32 accumulate, tf32 x tf32, M = 256 (the pair), N = 256.
__device__ __forceinline__ unsigned make_idesc(bool a_mn_major) {
    unsigned d = 0;
    d |= 1u << 4;                          // D format F32
    d |= 2u << 7;                          // A format TF32
    d |= 2u << 10;                         // B format TF32
    d |= (a_mn_major ? 1u : 0u) << 15;     // A major: 0 = K, 1 = MN
    d |= (unsigned)(256 >> 3) << 17;       // N
    d |= (unsigned)(256 >> 4) << 24;       // M
    return d;
}

// What the fused wgrad + Adam epilogue needs beside the GEMM operands.
struct AdamFuse {
    CUtensorMap p, m, v;               // [Hout][Hin] weight views of the flat parameter / moment buffers, boxes [32][32], 128B swizzle
    float *dW;                         // [Hout][ldw] weight gradient (still written: .grad stays valid after step())
    float *pb, *mb, *vb;               // the layer's bias slices [Hout]
    const float2 *bias_table;          // Adam bias corrections (see ddp_b200.h)
    const ddpb_opt_state *state;
    int bias_table_n, ldw;
    float w1, b2, w2, eps;             // 1 - beta1, beta2, 1 - beta2, eps as floats (torch's op order, csrc/optim.cu)
};

// torch's _single_tensor_adam order with explicit round-to-nearest intrinsics: the same arithmetic as
// adam_one() in optim.cu (grad_scale = 1), so the fused update is bit-identical to the separate pass.
__device__ __forceinline__ void adam_elem(float g, float &p, float &m, float &v, float w1, float b2, float w2, float eps,
                                          float neg_step, float bc2_sqrt) {
    m = __fmaf_rn(w1, __fsub_rn(g, m), m);
    v = __fadd_rn(__fmul_rn(v, b2), __fmul_rn(__fmul_rn(w2, g), g));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps);
    p = __fadd_rn(p, __fmul_rn(neg_step, __fdiv_rn(m, denom)));
}

__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap *map, unsigned bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

template <int EPI, bool A_MN, int SPLIT>
__global__ void __launch_bounds__(kThreads, Cfg<EPI>::kCtasPerSm)
pair_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_o /* neuron-major out [M][N]: boxes [32 m][32 n], 128B swizzle */,
                 const __grid_constant__ CUtensorMap map_t /* sample-major out [N][M]: boxes [32 n][32 m], no swizzle */,
                 const __grid_constant__ AdamFuse fuse, const EpiParams ep, int M, int N, int K) {
    using C = Cfg<EPI>;
    constexpr int kStages = C::kStages;
    constexpr int kMine = (kChunks + SPLIT - 1) / SPLIT;        // chunks this CTA reduces and finishes (c = z, z+S, ... < 8)
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem + C::kRing);
    // bars[0..S) full (used in the leader), bars[S..2S) empty, bars[2S] accumulator ready, then the TMEM address
    unsigned *tmem_slot = reinterpret_cast<unsigned *>(bars + 2 * kStages + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned x = blockIdx.x & 1u;                         // which half of the pair
    const unsigned z = SPLIT > 1 ? blockIdx.z : 0u;             // K-rank inside the cluster
    const unsigned leader = 2u * z;                             // cluster rank of this pair's leader (ranks: x + 2z)
    const int m0 = (int)(blockIdx.x >> 1) * 256 + (int)x * kRows;   // my 128 neurons
    const int n_tile0 = (int)blockIdx.y * 256;                  // the pair's 256 samples
    const int n_load0 = n_tile0 + (int)x * kRows;               // the 128 sample rows I load
    const int num_k = (K + BK - 1) / BK;
    const int kb_begin = (int)((long long)num_k * z / SPLIT), kb_count = (int)((long long)num_k * (z + 1) / SPLIT) - kb_begin;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 2 * kStages + 1; ++s) mbar_init(saddr(&bars[s]), 1);
        if (EPI == EPI_WGRAD_ADAM)                                   // {p, m, v} chunk landed: [epilogue warp][buffer]
            for (int s = 0; s < 8; ++s) mbar_init(saddr(&bars[2 * kStages + 2 + s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_o) : "memory");
        if (EPI < EPI_WGRAD) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_t) : "memory");
        if (EPI == EPI_WGRAD_ADAM) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&fuse.p) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&fuse.m) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&fuse.v) : "memory");
        }
    }
    if (warp == 2) tmem_alloc_pair(saddr(tmem_slot), kTmemCols);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();                                          // every barrier of the cluster is initialised, TMEM is allocated
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem_base = *tmem_slot;

    // ---- what the epilogue threads can fetch while the main loop runs
    const int q = warp & 3;                                      // epilogue warp -> TMEM lanes [32q, 32q+32)
    const int m = m0 + q * 32 + lane;                            // my neuron
    float bias_v = 0.f;
    unsigned act_mask[kMine];

    if (warp == 0) {
        // ===== TMA producer (both CTAs of the pair; every load reports to the leader's full barrier) =====
        if (lane == 0) {
            for (int it = 0; it < kb_count; ++it) {
                const int s = it % kStages, kb = kb_begin + it;
                if ((ep.debug & 1) && it >= kStages) break;
                if (it >= kStages) mbar_wait(saddr(&bars[kStages + s]), ((it / kStages) - 1) & 1);
                if (x == 0) mbar_expect_tx(saddr(&bars[s]), 2 * kStageBytes);              // both halves land on this barrier
                const unsigned full = map_to_rank(saddr(&bars[s]), leader);
                const unsigned a_dst = saddr(smem + s * kStageBytes), b_dst = a_dst + kTileBytes;
                if (!A_MN) {
                    tma_load_pair(a_dst, &map_a, full, kb * BK, m0);                         // [128 neurons][32 k]
                } else {
#pragma unroll
                    for (int a = 0; a < kRows / 32; ++a)                                     // boxes [32 k][32 neurons]
                        tma_load_pair(a_dst + a * 4096, &map_a, full, m0 + a * 32, kb * BK);
                }
                tma_load_pair(b_dst, &map_b, full, kb * BK, n_load0);                        // [128 samples][32 k]
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread of the pair's leader =====
        if (lane == 0 && x == 0) {
            const unsigned idesc = make_idesc(A_MN);
            const unsigned short pair_mask = (unsigned short)(3u << leader);
            for (int it = 0; it < kb_count; ++it) {
                const int s = it % kStages;
                if (!(ep.debug & 1) || it < kStages) mbar_wait(saddr(&bars[s]), (it / kStages) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned a_base = saddr(smem + s * kStageBytes), b_base = a_base + kTileBytes;
#pragma unroll
                for (int k = 0; k < ((ep.debug & 2) ? 0 : BK / 8); ++k) {                    // UMMA_K = 8 for tf32
                    const unsigned long long da = A_MN ? make_desc(a_base + k * 1024, 4096, 512, 1)
                                                       : make_desc(a_base + k * 32, 0, 1024, 2);
                    const unsigned long long db = make_desc(b_base + k * 32, 0, 1024, 2);
                    umma_pair_tf32(tmem_base, da, db, idesc, (it | k) != 0);
                }
                umma_commit_pair(saddr(&bars[kStages + s]), pair_mask);                      // frees the stage in both CTAs
            }
            umma_commit_pair(saddr(&bars[2 * kStages]), pair_mask);                          // accumulators complete in both CTAs
        }
    } else if (warp >= 4) {
        if (EPI == EPI_FWD) bias_v = m < M ? __ldg(ep.bias + m) : 0.f;
        if (EPI == EPI_DGRAD) {
            // LeakyReLU mask bits of my neuron for the samples of my chunks: act_prev is [sample][neuron], so
            // for one sample the lanes of a warp read consecutive floats
#pragma unroll
            for (int j = 0; j < kMine; ++j) {
                const int n_c0 = ((int)z + j * SPLIT) < kChunks ? n_tile0 + ((int)z + j * SPLIT) * 32 : N;
                unsigned bits = 0;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int n = n_c0 + i;
                    const float a = (m < M && n < N) ? __ldg(ep.act_prev + (long long)n * ep.ld_act + m) : 0.f;
                    bits |= (a > 0.f ? 1u : 0u) << i;
                }
                act_mask[j] = bits;
            }
        }
        mbar_wait(saddr(&bars[2 * kStages]), 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (SPLIT > 1) {
            // park the whole partial tile: [chunk][row][32 floats + pad] in the dead pipeline stages
#pragma unroll 1
            for (int c = 0; c < kChunks; ++c) {
                float v[32];
                tmem_ld32(tmem_base + ((unsigned)(q * 32) << 16) + (unsigned)(c * 32), v);
                float4 *pr = reinterpret_cast<float4 *>(smem + c * kParkChunk + (q * 32 + lane) * kParkRow);
#pragma unroll
                for (int i = 0; i < 8; ++i) pr[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
        }
    }
    if (SPLIT > 1) {
        __syncthreads();
        cluster_sync_all();                                      // every K-rank's partial tile is parked
    }
    if (warp >= 4 && EPI == EPI_WGRAD_ADAM) {
        // ===== wgrad with Adam in the epilogue (one rank) =====
        // Per warp and 32-column chunk: the {p, m, v} tiles of my 32 neurons arrive by TMA (two buffers per warp, so
        // the next chunk's loads fly while this one is updated), the gradient comes out of TMEM, the update runs in
        // registers in torch's op order, the new {p, m, v} leave as TMA stores from the same buffers and dW is stored
        // straight from registers.  HBM traffic per parameter: 12 B read + 16 B written (28 instead of 32: dW is never
        // read back), and no separate optimizer kernel for the layer.
        unsigned char *wstage = smem + q * kFuseWarpStage;
        const int sw = lane & 7;
        const int nvalid = min(kChunks, (N - n_tile0 + 31) / 32);                           // chunks that exist in this tile
        unsigned long long *lbar = bars + 2 * kStages + 2 + q * 2;
        float neg_step, bc2_sqrt;
        {
            const long long step = fuse.state->step;
            const float2 bias = fuse.bias_table[step < fuse.bias_table_n ? step : fuse.bias_table_n - 1];
            neg_step = bias.x; bc2_sqrt = bias.y;
        }
        auto issue = [&](int c) {                                                           // lane 0: request chunk c's tiles
            const unsigned bar = saddr(&lbar[c & 1]);
            const unsigned dst = saddr(wstage + (c & 1) * 12288);
            mbar_expect_tx(bar, 12288);
            tma_load_2d(dst, &fuse.p, bar, n_tile0 + c * 32, m0 + q * 32);
            tma_load_2d(dst + 4096, &fuse.m, bar, n_tile0 + c * 32, m0 + q * 32);
            tma_load_2d(dst + 8192, &fuse.v, bar, n_tile0 + c * 32, m0 + q * 32);
        };
        if (lane == 0) {
            if (nvalid > 0) issue(0);
            if (nvalid > 1) issue(1);
        }
#pragma unroll 1
        for (int c = 0; c < nvalid; ++c) {
            const int n_c0 = n_tile0 + c * 32;
            unsigned char *buf = wstage + (c & 1) * 12288;
            float g[32];
            tmem_ld32(tmem_base + ((unsigned)(q * 32) << 16) + (unsigned)(c * 32), g);
            mbar_wait(saddr(&lbar[c & 1]), (c >> 1) & 1);
            if (m < M) {                                                                     // dW and the bias (column Hin of the product)
                float *drow = fuse.dW + (long long)m * fuse.ldw + n_c0;
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (n_c0 + 4 * i < ep.n_real)
                        *reinterpret_cast<float4 *>(drow + 4 * i) = make_float4(g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]);
                const int bi = ep.n_real - n_c0;
                if (bi >= 0 && bi < 32) {
                    float gb = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; ++i) gb = (i == bi) ? g[i] : gb;
                    ep.bias_grad[m] = gb;
                    float pb = fuse.pb[m], mb = fuse.mb[m], vb = fuse.vb[m];
                    adam_elem(gb, pb, mb, vb, fuse.w1, fuse.b2, fuse.w2, fuse.eps, neg_step, bc2_sqrt);
                    fuse.pb[m] = pb; fuse.mb[m] = mb; fuse.vb[m] = vb;
                }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {                                                    // 16-byte piece i of my row sits at i ^ (row & 7)
                float4 *pp = reinterpret_cast<float4 *>(buf + lane * 128 + ((i ^ sw) << 4));
                float4 *mp = reinterpret_cast<float4 *>(buf + 4096 + lane * 128 + ((i ^ sw) << 4));
                float4 *vp = reinterpret_cast<float4 *>(buf + 8192 + lane * 128 + ((i ^ sw) << 4));
                float4 P4 = *pp, M4 = *mp, V4 = *vp;
                adam_elem(g[4 * i + 0], P4.x, M4.x, V4.x, fuse.w1, fuse.b2, fuse.w2, fuse.eps, neg_step, bc2_sqrt);
                adam_elem(g[4 * i + 1], P4.y, M4.y, V4.y, fuse.w1, fuse.b2, fuse.w2, fuse.eps, neg_step, bc2_sqrt);
                adam_elem(g[4 * i + 2], P4.z, M4.z, V4.z, fuse.w1, fuse.b2, fuse.w2, fuse.eps, neg_step, bc2_sqrt);
                adam_elem(g[4 * i + 3], P4.w, M4.w, V4.w, fuse.w1, fuse.b2, fuse.w2, fuse.eps, neg_step, bc2_sqrt);
                *pp = P4; *mp = M4; *vp = V4;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&fuse.p, saddr(buf), n_c0, m0 + q * 32);                        // rows >= M, columns >= Hin are clipped
                tma_store_2d(&fuse.m, saddr(buf + 4096), n_c0, m0 + q * 32);
                tma_store_2d(&fuse.v, saddr(buf + 8192), n_c0, m0 + q * 32);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                if (c + 2 < nvalid) {                                                        // this buffer takes chunk c + 2 next
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    issue(c + 2);
                }
            }
            __syncwarp();
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    } else if (warp >= 4) {
        // ===== reduce my chunks over the K-ranks, epilogue, TMA stores =====
        unsigned char *wstage = smem + (SPLIT > 1 ? kParkBytes : 0) + q * kWarpStage;
        unsigned char *so = wstage;                                                          // [32 neurons][32 samples], swizzled
        float *st = reinterpret_cast<float *>(wstage + 4096);                                // [32 samples][32 neurons]
        const int sw = lane & 7;
        const bool has_o = EPI == EPI_WGRAD || ep.has_t;
        int issued = 0;                                                                      // chunks handed to the TMA engine so far
#pragma unroll 1
        for (int j = 0; j < kMine; ++j) {
            const int c = (int)z + j * SPLIT;
            const int n_c0 = n_tile0 + c * 32;
            if (c >= kChunks || n_c0 >= N) continue;                                         // nothing of this chunk exists
            float v[32];
            if (SPLIT == 1) {
                tmem_ld32(tmem_base + ((unsigned)(q * 32) << 16) + (unsigned)(c * 32), v);
            } else {
                const unsigned local = saddr(smem + c * kParkChunk + (q * 32 + lane) * kParkRow);
                // every K-rank's row is requested before any is added (one DSMEM round trip instead of S - 1
                // in a row); the sum runs in ascending K-rank order, the same on every CTA
                float r1[32], r2[32], r3[32];
                ld_cluster_row(map_to_rank(local, x), v);
                if (SPLIT >= 2) ld_cluster_row(map_to_rank(local, x + 2u), r1);
                if (SPLIT >= 3) ld_cluster_row(map_to_rank(local, x + 4u), r2);
                if (SPLIT >= 4) ld_cluster_row(map_to_rank(local, x + 6u), r3);
                if (SPLIT >= 2) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += r1[i];
                }
                if (SPLIT >= 3) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += r2[i];
                }
                if (SPLIT >= 4) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += r3[i];
                }
            }
            if (EPI == EPI_FWD) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float zf = v[i] + bias_v;
                    v[i] = (ep.apply_act && zf <= 0.f) ? zf * kSlope : zf;
                }
            } else if (EPI == EPI_DGRAD) {
                const unsigned bits = act_mask[j];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = ((bits >> i) & 1u) ? v[i] : v[i] * kSlope;
            }
            if (EPI == EPI_WGRAD) {
                // wgrad stages only the neuron-major chunk, so the warp's 8 KB hold two of them: the TMA engine
                // drains one while the next is being written (8 chunks per tile; a single buffer serialised them)
                so = wstage + (issued & 1) * 4096;
                if (issued >= 2) {
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    __syncwarp();
                }
            } else if (issued > 0) {                                                         // the staging is about to be reused
                if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                __syncwarp();
            }
            ++issued;
            if (has_o) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    *reinterpret_cast<float4 *>(so + lane * 128 + ((i ^ sw) << 4)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            if (EPI != EPI_WGRAD) {
#pragma unroll
                for (int i = 0; i < 32; ++i) st[i * 32 + lane] = v[i];
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");                     // generic-proxy writes -> visible to the TMA engine
            __syncwarp();
            if (lane == 0) {
                if (has_o) tma_store_2d(&map_o, saddr(so), n_c0, m0 + q * 32);              // rows >= M and columns beyond the tensor are clipped
                if (EPI != EPI_WGRAD) tma_store_2d(&map_t, saddr(st), m0 + q * 32, n_c0);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            if (EPI == EPI_WGRAD) {                                                          // product with the row of ones = bias gradient
                const int bi = ep.n_real - n_c0;
                if (bi >= 0 && bi < 32 && m < M) {
                    float g = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; ++i) g = (i == bi) ? v[i] : g;
                    ep.bias_grad[m] = g;
                }
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");       // staging must outlive the bulk reads
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    cluster_sync_all();                 // nobody leaves while a peer may still read its parked tile or signal its barriers
    if (warp == 2) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// ---- host side ------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

struct MapKey {
    const void *base;
    long long rows, cols, ld;
    int box_rows, box_cols, swizzle;
    bool operator==(const MapKey &o) const {
        return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows &&
               box_cols == o.box_cols && swizzle == o.swizzle;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey &k) const {
        size_t h = reinterpret_cast<size_t>(k.base);
        auto mix = [&h](size_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
        mix((size_t)k.rows); mix((size_t)k.cols); mix((size_t)k.ld);
        mix((size_t)k.box_rows * 131 + (size_t)k.box_cols * 7 + (size_t)k.swizzle);
        return h;
    }
};

}  // namespace

// A tensor map depends only on (pointer, shape, pitch, box, swizzle): the training loop passes the same
// buffers every step, so each map is encoded once (cuTensorMapEncodeTiled costs ~1.5 us, 27 of them
// per step of the widened model before this cache).
int cached_map(CUtensorMap *out, const float *base, long long rows, long long cols, long long ld, int box_rows, int box_cols,
               int swizzle) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    const MapKey key{base, rows, cols, ld, box_rows, box_cols, swizzle};
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = cache.find(key);
        if (it != cache.end()) {
            *out = it->second;
            return 0;
        }
    }
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return fail(5, "cuTensorMapEncodeTiled is not available from this driver");
    DDPB_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 4) % 16 == 0, "TMA operand must be 16-byte aligned with a 16-byte row pitch");
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUtensorMap map;
    CUresult rc = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) return fail(3000 + (int)rc, "cuTensorMapEncodeTiled failed (CUresult %d)", (int)rc);
    {
        std::lock_guard<std::mutex> lock(mu);
        if (cache.size() > 4096) cache.clear();               // a caller that keeps changing buffers: start over
        cache.emplace(key, map);
    }
    *out = map;
    return 0;
}

namespace {

struct Maps { CUtensorMap a, b, o, t; };

const AdamFuse &no_fuse() {
    static AdamFuse zero = {};
    return zero;
}

template <int EPI, bool A_MN, int SPLIT>
int launch(const Maps &mp, const EpiParams &ep, int M, int N, int K, cudaStream_t stream, const AdamFuse &fuse = no_fuse()) {
    static bool configured = false;
    if (!configured) {
        DDPB_CUDA(cudaFuncSetAttribute(pair_gemm_kernel<EPI, A_MN, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<EPI>::kSmemBytes));
        configured = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * ((M + 255) / 256), (N + 255) / 256, SPLIT);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = Cfg<EPI>::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = SPLIT;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    DDPB_CUDA(cudaLaunchKernelEx(&cfg, pair_gemm_kernel<EPI, A_MN, SPLIT>, mp.a, mp.b, mp.o, mp.t, fuse, ep, M, N, K));
    count_launch();
    return 0;
}

// K split so that about one wave of SMs has work (DDPB_GEMM_SPLIT overrides): pairs * 2 * S CTAs.
int split_for(int M, int N, int K) {
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("DDPB_GEMM_SPLIT");
        forced = e ? atoi(e) : 0;
    }
    const long long pairs = (long long)((M + 255) / 256) * ((N + 255) / 256);
    const int num_k = (K + BK - 1) / BK;
    // clusters of 2 x S CTAs with ~194 KB of shared memory each: a B200 holds 74 / 33 / 15 of them for
    // S = 1 / 2 / 4 (cudaOccupancyMaxActiveClusters), so 16 pairs x S = 4 would need a second wave
    int s = forced ? forced : (pairs <= 15 ? 4 : pairs <= 22 ? 3 : pairs <= 33 ? 2 : 1);
    if (s < 1 || s > 4) s = 1;
    while (s > 1 && num_k < 4 * s) --s;                        // keep a few k-blocks per rank
    return s;
}

template <int EPI, bool A_MN>
int launch_split(const Maps &mp, const EpiParams &ep, int M, int N, int K, cudaStream_t stream) {
    switch (split_for(M, N, K)) {
        case 4: return launch<EPI, A_MN, 4>(mp, ep, M, N, K, stream);
        case 3: return launch<EPI, A_MN, 3>(mp, ep, M, N, K, stream);
        case 2: return launch<EPI, A_MN, 2>(mp, ep, M, N, K, stream);
        default: return launch<EPI, A_MN, 1>(mp, ep, M, N, K, stream);
    }
}

int debug_bits() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("DDPB_GEMM_DEBUG");
        v = e ? atoi(e) : 0;
    }
    return v;
}

constexpr int kSwz128 = (int)CU_TENSOR_MAP_SWIZZLE_128B;
constexpr int kSwzNone = (int)CU_TENSOR_MAP_SWIZZLE_NONE;
constexpr int kSwzAtom32 = (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;

}  // namespace

int linear_fwd(const float *act_in, const float *W, const float *bias, float *act_out, float *act_out_t, int B, int Hin,
               int Hout, int ld_t, int apply_act, cudaStream_t stream) {
    Maps mp;
    if (int rc = cached_map(&mp.a, W, Hout, Hin, Hin, kRows, BK, kSwz128)) return rc;            // A: [neuron][k]
    if (int rc = cached_map(&mp.b, act_in, B, Hin, Hin, kRows, BK, kSwz128)) return rc;          // B: [sample][k]
    if (int rc = cached_map(&mp.t, act_out, B, Hout, Hout, 32, 32, kSwzNone)) return rc;         // sample-major out
    mp.o = mp.t;
    if (act_out_t)
        if (int rc = cached_map(&mp.o, act_out_t, Hout, B, ld_t, 32, 32, kSwz128)) return rc;    // neuron-major out
    EpiParams ep = {};
    ep.bias = bias; ep.has_t = act_out_t != nullptr; ep.apply_act = apply_act; ep.n_real = B; ep.debug = debug_bits();
    return launch_split<EPI_FWD, false>(mp, ep, Hout, B, Hin, stream);
}

int linear_dgrad(const float *dz, const float *W, const float *act_prev, float *dz_prev, float *dz_prev_t, int B, int Hin,
                 int Hout, int ld_t, cudaStream_t stream) {
    Maps mp;
    // A(m = input neuron, k = output neuron) = W[k][m]: MN-major, boxes [32 k][32 m]
    if (int rc = cached_map(&mp.a, W, Hout, Hin, Hin, BK, 32, kSwzAtom32)) return rc;
    if (int rc = cached_map(&mp.b, dz, B, Hout, Hout, kRows, BK, kSwz128)) return rc;            // B: [sample][k]
    if (int rc = cached_map(&mp.t, dz_prev, B, Hin, Hin, 32, 32, kSwzNone)) return rc;
    mp.o = mp.t;
    if (dz_prev_t)
        if (int rc = cached_map(&mp.o, dz_prev_t, Hin, B, ld_t, 32, 32, kSwz128)) return rc;
    EpiParams ep = {};
    ep.act_prev = act_prev; ep.ld_act = Hin; ep.has_t = dz_prev_t != nullptr; ep.n_real = B; ep.debug = debug_bits();
    return launch_split<EPI_DGRAD, true>(mp, ep, Hin, B, Hout, stream);
}

int linear_wgrad(const float *dz_t, const float *act_t_ext, float *dW, float *db, int B, int Hin, int Hout, int ld_t,
                 cudaStream_t stream) {
    Maps mp;
    if (int rc = cached_map(&mp.a, dz_t, Hout, B, ld_t, kRows, BK, kSwz128)) return rc;          // A: [out neuron][k = sample]
    if (int rc = cached_map(&mp.b, act_t_ext, Hin + 1, B, ld_t, kRows, BK, kSwz128)) return rc;  // B: [in neuron | ones][k = sample]
    if (int rc = cached_map(&mp.o, dW, Hout, Hin, Hin, 32, 32, kSwz128)) return rc;              // dW rows into the flat bucket
    mp.t = mp.o;
    EpiParams ep = {};
    ep.bias_grad = db; ep.n_real = Hin; ep.debug = debug_bits();
    return launch<EPI_WGRAD, false, 1>(mp, ep, Hout, Hin + 1, B, stream);
}

int linear_wgrad_adam(const float *dz_t, const float *act_t_ext, float *dW, float *db, float *p, float *pb, float *m,
                      float *mb, float *v, float *vb, int B, int Hin, int Hout, int ld_t, double beta1, double beta2,
                      double eps, const float *bias_table, int bias_table_n, const void *state, cudaStream_t stream) {
    Maps mp;
    if (int rc = cached_map(&mp.a, dz_t, Hout, B, ld_t, kRows, BK, kSwz128)) return rc;
    if (int rc = cached_map(&mp.b, act_t_ext, Hin + 1, B, ld_t, kRows, BK, kSwz128)) return rc;
    AdamFuse fuse = {};
    if (int rc = cached_map(&fuse.p, p, Hout, Hin, Hin, 32, 32, kSwz128)) return rc;
    if (int rc = cached_map(&fuse.m, m, Hout, Hin, Hin, 32, 32, kSwz128)) return rc;
    if (int rc = cached_map(&fuse.v, v, Hout, Hin, Hin, 32, 32, kSwz128)) return rc;
    mp.o = mp.t = fuse.p;                                                                       // unused by this epilogue
    fuse.dW = dW; fuse.ldw = Hin;
    fuse.pb = pb; fuse.mb = mb; fuse.vb = vb;
    fuse.bias_table = reinterpret_cast<const float2 *>(bias_table);
    fuse.bias_table_n = bias_table_n;
    fuse.state = static_cast<const ddpb_opt_state *>(state);
    fuse.w1 = (float)(1.0 - beta1); fuse.b2 = (float)beta2; fuse.w2 = (float)(1.0 - beta2); fuse.eps = (float)eps;
    EpiParams ep = {};
    ep.bias_grad = db; ep.n_real = Hin; ep.debug = debug_bits();
    return launch<EPI_WGRAD_ADAM, false, 1>(mp, ep, Hout, Hin + 1, B, stream, fuse);
}

// How many clusters of the forward kernel (2 x 1 x split CTAs, ~194 KB of shared memory each) the device can
// hold at once (cudaOccupancyMaxActiveClusters): says whether a launch runs as one wave.
int max_active_clusters(int split, int *out) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2, 1, split);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = Cfg<EPI_FWD>::kSmemBytes;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = split;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    int n = 0;
    switch (split) {
        case 4:
            DDPB_CUDA(cudaFuncSetAttribute(pair_gemm_kernel<EPI_FWD, false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<EPI_FWD>::kSmemBytes));
            DDPB_CUDA(cudaOccupancyMaxActiveClusters(&n, pair_gemm_kernel<EPI_FWD, false, 4>, &cfg));
            break;
        case 2:
            DDPB_CUDA(cudaFuncSetAttribute(pair_gemm_kernel<EPI_FWD, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<EPI_FWD>::kSmemBytes));
            DDPB_CUDA(cudaOccupancyMaxActiveClusters(&n, pair_gemm_kernel<EPI_FWD, false, 2>, &cfg));
            break;
        default:
            DDPB_CUDA(cudaFuncSetAttribute(pair_gemm_kernel<EPI_FWD, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<EPI_FWD>::kSmemBytes));
            DDPB_CUDA(cudaOccupancyMaxActiveClusters(&n, pair_gemm_kernel<EPI_FWD, false, 1>, &cfg));
            break;
    }
    *out = n;
    return 0;
}

}  // namespace pairgemm
}  // namespace ddpb

// Development aid (tools/gemm_probe.py); not part of the public header.
DDPB_API int ddpb_gemm_max_active_clusters(int split, int *out) { return ddpb::pairgemm::max_active_clusters(split, out); }
