// Tiled tensor-core GEMMs for the WIDENED toy model (hidden width H a multiple of 128,
// BASELINE.json configs[4]): the three H x H linear layers run forward, dgrad and wgrad as
// tcgen05 (5th-generation tensor core) kind::tf32 MMAs fed by TMA.
//
// Reference: the same nn.Linear layers (toy_model_and_data.py:13-21, widths scaled up) that
// PyTorch runs as cuBLAS sgemm calls in forward and in autograd's backward (demo.py:104-109).
//
//   D[m][n] = sum_k A(m,k) * B(n,k)          one 128 x 128 output tile per CTA (or per cluster, see below)
//
//   forward   m = sample, n = output neuron, k = input neuron   A = act_in [B,Hin]   K-major
//                                                                B = W [Hout,Hin]      K-major
//             epilogue: + bias, LeakyReLU, store act_out [B,Hout] and its transpose [Hout,ld]
//   dgrad     m = sample, n = input neuron,  k = output neuron   A = dZ [B,Hout]      K-major
//                                                                B = W [Hout,Hin]      MN-major (n contiguous)
//             epilogue: x LeakyReLU'(act_prev), store dZ_prev [B,Hin] and its transpose
//   wgrad     m = output neuron, n = input neuron, k = sample    A = dZ^T [Hout,ld]    K-major
//                                                                B = act^T [Hin+..,ld] K-major
//             epilogue: store dW straight into the flat gradient bucket; the operand row
//             n = Hin is a row of ones, which makes column Hin of the product the bias gradient
//
// Measured (tools/gemm_timeline.py: clock64 stamps in the producer and MMA threads, globaltimer
// stamps per CTA and phase):
//   * main loop: a CTA retires one 32-wide k-block (16 KB of A + 16 KB of B) every ~0.31 us when 64
//     CTAs run and ~0.38 us when 128 run; a k-block is requested ~2600 cycles before it lands and the
//     6-stage ring hides that.  Multicast of the shared A slab, 2x2 clusters of neighbouring tiles, L2
//     prefetch of the weight tiles and 64-wide tiles were all tried: no gain.  Every operand byte
//     crosses shared memory twice (TMA write, tensor-core read) and every weight tile is fetched by
//     both batch halves, so L2 -> SM traffic is 4x the weight bytes.
//   * too few tiles at B <= 256: 2 x H/128 (64 at H = 4096) for 148 SMs.  So the K loop is split
//     across a cluster of SPLIT CTAs (blockIdx.z): ranks > 0 park their partial accumulator in their
//     own shared memory, rank 0 adds them through distributed shared memory in its epilogue (2-way:
//     49 -> 38 us at H = 4096; 4-way never wins, its 256 CTAs need two waves).
//   * epilogue: the first version stored straight from registers (each lane its own row: 32 rows x
//     16 B per instruction) and read the bias / activation rows the same way -- 11 us per launch, each
//     128-byte line a serialised cold miss of ~1.5 us.  Now the bias slice is fetched during the main
//     loop, the activation tile (dgrad) arrives by TMA, and the output tile and its transpose go
//     through shared memory (the dead pipeline stages) and leave as TMA bulk stores: 2.8 us.
//   * wgrad (K = B, thousands of tiles): 3-stage ring and two CTAs per SM, so one CTA's epilogue runs
//     under the other's main loop (54 -> 35 us at H = 4096).
// Structure (one CTA, 256 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected
// thread), warp 2 = TMEM allocator, warps 4..7 = epilogue (tcgen05.ld 32 lanes x 32 columns at a
// time).  A 6-stage ring of {A tile, B tile} (32 KB per stage, 128-byte swizzle) is filled by
// cp.async.bulk.tensor with mbarrier complete_tx and drained by tcgen05.commit.
// Operands are fp32 in HBM and shared memory; the tensor core reads them as tf32 (10-bit
// mantissa) and accumulates in fp32 in TMEM -- the tolerance of this path is stated in
// tests/test_gpu_wide.py.  At B <= 256 each weight byte feeds 128 FLOP per pass, below the tf32 ridge
// (~170 FLOP/B): the roofline is weight bytes / HBM bandwidth.
#include "common.cuh"
#include "gemm_pair.cuh"
#include <cuda.h>
#include <cstdlib>

namespace {

constexpr int BM = 128, BK = 32;                     // BK tf32 = 128 bytes = one swizzle atom row
constexpr int kThreads = 256;
constexpr int kATileBytes = BM * BK * 4;             // 16 KB
constexpr float kSlope = 0.01f;

enum { EPI_FWD = 0, EPI_DGRAD = 1, EPI_WGRAD = 2 };

// Tile configuration: bytes per pipeline stage, ring depth, dynamic shared memory.  Forward and dgrad
// stream a long K (= H) through a 6-stage ring, one CTA per SM.  Wgrad has K = B (8 k-blocks at
// B = 256) and thousands of tiles: a 3-stage ring lets two CTAs share an SM, so that one CTA's
// epilogue runs under the other's main loop.
template <int BN, int EPI> struct Cfg {
    static constexpr int kBTileBytes = BN * BK * 4;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
    static constexpr int kStages = EPI == EPI_WGRAD ? 3 : 6;
    static constexpr int kCtasPerSm = EPI == EPI_WGRAD ? 2 : 1;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/ + 512 /*bias*/;
    static constexpr int kTmemCols = BN;             // fp32 accumulator: 128 lanes x BN columns
    // epilogue staging per warp: {activation tile (DGRAD), output tile, transposed output tile} x 16 KB
    static constexpr int kWarpStage = EPI == EPI_WGRAD ? 16384 : 49152;
    static constexpr int kOutOffset = EPI == EPI_WGRAD ? 0 : 16384;
    static_assert(4 * kWarpStage <= kStages * kStageBytes, "epilogue staging must fit in the pipeline stages");
};

#ifdef DDPB_GEMM_TIMELINE
// Development aid (tools/gemm_timeline.py): clock64 stamps of CTA (0,0)'s producer and MMA threads.
__device__ long long g_gemm_stamps[3][512];
#define GEMM_STAMP(row, i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && (i) < 512) g_gemm_stamps[row][i] = clock64(); } while (0)
// Per-CTA wall-clock phases (globaltimer, ns): 0 entry, 1 set-up done, 2 accumulator ready, 3 epilogue done, 4 exit.
__device__ long long g_gemm_phase[1024][8];
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define GEMM_PHASE(i) do { const int cta_ = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x; if (cta_ < 1024) g_gemm_phase[cta_][i] = gtimer(); } while (0)
#else
#define GEMM_STAMP(row, i) do { } while (0)
#define GEMM_PHASE(i) do { } while (0)
#endif

// The output tile, its transposed copy and (DGRAD) the activation tile travel by TMA; what is left here
// is what the epilogue threads touch directly.
struct EpiParams {
    const float *bias;        // FWD: [N]
    float *bias_grad;         // WGRAD: [M], column n_real of the product
    int has_t;                // FWD / DGRAD: also store the transposed copy
    int apply_act;            // FWD: LeakyReLU after the bias
    int n_real;               // width of the stored tile (FWD / DGRAD: N; WGRAD: Hin, column Hin is the bias gradient)
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
__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap *map, unsigned bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
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
// v[0..32) += the 128 contiguous bytes at `local` in the shared memory of cluster rank `rank`.  One asm
// block, so that the eight 16-byte loads are in flight together (issued one by one, each waiting for
// the previous, this reduction cost ~9 us per tile).
__device__ __forceinline__ void add_dsmem_row(const void *local, unsigned rank, float (&v)[32]) {
    unsigned remote;
    float r[32];
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(saddr(local)), "r"(rank));
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
        : "r"(remote) : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += r[i];
}
constexpr int kParkRow = 144;                        // bytes per parked row of 32 floats: 128 + 16 of padding (conflict-free v4)
constexpr int kParkChunk = BM * kParkRow;            // one 32-column chunk of the partial tile
__device__ __forceinline__ void tmem_alloc(unsigned slot_smem, unsigned cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(unsigned taddr, unsigned cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_tf32(unsigned tmem_d, unsigned long long desc_a, unsigned long long desc_b,
                                          unsigned idesc, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
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

// Shared-memory matrix descriptor (tcgen05).
//   K-major tile  [rows][32 tf32], SWIZZLE_128B (type 2): rows 128 B apart, 8-row groups 1024 B
//     apart (SBO); LBO unused.
//   MN-major tile of 32-bit elements: the only legal layout is "128B swizzle with 32B atoms"
//     (type 1, TMA CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B): boxes [32 k][32 n] of 4 KB, k rows 128 B
//     apart, 4-row k groups 512 B apart (SBO), 32-wide n atoms 4096 B apart (LBO).  (Plain
//     SWIZZLE_128B on an MN-major tf32 operand silently multiplies by zero -- measured.)
__device__ __forceinline__ unsigned long long make_desc(unsigned smem_byte_addr, unsigned lbo_bytes, unsigned sbo_bytes,
                                                        unsigned layout_type) {
    unsigned long long d = 0;
    d |= (unsigned long long)((smem_byte_addr & 0x3FFFFu) >> 4);
    d |= (unsigned long long)(lbo_bytes >> 4) << 16;
    d |= (unsigned long long)(sbo_bytes >> 4) << 32;
    d |= 1ull << 46;                       // descriptor version (Blackwell)
    d |= (unsigned long long)layout_type << 61;
    return d;
}

// Instruction descriptor: fp32 accumulate, tf32 x tf32, M = 128, N = BN.
template <int BN>
__device__ __forceinline__ unsigned make_idesc(bool b_mn_major) {
    unsigned d = 0;
    d |= 1u << 4;                          // D format F32
    d |= 2u << 7;                          // A format TF32
    d |= 2u << 10;                         // B format TF32
    d |= (b_mn_major ? 1u : 0u) << 16;     // B major: 0 = K, 1 = MN
    d |= (unsigned)(BN >> 3) << 17;
    d |= (unsigned)(BM >> 4) << 24;
    return d;
}

template <int EPI, bool B_MN, int BN, int SPLIT>
__global__ void __launch_bounds__(kThreads, Cfg<BN, EPI>::kCtasPerSm)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_o /* out, boxes [32 m][32 n], 128B swizzle */,
                 const __grid_constant__ CUtensorMap map_t /* transposed out, boxes [32 n][32 m], no swizzle */,
                 const __grid_constant__ CUtensorMap map_c /* DGRAD: act_prev, same geometry as map_o */,
                 const EpiParams ep, int M, int N, int K) {
    using C = Cfg<BN, EPI>;
    constexpr int kStages = C::kStages, kStageBytes = C::kStageBytes, kTmemCols = C::kTmemCols;
    extern __shared__ unsigned char smem_raw[];
    // 1024-byte alignment for the swizzle atoms
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem + kStages * kStageBytes);
    // bars[0..S) full, bars[S..2S) empty, bars[2S] accumulator ready, bars[2S+1..2S+5) activation tile of
    // epilogue warp q landed (DGRAD), then the TMEM address slot; the bias slice follows the barrier block
    unsigned *tmem_slot = reinterpret_cast<unsigned *>(bars + 2 * kStages + 5);
    float *bias_s = reinterpret_cast<float *>(smem + kStages * kStageBytes + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
    const int num_k = (K + BK - 1) / BK;

    // SPLIT > 1: the SPLIT CTAs of a cluster (blockIdx.z) share one output tile and split its K loop.
    const unsigned krank = SPLIT > 1 ? cluster_ctarank() : 0u;
    const int kb_begin = (int)((long long)num_k * krank / SPLIT), kb_count = (int)((long long)num_k * (krank + 1) / SPLIT) - kb_begin;
    if (threadIdx.x == 0) {
        GEMM_PHASE(0);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(saddr(&bars[s]), 1);
            mbar_init(saddr(&bars[kStages + s]), 1);
        }
        for (int s = 0; s < 5; ++s) mbar_init(saddr(&bars[2 * kStages + s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_o) : "memory");
        if (EPI != EPI_WGRAD) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_t) : "memory");
        if (EPI == EPI_DGRAD) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_c) : "memory");
    }
    if (warp == 2) tmem_alloc(saddr(tmem_slot), kTmemCols);
    // The bias slice of this tile is fetched while the pipeline runs (a cold miss here costs ~1.5 us
    // per 128-byte line when it sits in the epilogue instead).
    if (EPI == EPI_FWD && warp >= 4) bias_s[threadIdx.x - 128] = (n0 + (int)threadIdx.x - 128 < N) ? __ldg(ep.bias + n0 + threadIdx.x - 128) : 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem_base = *tmem_slot;
    if (threadIdx.x == 0) GEMM_PHASE(1);

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int it = 0; it < kb_count; ++it) {
                const int s = it % kStages, kb = kb_begin + it;
                if (it >= kStages) mbar_wait(saddr(&bars[kStages + s]), ((it / kStages) - 1) & 1);
                const unsigned full = saddr(&bars[s]);
                const unsigned a_dst = saddr(smem + s * kStageBytes), b_dst = a_dst + kATileBytes;
                GEMM_STAMP(0, it);                                                    // producer: stage free, issuing loads
                mbar_expect_tx(full, kStageBytes);
                tma_load_2d(a_dst, &map_a, full, kb * BK, m0);                       // [128 rows][32 k]
                if (!B_MN) {
                    tma_load_2d(b_dst, &map_b, full, kb * BK, n0);                   // [BN rows][32 k]
                } else {
#pragma unroll
                    for (int a = 0; a < BN / 32; ++a)                                // boxes [32 k][32 n]
                        tma_load_2d(b_dst + a * 4096, &map_b, full, n0 + a * 32, kb * BK);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            const unsigned idesc = make_idesc<BN>(B_MN);
            for (int it = 0; it < kb_count; ++it) {
                const int s = it % kStages;
                mbar_wait(saddr(&bars[s]), (it / kStages) & 1);
                GEMM_STAMP(1, it);                                                    // consumer: data landed
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned a_base = saddr(smem + s * kStageBytes), b_base = a_base + kATileBytes;
#pragma unroll
                for (int k = 0; k < BK / 8; ++k) {                                   // UMMA_K = 8 for tf32
                    const unsigned long long da = make_desc(a_base + k * 32, 0, 1024, 2);
                    const unsigned long long db = B_MN ? make_desc(b_base + k * 1024, 4096, 512, 1)
                                                       : make_desc(b_base + k * 32, 0, 1024, 2);
                    umma_tf32(tmem_base, da, db, idesc, (it | k) != 0);
                }
                GEMM_STAMP(2, it);                                                    // consumer: MMAs issued
                umma_commit(saddr(&bars[kStages + s]));                              // frees the stage when the MMAs retire
            }
            umma_commit(saddr(&bars[2 * kStages]));                                  // accumulator complete
        }
    }
    // The pipeline stages are dead once the accumulator is complete: with SPLIT > 1 they become the
    // parking area of the partial tile: [chunk of 32 columns][row][32 floats + pad].
    unsigned char *park = smem;
    if (warp >= 4) {
        mbar_wait(saddr(&bars[2 * kStages]), 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (threadIdx.x == 128) GEMM_PHASE(2);
        if (SPLIT > 1 && krank > 0) {
            const int q = warp - 4;
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                float v[32];
                tmem_ld32(tmem_base + ((unsigned)(q * 32) << 16) + (unsigned)c0, v);
                float4 *pr = reinterpret_cast<float4 *>(park + (c0 >> 5) * kParkChunk + (q * 32 + lane) * kParkRow);
#pragma unroll
                for (int i = 0; i < 8; ++i) pr[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
        }
    }
    // Rank 0's own pipeline stages become the staging area of its epilogue: per epilogue warp
    // {activation tile (DGRAD), output tile, transposed output tile}, each 4 chunks of [32][32] floats
    // (WGRAD: the output tile only).
    unsigned char *wstage = smem + (warp >= 4 ? warp - 4 : 0) * C::kWarpStage;
    if (EPI == EPI_DGRAD && warp >= 4 && krank == 0 && lane == 0) {
        // LeakyReLU mask source: rows [m0+32q, +32) x columns [n0, +128) of act_prev, fetched by TMA as
        // soon as the stages are free (out-of-range parts arrive as zeros and are clipped on the store)
        const unsigned bar = saddr(&bars[2 * kStages + 1 + (warp - 4)]);
        mbar_expect_tx(bar, 16384);
#pragma unroll
        for (int c = 0; c < 4; ++c) tma_load_2d(saddr(wstage) + c * 4096, &map_c, bar, n0 + c * 32, m0 + (warp - 4) * 32);
    }
    if (SPLIT > 1) {
        __syncthreads();
        cluster_sync_all();                                                          // partial tiles are parked
    }
    if (warp >= 4 && krank == 0) {
        // ===== epilogue: TMEM (+ the other ranks' partials) -> registers -> shared -> TMA store =====
        const int q = warp - 4;                                                      // TMEM lanes [32q, 32q+32)
        const int m = m0 + q * 32 + lane;
        const int sw = lane & 7;                                                     // 128B swizzle: 16-byte chunk j of row r sits at j ^ (r & 7)
        if (EPI == EPI_DGRAD) mbar_wait(saddr(&bars[2 * kStages + 1 + q]), 0);
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            tmem_ld32(tmem_base + ((unsigned)(q * 32) << 16) + (unsigned)c0, v);
            if (SPLIT > 1) {
#pragma unroll 1
                for (unsigned r = 1; r < (unsigned)SPLIT; ++r)
                    add_dsmem_row(park + (c0 >> 5) * kParkChunk + (q * 32 + lane) * kParkRow, r, v);
            }
            const int nb = n0 + c0, chunk = c0 >> 5;
            if (EPI == EPI_FWD) {
                const float4 *bp = reinterpret_cast<const float4 *>(bias_s + c0);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 b = bp[i];
                    const float z0 = v[4 * i] + b.x, z1 = v[4 * i + 1] + b.y, z2 = v[4 * i + 2] + b.z, z3 = v[4 * i + 3] + b.w;
                    v[4 * i + 0] = (ep.apply_act && z0 <= 0.f) ? z0 * kSlope : z0;
                    v[4 * i + 1] = (ep.apply_act && z1 <= 0.f) ? z1 * kSlope : z1;
                    v[4 * i + 2] = (ep.apply_act && z2 <= 0.f) ? z2 * kSlope : z2;
                    v[4 * i + 3] = (ep.apply_act && z3 <= 0.f) ? z3 * kSlope : z3;
                }
            } else if (EPI == EPI_DGRAD) {
                const unsigned char *ap = wstage + chunk * 4096 + lane * 128;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 a = *reinterpret_cast<const float4 *>(ap + ((i ^ sw) << 4));
                    v[4 * i + 0] = a.x > 0.f ? v[4 * i + 0] : v[4 * i + 0] * kSlope;
                    v[4 * i + 1] = a.y > 0.f ? v[4 * i + 1] : v[4 * i + 1] * kSlope;
                    v[4 * i + 2] = a.z > 0.f ? v[4 * i + 2] : v[4 * i + 2] * kSlope;
                    v[4 * i + 3] = a.w > 0.f ? v[4 * i + 3] : v[4 * i + 3] * kSlope;
                }
            }
            unsigned char *so = wstage + C::kOutOffset + chunk * 4096;               // [32 m][32 n], swizzled
            float *st = reinterpret_cast<float *>(wstage + 32768 + chunk * 4096);    // [32 n][32 m]
            const bool store_o = nb < ep.n_real, store_t = EPI != EPI_WGRAD && ep.has_t && nb < N;
            if (store_o) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    *reinterpret_cast<float4 *>(so + lane * 128 + ((i ^ sw) << 4)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            if (store_t) {
#pragma unroll
                for (int i = 0; i < 32; ++i) st[i * 32 + lane] = v[i];
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");              // generic-proxy writes -> visible to the TMA engine
            __syncwarp();
            if (lane == 0) {
                if (store_o) tma_store_2d(&map_o, saddr(so), nb, m0 + q * 32);      // rows >= M and columns >= n_real are clipped
                if (store_t) tma_store_2d(&map_t, saddr(st), m0 + q * 32, nb);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            if (EPI == EPI_WGRAD && nb == ep.n_real && m < M) ep.bias_grad[m] = v[0];   // product with the row of ones
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // staging must outlive the bulk reads
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        if (threadIdx.x == 128) GEMM_PHASE(3);
    }
    __syncthreads();
    if (threadIdx.x == 0) GEMM_PHASE(4);
    if (SPLIT > 1) cluster_sync_all();                 // nobody leaves while rank 0 may still read its parked tile
    if (warp == 2) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        tmem_dealloc(tmem_base, kTmemCols);
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
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 2-D fp32 tensor [rows][cols] with row pitch ld (elements); box = [box_rows][box_cols], 128B swizzle.
int make_map(CUtensorMap *map, const float *base, long long rows, long long cols, long long ld, int box_rows, int box_cols,
             CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
    return ddpb::pairgemm::cached_map(map, base, rows, cols, ld, box_rows, box_cols, (int)swizzle);
}

// The five tensor maps of one launch: operands A and B, the output tile, its transposed copy and
// (DGRAD) the activation whose sign is the LeakyReLU mask.  Unused ones repeat the output map.
struct Maps { CUtensorMap a, b, o, t, c; };

template <int EPI, bool B_MN, int BN, int SPLIT>
int launch(const Maps &mp, const EpiParams &ep, int M, int N, int K, cudaStream_t stream) {
    static bool configured = false;
    if (!configured) {
        DDPB_CUDA(cudaFuncSetAttribute(gemm_tf32_kernel<EPI, B_MN, BN, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<BN, EPI>::kSmemBytes));
        configured = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((N + BN - 1) / BN, (M + BM - 1) / BM, SPLIT);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = Cfg<BN, EPI>::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 1; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = SPLIT;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    DDPB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tf32_kernel<EPI, B_MN, BN, SPLIT>, mp.a, mp.b, mp.o, mp.t, mp.c, ep, M, N, K));
    ddpb::count_launch();
    return 0;
}

// How many ways to split K so that about one wave of 148 SMs has a tile (DDPB_GEMM_SPLIT overrides).
int split_for(int M, int N, int K) {
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("DDPB_GEMM_SPLIT");
        forced = e ? atoi(e) : 0;
    }
    const long long tiles = (long long)((M + BM - 1) / BM) * ((N + 127) / 128);
    int s = forced ? forced : (tiles >= 100 ? 1 : 2);          // measured: 4-way never beats 2-way (H = 1024, 4096)
    if (s != 1 && s != 2 && s != 4) s = 1;
    while (s > 1 && (K + BK - 1) / BK < 4 * s) s >>= 1;       // keep a few k-blocks per rank
    return s;
}

template <int EPI, bool B_MN>
int launch_split(const Maps &mp, const EpiParams &ep, int M, int N, int K, cudaStream_t stream) {
    switch (split_for(M, N, K)) {
        case 4: return launch<EPI, B_MN, 128, 4>(mp, ep, M, N, K, stream);
        case 2: return launch<EPI, B_MN, 128, 2>(mp, ep, M, N, K, stream);
        default: return launch<EPI, B_MN, 128, 1>(mp, ep, M, N, K, stream);
    }
}

// Which kernel family runs a layer: the CTA-pair weight-stationary kernels of gemm_pair.cu or the
// 128 x 128 single-CTA kernels of this file.  DDPB_GEMM_KERNEL=pair|tile forces one (A/B measurements);
// the default picks per shape from what tools/gemm_probe.py measured on B200 (B = 256, L2 flushed;
// profiles/r02_gemm_probe.jsonl): the pair kernels halve the bytes that have to reach shared memory but
// have few, large tiles -- they win once a layer has >= 20 neuron blocks of 256 (H >= 5120: 49 vs 77 us
// at 5120, 59 vs 89 at 6144, 78 vs 94 at 8192) and lose a few us below that (42 vs 40 us at 4096).
// wgrad (K = batch) differs by a few percent either way; the pair form is taken where it measured faster.
bool use_pair_kernels(int epi, long long neurons, long long other) {
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("DDPB_GEMM_KERNEL");
        forced = !e ? 0 : (e[0] == 't' ? 1 : e[0] == 'p' ? 2 : 0);
    }
    if (forced) return forced == 2;
    const long long blocks = (neurons + 255) / 256;
    if (epi == EPI_WGRAD) return blocks >= 16 && blocks < 32 && other >= 4096;
    return blocks >= 20;
}

}  // namespace

#ifdef DDPB_GEMM_TIMELINE
DDPB_API int ddpb_gemm_stamps(long long *host_out /* [3*512] */) {
    DDPB_CUDA(cudaMemcpyFromSymbol(host_out, g_gemm_stamps, sizeof(long long) * 3 * 512));
    return 0;
}
DDPB_API int ddpb_gemm_phases(long long *host_out /* [1024*8] */) {
    DDPB_CUDA(cudaMemcpyFromSymbol(host_out, g_gemm_phase, sizeof(long long) * 1024 * 8));
    return 0;
}
#endif

DDPB_API int ddpb_linear_fwd(const float *act_in, const float *W, const float *bias, float *act_out, float *act_out_t,
                             int B, int Hin, int Hout, int ld_t, int apply_act, void *stream) {
    DDPB_REQUIRE(act_in && W && bias && act_out && B >= 1 && Hin >= 32 && Hout >= 32 && Hin % 4 == 0 && Hout % 4 == 0,
                 "ddpb_linear_fwd: bad arguments (B=%d Hin=%d Hout=%d)", B, Hin, Hout);
    DDPB_REQUIRE(!act_out_t || ld_t >= B, "ddpb_linear_fwd: transposed pitch %d < batch %d", ld_t, B);
    DDPB_REQUIRE(!act_out_t || ld_t % 4 == 0, "ddpb_linear_fwd: transposed pitch %d must be a multiple of 4", ld_t);
    if (use_pair_kernels(EPI_FWD, Hout, Hin))
        return ddpb::pairgemm::linear_fwd(act_in, W, bias, act_out, act_out_t, B, Hin, Hout, ld_t, apply_act, (cudaStream_t)stream);
    Maps mp;
    if (int rc = make_map(&mp.a, act_in, B, Hin, Hin, BM, BK)) return rc;
    if (int rc = make_map(&mp.b, W, Hout, Hin, Hin, 128, BK)) return rc;
    if (int rc = make_map(&mp.o, act_out, B, Hout, Hout, 32, 32)) return rc;
    mp.t = mp.c = mp.o;
    if (act_out_t)
        if (int rc = make_map(&mp.t, act_out_t, Hout, B, ld_t, 32, 32, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
    EpiParams ep = {bias, nullptr, act_out_t != nullptr, apply_act, Hout};
    return launch_split<EPI_FWD, false>(mp, ep, B, Hout, Hin, (cudaStream_t)stream);
}

DDPB_API int ddpb_linear_dgrad(const float *dz, const float *W, const float *act_prev, float *dz_prev, float *dz_prev_t,
                               int B, int Hin, int Hout, int ld_t, void *stream) {
    DDPB_REQUIRE(dz && W && act_prev && dz_prev && B >= 1 && Hin >= 32 && Hout >= 32 && Hin % 4 == 0 && Hout % 4 == 0,
                 "ddpb_linear_dgrad: bad arguments (B=%d Hin=%d Hout=%d)", B, Hin, Hout);
    DDPB_REQUIRE(!dz_prev_t || ld_t >= B, "ddpb_linear_dgrad: transposed pitch %d < batch %d", ld_t, B);
    DDPB_REQUIRE(!dz_prev_t || ld_t % 4 == 0, "ddpb_linear_dgrad: transposed pitch %d must be a multiple of 4", ld_t);
    if (use_pair_kernels(EPI_DGRAD, Hin, Hout))
        return ddpb::pairgemm::linear_dgrad(dz, W, act_prev, dz_prev, dz_prev_t, B, Hin, Hout, ld_t, (cudaStream_t)stream);
    Maps mp;
    if (int rc = make_map(&mp.a, dz, B, Hout, Hout, BM, BK)) return rc;               // A: [m = sample][k = out neuron]
    if (int rc = make_map(&mp.b, W, Hout, Hin, Hin, BK, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) return rc;   // B: boxes [32 k][32 n], n contiguous
    if (int rc = make_map(&mp.o, dz_prev, B, Hin, Hin, 32, 32)) return rc;
    if (int rc = make_map(&mp.c, act_prev, B, Hin, Hin, 32, 32)) return rc;
    mp.t = mp.o;
    if (dz_prev_t)
        if (int rc = make_map(&mp.t, dz_prev_t, Hin, B, ld_t, 32, 32, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
    EpiParams ep = {nullptr, nullptr, dz_prev_t != nullptr, 0, Hin};
    return launch_split<EPI_DGRAD, true>(mp, ep, B, Hin, Hout, (cudaStream_t)stream);
}

DDPB_API int ddpb_linear_wgrad(const float *dz_t, const float *act_t_ext, float *dW, float *db, int B, int Hin, int Hout,
                               int ld_t, void *stream) {
    DDPB_REQUIRE(dz_t && act_t_ext && dW && db && B >= 1 && Hin >= 32 && Hout >= 32 && Hin % 4 == 0 && ld_t % 4 == 0 && ld_t >= B,
                 "ddpb_linear_wgrad: bad arguments (B=%d Hin=%d Hout=%d ld_t=%d)", B, Hin, Hout, ld_t);
    if (use_pair_kernels(EPI_WGRAD, Hout, Hin))
        return ddpb::pairgemm::linear_wgrad(dz_t, act_t_ext, dW, db, B, Hin, Hout, ld_t, (cudaStream_t)stream);
    Maps mp;
    if (int rc = make_map(&mp.a, dz_t, Hout, B, ld_t, BM, BK)) return rc;             // A: [m = out neuron][k = sample]
    if (int rc = make_map(&mp.b, act_t_ext, Hin + 1, B, ld_t, 128, BK)) return rc;    // B: [n = in neuron | ones][k = sample]
    if (int rc = make_map(&mp.o, dW, Hout, Hin, Hin, 32, 32)) return rc;              // dW rows straight into the flat gradient bucket
    mp.t = mp.c = mp.o;
    EpiParams ep = {nullptr, db, 0, 0, Hin};
    return launch<EPI_WGRAD, false, 128, 1>(mp, ep, Hout, Hin + 1, B, (cudaStream_t)stream);
}

DDPB_API int ddpb_linear_wgrad_adam(const float *dz_t, const float *act_t_ext, float *dW, float *db, float *p, float *pb,
                                    float *m, float *mb, float *v, float *vb, int B, int Hin, int Hout, int ld_t,
                                    double beta1, double beta2, double eps, const float *bias_table, int bias_table_n,
                                    const ddpb_opt_state *state, void *stream) {
    DDPB_REQUIRE(dz_t && act_t_ext && dW && db && p && pb && m && mb && v && vb && state && bias_table && bias_table_n >= 1,
                 "ddpb_linear_wgrad_adam: null argument");
    DDPB_REQUIRE(B >= 1 && Hin >= 32 && Hout >= 32 && Hin % 4 == 0 && ld_t % 4 == 0 && ld_t >= B,
                 "ddpb_linear_wgrad_adam: bad arguments (B=%d Hin=%d Hout=%d ld_t=%d)", B, Hin, Hout, ld_t);
    DDPB_REQUIRE((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)dW) & 15u) == 0,
                 "ddpb_linear_wgrad_adam: dW, p, m and v must be 16-byte aligned");
    return ddpb::pairgemm::linear_wgrad_adam(dz_t, act_t_ext, dW, db, p, pb, m, mb, v, vb, B, Hin, Hout, ld_t, beta1, beta2,
                                             eps, bias_table, bias_table_n, state, (cudaStream_t)stream);
}
