// Host-side index stream of DistributedSampler (no device code in this file).
//
// The reference builds DistributedSampler(dataset, shuffle=True) at demo.py:144 and
// calls set_epoch at demo.py:96-98; the per-rank index stream must be bit-exact.
// torch draws the permutation with torch.randperm on a CPU generator seeded with
// seed+epoch (torch:utils/data/distributed.py:107-112): a 32-bit Mersenne Twister
// feeding a forward Fisher-Yates shuffle.  Doing it here (instead of calling
// torch.randperm + .tolist() + slicing every epoch) costs ~3 us per epoch and lets
// the host layer pre-compute thousands of epochs for the device-resident data path.
#include "common.cuh"
#include <vector>

namespace {

class Twister {
public:
    explicit Twister(uint64_t seed) {
        state_[0] = (uint32_t)seed;
        for (uint32_t i = 1; i < kN; ++i)
            state_[i] = 1812433253u * (state_[i - 1] ^ (state_[i - 1] >> 30)) + i;
        next_ = kN;
    }
    uint32_t draw() {
        if (next_ == kN) refill();
        uint32_t y = state_[next_++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        return y ^ (y >> 18);
    }

private:
    static constexpr uint32_t kN = 624, kM = 397;
    void refill() {
        for (uint32_t i = 0; i < kN; ++i) {
            uint32_t mixed = (state_[i] & 0x80000000u) | (state_[(i + 1) % kN] & 0x7fffffffu);
            uint32_t twist = (mixed >> 1) ^ ((mixed & 1u) ? 0x9908b0dfu : 0u);
            state_[i] = state_[(i + kM) % kN] ^ twist;
        }
        next_ = 0;
    }
    uint32_t state_[kN];
    uint32_t next_;
};

// torch's randperm_cpu keeps this algorithm for n < UINT32_MAX / 20 only.
constexpr int64_t kMaxPerm = (int64_t)(0xffffffffu / 20);

template <typename T>
void shuffled_range(uint64_t seed, int64_t n, T *out) {
    Twister gen(seed);
    for (int64_t i = 0; i < n; ++i) out[i] = (T)i;
    for (int64_t i = 0; i + 1 < n; ++i) {
        int64_t pick = i + (int64_t)(gen.draw() % (uint64_t)(n - i));
        T tmp = out[i];
        out[i] = out[pick];
        out[pick] = tmp;
    }
}

int64_t samples_per_rank(int64_t n, int world, int drop_last) {
    if (drop_last && n % world != 0) return (n - world + world - 1) / world;
    return (n + world - 1) / world;
}

}  // namespace

DDPB_API int ddpb_randperm(uint64_t seed, int64_t n, int64_t *out) {
    DDPB_REQUIRE(out != nullptr && n >= 0, "ddpb_randperm: bad arguments");
    DDPB_REQUIRE(n < kMaxPerm, "ddpb_randperm: n=%lld is beyond the 32-bit shuffle torch uses", (long long)n);
    shuffled_range(seed, n, out);
    return 0;
}

DDPB_API int64_t ddpb_sampler_indices(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                                      uint64_t epoch, int drop_last, int64_t *out, int64_t cap) {
    if (n <= 0 || n >= kMaxPerm || world <= 0 || rank < 0 || rank >= world || !out) {
        ddpb::fail(2, "ddpb_sampler_indices: bad arguments (n=%lld world=%d rank=%d)", (long long)n, world, rank);
        return -1;
    }
    const int64_t per_rank = samples_per_rank(n, world, drop_last);
    if (per_rank > cap) {
        ddpb::fail(2, "ddpb_sampler_indices: need room for %lld indices, got %lld", (long long)per_rank, (long long)cap);
        return -1;
    }
    std::vector<int64_t> order((size_t)n);
    if (shuffle) shuffled_range(seed + epoch, n, order.data());
    else for (int64_t i = 0; i < n; ++i) order[(size_t)i] = i;
    // rank r takes positions r, r+W, r+2W, ... of the list padded by wrap-around
    for (int64_t k = 0; k < per_rank; ++k) out[k] = order[(size_t)((rank + k * world) % n)];
    return per_rank;
}

DDPB_API int ddpb_sampler_epochs_i32(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                                     uint64_t epoch0, int nepochs, int drop_last, int32_t *out, int64_t stride) {
    DDPB_REQUIRE(n > 0 && n < kMaxPerm && n <= 0x7fffffff && world > 0 && rank >= 0 && rank < world && out && nepochs >= 0,
                 "ddpb_sampler_epochs_i32: bad arguments");
    const int64_t per_rank = samples_per_rank(n, world, drop_last);
    DDPB_REQUIRE(stride >= per_rank, "ddpb_sampler_epochs_i32: stride %lld < %lld indices per epoch",
                 (long long)stride, (long long)per_rank);
    std::vector<int32_t> order((size_t)n);
    for (int e = 0; e < nepochs; ++e) {
        if (shuffle) shuffled_range(seed + epoch0 + (uint64_t)e, n, order.data());
        else for (int64_t i = 0; i < n; ++i) order[(size_t)i] = (int32_t)i;
        int32_t *row = out + (int64_t)e * stride;
        for (int64_t k = 0; k < per_rank; ++k) row[k] = order[(size_t)((rank + k * world) % n)];
    }
    return 0;
}

// int64 rows for `nepochs` consecutive epochs in one call (the host loader refills its cache
// with this instead of paying one foreign-function call per epoch).
DDPB_API int ddpb_sampler_epochs_i64(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                                     uint64_t epoch0, int nepochs, int drop_last, int64_t *out, int64_t stride) {
    DDPB_REQUIRE(n > 0 && n < kMaxPerm && world > 0 && rank >= 0 && rank < world && out && nepochs >= 0,
                 "ddpb_sampler_epochs_i64: bad arguments");
    const int64_t per_rank = samples_per_rank(n, world, drop_last);
    DDPB_REQUIRE(stride >= per_rank, "ddpb_sampler_epochs_i64: stride %lld < %lld indices per epoch",
                 (long long)stride, (long long)per_rank);
    std::vector<int64_t> order((size_t)n);
    for (int e = 0; e < nepochs; ++e) {
        if (shuffle) shuffled_range(seed + epoch0 + (uint64_t)e, n, order.data());
        else for (int64_t i = 0; i < n; ++i) order[(size_t)i] = i;
        int64_t *row = out + (int64_t)e * stride;
        for (int64_t k = 0; k < per_rank; ++k) row[k] = order[(size_t)((rank + k * world) % n)];
    }
    return 0;
}

DDPB_API int ddpb_host_gather(const float *X, const float *Y, const int64_t *idx, int64_t B, int dx, int dy,
                              float *outX, float *outY) {
    DDPB_REQUIRE(X && Y && idx && outX && outY && B >= 0 && dx > 0 && dy > 0, "ddpb_host_gather: bad arguments");
    for (int64_t b = 0; b < B; ++b) {
        const float *sx = X + idx[b] * dx;
        const float *sy = Y + idx[b] * dy;
        for (int k = 0; k < dx; ++k) outX[b * dx + k] = sx[k];
        for (int k = 0; k < dy; ++k) outY[b * dy + k] = sy[k];
    }
    return 0;
}
