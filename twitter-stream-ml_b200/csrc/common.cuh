// common.cuh -- shared device helpers for libtwtml (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace twtml {

constexpr int kMaxShards = 8;      // virtual row shards per job: the fixed reduction order is over these
constexpr int kIterThreads = 512;  // CTA size of the dense iteration kernel
constexpr int kIterWarps = kIterThreads / 32;

// How a rank's rows are cut into shards and row blocks.  A block's partial sums are a function of
// its rows only and the blocks of a shard are added in block order, so the result does not depend
// on grid size; with n % 8 == 0 it does not depend on the number of GPUs (1/2/4/8) either.
struct Plan {
    int spr;                              // shards held by this rank
    int rb;                               // rows per block
    long long nblocks;                    // blocks on this rank
    long long shard_row0[kMaxShards + 1]; // first local row of each shard (+ end)
    long long shard_blk0[kMaxShards + 1]; // first block of each shard (+ end)
};

__device__ __forceinline__ void block_geom(const Plan& p, long long b, long long& row0, int& nrows)
{
    int s = 0;
#pragma unroll
    for (int i = 1; i < kMaxShards; ++i)
        if (i < p.spr && b >= p.shard_blk0[i]) s = i;
    row0 = p.shard_row0[s] + (b - p.shard_blk0[s]) * (long long)p.rb;
    long long rem = p.shard_row0[s + 1] - row0;
    nrows = rem < (long long)p.rb ? (int)rem : p.rb;
}

// Training-loop state kept on the device so that no iteration needs the host
// (GradientDescent.runMiniBatchSGD's i / converged / regVal / lossHistory [U]).
struct TrainState {
    int iter;        // next iteration number i (1-based)
    int done;        // converged, or iterations exhausted
    int nhist;       // entries in the loss history = iterations that saw m > 0
    int pad;
    double reg_val;  // regVal carried into the next loss entry
    double last_count;
};

// ---- Utils.round (Utils.scala:4-5): decimal HALF_UP == truncate, step away from zero when the
// exact remainder is >= 0.5; BigDecimal has no -0.0.
__device__ __forceinline__ double round_half_up(double x)
{
    if (!isfinite(x)) return x;
    double r = trunc(x);
    double rem = x - r;
    if (fabs(rem) >= 0.5) r += (x < 0.0) ? -1.0 : 1.0;
    return r + 0.0;
}

// ---- splitmix64 counter hash: synthetic generators and the Bernoulli row sampler.
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ double u01(uint64_t seed, uint64_t idx)
{
    return (double)(splitmix64(seed ^ idx) >> 11) * 0x1.0p-53;
}
// stand-in for RDD.sample(false, frac, 42 + i) [U]: keyed by the GLOBAL row index
__device__ __forceinline__ bool row_sampled(uint64_t iter_seed, uint64_t global_row, double frac)
{
    return u01(iter_seed, global_row) < frac;
}

// ---- StatCounter [U] as a mergeable 7-tuple: count, (mean, M2) of the labels, (mean, M2) of the
// rounded predictions, running mean of (label - pred)^2, count of non-finite predictions.
struct Stat7 {
    double n, mu_y, m2_y, mu_p, m2_p, mu_e, bad;
};
__device__ __forceinline__ void stat_zero(Stat7& s) { s.n = s.mu_y = s.m2_y = s.mu_p = s.m2_p = s.mu_e = s.bad = 0.0; }
__device__ __forceinline__ void stat_add(Stat7& s, double y, double p)
{
    s.n += 1.0;
    double dy = y - s.mu_y;
    s.mu_y += dy / s.n;
    s.m2_y += dy * (y - s.mu_y);
    double dp = p - s.mu_p;
    s.mu_p += dp / s.n;
    s.m2_p += dp * (p - s.mu_p);
    double e = (y - p) * (y - p);
    s.mu_e += (e - s.mu_e) / s.n;
    if (!isfinite(p)) s.bad += 1.0;
}
__device__ __forceinline__ void merge_moment(double n1, double& mu1, double& m21, double n2, double mu2, double m22)
{
    double delta = mu2 - mu1, n = n1 + n2;
    if (n2 * 10.0 < n1) mu1 = mu1 + (delta * n2) / n;
    else if (n1 * 10.0 < n2) mu1 = mu2 - (delta * n1) / n;
    else mu1 = (mu1 * n1 + mu2 * n2) / n;
    m21 += m22 + (delta * delta * n1 * n2) / n;
}
__device__ __forceinline__ void stat_merge(Stat7& a, const Stat7& b)
{
    if (b.n == 0.0) return;
    if (a.n == 0.0) { a = b; return; }
    merge_moment(a.n, a.mu_y, a.m2_y, b.n, b.mu_y, b.m2_y);
    merge_moment(a.n, a.mu_p, a.m2_p, b.n, b.mu_p, b.m2_p);
    double dummy = 0.0;
    merge_moment(a.n, a.mu_e, dummy, b.n, b.mu_e, 0.0);
    a.bad += b.bad;
    a.n += b.n;
}
constexpr int kStatStride = 8; // doubles per stored Stat7

__device__ __forceinline__ void stat_store(double* p, const Stat7& s)
{
    p[0] = s.n; p[1] = s.mu_y; p[2] = s.m2_y; p[3] = s.mu_p; p[4] = s.m2_p; p[5] = s.mu_e; p[6] = s.bad; p[7] = 0.0;
}
__device__ __forceinline__ void stat_load(Stat7& s, const double* p)
{
    s.n = p[0]; s.mu_y = p[1]; s.m2_y = p[2]; s.mu_p = p[3]; s.m2_p = p[4]; s.mu_e = p[5]; s.bad = p[6];
}

// ---- mbarrier + 1-D bulk TMA (cp.async.bulk, SASS UBLKCP) ---------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared, contiguous, completion counted in bytes on `bar`; streaming data: L2 evict-first
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ double shfl_xor_f64(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }

} // namespace twtml
