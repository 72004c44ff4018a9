// featurize.cuh -- tweets -> filtered CSR rows on the device (K1).
//
// Follows MllibHelper.scala: filtrate (:84-94), featurizeText (:42-56: text.sliding(2) over UTF-16
// code units, HashingTF [U] bucket = String.hashCode mod numFeatures, counts per bucket, indices
// ascending), featurizeNumbers (:58-71), featurize (:73-82: the four numeric slots always stored,
// label = retweet count).  Integer work is exact; the four numeric values are one IEEE multiply
// each, exactly as on the JVM.
//
// One warp per tweet.  The hashed terms go to shared memory, are rank-sorted there (a tweet has at
// most kMaxTerms terms, so the O(m^2/32) rank sort is a few hundred steps), and equal neighbours
// are run-length encoded into (bucket, count) pairs, written at the tweet's own offset of a
// scratch area that is as large as the text (a tweet yields at most as many pairs as it has code
// units).  A second kernel, after exclusive scans over the kept flags and the row lengths, packs
// the rows into CSR.
#pragma once
#include "common.cuh"

namespace twtml {

constexpr int kFeatWarps = 4;
constexpr int kMaxTerms = 1024; // terms per tweet (text of up to 1025 UTF-16 units)

struct TweetCols {
    const uint16_t* text;
    const long long* off; // n + 1
    const int* retweet_cnt;
    const int* followers;
    const int* favourites;
    const int* friends;
    const long long* created_ms;
    const uint8_t* is_retweet;
};

static __global__ void __launch_bounds__(kFeatWarps * 32) k_feat_terms(TweetCols tw, long long n, int rt_begin, int rt_end, int F,
                                                                const uint16_t* __restrict__ lower /* 65536 or null */,
                                                                int* __restrict__ pair_idx, int* __restrict__ pair_cnt,
                                                                long long* __restrict__ keep, long long* __restrict__ rownnz)
{
    __shared__ unsigned hbuf[kFeatWarps][kMaxTerms];
    __shared__ unsigned sbuf[kFeatWarps][kMaxTerms];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    unsigned* h = hbuf[wib];
    unsigned* s = sbuf[wib];
    const long long nw = (long long)gridDim.x * kFeatWarps;
    for (long long i = (long long)blockIdx.x * kFeatWarps + wib; i < n; i += nw) {
        const int rc = tw.retweet_cnt[i];
        const bool kept = tw.is_retweet[i] != 0 && rc >= rt_begin && rc <= rt_end; // filtrate
        if (!kept) {
            if (lane == 0) { keep[i] = 0; rownnz[i] = 0; }
            continue;
        }
        const long long o = tw.off[i];
        const int L = (int)(tw.off[i + 1] - o);
        const int m = L >= 2 ? L - 1 : L; // sliding(2): L-1 bigrams; a 1-unit text is one short term
        // The text of this tweet into shared memory first, with 16-byte loads over its aligned middle: tw.text may be
        // the caller's page-locked host buffer read in place over the host link (twtml_sparse.cu tweets_ingest), where
        // the size of a request is what the throughput depends on; no byte outside [o, o + L) is touched.
        unsigned short* txt = reinterpret_cast<unsigned short*>(s); // (s is free until the rank sort below)
        {
            const uint16_t* tp = tw.text + o;
            int head = (int)(((16u - (unsigned)(reinterpret_cast<uintptr_t>(tp) & 15u)) & 15u) >> 1);
            head = min(head, L);
            const int body = (L - head) >> 3, tail0 = head + 8 * body;
            if (lane < head) txt[lane] = tp[lane];
            for (int c = lane; c < body; c += 32) {
                const uint4 v = *reinterpret_cast<const uint4*>(tp + head + 8 * c);
                unsigned short* q = txt + head + 8 * c;
                q[0] = (unsigned short)(v.x & 0xffffu); q[1] = (unsigned short)(v.x >> 16);
                q[2] = (unsigned short)(v.y & 0xffffu); q[3] = (unsigned short)(v.y >> 16);
                q[4] = (unsigned short)(v.z & 0xffffu); q[5] = (unsigned short)(v.z >> 16);
                q[6] = (unsigned short)(v.w & 0xffffu); q[7] = (unsigned short)(v.w >> 16);
            }
            if (lane < L - tail0) txt[tail0 + lane] = tp[tail0 + lane];
        }
        __syncwarp();
        for (int p = lane; p < m; p += 32) {
            unsigned c0 = txt[p];
            if (lower) c0 = lower[c0];
            unsigned hv = c0; // String.hashCode of one unit
            if (L >= 2) {
                unsigned c1 = txt[p + 1];
                if (lower) c1 = lower[c1];
                hv = 31u * c0 + c1; // <= 2 097 120: never negative, so nonNegativeMod is a plain %
            }
            h[p] = hv % (unsigned)F;
        }
        __syncwarp();
        for (int p = lane; p < m; p += 32) { // rank sort (stable)
            const unsigned v = h[p];
            int rank = 0;
            for (int q = 0; q < m; ++q) {
                const unsigned u = h[q];
                rank += (u < v) || (u == v && q < p);
            }
            s[rank] = v;
        }
        __syncwarp();
        int nruns = 0;
        for (int base = 0; base < m; base += 32) { // run-length encode the sorted buckets
            const int p = base + lane;
            const bool valid = p < m;
            const unsigned v = valid ? s[p] : 0u;
            const bool head = valid && (p == 0 || s[p - 1] != v);
            const bool tail = valid && (p == m - 1 || s[p + 1] != v);
            const unsigned hb = __ballot_sync(0xffffffffu, head), tb = __ballot_sync(0xffffffffu, tail);
            const unsigned lt = (1u << lane) - 1u;
            // every run's head precedes (or is) its tail, so at a tail: tails before == heads before it - 1
            const int hk = nruns + __popc(hb & lt);
            if (head) { h[hk] = (unsigned)p; pair_idx[o + hk] = (int)v; }
            __syncwarp();
            if (tail) {
                const int tk = nruns + __popc(hb & (lt | (1u << lane))) - 1; // index of the run this tail closes
                pair_cnt[o + tk] = p - (int)h[tk] + 1;
            }
            nruns += __popc(hb);
            (void)tb;
            __syncwarp();
        }
        if (lane == 0) { keep[i] = 1; rownnz[i] = nruns + 4; }
        __syncwarp();
    }
}

// after the scans: krow[i] = kept rows before i, eoff[i] = entries before tweet i
static __global__ void __launch_bounds__(128) k_feat_pack(TweetCols tw, long long n, long long now_ms, int F,
                                                   const int* __restrict__ pair_idx, const int* __restrict__ pair_cnt,
                                                   const long long* __restrict__ keep, const long long* __restrict__ krow,
                                                   const long long* __restrict__ rownnz, const long long* __restrict__ eoff,
                                                   long long* __restrict__ rowptr, int* __restrict__ colidx,
                                                   double* __restrict__ vals, double* __restrict__ labels)
{
    const int lane = threadIdx.x & 31;
    const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += nw) {
        if (!keep[i]) continue;
        const long long o = tw.off[i], e = eoff[i], r = krow[i];
        const int k = (int)rownnz[i] - 4;
        for (int p = lane; p < k; p += 32) {
            colidx[e + p] = pair_idx[o + p];
            vals[e + p] = (double)pair_cnt[o + p];
        }
        if (lane == 0) {
            rowptr[r] = e;
            labels[r] = (double)tw.retweet_cnt[i];
            colidx[e + k + 0] = F + 0; vals[e + k + 0] = __dmul_rn((double)tw.followers[i], 1e-12);
            colidx[e + k + 1] = F + 1; vals[e + k + 1] = __dmul_rn((double)tw.favourites[i], 1e-12);
            colidx[e + k + 2] = F + 2; vals[e + k + 2] = __dmul_rn((double)tw.friends[i], 1e-12);
            colidx[e + k + 3] = F + 3; vals[e + k + 3] = __dmul_rn((double)(now_ms - tw.created_ms[i]), 1e-14);
        }
    }
}

static __global__ void k_feat_maxlen(const long long* __restrict__ off, long long n, int* __restrict__ maxlen)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicMax(maxlen, (int)(off[i + 1] - off[i]));
}

} // namespace twtml
