// msdlong.cu -- occurrence index for 17 <= k <= 31 without sorting 128-bit records.
//
// The 2k-bit code of a k-mer starts with the code of its first 11 residues, so the offsets-first
// MSD sort of msd3.cu, run on that 22-bit prefix (validity still decided by the whole k-mer),
// groups the positions into 4^11 buckets in three passes over 8-byte records.  A bucket holds
// ~n / 4^11 positions (120 at config-3 scale) in ascending order; what is left is a stable sort of
// every bucket on the remaining 2k - 22 bits: a warp reads the full code of each of the bucket's
// records from the packed store, sorts the bucket on (code, position) in shared memory and writes
// sorted {code, position} records, from which gbuild.cu forms the rows as before.  Buckets of more
// than 512 records get a whole CTA (up to 16 384 records).
// The LSD path this replaces (radix.cu: 8 passes over 16-byte records, 256 B/record of traffic)
// stays as the fallback for inputs whose buckets do not fit a window (long low-complexity repeats).
// Reference counterpart: the node_idx / srcnode_idx / tgtnode_idx B-trees of lemongraph.c:591-630.
#include "gindex.cuh"
#include "msd3.cuh"
#include <algorithm>

namespace pdg {

static constexpr int ML_THREADS = 256;
#ifndef ML_CFG_MINB
#define ML_CFG_MINB 4   /* CTAs per SM the bucket sort is compiled for (64 registers) */
#endif
#ifndef ML_CFG_LB
#define ML_CFG_LB 0     /* registers per batch of packed-word loads; 0 = one register at a time (faster: 16.4 vs 18.7 ms) */
#endif

__device__ __forceinline__ unsigned long long ml_code_at(const unsigned long long *__restrict__ packed, uint32_t p, int k) {
    const uint32_t w = p >> 5, off = p & 31u;
    const unsigned long long w0 = packed[w];
    unsigned long long hi = w0;
    if (off) hi = (w0 << (2 * off)) | (packed[w + 1] >> (64 - 2 * off));
    return hi >> (64 - 2 * k);
}

// first bucket b in [0, n_buckets] with rowtab[b] >= x (rowtab[n_buckets] = n)
__device__ __forceinline__ uint32_t ml_first_bucket(const uint32_t *__restrict__ rowtab, uint32_t n_buckets, uint32_t x) {
    uint32_t lo = 0, hi = n_buckets;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (rowtab[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// Sorts records [e0, e0 + cnt) of pos_in on (code, position) in shared memory (key / val hold at least the
// next power of two >= cnt entries) and writes them to out.  All threads of the CTA take part.
template <int THREADS>
__device__ __forceinline__ void ml_sort_range(const unsigned long long *__restrict__ packed,
                                              const uint32_t *__restrict__ pos_in, uint32_t e0, uint32_t cnt, int k,
                                              unsigned long long *key, uint32_t *val, uint4 *__restrict__ out,
                                              uint32_t *__restrict__ heads) {
    const int tid = threadIdx.x;
    uint32_t tn = 2;           // sort the next power of two >= cnt
    while (tn < cnt) tn <<= 1;
    for (uint32_t i = tid; i < tn; i += THREADS) {
        if (i < cnt) {
            const uint32_t p = pos_in[e0 + i];
            key[i] = ml_code_at(packed, p, k);
            val[i] = p;
        } else {
            key[i] = ~0ULL;
            val[i] = ~0u;
        }
    }
    // bitonic network on (code, position); positions are unique, so the order is total and stable
    for (uint32_t size = 2; size <= tn; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (uint32_t t = tid; t < tn / 2; t += THREADS) {
                const uint32_t i = 2 * t - (t & (stride - 1)), l = i + stride;
                const unsigned long long ka = key[i], kb = key[l];
                const uint32_t va = val[i], vb = val[l];
                const bool gt = ka > kb || (ka == kb && va > vb);
                const bool up = (i & size) == 0;
                if (gt == up) {
                    key[i] = kb; key[l] = ka;
                    val[i] = vb; val[l] = va;
                }
            }
        }
    }
    __syncthreads();
    for (uint32_t i = tid; i < cnt; i += THREADS) {
        const unsigned long long c = key[i];
        out[e0 + i] = make_uint4((uint32_t)c, (uint32_t)(c >> 32), val[i], 0u);
        heads[e0 + i] = (i == 0 || c != key[i - 1]) ? 1u : 0u;     // a bucket is a whole number of rows
    }
}

// Bitonic network over 32 * E keys held E per lane (key index = r * 32 + lane): strides >= 32 pair two
// registers of one lane, smaller strides pair two lanes through a shuffle.  No shared memory, no barrier.
// The (size, stride) loops stay rolled -- fully unrolled, the E = 16 network alone is ~6000 instructions
// and the kernel stalled on instruction fetch (ncu: no_instruction 33 cycles per issue) -- so register
// pairs are reached through one unrolled body per register stride.
template <int E, int RS>
__device__ __forceinline__ void ml_reg_stage(unsigned long long (&key)[E], uint32_t size) {
#pragma unroll
    for (int r = 0; r < E; ++r) {
        if (r & RS) continue;
        const bool up = (((uint32_t)r * 32u) & size) == 0u;       // size > stride >= 32: a register-index bit
        const unsigned long long a = key[r], b = key[r | RS];
        const bool sw = (a > b) == up;
        key[r] = sw ? b : a;
        key[r | RS] = sw ? a : b;
    }
}

template <int E>
__device__ __forceinline__ void ml_warp_bitonic(unsigned long long (&key)[E], int lane) {
#pragma unroll 1
    for (uint32_t size = 2; size <= 32u * E; size <<= 1) {
#pragma unroll 1
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            if (stride >= 32u) {
                switch (stride >> 5) {
                case 1: if (E > 1) ml_reg_stage<E, (E > 1 ? 1 : 0)>(key, size); break;
                case 2: if (E > 2) ml_reg_stage<E, (E > 2 ? 2 : 0)>(key, size); break;
                case 4: if (E > 4) ml_reg_stage<E, (E > 4 ? 4 : 0)>(key, size); break;
                default: if (E > 8) ml_reg_stage<E, (E > 8 ? 8 : 0)>(key, size); break;
                }
            } else if (size < 32u) {
                // direction and side depend on the lane only: one predicate for all registers
                const bool keep_min = (((uint32_t)lane & stride) == 0u) == (((uint32_t)lane & size) == 0u);
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[r], (int)stride);
                    key[r] = ((key[r] < o) == keep_min) ? key[r] : o;
                }
            } else {
                // direction depends on the register index only: bit r of dn is set where the run descends
                const bool lower = ((uint32_t)lane & stride) == 0u;
                uint32_t dn = 0;
#pragma unroll
                for (int r = 0; r < E; ++r) dn |= (((uint32_t)r * 32u) & size) ? (1u << r) : 0u;
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[r], (int)stride);
                    const bool keep_min = lower != (bool)((dn >> r) & 1u);
                    key[r] = ((key[r] < o) == keep_min) ? key[r] : o;
                }
            }
        }
    }
}

// Sorts bucket [e0, e0 + m) (m <= 32 * E) on (suffix, position) and writes its records.  Key = suffix of the
// code below the 22-bit prefix << 9 | index in the bucket: the index keeps equal codes in position order
// and finds the position again afterwards (the bucket's positions ascend).
template <int E>
__device__ __forceinline__ void ml_sort_bucket(const unsigned long long *__restrict__ packed,
                                               const uint32_t *__restrict__ pos_in, uint32_t e0, uint32_t m, int k,
                                               unsigned long long prefix, int lane, uint4 *__restrict__ out,
                                               uint32_t *__restrict__ heads) {
    const int sb = 2 * k - kMsd3MaxBits;                       // suffix bits, <= 40
    const unsigned long long smask = (1ULL << sb) - 1ULL;
    unsigned long long key[E];
#if ML_CFG_LB == 0
    // register by register: position, then its packed words.  The compiler keeps that order (2 E dependent trips to
    // memory per bucket, all of them long_scoreboard stalls in ncu's source page) -- and it is the FASTER form: see
    // the batched one below.
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint32_t i = r * 32 + lane;
#ifdef ML_FAKE_GATHER   /* experiment: no access to the packed store at all, keys from a hash of the position */
        key[r] = i < m ? (((mix64(__ldcs(pos_in + e0 + i)) & smask) << 9) | i) : ~0ULL;
#else
        key[r] = i < m ? (((ml_code_at(packed, __ldcs(pos_in + e0 + i), k) & smask) << 9) | i) : ~0ULL;
#endif
    }
#else
    // Loads in batches, every batch's loads independent of each other: first the positions of all E registers, then
    // -- ML_LB registers at a time -- both packed words of every position, then the keys.  Measured SLOWER at config 3
    // (18.4-19.4 ms for batches of 2, 4, 8 against 16.4 ms): more gathers in flight do not help a kernel whose
    // gathers already run at what the DRAM gives for random 64-byte lines (profiles/r02_k31_bucket_sort_experiments.txt).
    constexpr int ML_LB = E < ML_CFG_LB ? E : ML_CFG_LB;
    {
        uint32_t p[E];
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const uint32_t i = r * 32 + lane;
            p[r] = i < m ? pos_in[e0 + i] : 0u;                // position 0: a harmless, hot line for the padding lanes
        }
#pragma unroll
        for (int r0 = 0; r0 < E; r0 += ML_LB) {
            unsigned long long w0[ML_LB], w1[ML_LB];
#pragma unroll
            for (int q = 0; q < ML_LB; ++q) {
                const unsigned long long *src = packed + (p[r0 + q] >> 5);
                w0[q] = src[0];
                w1[q] = src[1];                                // the store has a word of slack behind the last residue
            }
#pragma unroll
            for (int q = 0; q < ML_LB; ++q) {
                const uint32_t i = (r0 + q) * 32 + lane, off = p[r0 + q] & 31u;
                // window of 32 residues at `off` of the 64 in (w0, w1); a shift by 64 is split in two
                const unsigned long long hi = (w0[q] << (2 * off)) | ((w1[q] >> 1) >> (63 - 2 * off));
                key[r0 + q] = i < m ? ((((hi >> (64 - 2 * k)) & smask) << 9) | i) : ~0ULL;
            }
        }
    }
#endif
#ifndef ML_NO_SORT      /* experiment: everything but the network (output unsorted) */
    ml_warp_bitonic<E>(key, lane);
#endif
    // the positions of the sorted records, all loads first (the bucket's slice of pos_in was read a moment ago)
    uint32_t pv[E];
#pragma unroll
    for (int r = 0; r < E; ++r) pv[r] = (uint32_t)(r * 32 + lane) < m ? pos_in[e0 + (uint32_t)(key[r] & 511u)] : 0u;
    // a record starts a row when its code differs from its predecessor's (the first of a bucket always does)
    unsigned long long carry = ~0ULL;                          // suffix of key index r * 32 - 1
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint32_t i = r * 32 + lane;
        const unsigned long long sfx = key[r] >> 9;
        unsigned long long prev = __shfl_up_sync(0xffffffffu, sfx, 1);
        if (lane == 0) prev = carry;
        carry = __shfl_sync(0xffffffffu, sfx, 31);
        if (i < m) {
            const unsigned long long c = (prefix << sb) | sfx;
            __stcs(out + e0 + i, make_uint4((uint32_t)c, (uint32_t)(c >> 32), pv[r], 0u));
            __stcs(heads + e0 + i, (i == 0 || sfx != prev) ? 1u : 0u);
        }
    }
}

// One WARP per bucket (strided over the 4^11 buckets), keys in registers.  Buckets of more than 512 records
// are appended to big[] for the CTA kernels below.  ctl[0] = overflow flag, ctl[1] = entries in big[].
// (Earlier versions, config-3 scale: windows of ~8 whole buckets sorted per CTA in shared memory, 40 ms --
// the network is log^2 in the window size and half of every window was padding; one warp per bucket in
// shared memory, 31.5 ms at 36 % issue -- every stage waits for a shared-memory round trip.)
template <bool PIPE>
__global__ void __launch_bounds__(ML_THREADS, ML_CFG_MINB) ml_bucket_sort_kernel(const unsigned long long *__restrict__ packed,
                                                                    const uint32_t *__restrict__ pos_in,
                                                                    const uint32_t *__restrict__ rowtab, uint32_t n_buckets,
                                                                    int k, uint4 *__restrict__ out,
                                                                    uint32_t *__restrict__ heads,
                                                                    unsigned int *__restrict__ ctl, uint2 *__restrict__ big,
                                                                    uint32_t big_cap) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t nw = gridDim.x * (ML_THREADS / 32);
    auto sort_one = [&](uint32_t b, uint32_t e0, uint32_t m) {
        if (m == 0) return;
        if (m <= 32) ml_sort_bucket<1>(packed, pos_in, e0, m, k, b, lane, out, heads);
        else if (m <= 64) ml_sort_bucket<2>(packed, pos_in, e0, m, k, b, lane, out, heads);
        else if (m <= 128) ml_sort_bucket<4>(packed, pos_in, e0, m, k, b, lane, out, heads);
        else if (m <= 256) ml_sort_bucket<8>(packed, pos_in, e0, m, k, b, lane, out, heads);
        else if (m <= 512) ml_sort_bucket<16>(packed, pos_in, e0, m, k, b, lane, out, heads);
        else if (lane == 0) {
            const unsigned int slot = atomicAdd(&ctl[1], 1u);
            if (slot < big_cap) big[slot] = make_uint2(e0, e0 + m); else atomicOr(&ctl[0], 1u);
        }
    };
    const uint32_t b0 = blockIdx.x * (ML_THREADS / 32) + w;
    if (!PIPE) {
        for (uint32_t b = b0; b < n_buckets; b += nw) {
            const uint32_t e0 = rowtab[b];
            sort_one(b, e0, rowtab[b + 1] - e0);
        }
        return;
    }
    // A warp's bucket costs three dependent trips to memory -- its bounds, its positions, the packed words its
    // positions point at (a random 64-byte line each) -- before the first compare; with 8 warps per scheduler that
    // chain showed as long_scoreboard, the top stall (ncu: issue slots 53 % used).  So the warp works three buckets
    // ahead of itself: while bucket i is sorted, the lines of bucket i + 1 are on their way to L2 (prefetch from the
    // positions loaded one iteration earlier), the first ML_PF x 32 positions of bucket i + 2 are being loaded and so
    // are the bounds of bucket i + 3.
    constexpr int ML_PF = 4;
    auto bounds = [&](uint32_t b, uint32_t &e, uint32_t &m) {
        e = 0; m = 0;
        if (b < n_buckets) { e = rowtab[b]; m = rowtab[b + 1] - e; }
    };
    auto fetch_pos = [&](uint32_t e, uint32_t m, uint32_t (&P)[ML_PF]) {
#pragma unroll
        for (int q = 0; q < ML_PF; ++q) P[q] = (uint32_t)(q * 32 + lane) < m ? pos_in[e + q * 32 + lane] : ~0u;
    };
    uint32_t e_c, m_c, e_1, m_1, e_2, m_2, P[ML_PF];
    bounds(b0, e_c, m_c);
    bounds(b0 + nw, e_1, m_1);
    bounds(b0 + 2 * nw, e_2, m_2);
    fetch_pos(e_1, m_1, P);
    for (uint32_t b = b0; b < n_buckets; b += nw) {
#pragma unroll
        for (int q = 0; q < ML_PF; ++q)
            if (P[q] != ~0u) asm volatile("prefetch.global.L2 [%0];" ::"l"(packed + (P[q] >> 5)));
        fetch_pos(e_2, m_2, P);
        uint32_t e_3, m_3;
        bounds(b + 3 * nw, e_3, m_3);
        sort_one(b, e_c, m_c);
        e_c = e_1; m_c = m_1;
        e_1 = e_2; m_1 = m_2;
        e_2 = e_3; m_2 = m_3;
    }
}

// one CTA per big bucket: up to ML_MID records in static shared memory (many CTAs per SM), or up to ML_BIG
// in dynamic shared memory (one CTA per SM) for the rare huge ones; each launch takes the buckets of its
// size class and skips the others
static constexpr int ML_MID = 2048, ML_BIG = 16384, ML_BIG_THREADS = 1024;
__global__ void __launch_bounds__(ML_THREADS) ml_mid_bucket_kernel(const unsigned long long *__restrict__ packed,
                                                                   const uint32_t *__restrict__ pos_in, int k,
                                                                   uint4 *__restrict__ out, uint32_t *__restrict__ heads,
                                                                   unsigned int *__restrict__ ctl, const uint2 *__restrict__ big) {
    __shared__ unsigned long long key[ML_MID];
    __shared__ uint32_t val[ML_MID];
    const uint2 r = big[blockIdx.x];
    const uint32_t cnt = r.y - r.x;
    if (cnt > (uint32_t)ML_MID) {   // left to ml_big_bucket_kernel
        if (threadIdx.x == 0) atomicAdd(&ctl[2], 1u);
        return;
    }
    ml_sort_range<ML_THREADS>(packed, pos_in, r.x, cnt, k, key, val, out, heads);
}

__global__ void __launch_bounds__(ML_BIG_THREADS) ml_big_bucket_kernel(const unsigned long long *__restrict__ packed,
                                                                       const uint32_t *__restrict__ pos_in, int k,
                                                                       uint4 *__restrict__ out, uint32_t *__restrict__ heads,
                                                                       unsigned int *__restrict__ ctl, const uint2 *__restrict__ big) {
    extern __shared__ __align__(16) unsigned char ml_raw[];
    unsigned long long *key = reinterpret_cast<unsigned long long *>(ml_raw);
    uint32_t *val = reinterpret_cast<uint32_t *>(ml_raw + (size_t)ML_BIG * 8);
    const uint2 r = big[blockIdx.x];
    const uint32_t cnt = r.y - r.x;
    if (cnt <= (uint32_t)ML_MID) return;   // done by ml_mid_bucket_kernel
    if (cnt > (uint32_t)ML_BIG) {          // the caller takes the LSD path
        if (threadIdx.x == 0) atomicOr(&ctl[0], 1u);
        return;
    }
    ml_sort_range<ML_BIG_THREADS>(packed, pos_in, r.x, cnt, k, key, val, out, heads);
}

// true: *sorted (= g->recA) holds the n records {code lo, code hi, position, 0} in (code, position) order and
// *head_flags (= g->recB) one word per record: 1 where a row (a new code) starts
// staged: a pipelined ingest already ran the prefix sort up to pass 2 (or 3) for exactly these contigs.
bool msdlong_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, bool staged, const uint4 **sorted,
                   const uint32_t **head_flags) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    const uint64_t n = hc.n_kmers;
    const uint32_t n_buckets = 1u << kMsd3MaxBits;
    static const bool off = getenv("PDG_NO_MSDLONG") != nullptr;
    if (off || 2 * k <= kMsd3MaxBits || n == 0 || n >= (1ULL << 32)) return false;
    if (!staged) {
        g->recA.reserve(n * 16, s);        // before the staging so that nothing is reallocated under it
        msd3_reserve(g, hc.pos0, hc.pos1, 1);
        msd3_stage_group(g, ct, hc.pos0, hc.pos1);
    }
    if (!g->m3_launched) msd3_finish_launch(g, /*rows=*/false);   // g_pos grouped by prefix, scratch0 = bucket starts (+ sentinel)
    if (staged) g->recA.reserve(n * 16, s);   // pass 3 no longer reads recA; reserve() drains the stream before it frees
    // room for EVERY bucket (33.5 MB): with a fixed 2^18 entries, 500 genomes per GPU (config 4 on 2 GPUs: ~600
    // records per bucket, nearly all of the 4^11 buckets above the warp path's 512) overflowed the list and sent the
    // whole build to the 8-pass LSD sort of 16-byte records -- 282 ms for 5x the records of a 22 ms sort
    const uint32_t big_cap = n_buckets;
    g->ml_flag.reserve(16 + (size_t)big_cap * sizeof(uint2), s);
    PDG_CUDA(cudaMemsetAsync(g->ml_flag.p, 0, 16, s));
    while (g->k_events.size() < 8) {
        cudaEvent_t ev;
        PDG_CUDA(cudaEventCreate(&ev));
        g->k_events.push_back(ev);
    }
    if (!(g->attr_mask & pdg_graph::ATTR_MSDLONG)) {
        PDG_CUDA(cudaFuncSetAttribute(ml_big_bucket_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ML_BIG * 12));
        g->attr_mask |= pdg_graph::ATTR_MSDLONG;
    }
    const unsigned grid = 148 * 16;                            // warps stride over the buckets
    const unsigned long long *packed = g->packed.as<unsigned long long>();
    const uint32_t *pos_in = g->g_pos.as<uint32_t>(), *rowtab = g->scratch0.as<uint32_t>();
    uint4 *out = g->recA.as<uint4>();
    uint32_t *heads = g->recB.as<uint32_t>();                  // free since pass 3; n * 4 <= n * 8 bytes
    unsigned int *ctl = g->ml_flag.as<unsigned int>();
    uint2 *big = reinterpret_cast<uint2 *>(g->ml_flag.as<unsigned char>() + 16);
    // (an L2 persistence window over the packed store was tried for the random code gathers: the kernel
    // did not get faster and the rest of the step lost 16 ms to the limit / reset calls)
    PDG_CUDA(cudaEventRecord(g->k_events[6], s));
    // the three-buckets-ahead loop is measured SLOWER (20.8 against 16.4 ms at config 3): off unless asked for, kept for the A/B
    static const bool pipe = getenv("PDG_ML_PIPE") && atoi(getenv("PDG_ML_PIPE")) != 0;
    if (pipe) ml_bucket_sort_kernel<true><<<grid, ML_THREADS, 0, s>>>(packed, pos_in, rowtab, n_buckets, k, out, heads, ctl, big, big_cap);
    else ml_bucket_sort_kernel<false><<<grid, ML_THREADS, 0, s>>>(packed, pos_in, rowtab, n_buckets, k, out, heads, ctl, big, big_cap);
    PDG_LAUNCHED(g);
    unsigned int h_ctl[4] = {0, 0, 0, 0};
    PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    if (!h_ctl[0] && h_ctl[1]) {   // buckets too large for a warp: one CTA each
        ml_mid_bucket_kernel<<<h_ctl[1], ML_THREADS, 0, s>>>(packed, pos_in, k, out, heads, ctl, big);
        PDG_LAUNCHED(g);
        PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 12, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        if (h_ctl[2]) {
            ml_big_bucket_kernel<<<h_ctl[1], ML_BIG_THREADS, ML_BIG * 12, s>>>(packed, pos_in, k, out, heads, ctl, big);
            PDG_LAUNCHED(g);
            PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 4, cudaMemcpyDeviceToHost, s));
        }
    }
    PDG_CUDA(cudaEventRecord(g->k_events[7], s));   // bucket sort + big buckets (with the host round trips between)
    PDG_CUDA(cudaStreamSynchronize(s));
    if (getenv("PDG_ML_DEBUG")) fprintf(stderr, "msdlong: %u buckets > %d records, %u > %d\n", h_ctl[1], 512, h_ctl[2], ML_MID);
    const unsigned int h_flag = h_ctl[0];
    const uint64_t n_groups = g->m3_groups.size();
    g->m3_groups.clear();
    g->m3_launched = false;
    if (h_flag) return false;
    {   // the bucket sort is timed on every path (its events bracket it here); the prefix passes only when they ran
        // back to back on this stream (a pipelined ingest interleaves them with the copies)
        float msw = 0.f;
        cudaEventElapsedTime(&msw, g->k_events[6], g->k_events[7]);
        g->stats.ms_sort_kernel_avg = msw;
        g->stats.sort_kernel_launches = 1;
    }
    if (g->m3_timed && !staged) {
        float ms[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 5; ++i) cudaEventElapsedTime(&ms[i], g->k_events[i], g->k_events[i + 1]);
        g->stats.ms_hist = ms[0];
        g->stats.ms_scans = ms[1];
        g->stats.ms_pass[0] = ms[2];
        g->stats.ms_pass[1] = ms[3];
        g->stats.ms_pass[2] = ms[4];
    }
    g->stats.sort_passes = 4;
    g->stats.sort_records = n;
    g->stats.sort_groups = n_groups;
    *sorted = out;
    *head_flags = heads;
    return true;
}

}  // namespace pdg
