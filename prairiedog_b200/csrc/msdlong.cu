// msdlong.cu -- occurrence index for 17 <= k <= 31 without sorting 128-bit records.
//
// The 2k-bit code of a k-mer starts with the code of its first 11 residues, so the offsets-first
// MSD sort of msd3.cu, run on that 22-bit prefix (validity still decided by the whole k-mer),
// groups the positions into 4^11 buckets in three passes over 8-byte records.  A bucket holds
// ~n / 4^11 positions (120 at config-3 scale) in ascending order; what is left is a stable sort of
// every bucket on the remaining 2k - 22 bits: a warp reads the full code of each of the bucket's
// records from the packed store, sorts the bucket on (code, position) in registers and writes the
// final position column plus, densely per bucket, the code and start of every row (distinct code) and
// the bucket's row count; ml_rows_kernel moves the rows to their ids after a scan over the bucket
// counts.  One launch per bucket-size class; buckets of more than 512 records get a whole CTA (warps
// sort runs in registers, places by binary search: up to 4096 records; a shared-memory network up to
// 16 384).
// The LSD path this replaces (radix.cu: 8 passes over 16-byte records, 256 B/record of traffic)
// stays as the fallback for inputs whose buckets do not fit a window (long low-complexity repeats).
// Reference counterpart: the node_idx / srcnode_idx / tgtnode_idx B-trees of lemongraph.c:591-630.
#include "gindex.cuh"
#include "msd3.cuh"
#include <algorithm>

namespace pdg {

static constexpr int ML_THREADS = 256;
#ifndef ML_CFG_CLASSES
#define ML_CFG_CLASSES 3   /* launches of the bucket sort by bucket-size class (see ml_class_sort_kernel); 1, 2 or 3 */
#endif
#ifndef ML_CFG_MIRROR
#define ML_CFG_MIRROR 1    /* class of 257-512 records: the network that skips its padding registers (ml_sort_bucket_c) */
#endif
#ifndef ML_CFG_SPOS
#define ML_CFG_SPOS 1      /* CTA sort: the bucket's positions are kept in shared memory instead of being read twice */
#endif
#ifndef ML_CFG_MINB_A
#define ML_CFG_MINB_A 8    /* CTAs per SM of the class of buckets <= 128 records */
#endif

// What a bucket sort writes.  Record i of the sorted order (i = bucket start + rank inside the bucket, which IS the
// record's final index: the buckets are in prefix order) gets its position in pos[i].  The HEADS of the bucket -- the
// records at which a row (a new code) starts -- are written densely from the bucket's start: hcode[e0 + j] and
// hidx[e0 + j] are the code and the record index of the bucket's j-th row, cnt[b] the number of rows of bucket b.
// ml_rows_kernel then copies them to their global row ids (an exclusive scan over the 4^11 bucket counts).
// (Until round 2 the kernels wrote 16-byte {code, position} records plus one 32-bit flag per record, and gbuild.cu
// made the rows with a scan over the 5e8 flags and a pass over the 8 GB of records: 5.5 ms of a 27 ms step at
// config 3, profiles/r02_launches_k31_classes.csv.  A first fix -- code of every record + one head BIT per record,
// rows by a popcount scan -- still spent 2.1 ms picking the heads' codes out of 4 GB, r02_launches_k31_final.csv.)
struct MlOut {
    uint32_t *pos;
    unsigned long long *hcode;
    uint32_t *hidx;
    uint32_t *cnt;
};

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
// next power of two >= cnt entries) and writes them.  All threads of the CTA take part.
template <int THREADS>
__device__ __forceinline__ void ml_sort_range(const unsigned long long *__restrict__ packed,
                                              const uint32_t *__restrict__ pos_in, uint32_t e0, uint32_t cnt, int k,
                                              unsigned long long *key, uint32_t *val, const MlOut &out,
                                              uint32_t bucket) {
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
    for (uint32_t i = tid; i < cnt; i += THREADS) out.pos[e0 + i] = val[i];
    if (tid < 32) {     // the rows of the bucket, in order, by one warp (these buckets are rare)
        uint32_t rows = 0;
        for (uint32_t i0 = 0; i0 < cnt; i0 += 32) {
            const uint32_t i = i0 + tid;
            const bool ok = i < cnt;
            const unsigned long long c = ok ? key[i] : 0ULL;
            const bool head = ok && (i == 0 || c != key[i - 1]);     // a bucket is a whole number of rows
            const uint32_t hm = __ballot_sync(0xffffffffu, head);
            if (head) {
                const uint32_t j = rows + __popc(hm & ((1u << tid) - 1u));
                out.hcode[e0 + j] = c;
                out.hidx[e0 + j] = e0 + i;
            }
            rows += __popc(hm);
        }
        if (tid == 0) out.cnt[bucket] = rows;
    }
}

// ---- one warp, one bucket: keys in registers -----------------------------------------------------------------
// Bitonic network over 32 * E keys held E per lane (key index = r * 32 + lane): strides >= 32 pair two registers of one
// lane, smaller strides pair two lanes through a shuffle.  No shared memory, no barrier.  (Earlier versions, config-3
// scale: windows of ~8 whole buckets sorted per CTA in shared memory, 40 ms; one warp per bucket in shared memory,
// 31.5 ms at 36 % issue -- every stage waits for a shared-memory round trip.)
//
// E is a RUN-TIME value.  Round 1 instantiated loads, network and stores for E = 1, 2, 4, 8, 16: 5184 SASS
// instructions, 83 KB, and since the warps of an SM work on buckets of different sizes at the same time the
// instruction caches thrashed (ncu: sm__icc_request_hit_rate 87.8 %, gcc__cache_requests_type_instruction at 91.6 % of
// peak, no_instruction 16 % of the stall samples) -- which is also why every variant that ADDED code (a
// three-buckets-ahead prefetch loop, batched loads) got slower.  Now the keys sit in EMAX register pairs, registers
// >= E are never touched, and the per-register work is skipped by warp-uniform branches over the register groups
// {0}, {1}, {2, 3}, {4..7}, {8..15}; the (size, stride) loops stay rolled.  2424 instructions for EMAX = 16;
// 17.5 -> 15.6 ms.  profiles/r02_k31_bucket_sort_experiments.txt has all of it.
template <int R> struct MlRc { static constexpr int value = R; };
template <int EMAX, typename F>
__device__ __forceinline__ void ml_for_regs(uint32_t E, F f) {
    f(MlRc<0>());
    if constexpr (EMAX > 1) if (E > 1) {
        f(MlRc<1>());
        if constexpr (EMAX > 2) if (E > 2) {
            f(MlRc<2>()); f(MlRc<3>());
            if constexpr (EMAX > 4) if (E > 4) {
                f(MlRc<4>()); f(MlRc<5>()); f(MlRc<6>()); f(MlRc<7>());
                if constexpr (EMAX > 8) if (E > 8) {
                    f(MlRc<8>()); f(MlRc<9>()); f(MlRc<10>()); f(MlRc<11>());
                    f(MlRc<12>()); f(MlRc<13>()); f(MlRc<14>()); f(MlRc<15>());
                }
            }
        }
    }
}

// stride 32 * RS pairs registers (r, r | RS); ascending where (r & s5) == 0, s5 = size / 32
template <int EMAX, int RS>
__device__ __forceinline__ void ml_reg_stage(unsigned long long (&key)[EMAX], uint32_t s5, uint32_t E) {
    ml_for_regs<EMAX>(E, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        if constexpr ((r & RS) != 0) {              // visited as the upper register of its pair: E > r > r ^ RS
            constexpr int lo = r ^ RS;
            const bool up = ((uint32_t)lo & s5) == 0u;
            const unsigned long long a = key[lo], b = key[r];
            const bool sw = (a > b) == up;
            key[lo] = sw ? b : a;
            key[r] = sw ? a : b;
        }
    });
}

template <int EMAX>
__device__ __forceinline__ void ml_warp_bitonic(unsigned long long (&key)[EMAX], uint32_t E, int lane) {
    const uint32_t n = 32u * E;
#pragma unroll 1
    for (uint32_t size = 2; size <= n; size <<= 1) {
        const uint32_t s5 = size >> 5;
        // bit r set: register r's block descends at this size (size < 32: the same for all registers of a lane)
        const uint32_t dn = size < 32u ? (((uint32_t)lane & size) ? 0xffffu : 0u)
                                       : (s5 == 1u ? 0xaaaau : s5 == 2u ? 0xccccu : s5 == 4u ? 0xf0f0u : s5 == 8u ? 0xff00u : 0u);
#pragma unroll 1
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            if (stride >= 32u) {
                const uint32_t rs = stride >> 5;
                if (rs == 1u) { if constexpr (EMAX > 1) ml_reg_stage<EMAX, 1>(key, s5, E); }
                else if (rs == 2u) { if constexpr (EMAX > 2) ml_reg_stage<EMAX, 2>(key, s5, E); }
                else if (rs == 4u) { if constexpr (EMAX > 4) ml_reg_stage<EMAX, 4>(key, s5, E); }
                else { if constexpr (EMAX > 8) ml_reg_stage<EMAX, 8>(key, s5, E); }
            } else {
                // keep the smaller key where (lower lane of the pair) != (descending block)
                const uint32_t mx = ((uint32_t)lane & stride) ? ~dn : dn;      // bit r set: keep the LARGER key
                ml_for_regs<EMAX>(E, [&](auto rc) {
                    constexpr int r = decltype(rc)::value;
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[r], (int)stride);
                    const bool keep_max = (mx >> r) & 1u;
                    key[r] = ((key[r] > o) == keep_max) ? key[r] : o;
                });
            }
        }
    }
}

// Sorts bucket [e0, e0 + m) (m <= 32 * EMAX) on (suffix, position) and writes it.  Key = suffix of the code below the
// 22-bit prefix << 9 | index in the bucket: the index keeps equal codes in position order and finds the position
// again afterwards (the bucket's positions ascend).
// The key build goes register by register -- position, then its packed words: 2 E dependent trips to memory, every
// one a long_scoreboard stall in ncu's source page.  Overlapping them inside the warp (all positions first, the packed
// words of four registers at a time) was measured twice and never won (13.3 vs 13.2-14.4 ms per class layout); what
// did win was more warps per SM -- see ml_class_sort_kernel.
template <int EMAX>
__device__ __forceinline__ void ml_sort_bucket(const unsigned long long *__restrict__ packed,
                                               const uint32_t *__restrict__ pos_in, uint32_t e0, uint32_t m, int k,
                                               unsigned long long prefix, int lane, const MlOut &out) {
    const uint32_t E = m <= 32u ? 1u : m <= 64u ? 2u : m <= 128u ? 4u : m <= 256u ? 8u : 16u;   // <= EMAX (caller)
    const int sb = 2 * k - kMsd3MaxBits;                       // suffix bits, <= 40
    const unsigned long long smask = (1ULL << sb) - 1ULL;
    unsigned long long key[EMAX];
    ml_for_regs<EMAX>(E, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        key[r] = i < m ? (((ml_code_at(packed, pos_in[e0 + i], k) & smask) << 9) | i) : ~0ULL;
    });
    ml_warp_bitonic<EMAX>(key, E, lane);
    // a record starts a row when its code differs from its predecessor's (the first of a bucket always does)
    unsigned long long carry = ~0ULL;                          // suffix of key index r * 32 - 1
    uint32_t rows = 0;                                         // rows of the bucket so far
    ml_for_regs<EMAX>(E, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        const unsigned long long sfx = key[r] >> 9;
        unsigned long long prev = __shfl_up_sync(0xffffffffu, sfx, 1);
        if (lane == 0) prev = carry;
        carry = __shfl_sync(0xffffffffu, sfx, 31);
        const bool ok = i < m, head = ok && (i == 0 || sfx != prev);
        const uint32_t hm = __ballot_sync(0xffffffffu, head);
        if (ok) __stcs(out.pos + e0 + i, pos_in[e0 + (uint32_t)(key[r] & 511u)]);
        if (head) {
            const uint32_t j = rows + __popc(hm & ((1u << lane) - 1u));
            __stcs(out.hcode + e0 + j, (prefix << sb) | sfx);
            __stcs(out.hidx + e0 + j, e0 + i);
        }
        rows += __popc(hm);
    });
    if (lane == 0) out.cnt[prefix] = rows;      // prefix = the bucket's number
}

// ---- the class of 257-512 records: a network that skips its padding ------------------------------------------
// A bucket of 300 records needs 10 key registers per lane, the bitonic network above a power of two: 16, of which 6
// hold nothing but +inf.  In the form of the network whose comparators ALL put the smaller key at the lower index (the
// first step of every merge pairs i with its mirror image i ^ (size - 1) instead of i ^ (size / 2)), a +inf at the top
// never moves, so every comparator that touches a register >= E can be left out: the 45 stages of n = 512 run over
// E2 = E rounded up to even (10, 12, 14, 16) registers.  The mirror step pairs register r with r ^ (size / 32 - 1) of
// the lane ^ 31; both keys of a pair are fetched before either is replaced.
__device__ __forceinline__ unsigned long long ml_min64(unsigned long long a, unsigned long long b) { return a < b ? a : b; }
__device__ __forceinline__ unsigned long long ml_max64(unsigned long long a, unsigned long long b) { return a < b ? b : a; }

// f(r) for the registers in use.  ANY = false: E2 >= 10 is known (257-512 records), registers 0..9 unguarded;
// ANY = true: E2 = 0, 2, ..., 16, every pair behind a warp-uniform branch
template <bool ANY, typename F>
__device__ __forceinline__ void ml_for_regs_c(uint32_t E2, F f) {
    if constexpr (ANY) {
        if (E2 > 0) {
            f(MlRc<0>()); f(MlRc<1>());
            if (E2 > 2) {
                f(MlRc<2>()); f(MlRc<3>());
                if (E2 > 4) {
                    f(MlRc<4>()); f(MlRc<5>());
                    if (E2 > 6) {
                        f(MlRc<6>()); f(MlRc<7>());
                        if (E2 > 8) { f(MlRc<8>()); f(MlRc<9>()); }
                    }
                }
            }
        }
    } else {
        f(MlRc<0>()); f(MlRc<1>()); f(MlRc<2>()); f(MlRc<3>()); f(MlRc<4>()); f(MlRc<5>()); f(MlRc<6>()); f(MlRc<7>());
        f(MlRc<8>()); f(MlRc<9>());
    }
    if (E2 > 10) {
        f(MlRc<10>()); f(MlRc<11>());
        if (E2 > 12) {
            f(MlRc<12>()); f(MlRc<13>());
            if (E2 > 14) { f(MlRc<14>()); f(MlRc<15>()); }
        }
    }
}

// mirror step of size 32 * S5 (S5 = 2, 4, 8, 16): (register a, lane) with (register a ^ (S5 - 1), lane ^ 31)
template <int S5, int A>
__device__ __forceinline__ void ml_mirror_pair(unsigned long long (&key)[16], uint32_t E2) {
    constexpr int B = A ^ (S5 - 1);
    if constexpr (A < B) {
        if ((uint32_t)B < E2) {
            const unsigned long long oa = __shfl_xor_sync(0xffffffffu, key[B], 31);
            const unsigned long long ob = __shfl_xor_sync(0xffffffffu, key[A], 31);
            key[A] = ml_min64(key[A], oa);
            key[B] = ml_max64(key[B], ob);
        }
    }
}
template <int S5>
__device__ __forceinline__ void ml_mirror_stage(unsigned long long (&key)[16], uint32_t E2) {
    ml_mirror_pair<S5, 0>(key, E2); ml_mirror_pair<S5, 1>(key, E2); ml_mirror_pair<S5, 2>(key, E2); ml_mirror_pair<S5, 3>(key, E2);
    ml_mirror_pair<S5, 4>(key, E2); ml_mirror_pair<S5, 5>(key, E2); ml_mirror_pair<S5, 6>(key, E2); ml_mirror_pair<S5, 7>(key, E2);
    ml_mirror_pair<S5, 8>(key, E2); ml_mirror_pair<S5, 9>(key, E2); ml_mirror_pair<S5, 10>(key, E2); ml_mirror_pair<S5, 11>(key, E2);
    ml_mirror_pair<S5, 12>(key, E2); ml_mirror_pair<S5, 13>(key, E2); ml_mirror_pair<S5, 14>(key, E2); ml_mirror_pair<S5, 15>(key, E2);
}

// stride 32 * RS: registers (a, a | RS) of one lane, smaller key to a
template <int RS, int A>
__device__ __forceinline__ void ml_up_pair(unsigned long long (&key)[16], uint32_t E2) {
    if constexpr ((A & RS) == 0) {
        constexpr int B = A | RS;
        if ((uint32_t)B < E2) {
            const unsigned long long a = key[A], b = key[B];
            key[A] = ml_min64(a, b);
            key[B] = ml_max64(a, b);
        }
    }
}
template <int RS>
__device__ __forceinline__ void ml_up_stage(unsigned long long (&key)[16], uint32_t E2) {
    ml_up_pair<RS, 0>(key, E2); ml_up_pair<RS, 1>(key, E2); ml_up_pair<RS, 2>(key, E2); ml_up_pair<RS, 3>(key, E2);
    ml_up_pair<RS, 4>(key, E2); ml_up_pair<RS, 5>(key, E2); ml_up_pair<RS, 6>(key, E2); ml_up_pair<RS, 7>(key, E2);
    ml_up_pair<RS, 8>(key, E2); ml_up_pair<RS, 9>(key, E2); ml_up_pair<RS, 10>(key, E2); ml_up_pair<RS, 11>(key, E2);
    ml_up_pair<RS, 12>(key, E2); ml_up_pair<RS, 13>(key, E2); ml_up_pair<RS, 14>(key, E2); ml_up_pair<RS, 15>(key, E2);
}

// all 512 (virtual) keys ascending; registers >= E2 do not exist
template <bool ANY>
__device__ __forceinline__ void ml_warp_sort_c(unsigned long long (&key)[16], uint32_t E2, int lane) {
    // lanes: exchange with lane ^ x, the lower lane of the pair (bit `lowbit` clear) keeps the smaller key
    auto lane_stage = [&](uint32_t x, uint32_t lowbit) {
        const bool lower = ((uint32_t)lane & lowbit) == 0u;
        ml_for_regs_c<ANY>(E2, [&](auto rc) {
            constexpr int r = decltype(rc)::value;
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[r], (int)x);
            key[r] = ((key[r] < o) == lower) ? key[r] : o;
        });
    };
    uint32_t nmax = 512u;                       // smallest power of two >= 32 * E2 (at least 32)
    if constexpr (ANY) { nmax = 32u; while (nmax < 32u * E2) nmax <<= 1; }
#pragma unroll 1
    for (uint32_t size = 2; size <= nmax; size <<= 1) {
        if (size <= 32u) lane_stage(size - 1u, size >> 1);
        else if (size == 64u) ml_mirror_stage<2>(key, E2);
        else if (size == 128u) ml_mirror_stage<4>(key, E2);
        else if (size == 256u) ml_mirror_stage<8>(key, E2);
        else ml_mirror_stage<16>(key, E2);
#pragma unroll 1
        for (uint32_t stride = size >> 2; stride > 0; stride >>= 1) {
            if (stride >= 32u) {
                if (stride == 32u) ml_up_stage<1>(key, E2);
                else if (stride == 64u) ml_up_stage<2>(key, E2);
                else ml_up_stage<4>(key, E2);
            } else {
                lane_stage(stride, stride);
            }
        }
    }
}

// bucket of 257..512 records (9 <= E <= 16 registers in use)
__device__ __forceinline__ void ml_sort_bucket_c(const unsigned long long *__restrict__ packed,
                                                 const uint32_t *__restrict__ pos_in, uint32_t e0, uint32_t m, int k,
                                                 unsigned long long prefix, int lane, const MlOut &out) {
    const uint32_t E2 = (((m + 31u) >> 5) + 1u) & ~1u;         // 10, 12, 14 or 16
    const int sb = 2 * k - kMsd3MaxBits;
    const unsigned long long smask = (1ULL << sb) - 1ULL;
    unsigned long long key[16];
    ml_for_regs_c<false>(E2, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        key[r] = i < m ? (((ml_code_at(packed, pos_in[e0 + i], k) & smask) << 9) | i) : ~0ULL;
    });
    ml_warp_sort_c<false>(key, E2, lane);
    unsigned long long carry = ~0ULL;
    uint32_t rows = 0;
    ml_for_regs_c<false>(E2, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        const unsigned long long sfx = key[r] >> 9;
        unsigned long long prev = __shfl_up_sync(0xffffffffu, sfx, 1);
        if (lane == 0) prev = carry;
        carry = __shfl_sync(0xffffffffu, sfx, 31);
        const bool ok = i < m, head = ok && (i == 0 || sfx != prev);
        const uint32_t hm = __ballot_sync(0xffffffffu, head);
        if (ok) __stcs(out.pos + e0 + i, pos_in[e0 + (uint32_t)(key[r] & 511u)]);
        if (head) {
            const uint32_t j = rows + __popc(hm & ((1u << lane) - 1u));
            __stcs(out.hcode + e0 + j, (prefix << sb) | sfx);
            __stcs(out.hidx + e0 + j, e0 + i);
        }
        rows += __popc(hm);
    });
    if (lane == 0) out.cnt[prefix] = rows;      // prefix = the bucket's number
}

// Buckets of (32 * ELO, 32 * EMAX] records, one WARP per bucket (strided over the 4^11 buckets); the other buckets are
// left to the launches of the other classes.  Splitting by class buys occupancy: the kernel is bound by the latency
// of its dependent loads, and what hides it is warps -- 3 CTAs per SM 19.0 ms, 4: 15.6, and a class of <= 256 records
// needs only 16 key registers = 40 registers per thread = 6 CTAs per SM, one of <= 128 records 31 = 8 CTAs.
// One launch for everything 15.6 ms, two classes 13.6, three 13.3 (config 3).  The class of the largest buckets
// (EMAX = 16) also lists the buckets above 512 records in big[] for the CTA kernels below.
// ctl[0] = overflow flag, ctl[1] = entries in big[].
template <int ELO, int EMAX, int MINB>
__global__ void __launch_bounds__(ML_THREADS, MINB) ml_class_sort_kernel(const unsigned long long *__restrict__ packed,
                                                                        const uint32_t *__restrict__ pos_in,
                                                                        const uint32_t *__restrict__ rowtab, uint32_t n_buckets,
                                                                        int k, MlOut out, unsigned int *__restrict__ ctl,
                                                                        uint32_t *__restrict__ big, uint32_t big_cap) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t nw = gridDim.x * (ML_THREADS / 32);
    for (uint32_t b = blockIdx.x * (ML_THREADS / 32) + w; b < n_buckets; b += nw) {
        const uint32_t e0 = rowtab[b], m = rowtab[b + 1] - e0;
        if (m <= 32u * ELO) continue;
        if (m <= 32u * EMAX) {
#if ML_CFG_MIRROR
            if constexpr (ELO == 8 && EMAX == 16) ml_sort_bucket_c(packed, pos_in, e0, m, k, b, lane, out);
            else
#endif
                ml_sort_bucket<EMAX>(packed, pos_in, e0, m, k, b, lane, out);
        }
        else if (EMAX == 16 && lane == 0) {
            const unsigned int slot = atomicAdd(&ctl[1], 1u);
            if (slot < big_cap) big[slot] = b; else atomicOr(&ctl[0], 1u);
        }
    }
}

// Buckets too large for a warp (big[] holds their numbers), one CTA each.
// 513..ML_MID records (ml_mid_bucket_kernel): the NW warps of the CTA each sort one NW-th of the bucket in registers
// with the network above, the sorted runs go to shared memory, and every record finds its place in the bucket as
// (index in its own run) + (records of the other runs below it) -- NW - 1 binary searches in shared memory; the
// keys are unique (the index in the bucket is part of them), so there are no ties to break.  Round 1 sorted these
// buckets with a bitonic network over shared memory, 55-66 block-wide stages: at 500 genomes per GPU, where most
// buckets are of this size, that was ~200 of a 272 ms step (config 4 on 2 GPUs).
// Up to ML_BIG records (ml_big_bucket_kernel, dynamic shared memory, one CTA per SM): the shared-memory network, for
// the rare huge bucket (low-complexity sequence); above that the caller takes the LSD path.
// NW = warps of the CTA = runs: 4 for buckets of up to 2048 records, 8 for up to 4096; a bucket above 512 NW records is
// appended to the list of the next tier (next[], its length in ctl[ctl_slot]) -- a launch over the FIRST list for every
// tier cost 1.4 M empty CTAs each at 500 genomes per GPU, 11 ms for the one-CTA-per-SM kernel alone.
static constexpr int ML_MID = 4096, ML_BIG = 16384, ML_BIG_THREADS = 1024;
template <int NW>
__global__ void __launch_bounds__(32 * NW) ml_mid_bucket_kernel(const unsigned long long *__restrict__ packed,
                                                                const uint32_t *__restrict__ pos_in,
                                                                const uint32_t *__restrict__ rowtab, int k, MlOut out,
                                                                unsigned int *__restrict__ ctl, const uint32_t *__restrict__ big,
                                                                uint32_t *__restrict__ next, int ctl_slot) {
    __shared__ unsigned long long skey[512 * NW];
    __shared__ uint32_t spos[512 * NW];        // the bucket's positions: read back by index after the sort (a second
                                               // read of pos_in missed L1 and L2 by then: ncu 186 B of DRAM reads per record)
    uint32_t *s_rows = spos;                   // rows per warp: spos is free again by then (48 KB of static shared memory at NW = 8)
    const uint32_t b = big[blockIdx.x], e0 = rowtab[b], cnt = rowtab[b + 1] - e0;
    if (cnt > 512u * NW) {          // passed on to the next tier's list
        if (threadIdx.x == 0) next[atomicAdd(&ctl[ctl_slot], 1u)] = b;
        return;
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int sb = 2 * k - kMsd3MaxBits;                       // suffix bits, <= 40; key = suffix << 12 | index in the bucket
    const unsigned long long smask = (1ULL << sb) - 1ULL;
    const uint32_t Q = (((cnt + NW - 1u) / NW) + 31u) & ~31u;  // run length, a multiple of 32, <= 512
    const uint32_t r0 = min((uint32_t)w * Q, cnt), m = min(Q, cnt - r0);   // this warp's run [r0, r0 + m)
    const uint32_t E2 = (((m + 31u) >> 5) + 1u) & ~1u;         // 0, 2, ..., 16 registers
    unsigned long long key[16];
    ml_for_regs_c<true>(E2, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        key[r] = ~0ULL;
        if (i < m) {
            const uint32_t p = pos_in[e0 + r0 + i];
#if ML_CFG_SPOS
            spos[r0 + i] = p;
#endif
            key[r] = ((ml_code_at(packed, p, k) & smask) << 12) | (r0 + i);
        }
    });
    ml_warp_sort_c<true>(key, E2, lane);
    ml_for_regs_c<true>(E2, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        if (i < m) skey[r0 + i] = key[r];
    });
    __syncthreads();
    // place in the bucket = index in the own run + records of every other run below the key (fixed-length searches:
    // no divergence, and the two registers of a pair search side by side)
    uint32_t place[16];
    ml_for_regs_c<true>(E2, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        place[r] = r * 32 + lane;
    });
#pragma unroll 1
    for (uint32_t ow = 0; ow < (uint32_t)NW; ++ow) {
        if (ow == (uint32_t)w) continue;
        const uint32_t o0 = min(ow * Q, cnt), om = min(Q, cnt - o0);
        if (om == 0) continue;
        const unsigned long long *run = skey + o0;
        ml_for_regs_c<true>(E2, [&](auto rc) {
            constexpr int r = decltype(rc)::value;
            uint32_t lo = 0;
#pragma unroll
            for (uint32_t step = 512; step; step >>= 1) {
                const uint32_t t = lo + step;
                if (t <= om && run[t - 1] < key[r]) lo = t;
            }
            place[r] += lo;
        });
    }
    __syncthreads();                                           // the runs have been read: skey becomes the sorted bucket
    ml_for_regs_c<true>(E2, [&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const uint32_t i = r * 32 + lane;
        if (i < m) {
            skey[place[r]] = key[r];
#if ML_CFG_SPOS
            out.pos[e0 + place[r]] = spos[(uint32_t)(key[r] & 4095u)];
#else
            out.pos[e0 + place[r]] = pos_in[e0 + (uint32_t)(key[r] & 4095u)];
#endif
        }
    });
    __syncthreads();
    // rows: warp w takes the sorted records [h0, h1); first their number, then -- after the warps before it -- the heads
    const uint32_t h0 = min((uint32_t)w * Q, cnt), h1 = min(h0 + Q, cnt);
    auto heads_of = [&](uint32_t i0, bool &head, unsigned long long &c, uint32_t &i) {
        i = i0 + lane;
        const bool ok = i < h1;
        c = ok ? (skey[i] >> 12) : 0ULL;
        head = ok && (i == 0 || c != (skey[i - 1] >> 12));
        return __ballot_sync(0xffffffffu, head);
    };
    uint32_t mine = 0;
    for (uint32_t i0 = h0; i0 < h1; i0 += 32) {
        bool head; unsigned long long c; uint32_t i;
        mine += __popc(heads_of(i0, head, c, i));
    }
    if (lane == 0) s_rows[w] = mine;
    __syncthreads();
    uint32_t rows = 0;
    for (int ww = 0; ww < w; ++ww) rows += s_rows[ww];
    for (uint32_t i0 = h0; i0 < h1; i0 += 32) {
        bool head; unsigned long long c; uint32_t i;
        const uint32_t hm = heads_of(i0, head, c, i);
        if (head) {
            const uint32_t j = rows + __popc(hm & ((1u << lane) - 1u));
            out.hcode[e0 + j] = ((unsigned long long)b << sb) | c;
            out.hidx[e0 + j] = e0 + i;
        }
        rows += __popc(hm);
    }
    if (w == NW - 1 && lane == 0) out.cnt[b] = rows;           // the last warp's running count is the bucket's
}

__global__ void __launch_bounds__(ML_BIG_THREADS) ml_big_bucket_kernel(const unsigned long long *__restrict__ packed,
                                                                       const uint32_t *__restrict__ pos_in,
                                                                       const uint32_t *__restrict__ rowtab, int k, MlOut out,
                                                                       unsigned int *__restrict__ ctl, const uint32_t *__restrict__ big) {
    extern __shared__ __align__(16) unsigned char ml_raw[];
    unsigned long long *key = reinterpret_cast<unsigned long long *>(ml_raw);
    uint32_t *val = reinterpret_cast<uint32_t *>(ml_raw + (size_t)ML_BIG * 8);
    const uint32_t b = big[blockIdx.x], e0 = rowtab[b], cnt = rowtab[b + 1] - e0;
    if (cnt > (uint32_t)ML_BIG) {          // the caller takes the LSD path
        if (threadIdx.x == 0) atomicOr(&ctl[0], 1u);
        return;
    }
    ml_sort_range<ML_BIG_THREADS>(packed, pos_in, e0, cnt, k, key, val, out, b);
}

// rows: eight lanes per bucket copy the bucket's heads to their global row ids; cbase = exclusive scan of cnt (+ total).
// (One WARP per bucket, bounds loaded after the count: 1.55 ms for 3 GB at config 3 -- ~30 rows per bucket do not keep a
// warp's loads busy; four buckets per warp and the three bounds loaded together: see r02_k31_bucket_sort_experiments.txt.)
__global__ void __launch_bounds__(256) ml_rows_kernel(const uint32_t *__restrict__ rowtab, const uint32_t *__restrict__ cnt,
                                                      const uint32_t *__restrict__ cbase, uint32_t n_buckets,
                                                      const unsigned long long *__restrict__ hcode,
                                                      const uint32_t *__restrict__ hidx, uint64_t *__restrict__ codes,
                                                      uint32_t *__restrict__ rowptr, uint32_t n) {
    const uint32_t b = blockIdx.x * 32 + (threadIdx.x >> 3), l = threadIdx.x & 7;
    if (b >= n_buckets) return;
    const uint32_t c = cnt[b], e0 = rowtab[b], base = cbase[b];
    if (b == n_buckets - 1 && l == 0) rowptr[cbase[n_buckets]] = n;
    for (uint32_t j0 = l; j0 < c; j0 += 32) {       // four rows per lane and trip: all eight loads before the first store
        unsigned long long hc[4];
        uint32_t hi[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t j = j0 + 8 * u;
            hc[u] = j < c ? __ldcs(hcode + e0 + j) : 0ULL;
            hi[u] = j < c ? __ldcs(hidx + e0 + j) : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t j = j0 + 8 * u;
            if (j < c) {
                codes[base + j] = hc[u];
                rowptr[base + j] = hi[u];
            }
        }
    }
}

// true: the occurrence index is complete -- g->g_pos holds the n positions in (code, position) order, g->g_codes the
// g->g_nodes distinct codes in ascending order, g->g_rowptr their row starts (+ sentinel)
// staged: a pipelined ingest already ran the prefix sort up to pass 2 (or 3) for exactly these contigs.
bool msdlong_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, bool staged) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    const uint64_t n = hc.n_kmers;
    const uint32_t n_buckets = 1u << kMsd3MaxBits;
    static const bool off = getenv("PDG_NO_MSDLONG") != nullptr;
    if (off || 2 * k <= kMsd3MaxBits || n == 0 || n >= (1ULL << 32)) return false;
    if (!staged) {
        msd3_reserve(g, hc.pos0, hc.pos1, 1);
        msd3_stage_group(g, ct, hc.pos0, hc.pos1);
    }
    if (!g->m3_launched) msd3_finish_launch(g, /*rows=*/false);   // g_pos grouped by prefix, scratch0 = bucket starts (+ sentinel)
    // recA (pass 1's output) and recB (pass 2's) are free once pass 3 is in the stream: the heads' codes go to recA, the
    // sorted positions to recB, which then changes places with g_pos.  (recA >= n * 8 and recB >= n * 8 bytes:
    // msd3_reserve.)  scratch1 (pass 2's offsets, not read after it): heads' record indices | rows per bucket | their scan.
    g->scratch1.reserve(((size_t)n + 2 * (size_t)n_buckets + 2) * 4, s);
    uint32_t *hidx = g->scratch1.as<uint32_t>(), *cnt = hidx + n, *cbase = cnt + n_buckets;
    PDG_CUDA(cudaMemsetAsync(cnt, 0, (size_t)n_buckets * 4, s));
    // room for EVERY bucket (33.5 MB): with a fixed 2^18 entries, 500 genomes per GPU (config 4 on 2 GPUs: ~600
    // records per bucket, nearly all of the 4^11 buckets above the warp path's 512) overflowed the list and sent the
    // whole build to the 8-pass LSD sort of 16-byte records -- 282 ms for 5x the records of a 22 ms sort
    const uint32_t big_cap = n_buckets;
    g->ml_flag.reserve(16 + 3 * (size_t)big_cap * sizeof(uint32_t), s);     // ctl | list > 512 | list > 2048 | list > 4096
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
    const unsigned long long *packed = g->packed.as<unsigned long long>();
    const uint32_t *pos_in = g->g_pos.as<uint32_t>(), *rowtab = g->scratch0.as<uint32_t>();
    const MlOut out{g->recB.as<uint32_t>(), g->recA.as<unsigned long long>(), hidx, cnt};
    unsigned int *ctl = g->ml_flag.as<unsigned int>();
    uint32_t *big = reinterpret_cast<uint32_t *>(g->ml_flag.as<unsigned char>() + 16);
    // (an L2 persistence window over the packed store was tried for the random code gathers: the kernel
    // did not get faster and the rest of the step lost 16 ms to the limit / reset calls)
    // PDG_ML_CLASSES: 1 = one launch for all bucket sizes (4 CTAs/SM), 2 = (<= 256 | <= 512), 3 = (<= 128 | <= 256 | <= 512)
    static const int classes = getenv("PDG_ML_CLASSES") ? atoi(getenv("PDG_ML_CLASSES")) : ML_CFG_CLASSES;
#define ML_LAUNCH(ELO, EMAX, MINB)                                                                                       \
    do {                                                                                                                 \
        ml_class_sort_kernel<ELO, EMAX, MINB><<<148 * MINB * 4, ML_THREADS, 0, s>>>(packed, pos_in, rowtab, n_buckets, k, \
                                                                                   out, ctl, big, big_cap);           \
        PDG_LAUNCHED(g);                                                                                                 \
    } while (0)
    PDG_CUDA(cudaEventRecord(g->k_events[6], s));
    if (classes >= 3) {
        ML_LAUNCH(0, 4, ML_CFG_MINB_A);
        ML_LAUNCH(4, 8, 6);
        ML_LAUNCH(8, 16, 4);
    } else if (classes == 2) {
        ML_LAUNCH(0, 8, 6);
        ML_LAUNCH(8, 16, 4);
    } else {
        ML_LAUNCH(0, 16, 4);
    }
#undef ML_LAUNCH
    g->stats.sort_kernel_launches = classes >= 3 ? 3 : classes == 2 ? 2 : 1;
    unsigned int h_ctl[4] = {0, 0, 0, 0};
    PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    if (!h_ctl[0] && h_ctl[1]) {   // buckets too large for a warp: one CTA each
        uint32_t *big2 = big + big_cap, *big3 = big2 + big_cap;
        ml_mid_bucket_kernel<4><<<h_ctl[1], 128, 0, s>>>(packed, pos_in, rowtab, k, out, ctl, big, big2, 2);
        PDG_LAUNCHED(g);
        PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 16, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        if (h_ctl[2]) {     // some above 2048 records
            ml_mid_bucket_kernel<8><<<h_ctl[2], 256, 0, s>>>(packed, pos_in, rowtab, k, out, ctl, big2, big3, 3);
            PDG_LAUNCHED(g);
            PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 16, cudaMemcpyDeviceToHost, s));
            PDG_CUDA(cudaStreamSynchronize(s));
        }
        if (h_ctl[3]) {     // some above 4096
            ml_big_bucket_kernel<<<h_ctl[3], ML_BIG_THREADS, ML_BIG * 12, s>>>(packed, pos_in, rowtab, k, out, ctl, big3);
            PDG_LAUNCHED(g);
            PDG_CUDA(cudaMemcpyAsync(h_ctl, ctl, 4, cudaMemcpyDeviceToHost, s));
        }
    }
    PDG_CUDA(cudaEventRecord(g->k_events[7], s));   // bucket sort + big buckets (with the host round trips between)
    PDG_CUDA(cudaStreamSynchronize(s));
    if (getenv("PDG_ML_DEBUG")) fprintf(stderr, "msdlong: %u buckets > 512 records, %u > 2048, %u > %d\n", h_ctl[1], h_ctl[2], h_ctl[3], ML_MID);
    const unsigned int h_flag = h_ctl[0];
    const uint64_t n_groups = g->m3_groups.size();
    g->m3_groups.clear();
    g->m3_launched = false;
    if (h_flag) return false;
    {   // the bucket sort is timed on every path (its events bracket it here); the prefix passes only when they ran
        // back to back on this stream (a pipelined ingest interleaves them with the copies)
        float msw = 0.f;
        cudaEventElapsedTime(&msw, g->k_events[6], g->k_events[7]);
        g->stats.ms_sort_kernel_avg = msw;        // all classes together
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
    // rows: scan over the rows per bucket, then the heads to their row ids
    exclusive_scan_u32_u32(cnt, cbase, n_buckets, g);
    uint32_t nn = 0;
    PDG_CUDA(cudaMemcpyAsync(&nn, cbase + n_buckets, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->g_nodes = nn;
    g->g_codes.reserve(std::max<size_t>((size_t)nn * 8, 16), s);
    g->g_rowptr.reserve(((size_t)nn + 1) * 4, s);
    ml_rows_kernel<<<(n_buckets + 31) / 32, 256, 0, s>>>(rowtab, cnt, cbase, n_buckets, out.hcode, hidx, g->g_codes.as<uint64_t>(),
                                                      g->g_rowptr.as<uint32_t>(), (uint32_t)n);
    PDG_LAUNCHED(g);
    // the sorted position column is the index's: the two buffers change places (both hold >= n * 4 bytes)
    std::swap(g->g_pos.p, g->recB.p);
    std::swap(g->g_pos.cap, g->recB.cap);
    return true;
}

}  // namespace pdg
