// comm.cu -- K5: multi-GPU node deduplication (one process per GPU, NCCL over NVLink/NVSwitch).
//
// The reference has no distributed path at all: every Snakemake job writes into
// one LMDB file behind a single-writer lock.  Here genomes shard by sample: every rank keeps the
// edges of its own genomes (an edge is unique on (type,value) and a type belongs to one genome,
// so edges never need to move), and only k-mer NODE keys are exchanged:
//   1. each rank deduplicates its own k-mer codes                         (bitmap / sort+RLE)
//   2. codes are range partitioned (range_owner below: equal slices of the key space) and exchanged with ONE
//      grouped ncclSend/ncclRecv all-to-all (counts first, W x W uint64 ncclAllGather)
//   3. each owner radix-sorts + run-length compacts what it received   -> its owned unique set
//   4. the owned sets (disjoint ranges, ascending with the rank) are all-gathered in rank order -> the
//      same sorted global node table, hence the same dense node ids, on every rank.
// (Round 1 partitioned by hash: the gathered table then needed a full 64-bit sort on every rank,
//  103 ms of a 160 ms step at k = 31 on 8 GPUs.)
// NCCL is resolved with dlopen("libnccl.so.2") so that the library has no link-time dependency
// on it and picks up the copy torch already loaded.
#include "common.cuh"
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>

namespace pdg {

struct NcclApi {
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi *nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        api.h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!api.h) api.h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (api.h) {
#define PDG_SYM(field, name) api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.h, name))
            PDG_SYM(GetUniqueId, "ncclGetUniqueId");
            PDG_SYM(CommInitRank, "ncclCommInitRank");
            PDG_SYM(CommDestroy, "ncclCommDestroy");
            PDG_SYM(AllGather, "ncclAllGather");
            PDG_SYM(Broadcast, "ncclBroadcast");
            PDG_SYM(Send, "ncclSend");
            PDG_SYM(Recv, "ncclRecv");
            PDG_SYM(GroupStart, "ncclGroupStart");
            PDG_SYM(GroupEnd, "ncclGroupEnd");
            PDG_SYM(GetErrorString, "ncclGetErrorString");
#undef PDG_SYM
        }
    }
    if (!api.h || !api.GetUniqueId || !api.CommInitRank || !api.AllGather || !api.Broadcast || !api.Send ||
        !api.Recv || !api.GroupStart || !api.GroupEnd) {
        set_error("libnccl.so.2 could not be loaded: %s", dlerror());
        return nullptr;
    }
    return &api;
}

struct CommError { int code; };
#define PDG_NCCL(api, expr)                                                                   \
    do {                                                                                      \
        ncclResult_t _r = (expr);                                                             \
        if (_r != ncclSuccess) {                                                              \
            set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                           \
                      (api)->GetErrorString ? (api)->GetErrorString(_r) : "nccl error");     \
            throw pdg::CudaError{PDG_E_COMM};                                                 \
        }                                                                                     \
    } while (0)

// Owner of a node key: the key space is cut into `world` equal ranges (opaque keys, bit 63 set, sort
// after every k-mer code and belong to the last rank).  Monotone in the key, so the owned sets are
// disjoint RANGES: a rank's sorted table is already grouped by owner, and the owned sets of ranks
// 0..W-1, concatenated, are the global sorted table -- no scatter before and no sort after the exchange.
// code_bits = 2k.  Host mirror: prairiedog_b200/distributed.py:owner.
__host__ __device__ __forceinline__ uint32_t range_owner(uint64_t key, int code_bits, uint32_t world) {
    if (key >> 63) return world - 1;
    const uint32_t t = code_bits >= 16 ? (uint32_t)(key >> (code_bits - 16)) : (uint32_t)(key << (16 - code_bits));
    return (uint32_t)(((uint64_t)(t > 0xffffu ? 0xffffu : t) * world) >> 16);
}

// off[p] = first index of the sorted table whose owner is >= p (p = 0..world)
__global__ void owner_bounds_kernel(const uint64_t *__restrict__ codes, uint64_t n, int code_bits, uint32_t world,
                                    unsigned long long *__restrict__ off) {
    const uint32_t p = threadIdx.x;
    if (p > world) return;
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (range_owner(codes[mid], code_bits, world) < p) lo = mid + 1; else hi = mid;
    }
    off[p] = p == world ? n : lo;
}

__global__ void head_flags_keys_kernel(const uint64_t *__restrict__ keys, uint64_t n, uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}
__global__ void compact_keys_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ flags,
                                    const uint64_t *__restrict__ pos, uint64_t n, uint64_t *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flags[i]) out[pos[i]] = keys[i];
}

// ---- owner side: union of the W sorted runs it received ---------------------------------------------------
// Every run is a slice of some rank's sorted duplicate-free table, so the owner's set is a MERGE, not a sort: a
// tree of two-way merges (log2 W rounds, 16 bytes of traffic per key and round) replaces the 8-pass LSD radix sort of
// round 1 (128 bytes per key whatever W).  Each CTA produces MG_TILE consecutive outputs of one pair: "merge path" --
// the split of the two inputs at an output diagonal is found by binary search (co-rank), the CTA's two input pieces
// are staged in shared memory, every thread finds its own split there and merges MG_ITEMS keys serially; the result
// goes back through shared memory so that the global stores are coalesced.  Equal keys of different runs stay
// adjacent and are removed by the run-length compaction that follows, as before.
static constexpr int MG_THREADS = 256, MG_ITEMS = 8, MG_TILE = MG_THREADS * MG_ITEMS;

// number of keys taken from a[0, na) among the first `diag` outputs of merge(a, b); ties take a first
template <typename T>
__device__ __forceinline__ uint32_t mg_corank(const T *a, uint32_t na, const T *b, uint32_t nb, uint32_t diag) {
    uint32_t lo = diag > nb ? diag - nb : 0u, hi = diag < na ? diag : na;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a[mid] <= b[diag - 1 - mid]) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ uint64_t mg_corank64(const uint64_t *__restrict__ a, uint64_t na, const uint64_t *__restrict__ b,
                                                uint64_t nb, uint64_t diag) {
    uint64_t lo = diag > nb ? diag - nb : 0ull, hi = diag < na ? diag : na;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (a[mid] <= b[diag - 1 - mid]) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(MG_THREADS) merge_pair_kernel(const uint64_t *__restrict__ a, uint64_t na,
                                                                const uint64_t *__restrict__ b, uint64_t nb,
                                                                uint64_t *__restrict__ out) {
    __shared__ uint64_t sk[MG_TILE];              // the CTA's piece of a, then its piece of b (together the tile's outputs, <= MG_TILE); later the merged tile
    __shared__ uint64_t s_split[2];
    const uint64_t d0 = (uint64_t)blockIdx.x * MG_TILE, total = na + nb;
    const uint64_t d1 = d0 + MG_TILE < total ? d0 + MG_TILE : total;
    if (threadIdx.x < 2) s_split[threadIdx.x] = mg_corank64(a, na, b, nb, threadIdx.x ? d1 : d0);
    __syncthreads();
    const uint64_t a0 = s_split[0], a1 = s_split[1], b0 = d0 - a0, b1 = d1 - a1;
    const uint32_t la = (uint32_t)(a1 - a0), lb = (uint32_t)(b1 - b0), n = la + lb;
    uint64_t *sa = sk, *sb = sk + la;
    for (uint32_t i = threadIdx.x; i < la; i += MG_THREADS) sa[i] = a[a0 + i];
    for (uint32_t i = threadIdx.x; i < lb; i += MG_THREADS) sb[i] = b[b0 + i];
    __syncthreads();
    const uint32_t t0 = min((uint32_t)threadIdx.x * MG_ITEMS, n), t1 = min(t0 + MG_ITEMS, n);
    uint32_t ia = mg_corank(sa, la, sb, lb, t0), ib = t0 - ia;
    uint64_t r[MG_ITEMS];
#pragma unroll
    for (int i = 0; i < MG_ITEMS; ++i) {
        const bool have_a = ia < la, have_b = ib < lb;
        const uint64_t ka = have_a ? sa[ia] : ~0ull, kb = have_b ? sb[ib] : ~0ull;
        const bool take_a = have_a && (!have_b || ka <= kb);
        r[i] = take_a ? ka : kb;
        if (t0 + i < t1) { if (take_a) ++ia; else ++ib; }
    }
    __syncthreads();                               // every thread has read its inputs: the buffer becomes the output tile
#pragma unroll
    for (int i = 0; i < MG_ITEMS; ++i)
        if (t0 + i < t1) sk[t0 + i] = r[i];
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += MG_THREADS) out[d0 + i] = sk[i];
}

// Merges the sorted runs [off[p], off[p + 1]) of `src` (p = 0..W-1) into one sorted array; `tmp` has the same size.
// Returns the buffer that holds the result (duplicates between runs are kept, adjacent).
static uint64_t *merge_sorted_runs(pdg_graph *g, uint64_t *src, uint64_t *tmp, std::vector<uint64_t> off) {
    cudaStream_t s = g->stream;
    uint64_t *cur = src, *oth = tmp;
    while (off.size() > 2) {
        std::vector<uint64_t> next;
        next.push_back(off[0]);
        size_t p = 0;
        for (; p + 2 < off.size(); p += 2) {
            const uint64_t na = off[p + 1] - off[p], nb = off[p + 2] - off[p + 1];
            if (na + nb) {
                merge_pair_kernel<<<(unsigned)((na + nb + MG_TILE - 1) / MG_TILE), MG_THREADS, 0, s>>>(
                    cur + off[p], na, cur + off[p + 1], nb, oth + off[p]);
                PDG_LAUNCHED(g);
            }
            next.push_back(off[p + 2]);
        }
        if (p + 1 < off.size()) {          // an odd run is carried over as it is
            const uint64_t nl = off[p + 1] - off[p];
            if (nl) PDG_CUDA(cudaMemcpyAsync(oth + off[p], cur + off[p], nl * 8, cudaMemcpyDeviceToDevice, s));
            next.push_back(off[p + 1]);
        }
        off.swap(next);
        std::swap(cur, oth);
    }
    return cur;
}

static inline unsigned gridf(uint64_t n) { return (unsigned)((n + 255) / 256); }

// Small k (every rank holds a presence bitmap over all 4^k codes and no explicit nodes): the global
// node set is the OR of the ranks' bitmaps -- one ncclAllGather of 4^k / 8 bytes per rank (512 KiB at
// k = 11), an OR + popcount pass, a prefix sum and an expansion back to sorted codes.
__global__ void or_bitmaps_kernel(const uint32_t *__restrict__ all, int world, uint64_t n_words,
                                  uint32_t *__restrict__ out, uint32_t *__restrict__ cnt) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    uint32_t m = 0;
    for (int p = 0; p < world; ++p) m |= all[(size_t)p * n_words + i];
    out[i] = m;
    cnt[i] = __popc(m);
}
__global__ void expand_bitmap_kernel(const uint32_t *__restrict__ bm, const uint32_t *__restrict__ prefix,
                                     uint64_t n_words, uint64_t *__restrict__ codes) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    uint32_t m = bm[i], at = prefix[i];
    while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        codes[at++] = (i << 5) | (uint64_t)b;
    }
}

static int comm_global_nodes_bitmap(pdg_graph *g, NcclApi *api, uint64_t *n_global) {
    cudaStream_t s = g->stream;
    ncclComm_t comm = reinterpret_cast<ncclComm_t>(g->nccl_comm);
    const int W = g->world;
    const uint64_t nw = g->bitmap_words;
    g->c_all.reserve((size_t)W * nw * 4, s);
    g->c_all2.reserve(nw * 4, s);
    g->c_tmp.reserve(nw * 4, s);
    g->c_own.reserve((nw + 1) * 4, s);
    PDG_NCCL(api, api->AllGather(g->bitmap.p, g->c_all.p, nw, ncclUint32, comm, s));
    or_bitmaps_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, s>>>(g->c_all.as<uint32_t>(), W, nw, g->c_all2.as<uint32_t>(),
                                                                  g->c_tmp.as<uint32_t>());
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u32(g->c_tmp.as<uint32_t>(), g->c_own.as<uint32_t>(), nw, g);
    uint32_t n32 = 0;
    PDG_CUDA(cudaMemcpyAsync(&n32, g->c_own.as<uint32_t>() + nw, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->u_codes_buf.reserve(std::max<size_t>((size_t)n32 * 8, 16), s);
    expand_bitmap_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, s>>>(g->c_all2.as<uint32_t>(), g->c_own.as<uint32_t>(), nw,
                                                                     g->u_codes_buf.as<uint64_t>());
    PDG_LAUNCHED(g);
    *n_global = n32;
    return PDG_OK;
}

int comm_materialize_nodes(pdg_graph *g);

// In: d_local holds this rank's sorted unique codes (n_loc of them; may alias g->u_codes_buf).
// Out: g->u_codes_buf holds the global sorted unique table, *n_global its size.
int comm_global_nodes(pdg_graph *g, const uint64_t *d_local, uint64_t n_loc, int key_bits, uint64_t *n_global) {
    NcclApi *api = nccl_api();
    if (!api) return PDG_E_COMM;
    if (g->world > 64) {
        set_error("world size %d > 64", g->world);
        return PDG_E_INVALID;
    }
    cudaStream_t s = g->stream;
    ncclComm_t comm = reinterpret_cast<ncclComm_t>(g->nccl_comm);
    const int W = g->world, R = g->rank;
    g->u_lazy = false;

    // 0. agree on the path: every rank reports {words of its presence bitmap or 0, key bits it needs}
    {
        g->c_counts.reserve(2 * 64 * 8, s);
        g->c_all_counts.reserve((size_t)64 * 64 * 8, s);
        const bool bm_ok = g->have_bitmap && g->x_n_nodes == 0 && d_local == g->g_codes.as<uint64_t>() &&
                           g->bitmap_words * 4 * (uint64_t)W <= (256ull << 20);
        const uint64_t mine[2] = {bm_ok ? g->bitmap_words : 0, (uint64_t)key_bits};
        PDG_CUDA(cudaMemcpyAsync(g->c_counts.p, mine, 16, cudaMemcpyHostToDevice, s));
        PDG_NCCL(api, api->AllGather(g->c_counts.p, g->c_all_counts.p, 2, ncclUint64, comm, s));
        std::vector<uint64_t> all(2 * W);
        PDG_CUDA(cudaMemcpyAsync(all.data(), g->c_all_counts.p, 2 * W * 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        bool use_bitmap = mine[0] != 0;
        for (int p = 0; p < W; ++p) {
            use_bitmap = use_bitmap && all[2 * p] == mine[0];
            key_bits = std::max<int>(key_bits, (int)all[2 * p + 1]);
        }
        // a rank that cannot take the bitmap path reports 0, so the decision is the same everywhere
        for (int p = 0; p < W; ++p) use_bitmap = use_bitmap && all[2 * p] != 0;
        if (use_bitmap) return comm_global_nodes_bitmap(g, api, n_global);
    }

    // 1. the local table is sorted, so the keys of owner p are the slice [send_off[p], send_off[p+1])
    DevBuf &d_counts = g->c_counts, &d_all_counts = g->c_all_counts, &d_recv = g->c_recv;
    d_counts.reserve((2 * W + 2) * 8, s);
    d_all_counts.reserve((size_t)W * W * 8, s);
    unsigned long long *bounds = d_counts.as<unsigned long long>();      // W + 1 offsets
    unsigned long long *counts = bounds + W + 1;                         // W counts
    const int code_bits = 2 * g->k;
    owner_bounds_kernel<<<1, 128, 0, s>>>(d_local, n_loc, code_bits, W, bounds);
    PDG_LAUNCHED(g);
    std::vector<uint64_t> send_off(W + 1), h_send(W), h_all((size_t)W * W);
    PDG_CUDA(cudaMemcpyAsync(send_off.data(), bounds, (W + 1) * 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    for (int p = 0; p < W; ++p) h_send[p] = send_off[p + 1] - send_off[p];
    PDG_CUDA(cudaMemcpyAsync(counts, h_send.data(), W * 8, cudaMemcpyHostToDevice, s));
    // 2. counts matrix, then ONE grouped all-to-all of the codes, straight out of the local table
    PDG_NCCL(api, api->AllGather(counts, d_all_counts.p, W, ncclUint64, comm, s));
    PDG_CUDA(cudaMemcpyAsync(h_all.data(), d_all_counts.p, (size_t)W * W * 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    std::vector<uint64_t> recv_off(W + 1, 0);
    for (int p = 0; p < W; ++p) recv_off[p + 1] = recv_off[p] + h_all[(size_t)p * W + R];   // p sends me this
    const uint64_t n_recv = recv_off[W];
    d_recv.reserve(std::max<size_t>(n_recv * 8, 16), s);
    PDG_NCCL(api, api->GroupStart());
    for (int p = 0; p < W; ++p) {
        if (h_send[p]) PDG_NCCL(api, api->Send(d_local + send_off[p], h_send[p], ncclUint64, p, comm, s));
        const uint64_t rc = h_all[(size_t)p * W + R];
        if (rc) PDG_NCCL(api, api->Recv(d_recv.as<uint64_t>() + recv_off[p], rc, ncclUint64, p, comm, s));
    }
    PDG_NCCL(api, api->GroupEnd());

    // 3. owner: merge + run-length compaction of what was received (W sorted runs of one key range)
    uint64_t n_own = 0;
    DevBuf &d_own = g->c_own, &d_tmp = g->c_tmp;
    d_own.reserve(std::max<size_t>(n_recv * 8, 16), s);
    if (n_recv) {
        d_tmp.reserve(n_recv * 8, s);
        static const bool lsd = getenv("PDG_EXCHANGE_LSD") != nullptr;     // round 1's owner-side radix sort, for A/B
        uint64_t pr = 0;
        uint64_t *sorted = lsd ? radix_sort_keys(g, d_recv.as<uint64_t>(), d_tmp.as<uint64_t>(), n_recv, 0, key_bits, &pr)
                               : merge_sorted_runs(g, d_recv.as<uint64_t>(), d_tmp.as<uint64_t>(), recv_off);
        g->scratch0.reserve(n_recv * 4, s);
        g->scratch1.reserve((n_recv + 1) * 8, s);
        head_flags_keys_kernel<<<gridf(n_recv), 256, 0, s>>>(sorted, n_recv, g->scratch0.as<uint32_t>());
        PDG_LAUNCHED(g);
        exclusive_scan_u32_u64(g->scratch0.as<uint32_t>(), g->scratch1.as<uint64_t>(), n_recv, g);
        PDG_CUDA(cudaMemcpyAsync(&n_own, g->scratch1.as<uint64_t>() + n_recv, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        compact_keys_kernel<<<gridf(n_recv), 256, 0, s>>>(sorted, g->scratch0.as<uint32_t>(),
                                                         g->scratch1.as<uint64_t>(), n_recv, d_own.as<uint64_t>());
        PDG_LAUNCHED(g);
    }
    // 4. sizes of the owned sets -> global node count.  The table itself (the owned sets in rank order:
    //    ranges ascend with the rank, so their concatenation IS the sorted global table) is gathered
    //    only when something asks for it (comm_materialize_nodes): its size grows with the whole job,
    //    not with this rank's share, and a build does not need it.
    PDG_CUDA(cudaMemcpyAsync(counts, &n_own, 8, cudaMemcpyHostToDevice, s));
    PDG_NCCL(api, api->AllGather(counts, d_all_counts.p, 1, ncclUint64, comm, s));
    g->c_own_counts.assign(W, 0);
    PDG_CUDA(cudaMemcpyAsync(g->c_own_counts.data(), d_all_counts.p, W * 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));   // the sends are done: d_local is no longer read
    uint64_t n_glob = 0;
    for (int p = 0; p < W; ++p) n_glob += g->c_own_counts[p];
    g->u_lazy = true;
    static const bool eager = getenv("PDG_EAGER_NODE_TABLE") != nullptr;
    if (eager) {
        const int rc = comm_materialize_nodes(g);
        if (rc != PDG_OK) return rc;
    }
    *n_global = n_glob;
    return PDG_OK;
}

// Collective: every rank of the job has to call it (pdg_export_nodes, pdg_node_ids and pdg_checksum
// do, on their first call after a multi-rank pdg_finalize).
int comm_materialize_nodes(pdg_graph *g) {
    if (!g->u_lazy) return PDG_OK;
    NcclApi *api = nccl_api();
    if (!api) return PDG_E_COMM;
    cudaStream_t s = g->stream;
    ncclComm_t comm = reinterpret_cast<ncclComm_t>(g->nccl_comm);
    const int W = g->world;
    std::vector<uint64_t> own_off(W + 1, 0);
    for (int p = 0; p < W; ++p) own_off[p + 1] = own_off[p] + g->c_own_counts[p];
    const uint64_t n_glob = own_off[W];
    PDG_CUDA(cudaStreamSynchronize(s));
    g->u_codes_buf.reserve(std::max<size_t>(n_glob * 8, 16), s);
    PDG_NCCL(api, api->GroupStart());
    for (int p = 0; p < W; ++p)
        if (g->c_own_counts[p])
            PDG_NCCL(api, api->Broadcast(g->c_own.p, g->u_codes_buf.as<uint64_t>() + own_off[p], g->c_own_counts[p],
                                         ncclUint64, p, comm, s));
    PDG_NCCL(api, api->GroupEnd());
    PDG_CUDA(cudaStreamSynchronize(s));
    g->u_codes = g->u_codes_buf.as<uint64_t>();
    g->u_lazy = false;
    return PDG_OK;
}

}  // namespace pdg

using namespace pdg;

namespace pdg {
// pdg_destroy: the communicator belongs to the handle
void comm_destroy(pdg_graph *g) {
    if (!g->nccl_comm) return;
    NcclApi *api = nccl_api();
    if (api && api->CommDestroy) api->CommDestroy(static_cast<ncclComm_t>(g->nccl_comm));
    g->nccl_comm = nullptr;
    g->world = 1;
    g->rank = 0;
}

}  // namespace pdg

extern "C" {

int pdg_comm_unique_id(uint8_t *h_id128) {
    NcclApi *api = nccl_api();
    if (!api) return PDG_E_COMM;
    ncclUniqueId id;
    ncclResult_t r = api->GetUniqueId(&id);
    if (r != ncclSuccess) {
        set_error("ncclGetUniqueId failed: %s", api->GetErrorString ? api->GetErrorString(r) : "?");
        return PDG_E_COMM;
    }
    memcpy(h_id128, id.internal, NCCL_UNIQUE_ID_BYTES);
    return PDG_OK;
}

int pdg_comm_init(pdg_graph *g, const uint8_t *h_id128, int rank, int world) {
    if (!g) {
        set_error("null graph handle");
        return PDG_E_INVALID;
    }
    NcclApi *api = nccl_api();
    if (!api) return PDG_E_COMM;
    if (cudaSetDevice(g->device) != cudaSuccess) {
        set_error("cudaSetDevice(%d) failed", g->device);
        return PDG_E_CUDA;
    }
    ncclUniqueId id;
    memcpy(id.internal, h_id128, NCCL_UNIQUE_ID_BYTES);
    ncclComm_t comm;
    ncclResult_t r = api->CommInitRank(&comm, world, id, rank);
    if (r != ncclSuccess) {
        set_error("ncclCommInitRank failed: %s", api->GetErrorString ? api->GetErrorString(r) : "?");
        return PDG_E_COMM;
    }
    if (g->nccl_comm && api->CommDestroy) api->CommDestroy(static_cast<ncclComm_t>(g->nccl_comm));
    g->nccl_comm = comm;
    g->rank = rank;
    g->world = world;
    return PDG_OK;
}

}  // extern "C"
