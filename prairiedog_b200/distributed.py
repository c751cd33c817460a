"""Host side of the multi-GPU build: one process per GPU, genomes sharded by sample.

The reference has no distributed path (every Snakemake job of ``Snakefile:46-111`` writes into one
LMDB file behind a single-writer lock), so this module has no counterpart there; it fixes the
conventions the CUDA side (csrc/comm.cu) relies on:

* ``shard``      rank r takes a contiguous block of the genome list (a genome never splits: an edge
                 label belongs to one genome, so edges never move between ranks);
* ``owner``      the rank that deduplicates a node key: the key space is cut into ``world`` equal
                 ranges (csrc/comm.cu:range_owner), so the owned sets concatenate to the sorted table;
* ``bind_host_to_device``  pins the process to the CPUs local to its GPU before pinned buffers are allocated;
* ``attach``     hands every rank the same 128-byte NCCL id through ``torch.distributed`` (NCCL on
                 the GPU box, gloo in the CPU tests) and opens the library's communicator;
* ``reduce_counts``  whole-job edge / k-mer counts (nodes are already global after ``save()``);
* ``type_bases`` / ``exchange_hits`` / ``type_strings``  the path query over an N-rank graph: every rank holds the
                 occurrence index of ITS genomes only, and a hit (edge_type, i, j) lives entirely inside one contig,
                 so the hits of a pair are the concatenation over the ranks of what each rank finds in its shard
                 (prairiedog/lemon_graph.py:275-334 answers over the whole pangenome; SURVEY.md 8(e) "Queries").

torch.distributed is plumbing only; the data path collective is the grouped ncclSend/ncclRecv inside
``pdg_finalize``.
"""
import numpy as np

_M1 = np.uint64(0xff51afd7ed558ccd)
_M2 = np.uint64(0xc4ceb9fe1a85ec53)
_S = np.uint64(33)


def shard(items, rank, world):
    """Contiguous block of ``items`` for ``rank`` (sizes differ by at most one)."""
    n = len(items)
    lo = (n * rank) // world
    hi = (n * (rank + 1)) // world
    return items[lo:hi]


def mix64(codes):
    """murmur3 finaliser on uint64 arrays (csrc/common.cuh:mix64)."""
    x = np.ascontiguousarray(codes, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> _S
        x *= _M1
        x ^= x >> _S
        x *= _M2
        x ^= x >> _S
    return x


def owner(codes, world, k):
    """Owner rank of every node key (csrc/comm.cu:range_owner): equal slices of the 2k-bit code space by
    the top 16 bits; opaque keys (bit 63) belong to the last rank."""
    x = np.ascontiguousarray(codes, dtype=np.uint64)
    bits = 2 * int(k)
    t = (x >> np.uint64(bits - 16)) if bits >= 16 else (x << np.uint64(16 - bits))
    t = np.minimum(t, np.uint64(0xffff))
    o = ((t * np.uint64(world)) >> np.uint64(16)).astype(np.int64)
    o[(x >> np.uint64(63)) != 0] = world - 1
    return o


def broadcast_bytes(buf, src=0, device=None):
    """Broadcasts a uint8 numpy array from ``src`` to every rank through torch.distributed."""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.ascontiguousarray(buf, dtype=np.uint8).copy())
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, src)
    return t.cpu().numpy()


def bind_host_to_device(index):
    """Restricts this process to the CPUs that are local to GPU ``index`` (NVML's affinity mask), so that
    pinned host buffers allocated afterwards land on the GPU's NUMA node (first touch) and host-to-device
    copies do not cross the socket link.  One process per GPU is the deployment model (torchrun), so this
    is the process-level analogue of `numactl --cpunodebind`.  Best effort: returns the number of CPUs
    bound to, or 0 when NVML is unavailable or the mask is empty / outside the allowed set."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        phys = index
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x.strip() for x in vis.split(",") if x.strip()]
            if index < len(ids) and ids[index].isdigit():
                phys = int(ids[index])
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 1) + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return 0
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0


def attach(graph, device=None):
    """Opens the NCCL communicator of ``graph`` (a B200Graph) for the current process group and makes its
    queries collective (B200Graph.path_many / query_many then answer over all ranks' shards).
    No-op for a single rank."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return 0, 1
    rank, world = dist.get_rank(), dist.get_world_size()
    uid = graph.engine.comm_unique_id() if rank == 0 else np.zeros(128, dtype=np.uint8)
    uid = broadcast_bytes(uid, 0, device)
    graph.engine.comm_init(uid, rank, world)
    graph._dist = (rank, world, device)
    return rank, world


# ------------------------------------------------------------------ queries over an N-rank graph
def _world():
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(), dist.get_world_size()


def _all_gather_i64(vec, device=None):
    """every rank's int64 vector (equal lengths) -> array [world, n]"""
    import torch
    import torch.distributed as dist
    rank, world = _world()
    t = torch.from_numpy(np.ascontiguousarray(vec, dtype=np.int64))
    if world == 1:
        return t.numpy()[None, :]
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return np.stack([o.cpu().numpy() for o in out])


def type_bases(n_types, device=None):
    """Global edge-type numbering: rank r's local type id t is the global id bases[r] + t, bases = exclusive prefix of
    every rank's type count (ranks hold contiguous blocks of the genome list, so global ids ascend with the genome
    order like the ids of a single-process graph that loaded the same files in order).  Collective.
    Returns bases[world + 1]."""
    counts = _all_gather_i64([int(n_types)], device)[:, 0]
    return np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


def merge_hits(counts, hits_per_rank):
    """Pure host step of exchange_hits.  counts[r][q]: hits rank r found for pair q; hits_per_rank[r]: its hit
    records, grouped by pair, already carrying GLOBAL type ids.  Returns (offsets[n_q + 1], hits, owner) with the
    hits of a pair in rank order -- the order a single-process graph returns them in (ascending genome)."""
    counts = np.asarray(counts, dtype=np.int64)
    world, n_q = counts.shape
    off = np.zeros(n_q + 1, dtype=np.int64)
    np.cumsum(counts.sum(axis=0), out=off[1:])
    total = int(off[-1])
    dtype = hits_per_rank[0].dtype
    hits = np.zeros(total, dtype=dtype)
    owner = np.zeros(total, dtype=np.int32)    # (returned as uint8: world <= 64)
    before = np.zeros(n_q, dtype=np.int64)          # hits of the same pair on lower ranks
    for r in range(world):
        c = counts[r]
        n_r = int(c.sum())
        if n_r:
            loc_off = np.concatenate([[0], np.cumsum(c)])[:-1]
            qof = np.repeat(np.arange(n_q), c)
            dest = off[qof] + before[qof] + (np.arange(n_r) - loc_off[qof])
            hits[dest] = hits_per_rank[r][:n_r]
            owner[dest] = r
        before += c
    return off.astype(np.uint64), hits, owner.astype(np.uint8)


def merge_hits_torch(counts, recv, totals):
    """merge_hits with torch ops on whatever device the tensors live on (the GPU under NCCL): counts [world, n_q]
    int64, recv[r] [cap, 4] int32 hit records of rank r, totals[r] how many of them are real.  Returns tensors
    (offsets [n_q + 1] int64, hits [total, 4] int32, owner [total] int32)."""
    import torch
    world, n_q = counts.shape
    dev = counts.device
    off = torch.zeros(n_q + 1, dtype=torch.int64, device=dev)
    torch.cumsum(counts.sum(dim=0), 0, out=off[1:])
    total = int(off[-1].item())
    hits = torch.zeros((total, 4), dtype=torch.int32, device=dev)
    owner = torch.zeros(total, dtype=torch.uint8, device=dev)         # world <= 64 (csrc/comm.cu)
    before = torch.cumsum(counts, 0) - counts             # hits of the same pair on lower ranks
    pairs = torch.arange(n_q, device=dev)
    for r in range(world):
        n_r = int(totals[r])
        if n_r == 0:
            continue
        c = counts[r]
        loc_off = torch.cumsum(c, 0) - c
        qof = torch.repeat_interleave(pairs, c, output_size=n_r)
        dest = off[qof] + before[r][qof] + (torch.arange(n_r, device=dev) - loc_off[qof])
        hits[dest] = recv[r][:n_r]
        owner[dest] = r
    return off, hits, owner


def exchange_hits(off, hits, type_base, device=None):
    """Collective.  Every rank ran the SAME pair batch against its own shard: off[n_q + 1] / hits are its local
    result (local type ids).  Returns (offsets, hits, owner) of the whole graph on every rank: global type ids
    (+ type_base), `aux` stays local to the owner rank (owner[h] says which).  Two all-gathers -- the per-pair counts,
    then the hit records padded to the largest rank -- and a merge into pair order on ``device`` (merge_hits_torch;
    the numpy merge_hits states the same step and is what the tests compare it with)."""
    import torch
    import torch.distributed as dist
    rank, world = _world()
    off = np.asarray(off).astype(np.int64)
    mine = hits.copy()
    mine["type_id"] += np.uint32(type_base)
    if world == 1:
        return merge_hits(np.diff(off)[None, :], [mine])
    c = torch.from_numpy(np.diff(off))
    if device is not None:
        c = c.to(device)
    cs = [torch.empty_like(c) for _ in range(world)]
    dist.all_gather(cs, c)
    counts = torch.stack(cs)
    totals = counts.sum(dim=1).cpu().tolist()
    cap = max(int(max(totals)), 1)
    send = torch.zeros((cap, 4), dtype=torch.int32, device=c.device)
    if mine.size:
        send[:mine.size] = torch.from_numpy(mine.view(np.int32).reshape(-1, 4)).to(c.device)
    recv = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(recv, send)
    goff, ghits, owner = merge_hits_torch(counts, recv, totals)
    return (goff.cpu().numpy().astype(np.uint64), ghits.cpu().numpy().reshape(-1).view(mine.dtype),
            owner.cpu().numpy())


def exchange_hits_device(off_t, hits_t, type_base):
    """exchange_hits for results that are already torch tensors on the rank's device (Engine.query_paths_device):
    nothing crosses the host before the merged answer is copied out once, through a pinned buffer.  Collective.
    off_t int64[n_q + 1], hits_t int32[n, 4] with LOCAL type ids in column 0."""
    import torch
    import torch.distributed as dist
    from .engine import HIT_DTYPE
    rank, world = _world()
    dev = off_t.device
    c = off_t[1:] - off_t[:-1]
    mine = hits_t.clone()
    if mine.shape[0]:
        mine[:, 0] += int(type_base)          # type ids stay below 2^31: int32 arithmetic is exact
    if world == 1:
        counts, recv, totals = c[None, :], [mine], [int(mine.shape[0])]
    else:
        cs = [torch.empty_like(c) for _ in range(world)]
        dist.all_gather(cs, c)
        counts = torch.stack(cs)
        totals = counts.sum(dim=1).cpu().tolist()
        cap = max(int(max(totals)), 1)
        send = torch.zeros((cap, 4), dtype=torch.int32, device=dev)
        send[:mine.shape[0]] = mine
        recv = [torch.empty_like(send) for _ in range(world)]
        dist.all_gather(recv, send)
    goff, ghits, owner = merge_hits_torch(counts, recv, totals)
    n = int(ghits.shape[0])
    if dev.type != "cuda":        # CPU tensors (the gloo tests): nothing to copy out
        return (goff.numpy().astype(np.uint64), ghits.contiguous().numpy().reshape(-1).view(HIT_DTYPE).copy(),
                owner.numpy())
    # one copy out, into pinned memory kept between calls.  The returned arrays are VIEWS of it, valid until the next
    # call (a private copy of 14 M hits = 227 MB cost more than the whole exchange).
    global _PINNED_HITS, _PINNED_OWNER
    if _PINNED_HITS is None or _PINNED_HITS.shape[0] < n:
        _PINNED_HITS = torch.empty((max(int(n * 1.25), 1 << 16), 4), dtype=torch.int32, pin_memory=True)
        _PINNED_OWNER = torch.empty(_PINNED_HITS.shape[0], dtype=torch.uint8, pin_memory=True)
    _PINNED_HITS[:n].copy_(ghits, non_blocking=True)
    _PINNED_OWNER[:n].copy_(owner, non_blocking=True)
    goff_h = goff.cpu()
    torch.cuda.synchronize(dev)
    return (goff_h.numpy().view(np.uint64), _PINNED_HITS[:n].numpy().reshape(-1).view(HIT_DTYPE), _PINNED_OWNER[:n].numpy())


_PINNED_HITS = None
_PINNED_OWNER = None


def gather_objects(obj):
    """all_gather_object (strings of the hits a rank owns, type labels): list indexed by rank"""
    import torch.distributed as dist
    rank, world = _world()
    if world == 1:
        return [obj]
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out


def reduce_counts(values, device=None, op="sum"):
    """All-reduce of a short list of integers (edge / k-mer counts, timings)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(v) for v in values], dtype=torch.float64)
    if device is not None:
        t = t.to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
    return [float(x) for x in t.cpu().tolist()]
