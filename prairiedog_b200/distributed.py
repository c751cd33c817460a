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
* ``reduce_counts``  whole-job edge / k-mer counts (nodes are already global after ``save()``).

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
    """Opens the NCCL communicator of ``graph`` (a B200Graph) for the current process group.
    No-op for a single rank."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return 0, 1
    rank, world = dist.get_rank(), dist.get_world_size()
    uid = graph.engine.comm_unique_id() if rank == 0 else np.zeros(128, dtype=np.uint8)
    uid = broadcast_bytes(uid, 0, device)
    graph.engine.comm_init(uid, rank, world)
    return rank, world


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
