#!/usr/bin/env python
"""bench.py -- genome bp/s into the pangenome graph (k=11) on N B200s (BASELINE.json metric).

A step = one pass of the hot path over one batch: FASTA images of G synthetic E. coli-like genomes
(per rank) -> device parse + 2-bit pack -> k-mer occurrence index (node dedup + labelled edges)
-> node table of the whole job (NCCL exchange when N > 1).  `value` times it with the FASTA images
already resident in HBM; `e2e` times the same call path from pinned HOST buffers, host->device
copies and the device->host read of the result (counts + checksums) inside the timed region.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--genomes G] [--bases B] [--impl reference]

N > 1: launched by torchrun, one rank per GPU (RANK / LOCAL_RANK / WORLD_SIZE from the env).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "genome_bp_per_sec_into_pangenome_graph_k11"      # k filled in from --k (default 11)
UNIT = "bp/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=3, choices=[3, 4],
                    help="BASELINE.json config: 3 = 100 genomes per GPU, k=11, weak scaling (default); 4 = 1000 genomes "
                         "in total, k=31, ceil(1000/N) per rank, strong scaling (N = 2, 4, 8: one GPU's 32-bit position "
                         "column holds 4.29e9 residues, the whole job has 5e9)")
    ap.add_argument("--genomes", type=int, default=None, help="genomes per GPU (config 3: 100) / in total (config 4: 1000)")
    ap.add_argument("--bases", type=int, default=5_000_000)
    ap.add_argument("--contigs", type=int, default=200)
    ap.add_argument("--k", type=int, default=None, help="default: 11 (config 3), 31 (config 4)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample-bases", type=int, default=1_500_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--iid", action="store_true",
                    help="every genome independent i.i.d. random (worst case for deduplication, SURVEY.md 8d) instead of "
                         "substitutions / indels of one base genome")
    ap.add_argument("--no-h2d-ceiling", action="store_true", help="skip the copy-only measurement (e2e.h2d_ceiling_gbs)")
    a = ap.parse_args()
    if a.k is None:
        a.k = 31 if a.config == 4 else 11
    if a.genomes is None:
        a.genomes = 1000 if a.config == 4 else 100
    return a


def rank_block(a, rank, world):
    """(first genome, number of genomes) of a rank: config 3 gives every rank its own `genomes` (weak scaling),
    config 4 cuts `genomes` into world blocks of ceil(genomes / world) (strong scaling)."""
    if a.config == 4:
        per = -(-a.genomes // world)
        first = min(rank * per, a.genomes)
        return first, min(per, a.genomes - first)
    return rank * a.genomes, a.genomes


def workload_config(a, world=1):
    kind = "independent i.i.d. random" if a.iid else "E. coli-like"
    if a.config == 4:
        per = -(-a.genomes // world)
        return {"workload": "config4: %d synthetic %s genomes in total (%.1f Mbp, %d contigs each), k=%d, one pangenome, "
                            "%d genomes per GPU" % (a.genomes, kind, a.bases / 1e6, a.contigs, a.k, per),
                "genomes_total": a.genomes, "genomes_per_gpu": per, "bases_per_genome": a.bases,
                "contigs_per_genome": a.contigs, "k": a.k,
                "l2": "inputs (%.0f MB of FASTA per GPU and step) exceed the 126 MB L2" % (per * a.bases * 1.0127 / 1e6)}
    return {"workload": "config3: %d synthetic %s genomes per GPU (%.1f Mbp, %d contigs each), k=%d, "
                        "one pangenome" % (a.genomes, kind, a.bases / 1e6, a.contigs, a.k),
            "genomes_per_gpu": a.genomes, "bases_per_genome": a.bases, "contigs_per_genome": a.contigs, "k": a.k,
            "l2": "inputs (%.0f MB of FASTA per step) exceed the 126 MB L2" % (a.genomes * a.bases * 1.0127 / 1e6)}


# ------------------------------------------------------------------ CPU reference arm
def cpu_reference(a, sample_bases):
    """Times the reference's own CPU implementation on a bounded sample of the same workload:
    oracle/_ref/lg_driver (the reference's native LemonGraph engine driven like
    subgraph_ref.py:51-105) when it was compiled, else the C port oracle/pdg_oracle.c."""
    from prairiedog_b200 import synth
    n_gen = max(1, -(-sample_bases // a.bases))
    tmp = tempfile.mkdtemp(prefix="pdg_cpu_")
    paths, total = [], 0
    for name, buf, lens in synth.genomes(n_gen, n_bases=a.bases, n_contigs=a.contigs, related=not a.iid):
        p = os.path.join(tmp, name)
        with open(p, "wb") as f:
            f.write(buf.tobytes())
        paths.append(p)
    drv = os.path.join(ROOT, "oracle", "_ref", "lg_driver")
    if os.path.exists(drv):
        db = os.path.join(tmp, "lg.db")
        out = subprocess.run([drv, db, str(a.k), str(sample_bases)] + paths, capture_output=True, text=True,
                             timeout=900)
        rec = json.loads(out.stdout.strip().splitlines()[-1])
        res = {"value": rec["bp_per_s"], "unit": UNIT, "cores": 1, "kind": "reference",
               "sample": "%d bp (%d edges) of the same synthetic genomes through the reference's native "
                         "LemonGraph engine (liblemongraph.so from /root/reference, single writer: 1 process), "
                         "%.2f s" % (rec["bases"], rec["edges"], rec["seconds"])}
    else:
        from oracle.c_oracle import COracle
        t0 = time.time()
        o = COracle(a.k)
        for p in paths:
            o.add_fasta(p)
        o.build()
        dt = time.time() - t0
        bases = o.counts()[2]
        res = {"value": bases / dt, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "%d k-mers through oracle/pdg_oracle.c (sort-based C port), %.2f s" % (bases, dt)}
    for p in paths:
        os.remove(p)
    return res


def cpu_reference_parallel(a, sample_bases, procs):
    """Snakemake-style upper bound: `procs` independent LemonGraph databases filled at the same
    time, one genome sample each.  NOT one pangenome -- LMDB has a single writer, so the reference itself
    cannot build one graph with more than one process; reported beside the single-writer number."""
    from prairiedog_b200 import synth
    drv = os.path.join(ROOT, "oracle", "_ref", "lg_driver")
    if not os.path.exists(drv) or procs < 2:
        return None
    tmp = tempfile.mkdtemp(prefix="pdg_cpu_par_")
    jobs = []
    for i, (name, buf, lens) in enumerate(synth.genomes(procs, n_bases=a.bases, n_contigs=a.contigs, related=not a.iid)):
        p = os.path.join(tmp, name)
        with open(p, "wb") as f:
            f.write(buf.tobytes())
        jobs.append((p, os.path.join(tmp, "lg%d.db" % i)))
    t0 = time.time()
    running = [subprocess.Popen([drv, db, str(a.k), str(sample_bases), p], stdout=subprocess.PIPE, text=True)
               for p, db in jobs]
    outs = [r.communicate(timeout=1800)[0] for r in running]
    wall = time.time() - t0
    bases = sum(json.loads(o.strip().splitlines()[-1])["bases"] for o in outs)
    for p, _ in jobs:
        os.remove(p)
    return {"processes": procs, "value": bases / wall, "unit": UNIT,
            "note": "%d independent LemonGraph databases filled concurrently (%d bp in %.2f s wall); an upper bound "
                    "for a Snakemake run, not one pangenome graph" % (procs, bases, wall)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for _ in range(a.warmup + a.steps):
        t0 = time.time()
        r = cpu_reference(a, a.cpu_sample_bases)
        vals.append((r, time.time() - t0))
    timed = vals[a.warmup:]
    v = sum(r["value"] for r, _ in timed) / len(timed)
    r = timed[-1][0]
    r["value"] = v
    par = cpu_reference_parallel(a, a.cpu_sample_bases, min(os.cpu_count() or 1, 8))
    if par:
        r["parallel_upper_bound"] = par
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * sum(t for _, t in timed) / len(timed),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": dict(workload_config(a, a.gpus), sample_bases=int(a.cpu_sample_bases),
                           note="every step of this arm is a %d bp SAMPLE of the workload, not the whole of it"
                                % a.cpu_sample_bases),
            "cpu_baseline": r,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------ algorithmic CPU baseline (the C port)
def cpu_port(a, genomes_1t=20, genomes_par=6):
    """oracle/pdg_oracle.c (load + sliding window + sort + dedup: the same graph, built the way a CPU would build it
    in bulk) on a bounded sample of the same genomes: one thread, then one process per host core (independent graphs,
    an upper bound for a sharded CPU build).  Genome generation is outside the timed region (oracle/port_bench.py)."""
    def run(first, count):
        out = subprocess.run([sys.executable, "-m", "oracle.port_bench", str(first), str(count), str(a.k), str(a.bases),
                              str(a.contigs), "0" if a.iid else "1"], cwd=ROOT, capture_output=True, text=True, timeout=900)
        return json.loads(out.stdout.strip().splitlines()[-1])
    one = run(0, genomes_1t)
    procs = max(1, min(os.cpu_count() or 1, 32))
    t0 = time.time()
    jobs = [subprocess.Popen([sys.executable, "-m", "oracle.port_bench", str(i * genomes_par), str(genomes_par), str(a.k),
                              str(a.bases), str(a.contigs), "0" if a.iid else "1"], cwd=ROOT, stdout=subprocess.PIPE,
                             text=True) for i in range(procs)]
    outs = [json.loads(j.communicate(timeout=900)[0].strip().splitlines()[-1]) for j in jobs]
    # the processes start together; the rate is all bases over the slowest process's timed region
    par_s = max(o["seconds"] for o in outs)
    return {"kind": "port", "what": "oracle/pdg_oracle.c, the sort-based C restatement of the path",
            "one_thread": {"value": one["bases"] / one["seconds"], "unit": UNIT, "cores": 1,
                           "sample": "%d genomes = %d bp in %.2f s" % (genomes_1t, one["bases"], one["seconds"])},
            "all_cores": {"value": sum(o["bases"] for o in outs) / par_s, "unit": UNIT, "cores": procs,
                          "sample": "%d processes x %d genomes = %d bp, slowest process %.2f s (independent graphs: an "
                                    "upper bound for a sharded CPU build, not one pangenome)"
                                    % (procs, genomes_par, sum(o["bases"] for o in outs), par_s)}}


# ------------------------------------------------------------------ parity gate
GOLDEN = os.path.join(ROOT, "tests", "golden", "bench_checksums.json")


def parity_verdict(a, world, n_nodes, nodes_sum, per_rank):
    """Compares what the timed graph holds with tests/golden/bench_checksums.json (written from oracle/pdg_oracle.c
    by tests/golden/make_bench_checksums.py).  per_rank: [(n_edges, n_kmers, edges_sum)] in rank order.
    Returns {"parity": "ok" | "MISMATCH" | "unchecked", ...}."""
    standard = (a.bases == 5_000_000 and a.contigs == 200 and not a.iid and a.k in (11, 31))
    if not standard or not os.path.exists(GOLDEN):
        return {"parity": "unchecked", "why": "no golden values for this shape" if standard or os.path.exists(GOLDEN)
                else "tests/golden/bench_checksums.json missing"}
    gold = json.load(open(GOLDEN))
    if a.config == 4 and a.genomes == 1000 and a.k == 31:
        bkey, nkey = "config4_n%d_k31" % world, "first_1000_k31"
    elif a.config == 3 and a.genomes == 100:
        bkey, nkey = "weak_k%d" % a.k, "first_%d_k%d" % (100 * world, a.k)
    else:
        return {"parity": "unchecked", "why": "no golden values for this shape"}
    blocks, nodes = gold["blocks"].get(bkey), gold["nodes"].get(nkey)
    if blocks is None or nodes is None or len(blocks) < world:
        return {"parity": "unchecked", "why": "golden file has no %s / %s" % (bkey, nkey)}
    bad = []
    if (int(n_nodes), int(nodes_sum)) != (nodes["n_nodes"], nodes["nodes_sum"]):
        bad.append("nodes: got (%d, %d) want (%d, %d)" % (n_nodes, nodes_sum, nodes["n_nodes"], nodes["nodes_sum"]))
    for r, (ne, nk, es) in enumerate(per_rank):
        b = blocks[r]
        if (int(ne), int(nk), int(es)) != (b["n_edges"], b["n_kmers"], b["edges_sum"]):
            bad.append("rank %d edges: got (%d, %d, %d) want (%d, %d, %d)" % (r, ne, nk, es, b["n_edges"], b["n_kmers"],
                                                                              b["edges_sum"]))
    return {"parity": "MISMATCH" if bad else "ok", "golden": "tests/golden/bench_checksums.json [%s, %s]" % (bkey, nkey),
            "checked": "global node count + node checksum; per rank edge count, k-mer count, edge checksum (%d ranks)" % world,
            **({"mismatches": bad} if bad else {})}


# ------------------------------------------------------------------ clocks
class ClockSampler(object):
    """SM clock and throttle reasons DURING the timed region.  NVML in a thread (a sample costs
    microseconds; the timed region of the default run is ~50 ms, shorter than one `nvidia-smi` start-up),
    `nvidia-smi -lms` only if the NVML binding is missing."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.rows, self.p, self.index = [], None, index
        self.nvml, self.h, self.stop_flag, self.sm, self.reasons, self.mx = None, None, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML counts physical devices; honour CUDA_VISIBLE_DEVICES
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [x.strip() for x in vis.split(",") if x.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def sample(self):
        n = self.nvml
        try:
            self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for name, bit in (("hw_slowdown", n.nvmlClocksThrottleReasonHwSlowdown),
                              ("hw_thermal_slowdown", n.nvmlClocksThrottleReasonHwThermalSlowdown),
                              ("sw_thermal_slowdown", n.nvmlClocksThrottleReasonSwThermalSlowdown),
                              ("sw_power_cap", n.nvmlClocksThrottleReasonSwPowerCap)):
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def _loop(self):
        while not self.stop_flag:
            self.sample()
            time.sleep(0.002)

    def start(self):
        if self.nvml:
            self.stop_flag = False
            self.t = threading.Thread(target=self._loop, daemon=True)
            self.t.start()
            return
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.p = None

    def _read(self):
        for ln in self.p.stdout:
            self.rows.append(ln.strip())

    def stop(self):
        if self.nvml:
            self.sample()            # at least one sample, taken before the last step has drained
            self.stop_flag = True
            self.t.join(timeout=2)
            sm = sorted(self.sm)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                    "samples": len(sm), "source": "nvml"}
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(self.NAMES, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvidia-smi"}


# ------------------------------------------------------------------ B200 arm
def run_b200(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    from prairiedog_b200 import B200Graph, synth, distributed

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    bound = distributed.bind_host_to_device(local)      # NUMA-local pinned buffers: one process per GPU
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    g = B200Graph(k=a.k, device=local)
    distributed.attach(g, dev)          # N > 1: every rank gets the NCCL id, opens the communicator

    # synthetic genomes of this rank: pinned host images and device-resident copies
    first, count = rank_block(a, rank, world)
    if count * a.bases >= (1 << 32):
        raise SystemExit("bench.py: %d genomes x %d bp on one GPU exceed the 32-bit position column of the occurrence "
                         "index (4.29e9 residues per GPU); config 4 runs on 2, 4 or 8 GPUs" % (count, a.bases))
    files, n_bases = [], 0
    for name, buf, lens in synth.genomes(count, n_bases=a.bases, n_contigs=a.contigs, first=first, related=not a.iid):
        pinned = torch.from_numpy(buf).pin_memory()
        files.append((name, pinned.numpy(), pinned.to(dev), pinned))
        n_bases += int(lens.sum())
    h2d_bytes = sum(int(f[1].size) for f in files)
    resident_files = [(f[0], f[1], f[2]) for f in files]
    host_files = [(f[0], f[1], None) for f in files]

    def step(resident):
        """One pass of the hot path over the batch.  resident: the FASTA images already sit in HBM.
        Otherwise the pinned HOST images cross the C ABI (pdg_ingest_fasta_batch with PDG_IN_HOST
        copies them, overlapped with the parse) and the result is read back to the host."""
        g.clear()
        g.add_fasta_batch(resident_files if resident else host_files)
        g.save()
        return g.engine.counts()                            # device -> host read of the result

    def timed(resident, steps, warmup, sampler=None):
        for _ in range(warmup):
            step(resident)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = step(resident)
        e1.record()
        clocks = sampler.stop() if sampler else None        # last sample while the last step is still running
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out, clocks

    ms, out, clocks = timed(True, a.steps, a.warmup, ClockSampler(local))
    st = g.engine.stats()
    ms_e2e, out_e2e, _ = timed(False, a.steps, a.warmup)
    st_e2e = g.engine.stats()
    # ---- parity gate (outside the timed region): checksums of the graph the last e2e step built, every rank's ----
    sums = g.engine.checksum()       # collective when world > 1: gathers the global node table on every rank
    mine = torch.tensor([int(out_e2e[1]), int(out_e2e[2]), int(sums[1]) & 0xffffffff, int(sums[1]) >> 32, bound],
                        device=dev, dtype=torch.int64)
    gathered = [torch.zeros_like(mine) for _ in range(world)]
    if world > 1:
        dist.all_gather(gathered, mine)
    else:
        gathered = [mine]
    per_rank = [(int(t[0]), int(t[1]), int(t[2]) | (int(t[3]) << 32)) for t in gathered]
    bound_all = [int(t[4]) for t in gathered]
    verdict = parity_verdict(a, world, int(out_e2e[0]), int(sums[0]), per_rank)

    # ---- what the node can deliver: every rank copies its pinned images to the device at once, no compute ----
    ceiling = None
    if not a.no_h2d_ceiling:
        for f in files:
            f[2].copy_(f[3], non_blocking=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3
        c0.record()
        for _ in range(reps):
            for f in files:
                f[2].copy_(f[3], non_blocking=True)
        c1.record()
        torch.cuda.synchronize()
        cms = torch.tensor([c0.elapsed_time(c1) / reps], device=dev)
        if world > 1:
            dist.all_reduce(cms, op=dist.ReduceOp.MAX)
        ceiling = float(cms.item())
    total_bases = torch.tensor([float(n_bases), float(h2d_bytes)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(total_bases)
    total_h2d = float(total_bases[1].item())
    total_bases = float(total_bases[0].item())

    if rank == 0:
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            peaks = json.load(open(pk))
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"
        kernels = kernel_table(st, a.k, peak)
        dom = max(kernels, key=lambda r: r["ms_per_launch"]) if kernels else None
        roof = {"bound": "hbm", "achieved": 0.0, "peak": peak, "unit": "GB/s", "frac": 0.0, "traffic": None}
        if dom:
            roof = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved"], "peak": peak, "unit": "GB/s",
                    "frac": dom["frac"], "peak_source": peak_src, "traffic": dom["traffic"],
                    "algorithmic_bytes_per_launch": dom["bytes"], "ms_per_launch": dom["ms_per_launch"],
                    "launches_per_step": 1, "units_per_launch": dom["units"],
                    "how": "CUDA events around each launch on the library's stream inside the last timed step "
                           "(pdg_get_build_stats); traffic = dram bytes/unit of the ncu --set full capture in "
                           "profiles/ x units", "all_kernels": kernels}
        line = {
            "metric": METRIC, "value": total_bases / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong" if a.config == 4 else "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic", "config": workload_config(a, world),
            "e2e": {"value": total_bases / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": 5 * 8 + int(st["n_contigs"]) * 32 + 48 * len(files), "ms_per_step": ms_e2e,
                    "how": "pinned host FASTA images -> pdg_ingest_fasta_batch(PDG_IN_HOST) -> pdg_finalize -> "
                           "pdg_counts + header table read back; copies overlap the parse inside the library",
                    # the same pinned buffers copied by all ranks at once with NO compute: the node's host->device
                    # ceiling for this job (max over ranks), against which e2e scaling has to be read
                    "h2d_ceiling_ms": ceiling,
                    "h2d_ceiling_gbs": (total_h2d / (ceiling * 1e-3) / 1e9) if ceiling else None,
                    "h2d_ceiling_gbs_per_gpu": (total_h2d / world / (ceiling * 1e-3) / 1e9) if ceiling else None},
            "gpu_launches": int(st["kernel_launches"]) * a.steps,
            "clocks": clocks,
            "host_binding": {"cpus_bound_per_rank": bound_all, "how": "NVML CPU affinity of the rank's GPU (0 = not bound)"},
            "roofline": roof,
            "result": dict(verdict, n_nodes=int(out_e2e[0]), nodes_sum=int(sums[0]),
                           per_rank=[{"n_edges": e, "n_kmers": k_, "edges_sum": s_} for e, k_, s_ in per_rank],
                           e2e_counts_equal_resident=list(out_e2e) == list(out)),
            "phases_ms": dict([(k, st[k]) for k in ("ms_upload", "ms_sort", "ms_explicit", "ms_nodes", "ms_total")] +
                              [("ms_ingest_a", st["ms_ingest_a"]), ("ms_ingest_b", st["ms_ingest_b"]),
                               ("ms_ingest_c", st["ms_ingest_c"]),
                               ("e2e_ms_ingest_call", st_e2e["ms_ingest_a"]), ("e2e_ms_h2d_copies", st_e2e["ms_ingest_b"]),
                               ("e2e_ms_finalize", st_e2e["ms_total"]), ("e2e_pipelined", st_e2e["pipelined"]),
                               ("e2e_sort_groups", st_e2e["sort_groups"])]),
        }
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference(a, a.cpu_sample_bases)
            line["cpu_baseline"]["port"] = cpu_port(a)     # the honest algorithmic CPU baseline beside the engine's
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if verdict["parity"] == "MISMATCH" or list(out_e2e) != list(out):
        sys.stderr.write("bench.py: PARITY MISMATCH against tests/golden/bench_checksums.json: %s\n" % json.dumps(verdict))
        raise SystemExit(3)


def kernel_table(st, k, peak):
    """Per-kernel roofline rows from the CUDA-event times of the last step.  Algorithmic bytes per
    unit are the ones DESIGN.md states; `traffic` scales the DRAM bytes/unit of the committed ncu
    capture (profiles/traffic.json) to this launch."""
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("dram_bytes_per_unit", {})
    n = float(st["sort_records"])
    nb = float(st["ingest_bytes"])
    rows = []

    def add(name, ms, units, bpu):
        if ms <= 0 or units <= 0:
            return
        ach = units * bpu / (ms * 1e-3) / 1e9
        t = traffic.get(name.split(" ")[0])
        rows.append({"kernel": name, "ms_per_launch": ms, "units": units, "bytes_per_unit": bpu, "bytes": units * bpu,
                     "achieved": ach, "frac": ach / peak, "traffic": (t * units) if t is not None else None})
    if st["ms_pass"][0] > 0:     # msd3 on the whole code (k <= 11) or on the 11-residue prefix (k > 16)
        add("m3_pass1_kernel (packed genome -> records grouped by digit 3)", st["ms_pass"][0], n, 8.25)
        add("m3_pass2_kernel (records -> grouped by digits 3,2)", st["ms_pass"][1], n, 16.0)
        add("m3_pass3_kernel (records -> row starts + position column)", st["ms_pass"][2], n, 12.0)
        add("m3_hist_kernel (joint digit histograms from the packed genome)", st["ms_hist"], n, 0.25)
    if k > 16 and st["sort_passes"] == 4 and st["ms_sort_kernel_avg"] > 0:
        # (also when the prefix passes ran inside a pipelined ingest and have no event times of their own)
        # per record: reads a position, gathers the two packed words of its code, writes the position; per ROW (distinct
        # code of this rank): writes the code and the record index of its head.  The launches of all bucket-size classes
        # (+ the CTA kernels of the few big buckets) are timed together.
        rows_per_record = float(st["n_nodes"]) / n if st.get("n_nodes") else 0.0
        add("ml_class_sort_kernel (prefix buckets -> position column + the rows' codes and starts; all size classes)",
            st["ms_sort_kernel_avg"], n, 24.0 + 12.0 * rows_per_record)
    elif st["ms_pass"][0] <= 0 and st["ms_sort_kernel_avg"] > 0:
        add("onesweep_kernel (one LSD partition pass)", st["ms_sort_kernel_avg"], n, 16.0 if k <= 16 else 32.0)
    add("fb_summ_kernel (FASTA bytes -> tile summaries)", st["ms_ingest_a"], nb, 1.0)
    add("fb_emit_kernel (FASTA bytes -> 2-bit store + header rows)", st["ms_ingest_c"], nb, 1.25)
    return rows


def main():
    global METRIC
    a = parse_args()
    METRIC = "genome_bp_per_sec_into_pangenome_graph_k%d" % a.k
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
