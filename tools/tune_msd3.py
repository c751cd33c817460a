#!/usr/bin/env python
"""Builds tuning variants of libpdgpu.so (the msd3 partition kernels) into prairiedog_b200/build/variants/ and, with
--run G, times each on the GPU: G genomes, 4 builds, per kernel the FASTEST of the last 3 (CUDA events inside the
library) plus the counts and checksums, which must agree between all variants.

  python tools/tune_msd3.py [--only a,b]            build (no GPU needed)
  python tools/tune_msd3.py --run 100 [--only a,b]  time (GPU box)
A variant is (threads, items, min CTAs per SM, extra -D flags...); "src:REV" builds csrc/msd3.cu + msd3.cuh as of git
revision REV (the rest of the tree as it is) -- e.g. the round-1 core for an A/B."""
import json
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VARIANTS = {
    "mlspos": (256, 8, 5),                               # k > 16 CTA sort (513-4096 records): positions in shared memory
    "mlnospos": (256, 8, 5, "-DML_CFG_SPOS=0"),          # ... read from pos_in a second time
}
OUT = os.path.join(ROOT, "prairiedog_b200", "build", "variants")


def run_variant(lib, n, k=11):
    code = r'''
import json, sys, torch
sys.path.insert(0, %r)
from prairiedog_b200 import B200Graph, synth
g = B200Graph(k=%d)
files = [(name, buf, torch.from_numpy(buf).cuda()) for name, buf, _ in synth.genomes(%d)]
best = {}
for it in range(4):
    g.clear(); g.add_fasta_batch(files); g.save(); torch.cuda.synchronize()
    st = g.engine.stats()
    if it:
        for key, v in (("hist", st["ms_hist"]), ("scans", st["ms_scans"]), ("p1", st["ms_pass"][0]), ("p2", st["ms_pass"][1]),
                       ("p3", st["ms_pass"][2])):
            best[key] = min(best.get(key, 1e9), v)
print(json.dumps({"counts": g.engine.counts(), "sums": g.engine.checksum(), "ms": best}))
''' % (ROOT, k, n)
    env = dict(os.environ, PDG_LIBPDGPU=lib)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    if r.returncode != 0:
        return None, r.stderr[-400:]
    return json.loads(r.stdout.strip().splitlines()[-1]), None


def main():
    from prairiedog_b200 import build
    os.makedirs(OUT, exist_ok=True)
    only = sys.argv[sys.argv.index("--only") + 1].split(",") if "--only" in sys.argv else sorted(VARIANTS)
    if "--run" in sys.argv:
        n = int(sys.argv[sys.argv.index("--run") + 1])
        first = None
        for name in only:
            lib = os.path.join(OUT, "libpdgpu_%s.so" % name)
            if not os.path.exists(lib):
                continue
            res, err = run_variant(lib, n)
            if res is None:
                print(name, "FAILED", err)
                continue
            same = "" if first is None or (res["counts"], res["sums"]) == first else "  ** RESULT DIFFERS **"
            first = first or (res["counts"], res["sums"])
            ms = res["ms"]
            print("%-10s hist %.3f scans %.3f p1 %.3f p2 %.3f p3 %.3f  sum %.3f  %s%s" % (
                name, ms["hist"], ms["scans"], ms["p1"], ms["p2"], ms["p3"], sum(ms.values()), res["counts"], same), flush=True)
        return
    for name in only:
        t, i, b = VARIANTS[name][:3]
        flags, csrc, tmp = [], None, None
        for x in VARIANTS[name][3:]:
            if x.startswith("src:"):
                tmp = tempfile.mkdtemp(prefix="pdg_csrc_")
                csrc = os.path.join(tmp, "csrc")
                shutil.copytree(build.CSRC, csrc)
                for f in ("msd3.cu", "msd3.cuh"):
                    src = subprocess.check_output(["git", "show", "%s:prairiedog_b200/csrc/%s" % (x[4:], f)], cwd=ROOT)
                    open(os.path.join(csrc, f), "wb").write(src)
                # common.cuh includes ../../include/pdgpu.h relative to csrc
                os.makedirs(os.path.join(os.path.dirname(tmp), "include"), exist_ok=True)
                shutil.copy(os.path.join(ROOT, "include", "pdgpu.h"), os.path.join(os.path.dirname(tmp), "include", "pdgpu.h"))
            else:
                flags.append(x)
        build.build_library(force=True, verbose=False, tag="_" + name, out=os.path.join(OUT, "libpdgpu_%s.so" % name),
                            extra_flags=["-DM3_CFG_THREADS=%d" % t, "-DM3_CFG_ITEMS=%d" % i, "-DM3_CFG_MINB=%d" % b] + flags,
                            csrc=csrc)
        if tmp:
            shutil.rmtree(tmp, ignore_errors=True)
        print("built", name, flush=True)


if __name__ == "__main__":
    main()
