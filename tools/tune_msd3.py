#!/usr/bin/env python
"""Builds tuning variants of libpdgpu.so (tile shape of the msd3 partition kernels) into
prairiedog_b200/build/variants/ and, with --run, times each on the GPU (tools/profile_step.py)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VARIANTS = {"t256i16b3": (256, 16, 3), "t256i16b4": (256, 16, 4), "t512i8b2": (512, 8, 2), "t512i8b3": (512, 8, 3),
            "t256i8b5": (256, 8, 5), "t512i16b1": (512, 16, 1), "t1024i8b1": (1024, 8, 1), "t128i16b6": (128, 16, 6),
            "cu1": (256, 8, 5, "-DM3_COUNT_UNROLL=1"), "cu2": (256, 8, 5, "-DM3_COUNT_UNROLL=2"), "cu8": (256, 8, 5, "-DM3_COUNT_UNROLL=8"),
            "reread": (256, 8, 5, "-DM3_REREAD_COUNTS")}
OUT = os.path.join(ROOT, "prairiedog_b200", "build", "variants")


def main():
    from prairiedog_b200 import build
    os.makedirs(OUT, exist_ok=True)
    if "--run" in sys.argv:
        n = sys.argv[sys.argv.index("--run") + 1]
        only = sys.argv[sys.argv.index("--only") + 1].split(",") if "--only" in sys.argv else sorted(VARIANTS)
        for name in only:
            lib = os.path.join(OUT, "libpdgpu_%s.so" % name)
            if not os.path.exists(lib):
                continue
            env = dict(os.environ, PDG_LIBPDGPU=lib)
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "profile_step.py"), n], env=env,
                               capture_output=True, text=True)
            out = r.stdout.strip()
            # counts first (every variant must print the same node / edge / k-mer counts), then the timings
            print(name, out[:40], out[out.find("'ms_hist'"):][:260] if r.returncode == 0 else "FAILED " + r.stderr[-300:])
        return
    only = sys.argv[sys.argv.index("--only") + 1].split(",") if "--only" in sys.argv else sorted(VARIANTS)
    for name in only:
        t, i, b = VARIANTS[name][:3]
        build.build_library(force=True, verbose=False, tag="_" + name, out=os.path.join(OUT, "libpdgpu_%s.so" % name),
                            extra_flags=["-DM3_CFG_THREADS=%d" % t, "-DM3_CFG_ITEMS=%d" % i, "-DM3_CFG_MINB=%d" % b] + list(VARIANTS[name][3:]))
        print("built", name)


if __name__ == "__main__":
    main()
