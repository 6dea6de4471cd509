"""SASS opcode histogram of libpaths_b200.so and the ptxas -v summary of its kernels (kept under profiles/).

    python tools/sass_histogram.py > profiles/sass_opcodes_r2.txt
    python tools/sass_histogram.py --ptxas > profiles/ptxas_r2.txt      (reads paths_b200/build/ptxas.log of the last build)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "paths_b200", "libpaths_b200.so")
WANT = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "UTMALDG.2CTA", "LDTM", "UTCBAR", "UTCBAR.2CTA", "UCGABAR_ARV", "UTCATOMSWS", "SYNCS", "ACQBULK",
        "PREEXIT", "FFMA", "MUFU"]


def sass():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kern, per = None, collections.OrderedDict()
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            per[kern] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and kern:
            op = m.group(1)
            base = op.split(".")[0]
            per[kern][base] += 1
            if ".2CTA" in op:
                per[kern][base + ".2CTA"] += 1
            per[kern]["_total"] += 1
    tot = collections.Counter()
    for c in per.values():
        tot.update(c)
    print("SASS opcode histogram of paths_b200/libpaths_b200.so (cuobjdump -sass, sm_100a), round 2 (final).")
    print("UTCHMMA = tcgen05.mma (kind::f16; .2CTA = cta_group::2), UTMALDG = TMA load (cp.async.bulk.tensor), LDTM = tcgen05.ld, UTCBAR = tcgen05.commit\n"
          "(.2CTA: multicast to the pair), UCGABAR_ARV = barrier.cluster, SYNCS = mbarrier ops, ACQBULK = griddepcontrol.wait (programmatic dependent\n"
          "launch: the first instruction of every kernel; PREEXIT = the early trigger, compiled out by default).\n")
    print("whole library: " + ", ".join(f"{k} {tot[k]}" for k in WANT if tot[k]) + f"; total instructions {tot['_total']}\n")
    print("kernels that issue tensor-core / TMEM / TMA instructions:")
    for k, c in per.items():
        if c["UTCHMMA"] or c["UTMALDG"] or c["LDTM"]:
            print("  " + k[:150])
            print("     " + ", ".join(f"{w} {c[w]}" for w in WANT if c[w]))


def ptxas():
    log = open(os.path.join(ROOT, "paths_b200", "build", "ptxas.log")).read()
    print("ptxas -v summary of every kernel of libpaths_b200.so (nvcc 12.9, sm_100a, -O3 -lineinfo), round 2 (final): registers, spills, static shared memory.\n")
    name = None
    for line in log.splitlines():
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
            spill = ""
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m:
            spill = f"spill st/ld {m.group(2)}/{m.group(3)}"
            continue
        m = re.search(r"Used (\d+) registers(.*)", line)
        if m and name:
            print(f"{name[:110]:112s} regs {int(m.group(1)):3d}  {spill}  {m.group(2).strip(', ')}")
            name = None


if __name__ == "__main__":
    ptxas() if "--ptxas" in sys.argv else sass()
