"""Per-kernel summary of the built library's SASS (cuobjdump): registers, shared memory, spills, and how
many bulk-copy (UBLKCP), mbarrier (SYNCS), tensor-map TMA (UTMA*) and tensor-core (MMA) instructions
each kernel holds -- the evidence that the string-staging and streaming kernels use the Blackwell copy
engine and that nothing on this path is a contraction.

    python tools/sass_summary.py > profiles/r2_sass_summary.txt
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "kubeflow-tfx-workshop_b200", "libtaxi_xf.so")


def demangle(name):
    out = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    return re.sub(r"\(.*", "", out)


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*(REG:\d+.*)", res):
        usage[m.group(1)] = m.group(2).strip()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            counts[cur]["total"] += 1
            for key in ("UBLKCP", "SYNCS", "UTMA", "HMMA", "IMMA", "QGMMA", "UTCMMA", "STL", "LDL", "ATOMS", "ATOMG", "RED"):
                if op.startswith(key):
                    counts[cur][key] += 1
    print(f"{'kernel':46s} {'SASS':>6s} {'UBLKCP':>6s} {'SYNCS':>6s} {'UTMA*':>6s} {'*MMA':>5s} {'STL/LDL':>8s}  resources")
    for fn, c in sorted(counts.items(), key=lambda kv: -kv[1]["total"]):
        mma = c["HMMA"] + c["IMMA"] + c["QGMMA"] + c["UTCMMA"]
        print(f"{demangle(fn)[:46]:46s} {c['total']:6d} {c['UBLKCP']:6d} {c['SYNCS']:6d} {c['UTMA']:6d} {mma:5d} {c['STL'] + c['LDL']:8d}  {usage.get(fn, '')}")
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print(f"\nall kernels: {tot['total']} SASS instructions, {tot['UBLKCP']} UBLKCP (1-D bulk copy), {tot['SYNCS']} SYNCS (mbarrier), "
          f"{tot['UTMA']} tensor-map TMA, {tot['HMMA'] + tot['IMMA'] + tot['QGMMA'] + tot['UTCMMA']} tensor-core instructions")


if __name__ == "__main__":
    main()
