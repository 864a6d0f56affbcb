"""Attribute executed SASS instructions / stall samples of an ncu capture to CUDA source lines.

    ncu -i prof.ncu-rep --page source --csv --kernel-name regex:<kernel> > sass.csv
    cuobjdump -xelf all build/<file>.o && nvdisasm -g <file>.sm_100a.cubin > dis.txt
    python tools/sass_lines.py sass.csv dis.txt <mangled-kernel-substring> [launch-index] [rows-per-launch]

ncu's CSV has per-instruction counters but no source lines; nvdisasm -g has lines but no counters;
both list the kernel's instructions in the same order, so they are joined by position.
"""
import collections
import csv
import re
import sys


def load_ncu(path, launch):
    rows = list(csv.reader(open(path)))
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = []
            blocks.append(cur)
            continue
        if cur is not None:
            cur.append(r)
    seen, uniq = set(), []
    for b in blocks:  # ncu repeats each launch twice in this page
        key = tuple(x[5] for x in b[1:] if len(x) > 5)
        if key not in seen:
            seen.add(key)
            uniq.append(b)
    blk = uniq[launch]
    hdr, data = blk[0], blk[1:]
    i_ex, i_s, i_src = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
    return [(r[i_src].strip(), int(r[i_ex]), int(r[i_s])) for r in data]


def load_dis(path, kernel):
    lines = open(path).read().split("\n")
    out, cur, active = [], None, False
    for l in lines:
        if l.startswith("\t.section\t.text."):
            active = kernel in l
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
            out.append(cur)
    return out


def main():
    ncu, dis, kernel = sys.argv[1], sys.argv[2], sys.argv[3]
    launch = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    nrows = float(sys.argv[5]) if len(sys.argv) > 5 else 100e6
    a, b = load_ncu(ncu, launch), load_dis(dis, kernel)
    print(f"ncu instructions {len(a)}, nvdisasm instructions {len(b)}")
    n = min(len(a), len(b))
    per = collections.OrderedDict()
    tot_ex = sum(x[1] for x in a)
    tot_s = sum(x[2] for x in a)
    for (src, ex, s), loc in zip(a[:n], b[:n]):
        d = per.setdefault(loc, [0, 0])
        d[0] += ex
        d[1] += s
    print(f"total warp-instr per 32 rows: {tot_ex / (nrows / 32):.1f}")
    for loc, (ex, s) in sorted(per.items(), key=lambda kv: -kv[1][0])[:45]:
        print(f"  {str(loc):34s} instr/32rows {ex / (nrows / 32):7.2f}  ({100 * ex / tot_ex:4.1f}%)   samples {100 * s / max(tot_s, 1):4.1f}%")


if __name__ == "__main__":
    main()
