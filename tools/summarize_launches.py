"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum ... --csv --log-file X.csv`) of a
bench.py run: kernels of the LAST step (from the launch after the previous transform_kernel to the
last transform_kernel), time and share each.

    python tools/summarize_launches.py profiles/r1_final_launches_100M.csv
"""
import collections
import csv
import sys


def main():
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
    h = rows[0]
    ik, im, iv, iid = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
    launches = collections.OrderedDict()
    for r in rows[1:]:
        if r[im] == "gpu__time_duration.sum":
            launches[int(r[iid])] = (r[ik].split("(")[0].replace("void ", "").strip(), float(r[iv].replace(",", "")) / 1e3)
    names = list(launches.values())
    tidx = [i for i, (k, _) in enumerate(names) if k.startswith("transform_")]
    print("launches", len(names))
    print("transform idx", tidx)
    lo = tidx[-2] + 1 if len(tidx) > 1 else 0
    step = names[lo:tidx[-1] + 1]
    agg = collections.OrderedDict()
    for k, us in step:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{us:10.1f} us  x {n:2d}  {100 * us / total:5.1f}%  {k}")
    print(f"step total us {total:.1f} launches/step {len(step)}")


if __name__ == "__main__":
    main()
