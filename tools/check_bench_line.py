"""Recompute the derived figures of bench.py JSON lines from their own primary fields:
value = global rows / ms_per_step, roofline.frac = achieved / peak, whole_pass.achieved = algorithmic
bytes per row x rows per GPU / ms_per_step, e2e.frac_of_ceiling = e2e bytes per second / measured ceiling.

    python tools/check_bench_line.py profiles/r2_bench_*.json        # prints one line per file, exit 1 on a mismatch
"""
import json
import sys


def check(line, tol=2e-3):
    """-> list of mismatches (empty = consistent)."""
    bad = []

    def near(name, got, want):
        if want == 0 or abs(got - want) > tol * abs(want):
            bad.append(f"{name}: line says {got:.6g}, recomputed {want:.6g}")

    cfg = line["config"]
    rows = cfg.get("global_rows") or cfg.get("rows_per_step")
    if rows:
        near("value", line["value"], rows / (line["ms_per_step"] * 1e-3))
    r = line.get("roofline")
    if r:
        near("roofline.frac", r["frac"], r["achieved"] / r["peak"])
        if r.get("algorithmic_bytes") and r.get("timing") and "phases_ms" in line:
            near("roofline.achieved", r["achieved"], r["algorithmic_bytes"] / (line["phases_ms"]["transform_alone"] * 1e-3) / 1e9)
        w = r.get("whole_pass")
        if w and cfg.get("rows_per_gpu"):
            near("whole_pass.achieved", w["achieved"], w["algorithmic_bytes_per_row"] * cfg["rows_per_gpu"] / (line["ms_per_step"] * 1e-3) / 1e9)
            near("whole_pass.frac", w["frac"], w["achieved"] / r["peak"])
    e = line.get("e2e")
    if e and e.get("frac_of_ceiling") and e.get("pcie_ceiling"):
        # bytes per step are per GPU; e2e.value / global_rows = steps per second of the whole job
        gbs = max(e["h2d_bytes_per_step"], e["d2h_bytes_per_step"]) * e["value"] / cfg["global_rows"] / 1e9 if cfg.get("global_rows") else None
        if gbs:
            near("e2e.frac_of_ceiling", e["frac_of_ceiling"], gbs / e["pcie_ceiling"]["gbs_per_gpu_each_way"])
    return bad


def main(paths):
    rc = 0
    for p in paths:
        line = json.loads(open(p).read().strip().splitlines()[-1])
        bad = check(line)
        print(p, "ok" if not bad else "; ".join(bad))
        rc |= bool(bad)
    return rc


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
