"""Turn one `ncu --set full` capture of tools/prof_passes.py into the per-kernel DRAM-traffic summary that
bench.py reports as `roofline.traffic`.

  # on the GPU box (after the same command has exited 0 without ncu):
  PROF_NCU=1 ncu --set full --clock-control none --import-source on --profile-from-start off \\
      -k regex:nb2_kernel -o gpurun_out/prof_c4_f32 python tools/prof_passes.py f32 c4 0 2> gpurun_out/prof_c4_f32.cases
  # here:
  ncu -i gpurun_out/prof_c4_f32.ncu-rep --page raw --csv > profiles/r2_ncu_full_c4_f32.csv
  python tools/ncu_dram_summary.py c4/f32 profiles/r2_ncu_full_c4_f32.csv gpurun_out/prof_c4_f32.cases

Every case of prof_passes.py launches ONE pass / vector kernel (plus, for cases with a reduction, the 1-CTA
`FinalReduce`, which is skipped here), in the order of the "case: <name>" lines, so the k-th remaining
profiled launch belongs to the k-th case.  Writes / updates profiles/dram_bytes.json:
  {"c4/f32": {"source": "<csv>", "cases": {"<case>": {"kernel": ..., "dram_bytes": read+write per launch,
                                                       "dram_read": ..., "dram_write": ..., "time_ms": ...}}}}
"""
import csv
import json
import os
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
        "ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3, "nsecond": 1e-6}


def main():
    key, csv_path, cases_path = sys.argv[1:4]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(cases_path) as f:
        cases = [ln.split(":", 1)[1].strip() for ln in f if ln.startswith("case:")]
    with open(csv_path, newline="") as f:
        rows = [r for r in csv.reader(f) if r]
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    head, units = rows[start], rows[start + 1]
    col = {n: i for i, n in enumerate(head)}

    def val(r, name):
        return float(r[col[name]].replace(",", "")) * UNIT[units[col[name]]]

    launches = [r for r in rows[start + 2:] if len(r) == len(head) and "FinalReduce" not in r[col["Kernel Name"]]]
    if len(launches) != len(cases):
        raise SystemExit(f"{len(launches)} profiled launches but {len(cases)} cases: {cases}")
    out = {}
    for name, r in zip(cases, launches):
        rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
        out[name] = {"kernel": r[col["Kernel Name"]][:120], "dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr,
                     "time_ms": val(r, "gpu__time_duration.sum")}
    path = os.path.join(root, "profiles", "dram_bytes.json")
    db = {}
    if os.path.exists(path):
        with open(path) as f:
            db = json.load(f)
    db[key] = {"source": os.path.relpath(os.path.abspath(csv_path), root), "cases": out}
    with open(path, "w") as f:
        json.dump(db, f, indent=1, sort_keys=True)
    for k, v in out.items():
        print(f"{k:28s} {v['dram_bytes'] / 1e9:8.3f} GB  {v['time_ms']:7.3f} ms   {v['kernel'][:70]}")


if __name__ == "__main__":
    main()
