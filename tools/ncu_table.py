#!/usr/bin/env python
"""Pivot an `ncu --metrics ... --csv` log (one row per launch and metric) into one row per launch and a per-kernel summary.
usage: ncu_table.py <ncu.csv> <out_launches.csv> <out_summary.md>"""
import csv, sys, collections, re
src, out_csv, out_md = sys.argv[1:4]
rows = [r for r in csv.reader(open(src)) if len(r) >= 15 and r[0].isdigit()]
launch = collections.OrderedDict()
for r in rows:
    d = launch.setdefault(int(r[0]), {"kernel": r[4], "block": r[7], "grid": r[8]})
    try:
        d[r[12]] = float(r[14].replace(",", ""))
    except ValueError:
        d[r[12]] = r[14]
    d["unit:" + r[12]] = r[13]
def short(k):
    k = re.sub(r"\(mtsb::[A-Za-z]+(, *mtsb::[A-Za-z]+|, *[a-z0-9 ]+|, *T\d)*\)|\(.*\)$", "", k.replace("void ", "").replace("mtsb::", ""))
    return k.replace("(bool)", "").replace("(int)", "")
metrics = [m for m in next(iter(launch.values())) if not m.startswith("unit:") and m not in ("kernel", "block", "grid")]
with open(out_csv, "w", newline="") as f:
    w = csv.writer(f); w.writerow(["id", "kernel", "grid", "block"] + metrics)
    for i, d in launch.items():
        w.writerow([i, short(d["kernel"]), d["grid"], d["block"]] + [d.get(m, "") for m in metrics])
def ms(d):
    v = d.get("gpu__time_duration.sum", 0.0); u = d.get("unit:gpu__time_duration.sum", "ns")
    return v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "msecond": 1.0, "ms": 1.0, "nsecond": 1e-6, "second": 1e3}.get(u, 1e-6)
def byts(d, m):
    v = d.get(m, 0.0); u = d.get("unit:" + m, "byte")
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
agg = collections.OrderedDict()
for d in launch.values():
    a = agg.setdefault(short(d["kernel"]), [])
    a.append(d)
tot = sum(ms(d) for d in launch.values())
with open(out_md, "w") as f:
    f.write("| kernel | launches | total ms | share | DRAM rd+wr MB | DRAM GB/s | L2 MB | L2 hit % | L1 hit % | warps active % | issue % | threads/inst | branch uniform % | regs |\n|---|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
    for k, ds in sorted(agg.items(), key=lambda kv: -sum(ms(d) for d in kv[1])):
        t = sum(ms(d) for d in ds)
        dr = sum(byts(d, "dram__bytes_read.sum") + byts(d, "dram__bytes_write.sum") for d in ds)
        l2 = sum(byts(d, "lts__t_bytes.sum") for d in ds)
        wavg = lambda m: sum(d.get(m, 0.0) * ms(d) for d in ds if isinstance(d.get(m, 0.0), float)) / max(t, 1e-12)
        f.write("| `%s` | %d | %.3f | %.1f %% | %.0f | %.0f | %.0f | %.0f | %.0f | %.0f | %.0f | %.1f | %.0f | %d |\n" % (
            k, len(ds), t, 100 * t / tot, dr / 1e6, dr / 1e9 / max(t * 1e-3, 1e-12), l2 / 1e6, wavg("lts__t_sector_hit_rate.pct"), wavg("l1tex__t_sector_hit_rate.pct"),
            wavg("sm__warps_active.avg.pct_of_peak_sustained_active"), wavg("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            wavg("smsp__thread_inst_executed_per_inst_executed.ratio"), wavg("smsp__sass_average_branch_targets_threads_uniform.pct"), int(ds[0].get("launch__registers_per_thread", 0))))
    f.write("\ntotal %.3f ms over %d launches\n" % (tot, len(launch)))
print(open(out_md).read())
