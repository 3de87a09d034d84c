"""Summarises an ncu capture of the conv launches of one step into profiles/<tag>_conv_ncu.json, the file bench.py reads
`roofline.traffic` from.  Two input formats:

  wide  (ncu -i X.ncu-rep --page raw --csv, from a `--set full` capture): one row per launch
  long  (ncu --metrics ... --csv --log-file X.csv, the metrics-only pass of tools/gpu_conv_metrics.sh): one row per
        launch and metric.  This is the pass used for the whole step: `--set full` never returned on the fused-upsample
        CTA-pair kernels (profiles/r02_conv.md), the metrics-only pass does.

  python tools/ncu_conv_summary.py <csv> <tag> [frames]

Per launch: duration, DRAM bytes read / written, tensor-pipe active %, DRAM throughput %; totals, the time-weighted
tensor-pipe figure, and the hash of the conv sources the capture belongs to (bench.py says "STALE" when they differ)."""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (conv_source_hash)

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
        "hz": 1.0, "Khz": 1e3, "Mhz": 1e6, "Ghz": 1e9,
        "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}
RT = "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"   # the figure VERDICT r01 quotes
TP = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"
NAMES = {"us": "gpu__time_duration.sum", "dram_read": "dram__bytes_read.sum", "dram_write": "dram__bytes_write.sum",
         "tensor_pipe_pct": TP, "tensor_pipe_realtime_pct": RT, "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
         "sm_hz": "sm__cycles_elapsed.avg.per_second"}
SCALED = ("us", "dram_read", "dram_write", "sm_hz")


def kname(s):
    return re.sub(r"^void ", "", re.sub(r"\(.*", "", s)).replace("<unnamed>::", "")


def num(v, unit, scaled):
    v = v.replace(",", "")
    if v in ("", "no data", "n/a"):
        return None
    return float(v) * (UNIT.get(unit, 1) if scaled else 1)


def read_wide(rd, head):
    units = next(rd)
    col = {h: i for i, h in enumerate(head)}
    out = []
    for row in rd:
        if len(row) < len(head):
            continue
        r = {"kernel": kname(row[col["Kernel Name"]]), "grid": row[col["Grid Size"]], "block": row[col["Block Size"]]}
        for k, m in NAMES.items():
            r[k] = num(row[col[m]], units[col[m]], k in SCALED) if m in col else None
        out.append(r)
    return out


def read_long(rd, head):
    col = {h: i for i, h in enumerate(head)}
    back = {m: k for k, m in NAMES.items()}
    rows, order = {}, []
    for row in rd:
        if len(row) < len(head):
            continue
        i = row[col["ID"]]
        if i not in rows:
            rows[i] = {"kernel": kname(row[col["Kernel Name"]]), "grid": row[col["Grid Size"]], "block": row[col["Block Size"]],
                       **{k: None for k in NAMES}}
            order.append(i)
        k = back.get(row[col["Metric Name"]])
        if k:
            rows[i][k] = num(row[col["Metric Value"]], row[col["Metric Unit"]], k in SCALED)
    return [rows[i] for i in order]


def main():
    path, tag = sys.argv[1], sys.argv[2]
    frames = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    lines = [l for l in open(path) if not l.startswith("==")]
    rd = csv.reader(lines)
    head = next(rd)
    out = read_long(rd, head) if "Metric Name" in head else read_wide(rd, head)
    out = [r for r in out if r["kernel"].startswith("k_conv")]
    if len(out) > 24:                       # two steps captured (tools/gpu_conv_metrics.sh): keep the second, warm one
        out = out[len(out) // 2:]
    tot_us = sum(r["us"] for r in out)

    def weighted(key):
        rows = [r for r in out if r[key] is not None]
        return sum(r[key] * r["us"] for r in rows) / max(sum(r["us"] for r in rows), 1e-9) if rows else None

    d = {"tag": tag, "frames": frames, "launches": len(out), "source_hash": bench.conv_source_hash(),
         "sum_us": tot_us, "dram_bytes_read": sum(r["dram_read"] or 0 for r in out),
         "dram_bytes_write": sum(r["dram_write"] or 0 for r in out),
         "tensor_pipe_pct_time_weighted": weighted("tensor_pipe_pct"),
         "tensor_pipe_realtime_pct_time_weighted": weighted("tensor_pipe_realtime_pct"),
         "per_launch": out, "from": os.path.basename(path)}
    dst = os.path.join(ROOT, "profiles", f"{tag}_conv_ncu.json")
    json.dump(d, open(dst, "w"), indent=1)
    for r in out:
        rt = f" (realtime {r['tensor_pipe_realtime_pct']:5.1f} %)" if r["tensor_pipe_realtime_pct"] is not None else ""
        print(f"{r['kernel'][:40]:40s} {r['us']:8.1f} us  rd {(r['dram_read'] or 0)/1e9:6.3f} GB  wr {(r['dram_write'] or 0)/1e9:6.3f} GB  "
              f"tensor {r['tensor_pipe_pct'] or 0:5.1f} %{rt}  dram {r['dram_pct'] or 0:5.1f} %  {(r['sm_hz'] or 0)/1e6:6.0f} MHz")
    print(f"{len(out)} launches, total {tot_us:.1f} us, {d['dram_bytes_read']/1e9:.2f} GB read + {d['dram_bytes_write']/1e9:.2f} GB written, "
          f"tensor pipe (time-weighted) {d['tensor_pipe_pct_time_weighted']:.1f} %")
    print("wrote", dst)


if __name__ == "__main__":
    main()
