"""Summarises an `ncu --set full` capture of the conv launches of one step into profiles/<tag>_conv_ncu.json, the file
bench.py reads `roofline.traffic` from.

  ncu -i gpurun_out/r02a_conv.ncu-rep --page raw --csv > profiles/r02a_conv_ncu_full.csv
  python tools/ncu_conv_summary.py profiles/r02a_conv_ncu_full.csv r02a [frames]

Per launch: duration, DRAM bytes read / written, tensor-pipe active %, DRAM throughput %; totals and the time-weighted
tensor-pipe figure; the hash of the conv sources the capture belongs to (bench.py says "STALE" when they differ)."""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (conv_source_hash)

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
        "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}


RT = "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"   # the figure VERDICT r01 quotes


def main():
    path, tag = sys.argv[1], sys.argv[2]
    frames = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    lines = [l for l in open(path) if not l.startswith("==")]
    rd = csv.reader(lines)
    head, units = next(rd), next(rd)
    col = {h: i for i, h in enumerate(head)}

    def get(row, name, scale=True):
        i = col[name]
        v = row[i].replace(",", "")
        if v in ("", "no data"):
            return 0.0
        return float(v) * (UNIT.get(units[i], 1) if scale else 1)

    out = []
    for row in rd:
        if len(row) < len(head):
            continue
        name = re.sub(r"^void ", "", re.sub(r"\(.*", "", row[col["Kernel Name"]])).replace("<unnamed>::", "")
        out.append({"kernel": name, "grid": row[col["Grid Size"]],
                    "us": get(row, "gpu__time_duration.sum"),
                    "dram_read": get(row, "dram__bytes_read.sum"), "dram_write": get(row, "dram__bytes_write.sum"),
                    "tensor_pipe_pct": get(row, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", False)
                    if "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed" in col else None,
                    "tensor_pipe_realtime_pct": get(row, RT, False) if RT in col else None,
                    "dram_pct": get(row, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", False),
                    "sm_mhz": get(row, "sm__cycles_elapsed.avg.per_second", False)})
    tot_us = sum(r["us"] for r in out)
    tp = [r for r in out if r["tensor_pipe_pct"] is not None]
    d = {"tag": tag, "frames": frames, "launches": len(out), "source_hash": bench.conv_source_hash(),
         "sum_us": tot_us, "dram_bytes_read": sum(r["dram_read"] for r in out),
         "dram_bytes_write": sum(r["dram_write"] for r in out),
         "tensor_pipe_pct_time_weighted": sum(r["tensor_pipe_pct"] * r["us"] for r in tp) / max(sum(r["us"] for r in tp), 1e-9) if tp else None,
         "tensor_pipe_realtime_pct_time_weighted": sum((r["tensor_pipe_realtime_pct"] or 0) * r["us"] for r in out) / max(tot_us, 1e-9),
         "per_launch": out, "from": os.path.basename(path)}
    dst = os.path.join(ROOT, "profiles", f"{tag}_conv_ncu.json")
    json.dump(d, open(dst, "w"), indent=1)
    for r in out:
        print(f"{r['kernel'][:44]:44s} {r['us']:8.1f} us  rd {r['dram_read']/1e9:6.3f} GB  wr {r['dram_write']/1e9:6.3f} GB  "
              f"tensor {r['tensor_pipe_pct'] or 0:5.1f} % (realtime {r['tensor_pipe_realtime_pct'] or 0:5.1f} %)  dram {r['dram_pct']:5.1f} %")
    print(f"total {tot_us:.1f} us, {d['dram_bytes_read']/1e9:.2f} GB read + {d['dram_bytes_write']/1e9:.2f} GB written, "
          f"tensor pipe (time-weighted) {d['tensor_pipe_pct_time_weighted']:.1f} %, realtime counter {d['tensor_pipe_realtime_pct_time_weighted']:.1f} %")
    print("wrote", dst)


if __name__ == "__main__":
    main()
