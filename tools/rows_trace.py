"""Reads the conv_rows pipeline traces written by a -DUB_ROWS_TRACE build (tools/build_trace_lib.sh) and prints, per
launch, where block 0's roles spent their time.  Usage: python tools/rows_trace.py gpurun_out/rows_trace_*.bin"""
import sys

import numpy as np

TR_N = 1024


def med(x):
    return float(np.median(x)) if len(x) else float("nan")


def report(path):
    raw = np.fromfile(path, dtype=np.int64)
    cout, kc, ncls, kbn, mt, stages, items, grid = raw[:8]
    t = raw[8:].reshape(8, TR_N, 8)
    groups = (items + grid - 1) // grid
    tiles = min(int(groups * mt), TR_N)
    boxes = min(int(groups * kbn), TR_N)
    print(f"== {path}: COUT={cout} KC={kc} NCLS={ncls} kbn={kbn} mt={mt} stages={stages} groups/CTA={groups}")
    p, m = t[0, :boxes], t[1, :tiles]
    t0 = p[0, 0]
    span = max(m[:, 2].max(), t[2:6, :, 3].max()) - t0
    print(f"  block 0 span {span} clk = {span / tiles:.0f} clk per M-tile")
    # producer
    print(f"  producer: wait-empty {med(p[2:, 1] - p[2:, 0]):.0f}  issue {med(p[2:, 2] - p[2:, 1]):.0f}  "
          f"box period {med(np.diff(p[2:, 2])):.0f}   (median clk per box)")
    # MMA
    w_empty = m[4:, 1] - m[4:, 0]
    issue = m[4:, 2] - m[4:, 1]
    first = m[::mt][1:]                                 # issuer 0 stamps the group's first tile, issuer 1 the second
    prev_last = m[mt - 2::mt] if mt >= 2 else m[mt - 1::mt]   # issuer 0's last tile of the previous group
    w_full = first[:, 3] - prev_last[: len(first), 2]
    print(f"  MMA: wait-tempty {med(w_empty):.0f} (mean {w_empty.mean():.0f})  issue+commit {med(issue):.0f}  "
          f"tile period {med(np.diff(m[4:, 2])):.0f}; per group: full-wait after previous group's last tile {med(w_full):.0f} "
          f"(mean {w_full.mean():.0f})")
    fw = m[::mt][1:]
    print(f"      issuer 0: previous group's last issue -> before full wait {med(fw[:, 4] - prev_last[: len(fw), 2]):.0f}  "
          f"waiting for the box {med(fw[:, 3] - fw[:, 4]):.0f}")
    # epilogue warpgroups
    nbuf = 2 if cout == 64 else 4
    for g in range(4):
        e = t[2 + g]
        n = int((e[:, 3] > 0).sum())
        if n < 4:
            continue
        e = e[2:n]
        print(f"  epilogue wg{g}: wait-tfull {med(e[:, 1] - e[:, 0]):.0f} (mean {(e[:, 1] - e[:, 0]).mean():.0f})  "
              f"tmem-read {med(e[:, 2] - e[:, 1]):.0f}  math+store {med(e[:, 3] - e[:, 2]):.0f}  "
              f"period {med(np.diff(e[:, 3])):.0f}")
        if mt == nbuf:
            print(f"      tile end -> next group top {med(e[1:, 4] - e[:-1, 3]):.0f}  group top -> before wait {med(e[:, 0] - e[:, 4]):.0f}  "
                  f"tfull seen -> first TMEM data {med(e[:, 5] - e[:, 1]):.0f}  first-half math + second load {med(e[:, 2] - e[:, 5]):.0f}")
    # interpolation warp 0 (fused upsample)
    u = t[6]
    nu = int((u[:, 3] > 0).sum())
    if nu > 4:
        u = u[2:nu]
        print(f"  interp warp 0: wait-low {med(u[:, 1] - u[:, 0]):.0f} (mean {(u[:, 1] - u[:, 0]).mean():.0f})  wait-empty {med(u[:, 2] - u[:, 1]):.0f} "
              f"(mean {(u[:, 2] - u[:, 1]).mean():.0f})  interpolate+store {med(u[:, 3] - u[:, 2]):.0f} (mean {(u[:, 3] - u[:, 2]).mean():.0f})  "
              f"stage period {med(np.diff(u[:, 3])):.0f} (mean {np.diff(u[:, 3]).mean():.0f})")
    lo = t[7]
    nl = int((lo[:, 1] > 0).sum())
    if nl > 8 and nu > 4:
        # interpolation warp 0 belongs to team 0 = even low boxes
        k = min(nl // 2, int((t[6][:, 3] > 0).sum())) - 2
        iss = lo[0:2 * k:2, 1][2:]; land = t[6][:k, 1][2:]; start = t[6][:k, 0][2:]
        print(f"  low box (team 0): producer wait-lempty {med(lo[2:nl, 1] - lo[2:nl, 0]):.0f}  issue -> landed-seen {med(land - iss):.0f} (mean {(land - iss).mean():.0f})  "
              f"issue relative to the team reaching its wait {med(iss - start):.0f}")
    # latency from MMA commit to epilogue start, from epilogue arrive to MMA resume
    if cout != 64:
        for g in range(1):
            e = t[2 + g]
            n = int((e[:, 3] > 0).sum())
            idx = np.arange(2, min(n, tiles // nbuf))
            commit = m[idx * nbuf + g, 2]
            print(f"  wg{g}: MMA issue-done -> tfull seen {med(e[idx, 1] - commit):.0f} clk "
                  f"(MMA execution + commit latency)")


if __name__ == "__main__":
    for f in sys.argv[1:]:
        report(f)
