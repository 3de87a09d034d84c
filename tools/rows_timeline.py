"""Event timeline of a few groups of block 0 from a conv_rows pipeline trace (see tools/rows_trace.py).
Usage: python tools/rows_timeline.py gpurun_out/rows_trace_06.bin [first_group] [n_groups]"""
import sys
import numpy as np
TR_N = 1024
raw = np.fromfile(sys.argv[1], dtype=np.int64)
cout, kc, ncls, kbn, mt, stages, items, grid = [int(v) for v in raw[:8]]
t = raw[8:].reshape(8, TR_N, 8)
g0 = int(sys.argv[2]) if len(sys.argv) > 2 else 20
ng = int(sys.argv[3]) if len(sys.argv) > 3 else 2
nbuf = 2 if cout == 64 else 4
ev = []
t0 = None
def add(time, what):
    if time > 0: ev.append((int(time), what))
nup = kbn - 1 if kbn > 1 else 0   # source-1 stages per group (decoder conv1: kb0 = 1 .. ) -- approximate for printing
for g in range(g0, g0 + ng):
    for m in range(mt):
        it = g * mt + m
        r = t[1, it]
        add(r[0], f"g{g} MMA tile{m}: before tempty wait"); add(r[1], f"g{g} MMA tile{m}: tempty ok"); add(r[2], f"g{g} MMA tile{m}: last commit issued")
        if m < 2: add(r[4], f"g{g} MMA issuer{m}: before first full wait"); add(r[3], f"g{g} MMA issuer{m}: first stage landed")
        wg = it % nbuf; use = it // nbuf
        e = t[2 + (wg if cout != 64 else wg * 2), use]
        add(e[0], f"g{g} EPI tile{m}: before tfull wait"); add(e[1], f"g{g} EPI tile{m}: tfull seen"); add(e[2], f"g{g} EPI tile{m}: accumulator released")
        add(e[3], f"g{g} EPI tile{m}: done")
    # producer skip boxes: index counts skip stages only
    kb0 = kbn - (2 if cout == 32 else 4) if kbn > 1 else kbn
    for k in range(kb0):
        p = t[0, g * kb0 + k]
        add(p[0], f"g{g} PROD skip{k}: before empty wait"); add(p[1], f"g{g} PROD skip{k}: slot free"); add(p[2], f"g{g} PROD skip{k}: issued")
    nu = kbn - kb0
    for k in range(nu):
        n = g * nu + k
        lo = t[7, n]
        add(lo[1], f"g{g} PROD low{k}: issued")
        if n % 2 == 0:
            u = t[6, n // 2]
            add(u[0], f"g{g} INTERP team0 up{k}: before low wait"); add(u[1], f"g{g} INTERP team0 up{k}: low landed"); add(u[2], f"g{g} INTERP team0 up{k}: slot free")
            add(u[3], f"g{g} INTERP team0 up{k}: stage written")
ev.sort()
base = ev[0][0]
print(f"COUT={cout} KC={kc} kbn={kbn} mt={mt} stages={stages}")
for tm, w in ev: print(f"{tm - base:8d}  {w}")
