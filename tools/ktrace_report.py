#!/usr/bin/env python
"""Summarise a kernel timeline written by B200_KTRACE=<file> (common.cuh: ktrace_event).

Events: "<ns> <gridid> <ev> <ctas> <threads>", ev 0 = kernel entry (block 0), 1 = block 0 passed its PDL wait, i.e. the
previous kernel has completed.  In a dependent chain the time a kernel really costs is  pass(k+1) - pass(k); how early it was
resident (and, for the GEMVs, streaming weights) is  pass(k) - entry(k).  Prints the steady-state pattern of one token:
kernels in order of their pass time, averaged over the last tokens of the trace."""
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    per_tok = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    cur, done = {}, []      # grid ids repeat at every replay of a CUDA graph: a new entry event opens a new instance
    for line in open(path):
        t, gid, e, ctas, thr = (int(x) for x in line.split())
        if e == 0:
            if gid in cur:
                done.append(cur.pop(gid))
            cur[gid] = {"entry": t, "pass": None, "ctas": ctas, "thr": thr, "marks": {}}
        elif e == 1:
            if gid in cur and cur[gid]["pass"] is None:
                cur[gid]["pass"] = t
        elif gid in cur and e not in cur[gid]["marks"]:
            cur[gid]["marks"][e] = t
    done += list(cur.values())
    ks = [k for k in done if k["pass"] is not None and k["entry"] is not None]
    ks.sort(key=lambda k: k["pass"])
    print("%d kernels with both events" % len(ks))
    if not per_tok:
        # period detection on the (ctas, thr) signature sequence, from the end
        sig = [(k["ctas"], k["thr"]) for k in ks]
        for per in range(50, 2000):
            if len(sig) > 3 * per and sig[-per:] == sig[-2 * per:-per] == sig[-3 * per:-2 * per]:
                per_tok = per
                break
    if not per_tok:
        print("no period found")
        return
    ntok = min(20, len(ks) // per_tok - 2)
    print("period %d launches/token, averaging the last %d tokens" % (per_tok, ntok))
    base = len(ks) - ntok * per_tok
    tot = (ks[-1]["pass"] - ks[base]["pass"]) / (ntok - 1 + (per_tok - 1) / per_tok) / 1e3
    print("token period %.1f us" % tot)
    rows, marks = [], []
    for i in range(per_tok):
        cost, lead, mk = [], [], defaultdict(list)
        for t in range(ntok):
            j = base + t * per_tok + i
            if j + 1 < len(ks):
                cost.append(ks[j + 1]["pass"] - ks[j]["pass"])
            lead.append(ks[j]["pass"] - ks[j]["entry"])
            for e, tm in ks[j]["marks"].items():
                mk[e].append(tm - ks[j]["pass"])
        rows.append((ks[base + i]["ctas"], ks[base + i]["thr"], sum(cost) / max(1, len(cost)) / 1e3, sum(lead) / len(lead) / 1e3))
        marks.append(" ".join("m%d=+%.2f" % (e, sum(v) / len(v) / 1e3) for e, v in sorted(mk.items())))
    # by signature
    agg = defaultdict(lambda: [0, 0.0, 0.0])
    for c, th, cost, lead in rows:
        a = agg[(c, th)]
        a[0] += 1; a[1] += cost; a[2] += lead
    print("\nby (ctas, threads): count, total us per token, avg cost us, avg lead us")
    for (c, th), a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("  ctas %6d thr %4d  x%4d  total %8.1f  cost %6.2f  lead %6.2f" % (c, th, a[0], a[1], a[1] / a[0], a[2] / a[0]))
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 24
    # find a rotation that starts at a stable place: print launches from the middle of the token
    start = per_tok // 2
    print("\nsequence (launch %d..%d of the token): ctas thr cost_us lead_us" % (start, start + n))
    for (c, th, cost, lead), mk in zip(rows[start:start + n], marks[start:start + n]):
        print("  %6d %4d %7.2f %7.2f   %s" % (c, th, cost, lead, mk))


if __name__ == "__main__":
    main()
