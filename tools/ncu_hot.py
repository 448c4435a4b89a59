"""Top stall sites of one kernel from `ncu -i X.ncu-rep --page source --csv --kernel-name K --print-source sass`.
usage: python tools/ncu_hot.py sass.csv [N]   -> N hottest SASS instructions (by stall samples) + stall-reason totals"""
import csv, sys
def main(path, n=25):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]; col = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[hi + 1:]:
        if r and r[0] == "Kernel Name": break          # first profiled instance only
        if len(r) == len(hdr) and r[0].startswith("0x"): data.append(r)
    f = lambda r, c: float(r[col[c]] or 0)
    tot = sum(f(r, "# Samples") for r in data) or 1
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = sorted(((sum(f(r, s) for r in data), s) for s in stalls), reverse=True)
    print("samples=%d  instr_exec=%d  " % (tot, sum(f(r, "Instructions Executed") for r in data)) + " ".join("%s=%.1f%%" % (s[6:], v / tot * 100) for v, s in agg[:8]))
    for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:n]:
        top = sorted(((f(r, s), s[6:]) for s in stalls), reverse=True)[:2]
        print("%5.1f%% %8d  %-70s %s" % (f(r, "# Samples") / tot * 100, f(r, "Instructions Executed"), r[col["Source"]][:70], " ".join("%s=%d" % (s, v) for v, s in top if v)))
if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
