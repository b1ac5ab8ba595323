"""Join an `ncu --page source --csv` export (SASS rows with 'Instructions Executed') with `nvdisasm -g -c` line info of the
same cubin, and print executed warp-instructions per source region.  Development aid.

    ncu -i X.ncu-rep --page source --csv > src.csv
    cuobjdump -xelf all coral_b200/libcoral_b200.so; nvdisasm -g -c scan_records.sm_100a.cubin > scan.dis
    python scripts/ncu_by_line.py src.csv scan.dis k_scan_records [--sass lo hi]
"""
import csv
import re
import sys


def load_dis(path, kernel):
    rows, cur, on = [], None, False
    for ln in open(path):
        if ln.startswith("//---") and ".text." in ln:
            on = kernel in ln
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            rows.append((int(m.group(1), 16), cur, m.group(2).strip()))
    return rows


def main():
    src, dis, kernel = sys.argv[1:4]
    d = load_dis(dis, kernel)
    rd = list(csv.reader(open(src)))
    hdr = next(i for i, r in enumerate(rd) if r and r[0] == "Address")
    col = {n: i for i, n in enumerate(rd[hdr])}
    rows = rd[hdr + 1:]
    base = int(rows[0][0], 16)
    ex = {int(r[0], 16) - base: (int(r[col["Instructions Executed"]]), int(r[col["Thread Instructions Executed"]]),
                                 int(r[col["# Samples"]] or 0)) for r in rows if r and r[0].startswith("0x")}
    total = sum(v[0] for v in ex.values())
    print("total warp instructions", total, "sass rows", len(ex), "dis rows", len(d))
    by = {}
    for off, loc, txt in d:
        e = ex.get(off, (0, 0, 0))
        by.setdefault(loc, [0, 0, 0])
        for k in range(3):
            by[loc][k] += e[k]
    if "--sass" in sys.argv:
        i = sys.argv.index("--sass")
        lo, hi = int(sys.argv[i + 1], 16), int(sys.argv[i + 2], 16)
        for off, loc, txt in d:
            if lo <= off <= hi:
                e = ex.get(off, (0, 0, 0))
                print("%05x %-22s %12d %5.1f  %s" % (off, "%s:%d" % loc if loc else "-", e[0], e[1] / max(e[0], 1), txt[:90]))
        return
    tot_samples = sum(v[2] for v in by.values())
    for loc, v in sorted(by.items(), key=lambda kv: (kv[0][0], kv[0][1]) if kv[0] else ("", 0)):
        if v[0] * 1000 >= total:
            print("%-28s %12d  %5.2f%%  lanes %5.1f  samples %5.2f%%" % ("%s:%d" % loc if loc else "-", v[0], 100.0 * v[0] / total,
                                                                        v[1] / max(v[0], 1), 100.0 * v[2] / max(tot_samples, 1)))


if __name__ == "__main__":
    main()
