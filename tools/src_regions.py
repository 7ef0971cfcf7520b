"""Scratch: per-region instruction / sample shares of one kernel from `ncu --page source --csv`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
for b in blocks:
    if want not in b["name"]:
        continue
    h = b["rows"][0]; data = b["rows"][1:]
    isrc, ie, isamp, itag = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples"), h.index("L1 Tag Requests Global")
    tot = sum(int(r[ie]) for r in data); tots = sum(int(r[isamp]) for r in data)
    print(b["name"][:80], "inst", tot, "samples", tots, "n", len(data))
    keys = sys.argv[3].split(",") if len(sys.argv) > 3 else ["LDG", "BAR.SYNC", "EXIT", "MUFU", "LDL", "STG"]
    acc_e = acc_s = 0; last = 0
    for i, r in enumerate(data):
        acc_e += int(r[ie]); acc_s += int(r[isamp])
        s = r[isrc]
        if any(k in s for k in keys):
            print("%5d %-58s exec %10s samp %6s tag %9s | since last mark: inst %5.1f%% samp %5.1f%%" % (
                i, s.strip()[:58], r[ie], r[isamp], r[itag], 100.0 * acc_e / tot, 100.0 * acc_s / tots))
            acc_e = acc_s = 0
    print("tail: inst %5.1f%% samp %5.1f%%" % (100.0 * acc_e / tot, 100.0 * acc_s / tots))
    break
