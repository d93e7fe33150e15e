"""Top stall sites of one kernel from `ncu -i rep --page source --csv --kernel-name regex:<k>` output.
    python tools/ncu_source_hot.py src.csv [top_n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
def f(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0
tot_s = sum(f(r, "# Samples") for r in data)
tot_i = sum(f(r, "Instructions Executed") for r in data)
print(f"total samples {tot_s:.0f}, warp instructions {tot_i:.0f}, shared wavefronts {sum(f(r,'L1 Wavefronts Shared') for r in data):.0f} (ideal {sum(f(r,'L1 Wavefronts Shared Ideal') for r in data):.0f})")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(f(r, s) for r in data) for s in stalls}
print("stall mix:", {k: round(100 * v / max(1, tot_s), 1) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:top]:
    main = max(stalls, key=lambda s: f(r, s))
    print(f"{100*f(r,'# Samples')/max(1,tot_s):5.1f}%  exec={f(r,'Instructions Executed'):9.0f} wf={f(r,'L1 Wavefronts Shared'):9.0f}/{f(r,'L1 Wavefronts Shared Ideal'):9.0f} {main:16s} {r[ix['Source']].strip()[:90]}")
