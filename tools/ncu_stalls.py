"""Top stalled SASS instructions of one launch in an ncu report (source page, SASS view).
python tools/ncu_stalls.py report.ncu-rep <launch index> [top N]"""
import csv, subprocess, sys
rep, idx = sys.argv[1], int(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(idx), "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
print(rows[0][1][:120])
hdr = rows[1]
i_src, i_s = hdr.index("Source"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr) and r[i_s].isdigit()]
tot = sum(int(r[i_s]) for r in body)
agg = {hdr[i]: sum(int(r[i]) for r in body) for i in stall_cols}
print("samples", tot, {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for n, r in sorted(enumerate(body), key=lambda nr: -int(nr[1][i_s]))[:top]:
    st = sorted(((int(r[i]), hdr[i]) for i in stall_cols), reverse=True)[:2]
    print(f"{n:5d} {int(r[i_s]):6d} {100*int(r[i_s])/max(tot,1):5.1f}%  {r[i_src].strip()[:70]:70s} {st}")
