"""Per-role stall summary of an am_project_tc_kernel ncu report (source page)."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; data = rows[2:]
isrc = hdr.index("Source"); isamp = hdr.index("# Samples"); iex = hdr.index("Instructions Executed")
stall_cols = [(n, hdr.index(n)) for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[isamp]) for r in data)
mma = [k for k, r in enumerate(data) if 'UTCHMMA' in r[isrc]]
ldtm = [k for k, r in enumerate(data) if 'LDTM' in r[isrc]]
def seg(a, b, name):
    s = sum(int(data[k][isamp]) for k in range(a, b)); ex = sum(int(data[k][iex]) for k in range(a, b))
    st = {n[6:]: sum(int(data[k][i] or 0) for k in range(a, b)) for n, i in stall_cols}
    print(name, f"{s/tot*100:.1f}% samples, exec {ex}", {k: v for k, v in sorted(st.items(), key=lambda x: -x[1])[:5]})
seg(0, mma[-1] + 40, "mma issuer(+setup)"); seg(mma[-1] + 40, ldtm[0] - 60, "producers"); seg(ldtm[0] - 60, len(data), "epilogue")
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.6
for k, r in enumerate(data):
    s = int(r[isamp])
    if s / tot * 100 >= thr:
        t2 = sorted(((int(r[i] or 0), n) for n, i in stall_cols), reverse=True)[:2]
        print(f"{k:5d} {s/tot*100:5.1f}% ex {int(r[iex]):9d} {r[isrc][:70]:70s}", " ".join(f"{n[6:]}:{v}" for v, n in t2 if v))
