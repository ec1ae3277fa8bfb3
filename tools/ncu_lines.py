"""Aggregate `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K` output:
hottest CUDA source lines with instruction share, average active lanes and stall-sample share."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = []; H = None; fname = "?"
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": H = r; iI = H.index("Instructions Executed"); iT = H.index("Thread Instructions Executed"); iS = H.index("# Samples"); continue
    if H is None or len(r) < len(H) or r[0] == "": continue          # rows with an empty line number are the SASS children
    try:
        ie = int(r[iI]); te = int(r[iT]); smp = int(r[iS])
    except ValueError:
        continue
    if ie > 0: out.append((ie, te, smp, fname, r[0], r[1].strip()[:90]))
tot = sum(o[0] for o in out); tots = max(1, sum(o[2] for o in out))
print(f"total warp-inst {tot}  avg lanes {sum(o[1] for o in out)/max(1,tot):.2f}  samples {tots}")
for ie, te, smp, f, ln, src in sorted(out, key=lambda x: -x[0])[:top]:
    print(f"{ie/tot*100:5.1f}% lanes={te/ie:5.1f} smp={smp/tots*100:5.1f}% {f}:{ln:>4} {src}")
