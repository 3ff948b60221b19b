"""Source-line hot-spot table of an .ncu-rep captured with --import-source on (ncu --page source --print-source cuda,sass):
per CUDA source line, the warp instructions executed and the stall samples of the SASS it became, top N by samples.
usage: python tools/ncu_hotspots.py file.ncu-rep [top=25]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
lines, fname, hdr, cur = {}, "", None, None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        i_ex, i_smp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        i_thr = hdr.index("Thread Instructions Executed") if "Thread Instructions Executed" in hdr else None
        i_cfl = next((i for i, h in enumerate(hdr) if "Shared" in h and "Excessive" in h), None)   # excess shared-memory wavefronts = bank conflicts
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0] != "":                      # a CUDA source line; the SASS rows that follow carry the numbers
        cur = (fname, int(r[0]), r[1].strip())
        lines.setdefault(cur, [0, 0, 0, 0])
        if r[2] == "-":
            continue
    if cur is not None and r[2] not in ("-", ""):
        try:
            lines[cur][0] += int(r[i_ex])
            lines[cur][1] += int(r[i_smp])
            if i_thr is not None:
                lines[cur][2] += int(r[i_thr])
            if i_cfl is not None and r[i_cfl] not in ("", "-"):
                lines[cur][3] += int(float(r[i_cfl]))
        except ValueError:
            pass
tot_ex = sum(v[0] for v in lines.values()) or 1
tot_s = sum(v[1] for v in lines.values()) or 1
tot_c = sum(v[3] for v in lines.values())
print("== %s: %d warp instructions, %d stall samples, %d excess shared-memory wavefronts (bank conflicts); top %d source lines by samples"
      % (rep, tot_ex, tot_s, tot_c, top))
print("%6s %6s %6s %6s  %-22s %s" % ("smp%", "inst%", "lanes", "cfl%", "file:line", "source"))
for (f, ln, src), (ex, smp, thr, cfl) in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%6.2f %6.2f %6.1f %6.2f  %-22s %s" % (100.0 * smp / tot_s, 100.0 * ex / tot_ex, thr / ex if ex else 0.0,
                                                  100.0 * cfl / tot_c if tot_c else 0.0, "%s:%d" % (f, ln), src[:110]))
