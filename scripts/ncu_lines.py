"""Per-source-line summary of an ncu report captured with --import-source on (kernels built with -lineinfo):
warp instructions, average active threads, stall samples.   python scripts/ncu_lines.py rep.ncu-rep [launch-skip] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]
skip = sys.argv[2] if len(sys.argv) > 2 else "0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif r[0] == "Kernel Name":
        print(r[1][:150])
    elif hdr and r[0].isdigit():
        d = dict(zip(hdr, r))
        # the first "Source" column is the CUDA line; duplicates in hdr: use positions
        i_inst = hdr.index("Instructions Executed")
        i_thr = hdr.index("Thread Instructions Executed")
        i_smp = hdr.index("# Samples")
        try:
            lines.append((fname, int(r[0]), r[1].strip()[:90], int(r[i_inst]), int(r[i_thr]), int(r[i_smp])))
        except ValueError:
            pass
import os
rev = os.environ.get("SRC_REV")  # the report shows today's files: take the line text from the profiled revision instead
if rev:
    cache = {}
    def text(f, n):
        if f not in cache:
            r = subprocess.run(["git", "show", f"{rev}:pbrt-v4_b200/csrc/{f}"], capture_output=True, text=True)
            cache[f] = r.stdout.splitlines() if r.returncode == 0 else None
        return cache[f][n - 1].strip()[:90] if cache[f] and n <= len(cache[f]) else "?"
    lines = [(l[0], l[1], text(l[0], l[1]), l[3], l[4], l[5]) for l in lines]
tot_i = sum(l[3] for l in lines) or 1
tot_s = sum(l[5] for l in lines) or 1
tot_t = sum(l[4] for l in lines)
print(f"total warp instructions {tot_i}, avg active threads {tot_t / tot_i:.2f}, samples {tot_s}")
print(f"{'file:line':28s} {'inst%':>6s} {'thr':>5s} {'smp%':>6s}  source")
for l in sorted(lines, key=lambda l: -l[5])[:top]:
    print(f"{l[0] + ':' + str(l[1]):28s} {100 * l[3] / tot_i:6.2f} {l[4] / max(l[3], 1):5.1f} {100 * l[5] / tot_s:6.2f}  {l[2]}")
