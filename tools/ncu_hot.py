"""Top stall instructions of the first kernel in an .ncu-rep, split at the consumer's setmaxnreg (producer | consumer)."""
import csv, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; data = rows[2:]
iS, iE, iSrc = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
S = lambda r: int(r[iS] or 0)
ia = [i for i, r in enumerate(data) if "USETMAXREG.TRY_ALLOC" in r[iSrc]]
if ia:
    ia = ia[0]
    print("producer samples", sum(S(r) for r in data[:ia]), "inst", sum(int(r[iE] or 0) for r in data[:ia]),
          "| consumer samples", sum(S(r) for r in data[ia:]), "inst", sum(int(r[iE] or 0) for r in data[ia:]))
top = sorted(range(len(data)), key=lambda i: -S(data[i]))[:ntop]
for i in sorted(top):
    print(i, S(data[i]), data[i][iE], data[i][iSrc][:100])
