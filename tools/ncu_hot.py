"""Top SASS instructions by stall samples from `ncu --page source --csv --print-source sass` output.
    python tools/ncu_hot.py sass.csv [N] [--around UBLKCP]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
H = rows[hi]
col = {n: i for i, n in enumerate(H)}
body = [r for r in rows[hi + 1:] if len(r) > col["# Samples"]]
def num(r, n):
    try: return float(r[col[n]].replace(",", ""))
    except Exception: return 0.0
tot = sum(num(r, "# Samples") for r in body)
N = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 40
stalls = [n for n in H if n.startswith("stall_") and "Not Issued" not in n]
if "--around" in sys.argv:
    pat = sys.argv[sys.argv.index("--around") + 1]
    for i, r in enumerate(body):
        if pat in r[col["Source"]]:
            for j in range(max(0, i - 3), min(len(body), i + 3)):
                rr = body[j]
                print(j, rr[col["Source"]][:90], int(num(rr, "# Samples")), int(num(rr, "Instructions Executed")))
            print("--")
    sys.exit()
print("total samples", tot, "instructions", sum(num(r, "Instructions Executed") for r in body))
idx = sorted(range(len(body)), key=lambda i: -num(body[i], "# Samples"))[:N]
for i in sorted(idx):
    r = body[i]
    top = sorted(((num(r, s), s) for s in stalls), reverse=True)[:2]
    print(f"{i:6d} {int(num(r, '# Samples')):6d} {100 * num(r, '# Samples') / tot:5.1f}% exec={int(num(r, 'Instructions Executed')):8d} {r[col['Source']][:80]:80s} {top[0][1]}:{int(top[0][0])} {top[1][1]}:{int(top[1][0])}")
