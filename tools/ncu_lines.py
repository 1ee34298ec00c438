"""Instructions executed / stall samples per SOURCE LINE: joins `ncu --page source --print-source sass --csv` (per-instruction
counters, program order) with `nvdisasm -g` of the same cubin (per-instruction source lines, program order).
    python tools/ncu_lines.py sass.csv all.dis <kernel-substring> [N]"""
import csv, re, sys, collections
sass, dis, kern = sys.argv[1], sys.argv[2], sys.argv[3]
N = int(sys.argv[4]) if len(sys.argv) > 4 else 40
lines = open(dis).read().split("\n")
start = [i for i, l in enumerate(lines) if l.startswith("//--------------------- .text.") and kern in l][0]
loc, cur = [], None
for l in lines[start + 1:]:
    if l.startswith("//--------------------- "):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]+\*/\s", l):
        loc.append(cur)
rows = list(csv.reader(open(sass)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
H = rows[hi]; col = {n: i for i, n in enumerate(H)}
body = [r for r in rows[hi + 1:] if len(r) > col["# Samples"]][:len(loc)]
def num(r, n):
    try: return float(r[col[n]].replace(",", ""))
    except Exception: return 0.0
ex, sm = collections.Counter(), collections.Counter()
for r, lc in zip(body, loc):
    ex[lc] += num(r, "Instructions Executed"); sm[lc] += num(r, "# Samples")
te, ts = sum(ex.values()), sum(sm.values())
print(f"instructions (SASS) {len(loc)}, executed {te:.0f}, samples {ts:.0f}")
src = {}
for lc, v in sorted(ex.items(), key=lambda kv: -kv[1])[:N]:
    if lc and lc[0] not in src:
        try: src[lc[0]] = open("/root/repo/sparkinfer_b200/csrc/" + lc[0]).read().split("\n")
        except Exception: src[lc[0]] = None
    text = src[lc[0]][lc[1] - 1].strip()[:70] if lc and src.get(lc[0]) else ""
    print(f"{100 * v / te:5.1f}% exec {int(v):9d}  samples {100 * sm[lc] / max(ts, 1):5.1f}%  {lc}  {text}")
