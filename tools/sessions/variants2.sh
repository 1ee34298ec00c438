show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1', 'tok/s %.1f e2e %.1f | ffn %.2f us (%.0f GB/s) pred %.2f us layer %.2f us' % (d['value'], d['e2e']['value'], r['us_per_launch'], r['achieved'], r['predictor']['us_per_layer'], r['layer']['us_per_layer']))"; }
B="python bench.py --steps 300 --no-cpu-baseline"
for pct in 45 62 80; do for split in 1 2; do
SPIF_GU_PCT=$pct SPIF_SMEM_SPLIT=$split $B | show "pct$pct split$split"
done; done
SPIF_GU_PCT=62 SPIF_DSLOTS=3 $B | show "pct62 d3 split2"
SPIF_GU_PCT=80 SPIF_DSLOTS=2 $B | show "pct80 d2 split2"
