show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1', 'tok/s %.1f e2e %.1f | ffn %.2f us (%.0f GB/s) pred %.2f us layer %.2f us' % (d['value'], d['e2e']['value'], r['us_per_launch'], r['achieved'], r['predictor']['us_per_layer'], r['layer']['us_per_layer']))"; }
B="python bench.py --steps 300 --no-cpu-baseline"
$B | show serial_early
$B --no-early-mask | show serial_noearly
$B --schedule two-streams | show two_streams
SPIF_SMEM_SPLIT=2 $B | show serial_split2
SPIF_SMEM_SPLIT=2 $B --schedule two-streams | show two_streams_split2
SPIF_CTAS_PER_SM=2 $B | show serial_2cta
SPIF_CTAS_PER_SM=2 $B --schedule two-streams | show two_streams_2cta
