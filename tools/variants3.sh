#!/bin/bash
# co-residency experiment: half-size FFN CTAs + small predictor CTAs, so that predictor(l+1) really overlaps FFN(l)
mkdir -p gpurun_out
run() {
    name=$1; shift
    env "$@" timeout 120 python bench.py --steps 1000 --warmup 10 --no-batched --no-cpu-baseline > gpurun_out/v3_$name.json 2> gpurun_out/v3_$name.err
    python - "$name" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads([l for l in open(f'gpurun_out/v3_{n}.json') if l.startswith('{')][0])
    r = d['roofline']
    print(f"{n:28s} tok/s {d['value']:7.1f} e2e {d['e2e']['value']:7.1f} | ffn {r['us_per_launch']:.2f} us pred {r['predictor']['us_per_layer']:.2f} us layer {r['layer']['us_per_layer']:.2f} us")
except Exception as e:
    print(n, 'FAILED', e)
PY
}
run small_w6 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=6
run small_w7 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=7
run small_w9 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=9
run small_w5 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=5
run small_w6_vl2 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=6 SPIF_DG_VL=2
run small_w8_vl2 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=8 SPIF_DG_VL=2
run small_w12_vl2 SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=12 SPIF_DG_VL=2
run small_w6_serial SPIF_DIRECT_SMALL=1 SPIF_DG_WARPS=6 SPIF_X=1
