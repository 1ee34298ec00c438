#!/bin/bash
# co-residency experiment: half-size FFN CTAs + small predictor CTAs, so that predictor(l+1) really overlaps FFN(l)
mkdir -p gpurun_out
run() {
    name=$1; shift
    env "$@" timeout 120 python bench.py --steps 600 --warmup 10 --no-batched --no-cpu-baseline > gpurun_out/v3_$name.json 2> gpurun_out/v3_$name.err
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
run p132 SPIF_FFN_CTAS=132
run p120 SPIF_FFN_CTAS=120
run p104 SPIF_FFN_CTAS=104
run p120_w12 SPIF_FFN_CTAS=120 SPIF_DG_WARPS=12
