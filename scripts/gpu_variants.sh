#!/bin/bash
# dev aid: kernel_bench over the regular library and every variant under _build/variants; $1 = --only filter
mkdir -p gpurun_out
f=${1:-mxfp8_quantize}
for so in transformerengine_b200/libtransformer_engine.so transformerengine_b200/_build/variants/libte_*.so; do
  n=$(basename $so .so)
  timeout 300 baseline/_ref/kernel_bench $so --name $n --only $f 2>&1 | grep -v checksums | python -c "
import sys, json
for ln in sys.stdin:
    try: d = json.loads(ln)
    except Exception: print(ln.strip()[:200]); continue
    print(f\"{d['lib'][:28]:28s} {d['case']:46s} {d.get('us', -1):9.2f} us  {d.get('rate', -1):8.1f} {d.get('unit','')} {d.get('error','')[:80]}\")
"
done | tee gpurun_out/r02_variants.txt
