#!/bin/bash
# bench several builds of the library: librsp_b200_<tag>.so next to the default one
mkdir -p gpurun_out
cd rsp_chains_b200
cp librsp_b200.so /tmp/base.so
for f in librsp_b200_*.so; do
  tag=${f#librsp_b200_}; tag=${tag%.so}
  cp $f librsp_b200.so
  (cd .. && timeout 300 python bench.py --no-cpu-baseline --e2e-steps 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$tag', round(d['value']), round(d['roofline']['frac'],4), d['config']['variant'])")
done
cp /tmp/base.so librsp_b200.so
(cd .. && timeout 300 python bench.py --no-cpu-baseline --e2e-steps 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('base', round(d['value']), round(d['roofline']['frac'],4), d['config']['variant'])")
