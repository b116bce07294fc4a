#!/bin/bash
# A/B: bench one configuration (scripts/bench_one.py key $1) with the default library and with every librsp_b200_<tag>.so beside it
KEY=${1:-c3}
cd rsp_chains_b200; cp librsp_b200.so /tmp/base.so
(cd .. && python scripts/bench_one.py $KEY | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('base', d['variant'], d['Msamples_per_s'], d['frac_of_measured_hbm_peak'])")
for f in librsp_b200_*.so; do
  tag=${f#librsp_b200_}; tag=${tag%.so}
  cp $f librsp_b200.so
  (cd .. && python scripts/bench_one.py $KEY | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$tag', d['variant'], d['Msamples_per_s'], d['frac_of_measured_hbm_peak'])")
done
cp /tmp/base.so librsp_b200.so
