#!/bin/bash
# development aid (under gpurun): full-length dbmixer transient for several (CSIM_TPI_OPTS mask, instances) pairs
# usage: tools/tpi_cta_sweep.sh mask:instances ...
for mi in "$@"; do
  m=${mi%%:*}; n=${mi##*:}
  if [ "$m" = d ]; then unset CSIM_TPI_OPTS; m=-1; else export CSIM_TPI_OPTS=$m; fi
  timeout 200 python bench.py --steps 1 --warmup 1 --instances-per-gpu $n --no-cpu-baseline --no-e2e --stamp-instances 65536 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('opts', $m, 'instances', $n, 'value %.1f M' % (d['value']/1e6), 'ms %.0f' % d['ms_per_step'])"
done
