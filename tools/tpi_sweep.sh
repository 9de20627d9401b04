#!/bin/bash
# development aid (under gpurun): short dbmixer bench for several CSIM_TPI_OPTS masks
for m in "$@"; do
  CSIM_TPI_OPTS=$m timeout 120 python bench.py --steps 2 --warmup 2 --tran-steps 2000 --no-cpu-baseline --no-e2e --stamp-instances 65536 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('opts', $m, 'value %.1f M' % (d['value']/1e6), 'ms', d['ms_per_step'])"
done
