for caps in "8 16" "4 16" "8 8" "12 16"; do set -- $caps; echo "tran cap $1 dc cap $2"; CSIM_TPI_MAX_VARIANTS=$1 CSIM_DC_MAX_VARIANTS=$2 python bench.py --netlist buffer --instances-per-gpu 65536 --steps 3 --warmup 14 --no-cpu-baseline --no-e2e --strong-total 0 2>/dev/null | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.read()); print('   value %.1f M'%(l['value']/1e6), 'ms %.1f'%l['ms_per_step'], 'dc %.1f'%l['roofline']['dc_kernel_ms'], 'tran %.1f'%l['roofline']['ms_per_launch'], l['parity_check']['ok'])"; done
