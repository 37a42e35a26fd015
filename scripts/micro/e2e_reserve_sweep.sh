for r in 2 16 32; do
  DLAP_E2E_RESERVE=$r timeout 200 python bench.py --steps 50 --no-cpu 2>/dev/null > gpurun_out/sw_$r.json
  python -c "import json; d=json.loads(open('gpurun_out/sw_$r.json').read().strip().splitlines()[-1]); print($r, d['e2e']['value']/1e6, d['e2e']['ms_per_step'])"
done
