for c in 2 4 6 8 12 16; do
  python bench.py --no-pt --no-cpu-baseline --steps 5 --opt trace.host_chunks=$c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('chunks',$c,'value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['e2e']['ms_per_step'],2),'soa',round(d['e2e']['soa_sync']['value']))"
done
