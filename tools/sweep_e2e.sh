# e2e (host buffers, record form, trace.async = 1) against the number of ray ranges a call is cut into; run under gpurun
STEPS=${STEPS:-10}
for c in ${CHUNKS:-1 2 3 4 6 8}; do
  python bench.py --no-pt --no-cpu-baseline --steps $STEPS --opt trace.host_chunks=$c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('chunks',$c,'steps',d['e2e']['steps'],'value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['e2e']['ms_per_step'],2),'soa',round(d['e2e']['soa_sync']['value']))"
done
