for i in 1 2; do python bench.py --no-classroom --no-dragon --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); v=d['path_tracer_cbox_dispersion']; print('disp ms',round(v['ms'],1),'sum stages',round(sum(v['roofline']['stage_ms'].values()),1), 'cbox ms', round(d['path_tracer']['ms'],2))"; done
