for cfg in "33554432 148" "33554432 296" "50331648 148" "67108864 296" "16777216 148" "33554432 74"; do set -- $cfg; echo "group $1 minchunks $2"; PDG_FB_GROUP_BYTES=$1 PDG_M3_MIN_CHUNKS=$2 python bench.py --steps 5 --warmup 3 --no-cpu-baseline | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); p=d['phases_ms']
print('resident',round(d['ms_per_step'],2),'e2e',round(d['e2e']['ms_per_step'],2),'call',round(p['e2e_ms_ingest_call'],2),'copies',round(p['e2e_ms_h2d_copies'],2),'fin',round(p['e2e_ms_finalize'],2),'groups',p['e2e_sort_groups'])"; done
