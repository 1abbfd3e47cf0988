for f in "5,2.5,2.5" "5,2.5,3" "5,2.5,2" "5,3,2.5" "5,2,2.5" "4,2.5,2.5" "6,2.5,2.5" "5,3,3" "4,2,2"; do
  LSB_KNN_CELL_FACTORS=$f timeout 120 python bench.py --no-cpu-baseline --multi-seq 0 --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$f', round(d['value'],1), round(d['ms_per_step'],4), {k:round(v,3) for k,v in d['phase_ms_per_frame'].items()})"
done
