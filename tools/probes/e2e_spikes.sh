# Probe: hunts host-side stalls in the e2e arm.  For each pool head-room setting it runs bench.py several
# times with LSB_TRACE=1 and prints the trace line of every frame holding a span above 5 ms, plus the
# state of the device's stream-ordered memory pool at the end of the run.
for mb in ${POOL_MBS:-4096 2048 0}; do
echo "== LSB_POOL_PREWARM_MB=$mb"
for i in 1 2 3 4 5 6; do
  LSB_POOL_PREWARM_MB=$mb LSB_TRACE=1 timeout 200 python bench.py --no-cpu-baseline --multi-seq 0 --steps 40 --warmup 5 2> /tmp/trace_$i.txt > /tmp/out_$i.json
  python - <<PY
import re,json
d=json.load(open('/tmp/out_$i.json'))
print("run $i: e2e", round(d['e2e']['value'],1), "max", round(d['e2e']['max_ms_per_frame'],2))
for line in open('/tmp/trace_$i.txt'):
    if 'memory pool' in line: print(line.strip())
    if line.startswith('[lsb trace]'):
        nums=[int(x) for x in re.findall(r'[=(h](\d+)', line)]
        if any(n>5000 for n in nums): print(line.strip()[:600])
PY
done
done
