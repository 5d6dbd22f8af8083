python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for w in "cornell_smoke 64" "final 64"; do python tools/bvh_probe.py $w; done 2>&1 | grep -v Warn
python bench.py --workload cornell_smoke --spp 64 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['roofline']['kernel_ms_per_step'])"
