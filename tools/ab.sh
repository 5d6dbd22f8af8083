python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python tools/bvh_probe.py cornell_smoke 64 2>&1 | grep -v Warn
python bench.py --spp 256 --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],1), d['e2e']['value'])"
