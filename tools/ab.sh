run() { python bench.py --spp 128 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --pool 4194304 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']; print('$1', round(d['value']), round(d['ms_per_step'],1), round(k['extend'],1), round(k['shade'],1))"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do RT_EXP=0 run base; RT_EXP=8 run shade128; RT_EXP=4 run parked; RT_EXP=12 run parked+shade128; done
