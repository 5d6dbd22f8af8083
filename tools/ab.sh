run() { python bench.py --spp 128 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']; print('$1', round(d['value']), round(d['ms_per_step'],1), round(k['extend'],1), round(k['shade'],1))"; }
for i in 1 2; do run nb8; RT_B200_LIB=$PWD/raytracing_gpu_b200/librt_b200_nb6.so run nb6; RT_B200_LIB=$PWD/raytracing_gpu_b200/librt_b200_nb4.so run nb4; done
