# Same-box A/B of library variants: tools/ab.sh label[=lib.so[,RT_EXP]] ...   (run inside ONE gpurun call; box-to-box
# variance is +-5 %).  Prints Msamples/s, ms/step, extend ms, shade ms for two interleaved rounds.
run() { python bench.py --spp ${SPP:-128} --steps 2 --warmup 3 --no-cpu-baseline --no-e2e ${WORKLOAD:+--workload $WORKLOAD} 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']; print('$1', round(d['value']), round(d['ms_per_step'],1), round(k['extend'],1), round(k['shade'],1), d.get('roofline_fp32',{}).get('warp_lane_utilisation_node_loop'))"; }
for i in 1 2; do
  for spec in "$@"; do
    label=${spec%%=*}; rest=${spec#*=}; [ "$rest" = "$spec" ] && rest=""
    lib=${rest%%,*}; exp=${rest#*,}; [ "$exp" = "$rest" ] && exp=""
    ( [ -n "$lib" ] && export RT_B200_LIB=$PWD/raytracing_gpu_b200/$lib; [ -n "$exp" ] && export RT_EXP=$exp; run $label )
  done
done
