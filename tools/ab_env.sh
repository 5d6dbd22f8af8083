# Same-box A/B of run-time variants: tools/ab_env.sh "label:VAR=val VAR=val" ...   (inside ONE gpurun call; box-to-box
# variance is +-5 %).  Prints Msamples/s, ms/step, extend ms, shade ms, node-loop lane utilisation for two interleaved rounds.
run() { python bench.py --spp ${SPP:-128} --steps 2 --warmup 3 --no-cpu-baseline --no-e2e ${WORKLOAD:+--workload $WORKLOAD} ${POOL:+--pool $POOL} 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']; print('$1', round(d['value']), round(d['ms_per_step'],1), round(k['extend'],1), round(k['shade'],1), d.get('roofline_fp32',{}).get('warp_lane_utilisation_node_loop'))"; }
for i in 1 2; do
  for spec in "$@"; do
    label=${spec%%:*}; envs=${spec#*:}; [ "$envs" = "$spec" ] && envs=""
    ( for kv in $envs; do export "$kv"; done; run "$label" )
  done
done
