# usage: bash tools/ab_libs.sh "<workload spp>" ... -- <variant names>   (variant "" = the default library)
# e.g.   bash tools/ab_libs.sh "mesh 32" "final 128" -- "" wmagic wnosort
ws=(); vs=(); seen=0
for a in "$@"; do if [ "$a" = "--" ]; then seen=1; elif [ $seen = 0 ]; then ws+=("$a"); else vs+=("$a"); fi; done
for w in "${ws[@]}"; do
  for v in "${vs[@]}"; do
    if [ -z "$v" ]; then unset RT_B200_LIB; else export RT_B200_LIB=$PWD/raytracing_gpu_b200/librt_b200_$v.so; fi
    AB_ONLY_DEFAULT=1 python tools/ab_variants.py $w 3 2>&1 | grep -E "variant|n_wide" | grep -v warp_shade
  done
done
