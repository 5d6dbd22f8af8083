for w in "final 128" "big1 16" "cornell_smoke 128" "mesh 32"; do
  for v in "" t256b3 t256b4 t128b2 t128b3 t128b4 t64b3 t64b4; do
    if [ -z "$v" ]; then unset RT_B200_LIB; else export RT_B200_LIB=$PWD/raytracing_gpu_b200/librt_b200_$v.so; fi
    AB_ONLY_DEFAULT=1 python tools/ab_variants.py $w 3 2>&1 | grep variant
  done
done
