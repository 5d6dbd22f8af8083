for top in 1 2048 8192 32768; do RT_PLOC_TOP=$top python tools/bvh_probe.py mesh 16 2>/dev/null | sed "s/^/top=$top /"; done
