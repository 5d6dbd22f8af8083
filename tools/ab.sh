python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for w in "final 64" "big1 16" "cornell_smoke 16" "mesh 16"; do RT_BVH=karras python tools/bvh_probe.py $w; python tools/bvh_probe.py $w; done 2>&1 | grep -v Warn
