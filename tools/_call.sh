O=gpurun_out/call20; mkdir -p $O
timeout 90 python bench.py --impl reference --gpus 1 --scale 0.1 --steps 5 --warmup 3 > $O/ref_small.json 2> $O/ref_small.err; echo "ref rc=$?"; cut -c1-700 $O/ref_small.json; tail -3 $O/ref_small.err | cut -c1-300
