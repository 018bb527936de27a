set -u
O=gpurun_out/call5; mkdir -p $O
python -m dgraph_b200.build > $O/build.log 2>&1 || { echo BUILD FAILED; exit 1; }
timeout 600 python -m pytest tests -q -m gpu -x -p no:cacheprovider > $O/tests.log 2>&1; echo "tests rc=$? $(tail -1 $O/tests.log)"
timeout 300 python benchmarks/bench_segment.py > $O/seg.json 2> $O/seg.err; echo "seg rc=$?"; cat $O/seg.json
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/one.json 2> $O/one.err; echo "one rc=$?"; python -c "import json; d=json.loads(open('$O/one.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_first'], d['final_loss'], d['gpu_launches'])"
timeout 300 env DGRAPH_B200_GEMM_BRES=1 python bench.py --gpus 1 --steps 20 --warmup 5 --no-e2e > $O/one_bres.json 2> $O/one_bres.err; python -c "import json; d=json.loads(open('$O/one_bres.json').read().strip().splitlines()[-1]); print('bres', d['ms_per_step'], d['final_loss'])"
