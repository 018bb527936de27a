set -u
O=gpurun_out/call16; mkdir -p $O
python -m dgraph_b200.build > $O/build.log 2>&1 || { echo BUILD FAILED; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
show() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print(d.get('value'), d.get('unit'), 'ms', d.get('ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_step'), 'loss', d.get('loss_first'), d.get('final_loss'), 'launches', d.get('gpu_launches'), 'pl', ((d.get('config') or {}).get('powerlaw_variant') or {}).get('ms_per_step'), d.get('max_mem_GB_rank0'))" 2>&1 | tail -1; }
timeout 300 $TR --master-port 29681 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench8.json 2> $O/bench8.err; echo "bench8 rc=$? $(show $O/bench8.json)"
timeout 600 $TR --master-port 29683 bench.py --gpus 8 --config papers100m --steps 10 --warmup 3 > $O/papers8.json 2> $O/papers8.err; echo "papers8 rc=$? $(show $O/papers8.json)"; [ -s $O/papers8.json ] || grep -i "error" $O/papers8.err | head -5
timeout 400 $TR --master-port 29686 bench.py --gpus 8 --config halo --halo-min-rows 100000 --steps 30 > $O/halo8.json 2> $O/halo8.err; echo "halo8 rc=$? $(show $O/halo8.json)"; [ -s $O/halo8.json ] || grep -i "error" $O/halo8.err | head -5
timeout 300 python -m pytest tests/test_peer_engine_multigpu.py -q -m gpu -x -p no:cacheprovider -k "8" > $O/mgpu_test.log 2>&1; echo "mgpu tests rc=$? $(tail -1 $O/mgpu_test.log)"; grep "AssertionError" $O/mgpu_test.log | head -3 | cut -c1-900
