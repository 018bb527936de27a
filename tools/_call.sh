set -u
O=gpurun_out/call4; mkdir -p $O
python -m dgraph_b200.build > $O/build.log 2>&1 || { echo BUILD FAILED; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 240 $TR --nproc-per-node 8 --master-port 29631 benchmarks/scale_modes.py --modes default,pull_kernel,pull_buffer,exchange_x,overlap_gemm --out $O/modes8.json > $O/modes8.log 2>&1; echo "modes8 rc=$?"; grep "ms/step" $O/modes8.log
timeout 240 $TR --nproc-per-node 8 --master-port 29632 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench8.json 2> $O/bench8.err; echo "bench8 rc=$?"; cut -c1-300 $O/bench8.json
timeout 240 $TR --nproc-per-node 8 --master-port 29633 bench.py --gpus 8 --steps 20 --warmup 5 --cuda-graph --no-e2e > $O/bench8_graph.json 2> $O/bench8_graph.err; echo "bench8graph rc=$?"; cut -c1-300 $O/bench8_graph.json
timeout 240 $TR --nproc-per-node 4 --master-port 29634 benchmarks/scale_modes.py --modes default,pull_kernel,pull_buffer --out $O/modes4.json > $O/modes4.log 2>&1; echo "modes4 rc=$?"; grep "ms/step" $O/modes4.log
timeout 400 python -m pytest tests/test_peer_engine_multigpu.py -q -m gpu -x -p no:cacheprovider > $O/mgpu_test.log 2>&1; echo "mgpu tests rc=$? $(tail -1 $O/mgpu_test.log)"
