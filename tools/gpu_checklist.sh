#!/usr/bin/env bash
# First GPU call of a round: run every validation / measurement that is waiting for hardware, each step in its own process with its
# own timeout (a kernel fault or a hang in one experimental path must not take the others down), logs under gpurun_out/checklist/.
#   tools/gpu_checklist.sh            # one GPU
#   tools/gpu_checklist.sh multi      # additionally the multi-GPU steps (uses every visible GPU, at least 2)
# Nothing here is a bench value for the judge (bench.py alone is); it tells which experimental switches are correct and which win.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/checklist
mkdir -p "$OUT"
SUMMARY="$OUT/summary.txt"
: > "$SUMMARY"

step() {  # step <name> <timeout_s> <command...>
  local name=$1 t=$2; shift 2
  local t0=$(date +%s)
  timeout --kill-after=10 "$t" "$@" > "$OUT/$name.log" 2>&1
  local rc=$?
  echo "$name rc=$rc $(( $(date +%s) - t0 ))s | $(grep -v '^\s*$' "$OUT/$name.log" | tail -1 | cut -c1-220)" | tee -a "$SUMMARY"
}

python -m dgraph_b200.build > "$OUT/build.log" 2>&1 || { echo "build failed" | tee -a "$SUMMARY"; exit 1; }

# ---- correctness, one GPU -----------------------------------------------------------------------------------------------------
step tests_gpu            900 python -m pytest tests -q -m gpu -p no:cacheprovider
step smoke                180 python -c "import __graft_entry__ as g; g.smoke()"

# ---- timing, one GPU ----------------------------------------------------------------------------------------------------------
step bench_1gpu           600 python bench.py --gpus 1 --steps 10 --warmup 3
step bench_1gpu_nobres    600 env DGRAPH_B200_GEMM_BRES=0 python bench.py --gpus 1 --steps 10 --warmup 3 --no-e2e --no-powerlaw
step gemm                 300 python benchmarks/bench_gemm.py
step gemm_bres            300 env DGRAPH_B200_GEMM_BRES=1 python benchmarks/bench_gemm.py
step graphcast            600 python bench.py --config graphcast --steps 10 --warmup 3 --no-e2e
step graphcast_unfused    600 env DGRAPH_B200_FUSED_MLP=0 python bench.py --config graphcast --steps 10 --warmup 3 --no-e2e
step graphcast_cudagraph  600 python bench.py --config graphcast --steps 10 --warmup 3 --no-e2e --cuda-graph
step rgat                 300 python bench.py --config rgat --rgat-relations first --steps 20 --warmup 5 --cuda-graph
step rgat_reference       600 python bench.py --config rgat --impl reference --steps 20 --warmup 5

if [ "${1:-}" = "multi" ]; then
  N=$(nvidia-smi -L | wc -l)
  if [ "$N" -ge 2 ]; then
    TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
    step test_multigpu        900 python -m pytest tests/test_peer_engine_multigpu.py tests/test_peer_hybrid_multigpu.py -q -m gpu -p no:cacheprovider
    step test_graph_multigpu  300 env DGRAPH_B200_TEST_DEVICE_EPOCH=1 python -m pytest tests/test_peer_graph_multigpu.py -q -m gpu -p no:cacheprovider
    for n in 2 4 8; do
      [ "$n" -le "$N" ] || continue
      step bench_${n}gpu            600 $TR --nproc-per-node $n --master-port 29601 bench.py --gpus $n --steps 20 --warmup 3
      step bench_${n}gpu_xchg       600 env DGRAPH_B200_GCN_EXCHANGE=x $TR --nproc-per-node $n --master-port 29602 bench.py --gpus $n --steps 20 --warmup 3 --no-e2e
      step bench_${n}gpu_eager      600 $TR --nproc-per-node $n --master-port 29603 bench.py --gpus $n --steps 20 --warmup 3 --no-e2e --no-cuda-graph --no-powerlaw
      step modes_${n}gpu            600 $TR --nproc-per-node $n --master-port 29607 benchmarks/scale_modes.py --out gpurun_out/checklist/modes_${n}gpu.json
    done
    step graphcast_${N}gpu            900 $TR --nproc-per-node $N --master-port 29604 experiments/GraphCast/train_graphcast.py --is_distributed True --backend nccl --procs_per_graph $N --iters 8
    step graphcast_${N}gpu_cudagraph  900 $TR --nproc-per-node $N --master-port 29605 experiments/GraphCast/train_graphcast.py --is_distributed True --backend nccl --procs_per_graph $N --iters 8 --cuda_graph True
    step papers100m_${N}gpu           900 $TR --nproc-per-node $N --master-port 29606 bench.py --gpus $N --config papers100m --steps 10 --warmup 3
    step halo_${N}gpu                 900 $TR --nproc-per-node $N --master-port 29608 bench.py --gpus $N --config halo --steps 30
    [ -d baseline/_ref/tests ] && step reference_dist 600 tools/run_tiers.sh reference-dist
  else
    echo "multi: only $N GPU visible" | tee -a "$SUMMARY"
  fi
fi
echo "---- summary ----"; cat "$SUMMARY"
