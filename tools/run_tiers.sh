#!/usr/bin/env bash
# Test tiers in one place (see tests/README.md).  Usage: tools/run_tiers.sh cpu | gpu | multigpu | reference-compat | reference-mpi | reference-dist | graphcast
set -euo pipefail
cd "$(dirname "$0")/.."
case "${1:-cpu}" in
  cpu)       python -m pytest tests -x -q -m "not gpu" ;;
  gpu)       python -m dgraph_b200.build && python -m pytest tests -x -q -m gpu ;;
  multigpu)  python -m pytest tests/test_peer_engine_multigpu.py -x -q -m gpu ;;
  reference-compat)  # the reference's own single-process tests against this package (needs /root/reference)
             python -m pytest tests/test_vs_reference_oracle.py -q -m "not gpu" ;;
  reference-mpi)     # the reference's own MPI back-end tests (written for `mpirun -n 2`), unmodified, as two CPU ranks on the
             # shared-memory window engine (no GPU, no MPI needed)
             PYTHONPATH="$PWD" python tools/run_ranks.py -n 2 -- python -m pytest /root/reference/tests/test_mpi_backend.py -q -p no:cacheprovider ;;
  reference-dist)    # the reference's own torchrun tests (2 GPUs, NCCL) against this package.  The reference tree is not on a GPU box:
             # copy its tests next to the installed reference first (here): cp -r /root/reference/tests baseline/_ref/tests
             T=/root/reference/tests; [ -d "$T" ] || T=baseline/_ref/tests
             PYTHONPATH="$PWD" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
               -m pytest "$T/test_comm_info.py" "$T/test_NCCLCommPlan.py" -q -p no:cacheprovider ;;
  graphcast) (cd experiments/GraphCast && python -m pytest tests -q) ;;
  *) echo "unknown tier $1" >&2; exit 2 ;;
esac
