#!/usr/bin/env bash
# Test tiers in one place (see tests/README.md).  Usage: tools/run_tiers.sh cpu | gpu | multigpu | reference-compat
set -euo pipefail
cd "$(dirname "$0")/.."
case "${1:-cpu}" in
  cpu)       python -m pytest tests -x -q -m "not gpu" ;;
  gpu)       python -m dgraph_b200.build && python -m pytest tests -x -q -m gpu ;;
  multigpu)  python -m pytest tests/test_peer_engine_multigpu.py -x -q -m gpu ;;
  reference-compat)  # the reference's own single-process tests against this package (needs /root/reference)
             python -m pytest tests/test_vs_reference_oracle.py -q -m "not gpu" ;;
  graphcast) (cd experiments/GraphCast && python -m pytest tests -q) ;;
  *) echo "unknown tier $1" >&2; exit 2 ;;
esac
