#!/usr/bin/env bash
# Re-creates baseline/_ref (git-ignored): the UNMODIFIED reference installed offline from /root/reference, plus unmodified copies of
# its experiment scripts and tests (they are not part of its wheel).  bench.py --impl reference and tools/run_tiers.sh reference-dist
# need it; run this after a fresh checkout or when the machine was replaced.  Takes about six minutes (torch headers).
set -euo pipefail
cd "$(dirname "$0")/.."
REF=${1:-/root/reference}
PY=${PYTHON:-python}
SITE=$($PY -c "import torch, os; print(os.path.dirname(os.path.dirname(torch.__file__)))")
rm -rf /tmp/ref_src baseline/_ref
cp -r "$REF" /tmp/ref_src          # the build writes into its source tree; the reference itself is read-only
export CMAKE_ARGS="-DDGRAPH_ENABLE_NVSHMEM=OFF -DTORCH_CUDA_ARCH_LIST=10.0 -DCMAKE_CUDA_ARCHITECTURES=100 -DTorch_DIR=$SITE/torch/share/cmake/Torch -Dpybind11_DIR=$SITE/pybind11/share/cmake/pybind11"
$PY -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref /tmp/ref_src
mkdir -p baseline/_ref/experiments
for d in OGB OGB-LSC GraphCast; do cp -r "$REF/experiments/$d" "baseline/_ref/experiments/$d"; done
cp -r "$REF/tests" baseline/_ref/tests
ls baseline/_ref/DGraph/torch_local*.so
echo "reference installed under baseline/_ref"
