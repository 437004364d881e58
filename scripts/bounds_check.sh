#!/bin/bash
# Builds libdli.so with -DDLI_BOUNDS_CHECK (every index the kernels form into a mesh buffer is checked on the device:
# printf + trap) and runs the small GPU parity tests against it.  Stands in for compute-sanitizer memcheck, which is
# not available on the build pool.  Usage (on a GPU box): bash scripts/bounds_check.sh [log]
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
LOG="${1:-$ROOT/gpurun_out/bounds_check.log}"
cd "$ROOT"
[ -f build/libdli_bounds.so ] || python scripts/build_variants.py bounds=-DDLI_BOUNDS_CHECK || exit 1
DLI_LIB="$ROOT/build/libdli_bounds.so" python -m pytest -m gpu -q -x \
  tests/test_gpu_parity_2d.py tests/test_gpu_parity_dd.py tests/test_gpu_gradient.py tests/test_gpu_update.py \
  tests/test_gpu_snapshot.py tests/test_gpu_round2.py tests/test_plot_data.py > "$LOG" 2>&1
rc=$?
echo "bounds-check build: pytest rc=$rc; DLI_BOUNDS_CHECK reports: $(grep -c 'DLI_BOUNDS_CHECK:' "$LOG")" | tee -a "$LOG"
tail -3 "$LOG"
exit $rc
