#!/usr/bin/env bash
# Sanitizer targets (the reference has none).
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck   - compute-sanitizer over the kernel tests
#                                                                 (ONE tool per gpurun call, see B200_PROFILING.md)
#   tools/sanitize.sh tsan                                      - ThreadSanitizer build of the host controller core
#                                                                 + the protocol tests (CPU only)
set -euo pipefail
cd "$(dirname "$0")/.."
case "${1:-memcheck}" in
  memcheck|racecheck|synccheck|initcheck)
    python build.py kernels
    exec compute-sanitizer --tool "$1" --error-exitcode 1 \
      python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "weighted_sum_fast or quantifier or optimizer_step or rolling"
    ;;
  tsan)
    mkdir -p build/tsan
    CORE=$(ls csrc/core/*.cc csrc/ckks/*.cc)
    g++ -O1 -g -std=c++17 -shared -fPIC -fopenmp -fsanitize=thread \
      -I "$(python -c 'import pybind11;print(pybind11.get_include())')" \
      -I "$(python -c 'import sysconfig;print(sysconfig.get_paths()["include"])')" -I include -I csrc $CORE \
      -o "build/tsan/_core$(python -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
    # run the protocol tests against the instrumented module (shadowing the regular build through PYTHONPATH order)
    mkdir -p build/tsan/metisfl_b200 && cp -r metisfl_b200/*.py build/tsan/metisfl_b200/ 2>/dev/null || true
    LD_PRELOAD="$(g++ -print-file-name=libtsan.so)" TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" \
      METISFL_CORE_OVERRIDE="$PWD/build/tsan" python -m pytest tests/test_controller_core.py -x -q
    ;;
  *) echo "usage: $0 memcheck|racecheck|synccheck|initcheck|tsan"; exit 2;;
esac
