#!/usr/bin/env bash
# Sanitizer targets (the reference has none).
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck [core|train] - compute-sanitizer over the kernel tests
#                                                                 (ONE tool per gpurun call, see B200_PROFILING.md)
#   tools/sanitize.sh asan                                      - AddressSanitizer + UBSan build of the host core under the
#                                                                 wire-codec property tests and the aggregation rules (CPU only)
#   tools/sanitize.sh tsan                                      - ThreadSanitizer build of the host controller core (OpenMP off:
#                                                                 libgomp is not instrumented and only yields false positives)
#                                                                 + the protocol tests (CPU only)
set -euo pipefail
cd "$(dirname "$0")/.."
case "${1:-memcheck}" in
  memcheck|racecheck|synccheck|initcheck)
    python build.py kernels
    if [ "${2:-core}" = "train" ]; then  # the training-side kernels: K14 BatchNorm, K15 conv, K16 loss, K17 / MN-major GEMM
      exec compute-sanitizer --tool "$1" --error-exitcode 1 \
        python -m pytest tests/test_bn_gpu.py tests/test_loss_gpu.py tests/test_gemm_gpu.py tests/test_conv_gpu.py -x -q -m gpu \
          -k "(three_launches and (shape2 or shape4 or eval or affine)) or lm_cross_entropy or colsum or (mn_major and 200) or (weight_stationary and 5-6)"
    fi
    exec compute-sanitizer --tool "$1" --error-exitcode 1 \
      python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "weighted_sum_fast or quantifier or optimizer_step or rolling"
    ;;
  tsan)
    mkdir -p build/tsan
    CORE=$(ls csrc/core/*.cc csrc/ckks/*.cc)
    g++ -O1 -g -std=c++17 -shared -fPIC -ffp-contract=off -Wno-unknown-pragmas -fsanitize=thread \
      -I "$(python -c 'import pybind11;print(pybind11.get_include())')" \
      -I "$(python -c 'import sysconfig;print(sysconfig.get_paths()["include"])')" -I include -I csrc $CORE \
      -o "build/tsan/_core$(python -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
    # swap the instrumented module in for the duration of the protocol tests, then restore the regular build
    EXT="$(python -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
    cp "metisfl_b200/_core$EXT" "build/tsan/_core_regular$EXT"
    trap 'cp "build/tsan/_core_regular$EXT" "metisfl_b200/_core$EXT"' EXIT
    cp "build/tsan/_core$EXT" "metisfl_b200/_core$EXT"
    LD_PRELOAD="$(g++ -print-file-name=libtsan.so)" TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" \
      python -m pytest tests/test_controller_core.py tests/test_policies_and_stores.py -x -q -p no:cacheprovider
    ;;
  asan)
    # AddressSanitizer + UBSan build of the host core under the wire-codec property tests (untrusted learner bytes)
    # and the aggregation rules (payload-length checks)
    mkdir -p build/asan
    CORE=$(ls csrc/core/*.cc csrc/ckks/*.cc)
    EXT="$(python -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
    g++ -O1 -g -std=c++17 -shared -fPIC -fopenmp -ffp-contract=off -fsanitize=address,undefined -fno-omit-frame-pointer \
      -I "$(python -c 'import pybind11;print(pybind11.get_include())')" \
      -I "$(python -c 'import sysconfig;print(sysconfig.get_paths()["include"])')" -I include -I csrc $CORE \
      -o "build/asan/_core$EXT"
    cp "metisfl_b200/_core$EXT" "build/asan/_core_regular$EXT"
    trap 'cp "build/asan/_core_regular$EXT" "metisfl_b200/_core$EXT"' EXIT
    cp "build/asan/_core$EXT" "metisfl_b200/_core$EXT"
    LD_PRELOAD="$(g++ -print-file-name=libasan.so):$(g++ -print-file-name=libubsan.so)" \
      ASAN_OPTIONS="detect_leaks=0:halt_on_error=1" UBSAN_OPTIONS="print_stacktrace=1" \
      python -m pytest tests/test_wire_fuzz.py tests/test_proto_codec.py tests/test_aggregation_rules.py tests/test_ckks_host.py \
        tests/test_controller_core.py tests/test_policies_and_stores.py tests/test_envelope.py -x -q -p no:cacheprovider
    ;;
  *) echo "usage: $0 memcheck|racecheck|synccheck|initcheck|tsan|asan"; exit 2;;
esac
