#!/usr/bin/env bash
# Regenerate the default self-signed pair (counterpart of the reference's resources/ssl_config/gen_certificates.sh,
# which shells out to openssl; here the generator is the `cryptography`-based SSLConfigurator so no CLI tool is needed).
set -euo pipefail
out="${1:-$(dirname "$0")/default}"
mkdir -p "$out"
python -m metisfl_b200.utils.ssl_configurator --out "$out" "${@:2}"
