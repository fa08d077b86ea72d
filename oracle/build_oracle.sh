#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY: builds the plain-C oracle into oracle/_build/liboracle_c.so.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$HERE/_build"
gcc -O2 -fPIC -shared -fopenmp -o "$HERE/_build/liboracle_c.so" "$HERE/oracle_c.c" -lm
echo "[build_oracle] built $HERE/_build/liboracle_c.so"
