#!/bin/sh
# A/B builds of kernel variants: tools/build_variant.sh NAME "-DFLAG=1 ..." -> rslqr_b200/variants/librslqr_NAME.so
# (git-ignored, travels to the GPU box); select it with RSLQR_LIB=rslqr_b200/variants/librslqr_NAME.so.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$ROOT/rslqr_b200/variants"
make -C "$ROOT/rslqr_b200/csrc" -j8 OUT="$ROOT/rslqr_b200/variants/librslqr_$1.so" BUILD="$ROOT/rslqr_b200/csrc/build/variant_$1" EXTRA="$2" >/dev/null
echo "built rslqr_b200/variants/librslqr_$1.so ($2)"
