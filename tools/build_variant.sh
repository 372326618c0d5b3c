#!/bin/bash
# Usage: tools/build_variant.sh <name> [nvcc flags, e.g. -DSB_X=1]  -> variants/<name>.so (git-ignored; travels with gpurun)
# For A/B timing of compile-time variants in ONE gpurun call: cp variants/<name>.so scran_b200/libscranb200.so between runs.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p variants
tmp=$(mktemp -d)
mkdir -p "$tmp/pkg" "$tmp/include"
cp -r scran_b200/csrc "$tmp/pkg/csrc"; rm -rf "$tmp/pkg/csrc/build"
cp include/scranb200.h "$tmp/include/"
make -s -C "$tmp/pkg/csrc" EXTRA="$*" -j8 >/dev/null
cp "$tmp/pkg/libscranb200.so" "variants/$name.so"
rm -rf "$tmp"
echo "variants/$name.so"
