#!/bin/bash
# Installs the reference's own Blackwell kernel (CuTe-DSL FlashAttentionForwardSm100, flash_attn/cute/) UNMODIFIED into the
# git-ignored baseline/_ref/ so that it travels to the GPU box with gpurun (where /root/reference does not exist).
# Library-side comparison only (tools/compare_libs.py): nothing under baseline/_ref is on the product path.
# The package pins nvidia-cutlass-dsl==4.3.0.dev0 (flash_attn/cute/pyproject.toml:25); the image has 4.5.0 -- whether the
# kernel JIT-compiles under it is what compare_libs.py finds out on the box (a number, or the captured traceback).
set -eu
SRC=${1:-/root/reference/flash_attn/cute}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
DST=$ROOT/baseline/_ref/flash_attn
mkdir -p "$DST"
rm -rf "$DST/cute"
cp -r "$SRC" "$DST/cute"
# a bare namespace marker: the stock flash_attn/__init__.py imports the FA2 CUDA extension, which the cute package does not need
: > "$DST/__init__.py"
echo "installed $(ls "$DST/cute" | wc -l) files into $DST/cute"
