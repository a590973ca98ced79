#!/bin/sh
# Puts the UNMODIFIED reference package where it can travel to the GPU box: baseline/_ref/ is
# git-ignored (never part of this repository's history) but not gpurun-ignored.
# `pip install --no-index --no-build-isolation --target baseline/_ref /root/reference` fails in
# this image (the build backend, hatchling, is not installed); the reference is pure Python,
# so installing it amounts to copying its package directory.
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
rm -rf "$HERE/_ref"
mkdir -p "$HERE/_ref"
cp -r /root/reference/src/aiko_services "$HERE/_ref/aiko_services"
find "$HERE/_ref" -name '__pycache__' -type d -exec rm -rf {} + 2>/dev/null || true
echo "reference installed into $HERE/_ref"
