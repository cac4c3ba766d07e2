#!/bin/sh
# Binary wheel with the sm_100a library inside (the reference shipped prebuilt armv7 wheels for its Raspberry Pi nodes,
# armv7_builds/*/Dockerfile; the B200 counterpart is one x86-64 / aarch64 wheel carrying libb200rnn.so).
#   deploy/build_wheel.sh [outdir]        -> outdir/pytorch_distributed_rnn_b200-*.whl   (default: dist/)
set -e
cd "$(dirname "$0")/.."
out="${1:-dist}"
python -m pytorch_distributed_rnn_b200._build            # compiles csrc/ for sm_100a if the library is stale
tmp="$(mktemp -d)"
# build from a copy: setuptools writes build/ and *.egg-info next to the sources
mkdir -p "$tmp/src" && cp -r pyproject.toml pytorch_distributed_rnn_b200 "$tmp/src/"
find "$tmp/src" -name __pycache__ -type d -prune -exec rm -rf {} +
python -m pip wheel --no-deps --no-build-isolation --no-index -w "$out" "$tmp/src"
rm -rf "$tmp"
ls -l "$out"
