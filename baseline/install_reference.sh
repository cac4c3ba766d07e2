#!/bin/sh
# Offline install of the UNMODIFIED reference into baseline/_ref (git-ignored).
# Upstream ships no setup.py / pyproject.toml, so `pip install /root/reference` fails with
# "neither setup.py nor pyproject.toml found"; a packaging-only shim is added to a COPY under /tmp
# (source files are installed byte-for-byte: `diff -r baseline/_ref/motion /root/reference/src/motion`).
set -e
here="$(cd "$(dirname "$0")" && pwd)"
rm -rf /tmp/ref_copy && cp -r /root/reference /tmp/ref_copy
cat > /tmp/ref_copy/setup.py <<'PY'
from setuptools import setup
setup(name="pytorch-distributed-rnn-reference", version="0.0.0", package_dir={"": "src"},
      packages=["motion", "motion.trainer", "motion.param_server", "example"])
PY
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$here/_ref" --upgrade /tmp/ref_copy
