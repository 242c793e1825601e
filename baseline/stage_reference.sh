#!/usr/bin/env bash
# Installs the UNMODIFIED reference library into baseline/_ref (git-ignored, shipped to the
# GPU box by gpurun) and stages its benchmark model builders next to it.
#
# The reference's setup.py lists `pytest-runner` in setup_requires, which is not in the offline
# wheelhouse; an empty placeholder wheel of that name lets metadata generation succeed without
# touching the reference sources. `--no-deps` because torch is already in the image and the
# wheelhouse has no torch wheel to "resolve" against.
set -euo pipefail
cd "$(dirname "$0")/.."
WORK=$(mktemp -d)
cp -r /root/reference "$WORK/refsrc"            # /root/reference is read-only; build writes egg-info
mkdir -p "$WORK/ptr" && cd "$WORK/ptr"
printf "from setuptools import setup\nsetup(name='pytest-runner', version='6.0.1', py_modules=['ptr'])\n" > setup.py
echo "# placeholder" > ptr.py
python -m pip wheel --no-index --no-build-isolation --no-deps -w "$WORK/wheels" . >/dev/null
cd - >/dev/null
rm -rf baseline/_ref
PIP_NO_INDEX=1 PIP_FIND_LINKS="/opt/wheelhouse $WORK/wheels" \
python -m pip install --no-index --no-build-isolation --no-deps \
    --find-links /opt/wheelhouse --find-links "$WORK/wheels" \
    --target baseline/_ref "$WORK/refsrc"
# Model builders are not part of the pip package (reference setup.py:69); stage them verbatim.
mkdir -p baseline/_ref/benchmarks_models
cp -r /root/reference/benchmarks/models/. baseline/_ref/benchmarks_models/
# The fork's distributed mode is not packaged either (no __init__.py); stage verbatim.
cp -r /root/reference/torchgpipe/distributed baseline/_ref/torchgpipe/distributed
rm -rf "$WORK"
echo "reference staged under baseline/_ref"
