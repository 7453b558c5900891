#!/bin/sh
# "Installs" the UNMODIFIED reference into the git-ignored baseline/_ref/ so that it travels to the GPU box
# with the gpurun snapshot (like the built .so files).  The reference has no setup.py / pyproject.toml, so
# `pip install --target baseline/_ref /root/reference` has nothing to build: the install is a plain copy of the
# package, its CLI and its shipped config.  Nothing under baseline/_ref is product source or enters git history;
# it is used only (a) by the -m gpu tests that run the reference's own `main.py --train/--eval` against the CUDA
# facades, (b) by bench.py's reference-timing legs (cpu_baseline_reference).
set -e
SRC="${1:-/root/reference}"
HERE="$(cd "$(dirname "$0")/.." && pwd)"
DST="$HERE/baseline/_ref"
if [ ! -d "$SRC/handyrl" ]; then
    echo "install_ref: no reference tree at $SRC (GPU box?) - keeping $DST as is"
    exit 0
fi
rm -rf "$DST"
mkdir -p "$DST"
cp -r "$SRC/handyrl" "$DST/handyrl"
cp "$SRC/main.py" "$SRC/config.yaml" "$DST/"
[ -f "$SRC/LICENSE" ] && cp "$SRC/LICENSE" "$DST/"
find "$DST" -name __pycache__ -type d -prune -exec rm -rf {} +
( cd "$SRC" && find handyrl main.py config.yaml -type f -name '*.py' -o -name '*.yaml' | sort | xargs sha256sum ) > "$DST/SHA256SUMS"
echo "install_ref: copied $(find "$DST" -name '*.py' | wc -l) python files into $DST"
