#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- builds the UNMODIFIED reference (gabrielfougeron/Choreographies2,
# `choreo`) from the sources where they lie under /root/reference into oracle/_ref/ so that
# tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs can use the
# real reference as the checker / CPU baseline.  Nothing in choreographies2_b200/ may import it.
#
# Recipe = SURVEY.md section 8(c):
#   * the reference's setup.py writes choreo/cython/optional_pyfftw.pxd into its own tree, and
#     /root/reference is read-only, so the build happens in a scratch copy under /tmp;
#   * CC=/usr/bin/gcc (the venv's gcc cannot link -fopenmp), --no-fftw (pyfftw is not installed,
#     the reference then uses its own pyfftw_fake stub and the scipy FFT backend);
#   * -march=native is replaced by -march=$MARCH (default x86-64-v3) because the built .so files
#     travel to the GPU box whose host CPU may differ from this container's; no source is changed;
#   * only the INSTALLED package (the .py files + built .so, what `pip install --target` would
#     put there) is copied to oracle/_ref/ (git-ignored); .pyx/.c sources never enter the repo.
#   * a 2-file matplotlib stub is written next to it (choreo/NBodySyst_build.py imports
#     matplotlib unconditionally; plotting is never reached on the hot path).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${CHOREO_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
SCRATCH="${CHOREO_REF_SCRATCH:-/tmp/choreo_ref_build}"
MARCH="${MARCH:-x86-64-v3}"
PY="${PYTHON:-python}"

if [ ! -d "$REF/choreo" ]; then
  echo "build_ref.sh: $REF not present (GPU box?) -- keeping prebuilt oracle/_ref as is" >&2
  exit 0
fi
if [ -f "$OUT/.built" ] && [ "${FORCE:-0}" != "1" ]; then
  echo "build_ref.sh: oracle/_ref already built (FORCE=1 to rebuild)"; exit 0
fi

rm -rf "$SCRATCH"; mkdir -p "$SCRATCH"
for f in choreo setup.py pyproject.toml include README.md LICENSE LICENSES requirements; do
  cp -r "$REF/$f" "$SCRATCH/"
done
chmod -R u+w "$SCRATCH"
sed -i "s/-march=native/-march=$MARCH/g" "$SCRATCH/setup.py"
( cd "$SCRATCH" && CC=/usr/bin/gcc LDSHARED="/usr/bin/gcc -shared" \
    "$PY" setup.py build_ext --inplace --no-fftw > "$SCRATCH/build.log" 2>&1 ) \
  || { tail -50 "$SCRATCH/build.log"; exit 1; }

rm -rf "$OUT"; mkdir -p "$OUT"
( cd "$SCRATCH" && find choreo \( -name '*.py' -o -name '*.so' \) -not -path '*/legacy/*' -print0 \
    | xargs -0 -I{} cp --parents {} "$OUT/" )
mkdir -p "$OUT/matplotlib"
printf 'colormaps = {}\n' > "$OUT/matplotlib/__init__.py"
: > "$OUT/matplotlib/pyplot.py"
echo "march=$MARCH built $(date -u +%FT%TZ)" > "$OUT/.built"
echo "build_ref.sh: reference installed into $OUT"
