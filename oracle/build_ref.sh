#!/usr/bin/env bash
# oracle/build_ref.sh -- TEST INFRASTRUCTURE.  Builds the real reference (luk-zm/cuda_raytracing)
# from the sources where they lie under $REF (default /root/reference) into oracle/_ref/:
#
#   raytracing_ref       the untouched reference binary  (nvcc -arch sm_100 -O3 src/main.cu; Makefile:2 with
#                        `-arch native` replaced because this container has no GPU)
#   raytracing_ref_ext   same sources through a sed-patched TEMP copy (deleted afterwards, never committed):
#                        env RT_WIDTH overrides settings.img_width (the reference has no width flag, needed
#                        for the 4K config) and env RT_DUMP_F32 names a file that receives the raw f32 image
#                        that pixels_to_ppm() is given (bit-exact tier-1 comparisons)
#   libref_host.so       the reference's __host__ __device__ code compiled for the host cores through
#                        oracle/ref_host_driver.cu (CPU oracle of record / CPU baseline kind "reference")
#
# oracle/_ref/ is git-ignored but NOT gpurun-ignored: the binaries travel to the GPU box, the sources do not.
set -euo pipefail
REF="${REF:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
  echo "build_ref.sh: $REF/src not present; keeping prebuilt files in $OUT (if any)"; exit 0
fi
mkdir -p "$OUT"
NVCC="${NVCC:-nvcc}"
ARCH="-gencode arch=compute_100,code=sm_100"
SUPP="-diag-suppress 550 -diag-suppress 177 -w"

# 1. untouched reference binary
$NVCC $ARCH -O3 -lm "$REF/src/main.cu" -I "$REF/include" $SUPP -o "$OUT/raytracing_ref"

# 2. width / float-dump extension through a temporary patched copy
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
cp "$REF"/src/*.cu "$REF"/src/*.cpp "$REF"/src/*.h "$TMP"/
sed -i 's|  render(&world, &settings);|  { const char *w_ = getenv("RT_WIDTH"); if (w_) settings.img_width = atoi(w_); }\n  render(\&world, \&settings);|' "$TMP/main.cu"
sed -i 's|  pixels_to_ppm(cpu_pixels_colors, img_width, img_height, settings->img_name);|  { const char *d_ = getenv("RT_DUMP_F32"); if (d_) { FILE *f_ = fopen(d_, "wb"); if (f_) { fwrite(cpu_pixels_colors, sizeof(vec3), n, f_); fclose(f_); } } }\n  if (!getenv("RT_NO_PPM")) pixels_to_ppm(cpu_pixels_colors, img_width, img_height, settings->img_name);|' "$TMP/raytracing.cu"
grep -q RT_WIDTH "$TMP/main.cu" && grep -q RT_DUMP_F32 "$TMP/raytracing.cu"
$NVCC $ARCH -O3 -lm "$TMP/main.cu" -I "$REF/include" $SUPP -o "$OUT/raytracing_ref_ext"

# 3. host build of the reference's __host__ __device__ code
$NVCC $ARCH -O3 -shared -Xcompiler -fPIC,-fopenmp -DREF_SRC="$REF/src" -I "$REF/include" $SUPP \
  "$HERE/ref_host_driver.cu" -o "$OUT/libref_host.so" -lgomp
echo "build_ref.sh: built $(ls "$OUT" | tr '\n' ' ')"
