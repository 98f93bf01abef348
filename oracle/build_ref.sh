#!/usr/bin/env bash
# Build the reference's own C for the host -> oracle/_ref/*.so   (TEST INFRASTRUCTURE)
#
# Sources are compiled where they lie under $REF (default /root/reference); the only
# files copied are audioprocessing.{c,h} + audio/mel_constants.h, into a throw-away
# directory under /tmp, because audioprocessing.h quote-includes "main.h" and would
# otherwise pick up the STM32 HAL header sitting next to it.  Nothing from $REF is
# written into the repo; outputs go to oracle/_ref/ only (git-ignored, shipped to the
# GPU box by gpurun like our own .so files).
#
# Libraries produced:
#   libref_mfcc_simd.so       audioprocessing.c -O0 (F6) + CMSIS-DSP, M4 SIMD branches on
#                             host-emulated intrinsics  == what the firmware runs (Makefile:30)
#   libref_mfcc_c.so          same, plain-C CMSIS branches (+-1 LSB different, F4)
#   libref_nnom_local.so      NNoM runtime, local C backend
#   libref_nnom_cmsis.so      NNoM + CMSIS-NN, reference (non-DSP) branches
#   libref_nnom_cmsis_simd.so NNoM + CMSIS-NN, M4 SIMD branches (firmware-true, nnom_port.h:42)
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
FW="$REF/firmware/src"
DSP="$FW/lib/CMSIS/DSP"
NN="$FW/lib/CMSIS/NN"
if [ ! -d "$FW" ]; then echo "build_ref: $REF not present, keeping prebuilt oracle/_ref" >&2; exit 0; fi
mkdir -p "$OUT"
TMP="$(mktemp -d /tmp/edison_ref_build.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT

python3 "$HERE/extract_tables.py" "$FW/lib/CMSIS/Lib/GCC/libarm_cortexM4lf_math.a" "$TMP/cmsis_tables.c" "$OUT/tables_manifest.json"

mkdir -p "$TMP/fw/audio"
cp "$FW/audioprocessing.c" "$FW/audioprocessing.h" "$TMP/fw/"
cp "$FW/audio/mel_constants.h" "$TMP/fw/audio/"
cp "$HERE"/shim/*.h "$TMP/fw/"

CC="${CC:-gcc}"
COMMON="-fPIC -w -D__GNUC_PYTHON__ -ffunction-sections -fdata-sections -I$DSP/Include -I$TMP/fw"
DSP_SRCS="TransformFunctions/arm_cfft_q15.c TransformFunctions/arm_cfft_radix4_q15.c TransformFunctions/arm_rfft_q15.c
          TransformFunctions/arm_rfft_init_q15.c TransformFunctions/arm_bitreversal2.c TransformFunctions/arm_bitreversal.c
          FastMathFunctions/arm_sqrt_q31.c SupportFunctions/arm_copy_q15.c CommonTables/arm_const_structs.c"

build_mfcc () {  # $1 = variant name, $2 = extra flags for the CMSIS sources, $3 = optimisation of the CMSIS sources, $4 = of audioprocessing.c
  local v="$1" extra="$2" opt="${3:--O2}" uopt="${4:--O0}" objs=""
  mkdir -p "$TMP/$v"
  for s in $DSP_SRCS; do
    o="$TMP/$v/$(basename "$s" .c).o"
    $CC $COMMON $opt $extra -c "$DSP/Source/$s" -o "$o"; objs="$objs $o"
  done
  # user code: -O0, like the firmware (Makefile:42) -- and required, see SURVEY F6 (the B2 variant below is the exception)
  $CC $COMMON $uopt $extra -c "$TMP/fw/audioprocessing.c" -o "$TMP/$v/audioprocessing.o"
  $CC $COMMON -O2 -c "$TMP/cmsis_tables.c" -o "$TMP/$v/cmsis_tables.o"
  $CC $COMMON -O2 -c "$HERE/drivers/ref_mfcc_driver.c" -o "$TMP/$v/driver.o"
  $CC -shared -o "$OUT/libref_mfcc_$v.so" $objs "$TMP/$v/audioprocessing.o" "$TMP/$v/cmsis_tables.o" "$TMP/$v/driver.o" \
      -Wl,--gc-sections -Wl,--version-script="$HERE/drivers/exports.map" -Wl,-z,defs -lm
}
build_mfcc simd "-include arm_math.h -include $HERE/shim/define_dsp.h"
build_mfcc c ""
# B2 of SURVEY 8(d), the "fair CPU" baseline: the reference's plain-C CMSIS path at -O3 for AVX2 hosts (x86-64-v3 rather than
# -march=native: the .so is built in one container and timed on another host).  audioprocessing.c goes to -O2 here: the
# out-of-bounds tail of cmpl_mag_q15's 2048-element call (F6) lands in the neighbouring static buffer either way and is never
# read; tests/test_oracle_cpu.py checks that this build's outputs equal libref_mfcc_c.so's.
build_mfcc c_o3 "" "-O3 -march=x86-64-v3" "-O2 -march=x86-64-v3"

# ---- float MFCC of firmware/src/audio/mfcc.c: mfcc.c + CMSIS f32 real FFT, compiled where they lie
build_mfccf () {
  local d="$TMP/mfccf" objs=""
  mkdir -p "$d"
  local F32_SRCS="TransformFunctions/arm_rfft_fast_f32.c TransformFunctions/arm_rfft_fast_init_f32.c TransformFunctions/arm_cfft_f32.c TransformFunctions/arm_cfft_init_f32.c
                  TransformFunctions/arm_cfft_radix8_f32.c TransformFunctions/arm_bitreversal2.c CommonTables/arm_const_structs.c"
  for s in $F32_SRCS; do
    o="$d/$(basename "$s" .c).o"
    $CC $COMMON -O2 -ffp-contract=off -c "$DSP/Source/$s" -o "$o"; objs="$objs $o"
  done
  $CC $COMMON -O2 -ffp-contract=off -Dmalloc=ref_pad_malloc -U_FORTIFY_SOURCE -D_FORTIFY_SOURCE=0 -I"$FW/audio" -c "$FW/audio/mfcc.c" -o "$d/mfcc.o"
  $CC $COMMON -O2 -c "$TMP/cmsis_tables.c" -o "$d/cmsis_tables.o"
  $CC $COMMON -O2 -I"$FW/audio" -c "$HERE/drivers/ref_mfccf_driver.c" -o "$d/driver.o"
  $CC -shared -o "$OUT/libref_mfcc_f32.so" $objs "$d/mfcc.o" "$d/cmsis_tables.o" "$d/driver.o" \
      -Wl,--gc-sections -Wl,--version-script="$HERE/drivers/exports.map" -Wl,-z,defs -lm
}
build_mfccf

NNOM="$FW/ai/nnom"
NNOM_SRCS="$(ls "$NNOM"/src/core/*.c "$NNOM"/src/layers/*.c "$NNOM"/src/backends/*.c)"
NN_SRCS="ConvolutionFunctions/arm_convolve_HWC_q7_basic_nonsquare.c ConvolutionFunctions/arm_convolve_HWC_q7_fast_nonsquare.c
         ConvolutionFunctions/arm_convolve_HWC_q7_basic.c ConvolutionFunctions/arm_convolve_HWC_q7_fast.c
         ConvolutionFunctions/arm_convolve_HWC_q7_RGB.c
         ConvolutionFunctions/arm_convolve_1x1_HWC_q7_fast_nonsquare.c
         ConvolutionFunctions/arm_depthwise_separable_conv_HWC_q7.c ConvolutionFunctions/arm_depthwise_separable_conv_HWC_q7_nonsquare.c
         ConvolutionFunctions/arm_nn_mat_mult_kernel_q7_q15.c ConvolutionFunctions/arm_nn_mat_mult_kernel_q7_q15_reordered.c
         FullyConnectedFunctions/arm_fully_connected_q7.c FullyConnectedFunctions/arm_fully_connected_q7_opt.c
         SoftmaxFunctions/arm_softmax_q7.c ActivationFunctions/arm_relu_q7.c ActivationFunctions/arm_nn_activations_q7.c
         PoolingFunctions/arm_pool_q7_HWC.c
         NNSupportFunctions/arm_q7_to_q15_no_shift.c NNSupportFunctions/arm_q7_to_q15_reordered_no_shift.c
         NNSupportFunctions/arm_nntables.c NNSupportFunctions/arm_nn_mult_q7.c NNSupportFunctions/arm_nn_add_q7.c"

build_nnom () {  # $1 = variant, $2 = port dir (selects NNOM_USING_CMSIS_NN), $3 = extra flags, $4 = with CMSIS-NN sources (0/1)
  local v="$1" port="$2" extra="$3" with_nn="$4" opt="${5:--O2}" objs=""
  mkdir -p "$TMP/nn_$v"
  local inc="-fPIC -w $opt -ffunction-sections -fdata-sections -I$NNOM -I$port -I$NNOM/inc -D__GNUC_PYTHON__ -D__RESTRICT=__restrict -I$NN/Include -I$DSP/Include"
  local i=0
  for s in $NNOM_SRCS; do
    o="$TMP/nn_$v/n$i.o"; i=$((i+1))
    $CC $inc $extra -c "$s" -o "$o"; objs="$objs $o"
  done
  if [ "$with_nn" = 1 ]; then
    for s in $NN_SRCS; do
      [ -f "$NN/Source/$s" ] || continue
      o="$TMP/nn_$v/c$i.o"; i=$((i+1))
      $CC $inc $extra -c "$NN/Source/$s" -o "$o"; objs="$objs $o"
    done
  fi
  $CC $inc $extra -c "$HERE/drivers/ref_nnom_driver.c" -o "$TMP/nn_$v/driver.o"
  $CC -shared -o "$OUT/libref_nnom_$v.so" $objs "$TMP/nn_$v/driver.o" -Wl,--gc-sections -Wl,--version-script="$HERE/drivers/exports.map" -Wl,-z,defs -lm
}
# local backend: the submodule's port header (NNOM_USING_CMSIS_NN commented out)
build_nnom local "$REF/audio/edison/deploy/nnom/port" "" 0
build_nnom local_o3 "$REF/audio/edison/deploy/nnom/port" "" 0 "-O3 -march=x86-64-v3"   # B2
# CMSIS-NN reference branches: firmware's port header (NNOM_USING_CMSIS_NN on), ARM_MATH_DSP unset
build_nnom cmsis "$NNOM/port" "" 1
# CMSIS-NN SIMD branches == firmware build
build_nnom cmsis_simd "$NNOM/port" "-include arm_math.h -include $HERE/shim/dsp_shim_nn.h" 1
# more real models for the generic executor: upstream NNoM examples inside the reference tree, local backend.
# (mnist-densenet and octave-conv are left out: this NNoM version's concat_build/concat_run size every input like the first
#  one (nnom_concat.c:104, :208) and local_up_sampling_q7_HWC never writes the odd rows (nnom_local.c:383), so their
#  outputs depend on stale buffer contents.)
build_nnom_example () {  # $1 = example directory name, $2 = output library name
  local d="$TMP/nn_example_$1" objs="" i=0
  mkdir -p "$d"
  local port="$REF/audio/edison/deploy/nnom/port"
  local inc="-fPIC -w -O2 -ffunction-sections -fdata-sections -I$NNOM -I$port -I$NNOM/inc -I$REF/audio/edison/deploy/nnom/examples/$1/mcu"
  for s in $NNOM_SRCS; do
    o="$d/n$i.o"; i=$((i+1))
    $CC $inc -c "$s" -o "$o"; objs="$objs $o"
  done
  $CC $inc -c "$HERE/drivers/ref_nnom_example_driver.c" -o "$d/driver.o"
  $CC -shared -o "$OUT/$2" $objs "$d/driver.o" -Wl,--gc-sections -Wl,--version-script="$HERE/drivers/exports.map" -Wl,-z,defs -lm
}
build_nnom_example keyword_spotting libref_nnom_example_local.so
build_nnom_example mnist-simple libref_nnom_ex_mnist_simple.so
build_nnom_example uci-inception libref_nnom_ex_uci_inception.so
ls -la "$OUT"
