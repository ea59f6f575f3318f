#!/bin/bash
# Builds, from the reference sources where they lie under /root/reference (nothing
# is copied into the repository):
#   oracle/_ref/libpga_ref.a        the reference library itself, for sm_100a
#   oracle/_ref/ref/test{1,2,3}     the reference's demo programs linked against it
#   oracle/_ref/dropin/test{1,2,3}  the SAME unmodified demo sources linked against
#                                   THIS repository's libpga.a (the drop-in check)
#   oracle/_ref/gen                 the reference's test3 instance generator
#   oracle/_ref/pin/pin{1,2,3}      the reference library + its demo hooks, cuRAND replaced at
#                                   link time by oracle/curand_shim.c, compared lock-step with
#                                   the oracle (oracle/ref_pin.cu)
#   oracle/_ref/time/{ref1,ours1}   oracle/ref_time.cu: pga_run timed on the test/ workload with
#                                   the reference library and with this repository's libpga.a
# Only the architecture flag differs from the reference's Makefiles (-arch=sm_20 is
# rejected by nvcc 12.9; Makefile:2, test/Makefile:2-3).  Outputs are git-ignored but
# travel to the GPU box.  Test infrastructure, not product.
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
OUT="$HERE/_ref"
ARCH="-gencode arch=compute_100a,code=sm_100a"
[ -d "$REF/src" ] || { echo "no reference checkout at $REF; keeping prebuilt oracle/_ref"; exit 0; }
[ -f "$ROOT/libpga.a" ] || make -C "$ROOT" libpga.a
mkdir -p "$OUT/ref" "$OUT/dropin" "$OUT/obj"

# the reference library (Makefile:11-17)
nvcc -dc $ARCH -w -I"$REF/include" "$REF/src/pga.cu" -o "$OUT/obj/pga_ref.o"
rm -f "$OUT/libpga_ref.a"
nvcc $ARCH -lib "$OUT/obj/pga_ref.o" -o "$OUT/libpga_ref.a"

i=1
for t in test test2 test3; do
    # reference program + reference library
    nvcc -dc $ARCH -w "$REF/$t/test.cu" -I"$REF/include" -o "$OUT/obj/${t}_ref.o"
    nvcc $ARCH "$OUT/obj/${t}_ref.o" -o "$OUT/ref/test$i" "$OUT/libpga_ref.a" -lcurand
    # the same source + this repository's header and archive (test/Makefile:2-3 with -L<repo>)
    nvcc -dc $ARCH -w "$REF/$t/test.cu" -I"$ROOT/include" -o "$OUT/obj/${t}_dropin.o"
    nvcc $ARCH "$OUT/obj/${t}_dropin.o" -o "$OUT/dropin/test$i" -L"$ROOT" -lpga -lcurand
    i=$((i + 1))
done
/usr/bin/gcc -O2 -w -o "$OUT/gen" "$REF/test3/gen.c" 2>/dev/null || gcc -O2 -w -o "$OUT/gen" "$REF/test3/gen.c"

# ---- the pin: the reference's own kernels against the oracle, bit for bit ----------------
# oracle/_ref/pin/pinN = oracle/ref_pin.cu (-DPIN_CASE=N) + the reference's unmodified
# test*/test.cu (main renamed, so its device hooks are the reference's own code) +
# libpga_ref.a + oracle/curand_shim.c IN PLACE OF -lcurand + oracle/pga_oracle.c.
HOSTCC=/usr/bin/gcc; [ -x $HOSTCC ] || HOSTCC=gcc
CUDA_INC="$(dirname "$(dirname "$(command -v nvcc)")")/include"
mkdir -p "$OUT/pin"
$HOSTCC -O2 -ffp-contract=off -fopenmp -fPIC -c "$HERE/pga_oracle.c" -o "$OUT/obj/oracle.o"
$HOSTCC -O2 -fopenmp -fPIC -I"$CUDA_INC" -I"$HERE" -c "$HERE/curand_shim.c" -o "$OUT/obj/shim.o"
i=1
for t in test test2 test3; do
    nvcc -dc $ARCH -w -Dmain=reference_demo_main "$REF/$t/test.cu" -I"$REF/include" -o "$OUT/obj/${t}_hooks.o"
    nvcc -dc $ARCH -w -DPIN_CASE=$i "$HERE/ref_pin.cu" -I"$REF/include" -I"$HERE" -o "$OUT/obj/pin$i.o"
    nvcc $ARCH "$OUT/obj/pin$i.o" "$OUT/obj/${t}_hooks.o" "$OUT/obj/shim.o" "$OUT/obj/oracle.o" "$OUT/libpga_ref.a" \
        -Xcompiler -fopenmp -lgomp -lm -o "$OUT/pin/pin$i"
    i=$((i + 1))
done

# ---- timing: pga_run on the reference's test/ workload, reference library vs this one ------
# oracle/ref_time.cu + the reference demo's objective (main renamed); the same source is
# linked once against libpga_ref.a + cuRAND and once against this repository's libpga.a.
mkdir -p "$OUT/time"
nvcc -dc $ARCH -w -Dmain=reference_demo_main "$REF/test/test.cu" -I"$REF/include" -o "$OUT/obj/test_hooks_ref.o"
nvcc -dc $ARCH -w "$HERE/ref_time.cu" -I"$REF/include" -o "$OUT/obj/time_ref.o"
nvcc $ARCH "$OUT/obj/time_ref.o" "$OUT/obj/test_hooks_ref.o" "$OUT/libpga_ref.a" -lcurand -o "$OUT/time/ref1"
nvcc -dc $ARCH -w -Dmain=reference_demo_main "$REF/test/test.cu" -I"$ROOT/include" -o "$OUT/obj/test_hooks_ours.o"
nvcc -dc $ARCH -w "$HERE/ref_time.cu" -I"$ROOT/include" -o "$OUT/obj/time_ours.o"
nvcc $ARCH "$OUT/obj/time_ours.o" "$OUT/obj/test_hooks_ours.o" -L"$ROOT" -lpga -lcurand -o "$OUT/time/ours1"

rm -rf "$OUT/obj"
ls -la "$OUT" "$OUT/ref" "$OUT/dropin" "$OUT/pin" "$OUT/time"
