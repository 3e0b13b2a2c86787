#!/bin/bash
# Tuning sweep helper: builds variants of libsctl_b200 (Stokes3D_FxU only, for speed) with different compile-time knobs
# into sctl_b200/lib/variants/. Run scripts/sweep_variants.sh on the GPU box afterwards.
set -e
cd "$(dirname "$0")/.."
build() { # name flags...
  local name=$1; shift
  make -s -j8 -C sctl_b200/csrc VARIANT=$name EXTRA_NVFLAGS="-DSCTL_B200_ONLY_KERNEL=3 $*" && echo "built $name: $*" && \
  grep -A2 "KStokesFxUEdLi[0-9]*ELi2ELi" build/csrc_$name/eval_f64.ptxas.log | grep -E "registers|spill" | tr '\n' ' '; echo
}
for v in "$@"; do
  case $v in
    base) build base ;;
    r2) build r2 -DSCTL_B200_RBIG=2 -DSCTL_B200_MINB_BIG=8 ;;
    r2t256) build r2t256 -DSCTL_B200_RBIG=2 -DSCTL_B200_MINB_BIG=4 -DSCTL_B200_EVAL_THREADS=256 ;;
    r3) build r3 -DSCTL_B200_RBIG=3 -DSCTL_B200_MINB_BIG=5 ;;
    r6) build r6 -DSCTL_B200_RBIG=6 -DSCTL_B200_MINB_BIG=2 ;;
    r8) build r8 -DSCTL_B200_RBIG=8 -DSCTL_B200_MINB_BIG=2 ;;
    r4m3) build r4m3 -DSCTL_B200_MINB_BIG=3 ;;
    t256) build t256 -DSCTL_B200_EVAL_THREADS=256 -DSCTL_B200_MINB_BIG=2 ;;
    t64) build t64 -DSCTL_B200_EVAL_THREADS=64 -DSCTL_B200_MINB_BIG=8 ;;
    u1) build u1 -DSCTL_B200_UNROLL=1 ;;
    u4) build u4 -DSCTL_B200_UNROLL=4 ;;
    tile256) build tile256 -DSCTL_B200_EVAL_TILE=256 ;;
    repair:*) # repair:MODE  (SCTL_B200_REPAIR_MODE sweep of eval_impl.cuh)
      build repair${v#repair:} -DSCTL_B200_REPAIR_MODE=${v#repair:} ;;
    custom:*) # custom:name:R:MINB:THREADS:UNROLL[:TILE[:STAGES]]
      IFS=: read -r _ n R M T U TL ST <<< "$v"
      build $n -DSCTL_B200_RBIG=$R -DSCTL_B200_MINB_BIG=$M -DSCTL_B200_EVAL_THREADS=$T -DSCTL_B200_UNROLL=$U -DSCTL_B200_EVAL_TILE=${TL:-128} -DSCTL_B200_EVAL_STAGES=${ST:-4} ;;
    *) echo "unknown variant $v"; exit 1 ;;
  esac
done
