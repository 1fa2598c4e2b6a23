#!/bin/bash
# Round 2, session 2: localize A/B combos, then compute-sanitizer memcheck per kernel family.
O=gpurun_out; T=r02b; mkdir -p $O
ab() {  # name, mode, flags
  touch ransaclib_b200/csrc/localize_kernels.cu
  make -s -C ransaclib_b200/csrc EXTRA="$3" 2>&1 | grep -E "localize(_cta)?_kernelENS" -A2 | grep "Used" | tr '\n' ' '; echo
  for Q in 512 4096; do echo -n "$1 [$3] mode $2: "; RLB_LOC_MODE=$2 timeout 100 python tools/time_localize.py $Q dev 3 2>&1 | tail -1; done
}
( ab base 1 ""
  ab smem+3 1 "-DRLB_LM_SMEM=1 -DRLB_LOC_CTA_MINB=3"
  ab smem+roll 1 "-DRLB_LM_SMEM=1 -DRLB_ROLL_LM=1"
  ab smem+slices 1 "-DRLB_LM_SMEM=1 -DRLB_FULL_SLICES=1"
  ab warp-smem 0 "-DRLB_LM_SMEM=1"
  ab warp-smem-4 0 "-DRLB_LM_SMEM=1 -DRLB_LOC_MINB=4"
) 2>&1 | tee $O/${T}_ab.txt
touch ransaclib_b200/csrc/localize_kernels.cu; make -s -C ransaclib_b200/csrc > /dev/null 2>&1
python tools/sanitize_cases.py all > $O/${T}_sanitize_plain.log 2>&1; echo "plain rc=$?"; tail -3 $O/${T}_sanitize_plain.log
for fam in tc ffma2 f64 line inlier localize localize_warp line_lomsac; do
  timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_cases.py $fam > $O/${T}_memcheck_$fam.log 2>&1
  echo "memcheck $fam rc=$?"; grep -E "ERROR SUMMARY|done" $O/${T}_memcheck_$fam.log | tail -2
done
