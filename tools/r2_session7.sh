#!/bin/bash
# f64 sweep A/B (RLB_F64_TRIM 0/1/2) with timing, then ncu --set full of the best and of variant 0
O=gpurun_out; T=r02g; mkdir -p $O
for v in 0 1 2; do
  touch ransaclib_b200/csrc/score_kernels.cu
  make -s -C ransaclib_b200/csrc EXTRA="-DRLB_F64_TRIM=$v" 2>&1 | grep -A2 "sweep_kernelINS_5PoseFEdLi4E" | grep Used
  echo -n "TRIM=$v: "; timeout 200 python bench.py --precision f64 --profile --steps 3 --warmup 2 2>/dev/null | head -c 300; echo
  cp ransaclib_b200/lib/libransaclib_b200.so /tmp/lib_trim$v.so
done 2>&1 | tee $O/${T}_f64_ab.txt
for v in 0 2; do
  cp /tmp/lib_trim$v.so ransaclib_b200/lib/libransaclib_b200.so
  timeout 300 python bench.py --precision f64 --profile --steps 1 --warmup 1 --k-override 16384 > /dev/null 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -c 1 -f -o $O/${T}_f64sweep_trim$v \
     python bench.py --precision f64 --profile --steps 1 --warmup 1 --k-override 16384 > $O/${T}_ncu_trim$v.log 2>&1; echo "ncu trim$v rc=$?"
done
