#!/bin/bash
# A/B of k_recognize_reg build variants on the box (needs nvcc there: the image has it)
cd speechrecognition_b200/csrc
for v in "-DSGMM_REC_LEAN=0 -DSGMM_REC_MINB=4" "-DSGMM_REC_LEAN=1 -DSGMM_REC_MINB=3" "-DSGMM_REC_LEAN=0 -DSGMM_REC_MINB=3" "-DSGMM_REC_LEAN=1 -DSGMM_REC_MINB=2"; do
  touch k_recognize.cu
  make EXTRA_NVFLAGS="$v" > /dev/null 2>&1 || { echo "build failed $v"; continue; }
  echo "== $v"; grep -E "Used" build/k_recognize.ptxas.log | sed -n 3p
  (cd ../..; timeout 200 python tools/rec_ab.py all279 all1399_148 bench 2>&1 | tail -3)
done
touch k_recognize.cu; make > /dev/null 2>&1
