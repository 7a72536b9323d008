#!/bin/bash
# usage (on the GPU box, via gpurun): tools/checked.sh <tag>  — the stand-in for the compute-sanitizer pass (closed on this pool): the B200PT_CHECKED build
# (build it first, here: tools/build_variant.sh checked -DB200PT_CHECKED) bounds-checks every queue / pool / stack / node / triangle / mesh / texel index in
# the kernels; run over tools/sanitize_target.py (every scheduling mode, results compared with the oracle and across modes) and over the GPU parity tests.
tag=$1
out=gpurun_out/checked_$tag.txt
{ echo "# B200PT_LIB=build/variants/libb200pt_checked.so (nvcc -DB200PT_CHECKED): violations counted per .cu file, printed by every b200pt_destroy"
  echo "## python tools/sanitize_target.py"
  B200PT_LIB=build/variants/libb200pt_checked.so python tools/sanitize_target.py 2>&1 | grep -E "B200PT_CHECKED|sanitize target|smoke:|Error|error" | sort | uniq -c
  echo "## python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -m gpu -q"
  B200PT_LIB=build/variants/libb200pt_checked.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -m gpu -q -s 2>&1 | grep -E "B200PT_CHECKED|passed|failed" | sed 's/^[.F]*//' | sort | uniq -c
} > $out
cat $out
