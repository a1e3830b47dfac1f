#!/bin/bash
# A/B: split-phase barriers vs one __syncthreads per step (same K0 plan)
mkdir -p gpurun_out
for v in "" "-DSV_LS_NO_OVL"; do
SV_NVCC_EXTRA="$v" python -m spec_viterbi_b200.build --force > gpurun_out/ab_build.log 2>&1
SV_LS_SA_ITERS=6000000 timeout 600 python bench.py --kernel ls --steps 3 --no-cpu > gpurun_out/ab.log 2>&1
echo "variant [$v]"; grep -o '"step_ms": \[[^]]*\]' gpurun_out/ab.log
done
python -m spec_viterbi_b200.build --force > /dev/null 2>&1
