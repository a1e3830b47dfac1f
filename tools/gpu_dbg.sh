#!/bin/bash
mkdir -p gpurun_out
SV_NVCC_EXTRA=-DSV_LS_DEBUG python -m spec_viterbi_b200.build --force > gpurun_out/dbg_build.log 2>&1
SV_LS_SA_ITERS=2000000 timeout 600 python tools/dbg_ls.py > gpurun_out/dbg_ls.log 2>&1
tail -6 gpurun_out/dbg_ls.log
