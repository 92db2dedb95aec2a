#!/usr/bin/env bash
# One-GPU profiling pass (run through gpurun): CUDA-event timings without a profiler first, then one
# `ncu --set full` capture per kernel of interest.  Reports land in gpurun_out/ (scratch); summarise them
# with tools/ncu_summarise.py into profiles/.
set -u
mkdir -p gpurun_out
python tools/prof_kernels.py all > gpurun_out/plain_all.log 2>&1 || { tail -5 gpurun_out/plain_all.log; exit 1; }
tail -1 gpurun_out/plain_all.log
capture() { # <report name> <kernel regex> <prof_kernels mode> <skip>
    timeout 170 ncu --set full --clock-control none --import-source on -k "regex:$2" -s "$4" -c 1 -f \
        -o "gpurun_out/prof_$1" python tools/prof_kernels.py "$3" > "gpurun_out/ncu_$1.log" 2>&1 || tail -3 "gpurun_out/ncu_$1.log"
}
capture crc32c_kernel crc32c_kernel hash 8
capture allreduce_two_shot_quant_kernel allreduce_two_shot_quant_kernel allreduce 2
capture quantize_kernel quantize_kernel quant 2
capture outer_step_two_shot_kernel outer_step_two_shot_kernel outer 2
ls -la gpurun_out/*.ncu-rep
