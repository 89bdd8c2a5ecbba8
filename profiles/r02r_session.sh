mkdir -p gpurun_out
for lg in 11 12 13 14; do timeout 300 python tools/trace_case.py msm $lg tables > gpurun_out/r02r_trace_msm_${lg}_tables.txt 2>&1; cat gpurun_out/r02r_trace_msm_${lg}_tables.txt; done
