#!/bin/bash
# re-capture the round's ncu evidence after a kernel change (development aid): step-level per-kernel metrics + the NTT kernel's full set
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
LSB_DUAL_STREAM=0 timeout 600 ncu --metrics $M --clock-control none -k 'regex:k_ntt|k_bex|k_ks_inner|k_tensor|k_rescale_lift|k_basis' -c 60 --csv --log-file gpurun_out/r2_step_kernels.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-api --no-secondary > gpurun_out/r2_step_ncu.log 2>&1
tail -2 gpurun_out/r2_step_ncu.log | cut -c1-300
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ntt -s 4 -c 2 -o /tmp/ntt14 -f python tools/prof_ntt.py 14 8192 3 > gpurun_out/r2_ntt_ncu.log 2>&1
python tools/ncu_summary.py /tmp/ntt14.ncu-rep > gpurun_out/r2_ntt14_cluster_ncu.txt 2>&1
tail -5 gpurun_out/r2_ntt14_cluster_ncu.txt | cut -c1-200
