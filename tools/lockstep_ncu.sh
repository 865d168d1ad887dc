M=gcc__cache_requests_type_instruction.sum,gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed,sm__icc_request_hit_rate.pct,sm__icc_requests.sum,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,gpu__time_duration.sum,smsp__inst_executed.sum,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio
for w in 1 4 13; do
  URGYM_WARPS=$w ncu --metrics $M --clock-control none -k regex:env_kernel -s 2 -c 1 --csv python tools/lockstep_probe.py BallBalance 16384 2>/dev/null | grep -E "env_kernel" | awk -F'","' -v w=$w '{print "W=" w, $(NF-2), $(NF)}'
done
