#!/bin/bash
# round-2 ncu --set full captures of the third step of profiles/run_step.py (each only after the same command ran clean)
mkdir -p gpurun_out
cap() {  # name, skip, count, regex, run_step args...
  local name=$1 skip=$2 count=$3 regex=$4; shift 4
  python profiles/run_step.py "$@" > gpurun_out/${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$regex" -s $skip -c $count -f -o gpurun_out/$name \
      python profiles/run_step.py "$@" > gpurun_out/${name}_ncu.log 2>&1
  tail -1 gpurun_out/${name}_plain.log
}
CLAIM='fused_score_scatter|fused_adagrad_claim|parse_count_kernel|parse_fill_kernel'
SORTED='fused_score_scatter|sorted_update|sorted_finish|scatter_kernel|count_kernel|expand_kernel'
cap r2_claim_k16 8 4 "$CLAIM" 3 uniform claim 16 24 65536
cap r2_sorted_k16 20 10 "$SORTED" 3 uniform sorted 16 24 65536
cap r2_claim_k16_zipf 8 4 "$CLAIM" 3 zipf claim 16 24 65536
cap r2_claim_k64 8 4 "$CLAIM" 3 uniform claim 64 24 32768
cap r2_claim_k32 8 4 "$CLAIM" 3 uniform claim 32 24 65536
