#!/bin/bash
# compute-sanitizer over the kernels with hand-written synchronisation protocols (SURVEY.md section 5 "race detection"):
#   rnn_persistent_kernel  -- relaxed-load sentinel polling of the L2 exchange tensor (cooperative launch)
#   rnn_cluster_kernel     -- st.async into peer shared memory, step-parity tiles, mbarrier complete_tx
#   conv_gemm_tc(_pair)    -- TMA/mbarrier ring, tcgen05 commits (multicast in the pair form), cross-CTA tmem_empty arrives
#   lr_scan / lr_gather    -- warp scans, atomics on meta
# Run on the B200 box:  bash scratch/sanitize.sh  > gpurun_out/sanitize.txt 2>&1   (copy the summary to profiles/)
# CUDA graphs are off (FTB_CUDA_GRAPHS=0) so every launch is seen as a launch; sizes are the tests' own (small).
set -u
SEL='test_recurrence or test_gemm_cta_pair or test_length_regulator_bit_exact or test_conv_split or test_program_set_seq_len or test_highway'
for tool in memcheck racecheck synccheck; do
  echo "=== compute-sanitizer --tool $tool"
  FTB_CUDA_GRAPHS=0 timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
      python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "$SEL" 2>&1 | grep -v "^$" | tail -25
  echo "=== $tool exit code: ${PIPESTATUS[0]}"
done
