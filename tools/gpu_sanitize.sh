#!/bin/bash
# compute-sanitizer over a reduced selection of the GPU parity tests (small shapes; every kernel family once):
# memcheck, racecheck (shared-memory hazards: mbarrier rings, TMEM-fed epilogues staging through smem), synccheck,
# initcheck.  Summaries go to gpurun_out/sanitizer_<tool>.txt.
set -u
mkdir -p gpurun_out
SEL='tests/test_gpu_parity.py::test_fftmix_fwd_bwd tests/test_gpu_parity.py::test_dwtmix_fwd_bwd tests/test_gpu_parity.py::test_dwtmix_multilevel_fwd_bwd tests/test_gpu_fused.py::test_norm_act_fwd_bwd tests/test_gpu_fused.py::test_norm_act_pooled_second_output tests/test_gpu_fused.py::test_encoder_layer_fused_matches_torch tests/test_gpu_fused.py::test_encoder_layer_c4_widths_pooled_handover tests/test_gpu_fused.py::test_stacked_layers_hand_the_gradient_over_in_two_parts tests/test_gpu_fused.py::test_patch_embed_kernels_match_reference_module tests/test_gpu_fused.py::test_permut_mix_gather'
KEXPR='not 300-14 and not 700-8 and not 67-7'
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ "$tool" = "memcheck" ] && extra="--leak-check no"
  timeout 2400 compute-sanitizer --tool $tool $extra --print-limit 20 --error-exitcode 0 \
    python -m pytest $SEL -q -x -k "$KEXPR" -p no:cacheprovider > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" gpurun_out/sanitizer_$tool.log | tail -5
  { echo "== compute-sanitizer --tool $tool: $(grep -cE '^========= (Invalid|Race|Error|Uninitialized|Barrier|Hazard)' gpurun_out/sanitizer_$tool.log) report lines"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitizer_$tool.log | tail -3; grep -E "^========= (Invalid|Race|Error|Uninitialized|Barrier|Hazard|Warning)" gpurun_out/sanitizer_$tool.log | sort | uniq -c | sort -rn | head -20; } > gpurun_out/sanitizer_$tool.txt
  head -c 200000 gpurun_out/sanitizer_$tool.log > gpurun_out/sanitizer_$tool.head.log; rm -f gpurun_out/sanitizer_$tool.log
done
