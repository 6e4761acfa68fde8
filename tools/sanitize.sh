#!/bin/bash
# compute-sanitizer over small-size runs of every kernel family ("race detection" row):
#   memcheck  -- out-of-bounds / misaligned accesses, leaks of device memory at exit
#   racecheck -- shared-memory data races (the round-1 K2 race was found only by accident; this is the net for it)
#   synccheck -- divergent barriers / invalid mbarrier use
# Run on the GPU box:  bash tools/sanitize.sh [outdir]   (about 6-10 minutes; writes one log per tool and a summary)
# The targets are the GPU tests themselves at their smallest sizes (-k selections below), so what is checked is what
# is tested. ASLAM_LARGE_NO_PDL=1: compute-sanitizer serialises kernels; programmatic dependent launch adds nothing.
set -u
OUT=${1:-gpurun_out/sanitize}
mkdir -p "$OUT"
export ASLAM_LARGE_NO_PDL=1
SEL='cfg1_single_ekf or gated or ragged_observations or split_launch or checkpoint or small_map_q19 or ragged_sizes or row_sharded or fused_peer_exchange_one_gpu or gives_up or laser_kernel_matches or synthetic_scan or reproducible_through_growth or ukf_stationary or dense_gemm_lower_only or dense_cholesky or large_ukf_step_matches'
for tool in memcheck racecheck synccheck; do
  echo "== $tool"
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 86 --print-limit 20 \
      python -m pytest tests -m gpu -x -q -k "$SEL" -p no:cacheprovider > "$OUT/$tool.log" 2>&1
  echo "exit=$?" >> "$OUT/$tool.log"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|exit=" "$OUT/$tool.log" | tail -4
done
{
  echo "# compute-sanitizer summary ($(date -u +%FT%TZ))"
  for tool in memcheck racecheck synccheck; do
    echo "## $tool"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|deselected|exit=" "$OUT/$tool.log" | tail -5
  done
} > "$OUT/summary.md"
cat "$OUT/summary.md"
