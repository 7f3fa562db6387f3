#!/bin/bash
# compute-sanitizer over the kernel library: one tool per invocation, as the
# profiling recipe asks (several tools in one GPU lease have left a B200 unusable).
#
#   gpurun --timeout 900 -- 'bash tools/sanitize.sh memcheck'     # or racecheck | synccheck | initcheck
#
# Targets: tools/smoke_all_kernels.py (one launch of every kernel family on odd shapes: partial
# tiles, short row groups, item switches of the persistent attention kernels), the multi-rank
# protocol emulated on one GPU (tools/gpu_check_comm.py without the dead-peer case, whose trap is
# the point of that case) and one tiny fused MLP engine step. Each target first runs WITHOUT the tool and
# must exit 0. Logs: gpurun_out/sanitize_<tool>.log; copy the summary lines into profiles/.
set -u
TOOL=${1:-memcheck}
OUT=gpurun_out/sanitize_${TOOL}.log
mkdir -p gpurun_out
: > "$OUT"
export B200DIST_NO_PDL=1            # the sanitizer serialises kernels anyway; keep launches plain
export B200DIST_COMM_TIMEOUT_MS=600000   # instrumented kernels are slow: do not mistake that for a dead peer
run() {
  local name=$1; shift
  echo "== $name (plain run)" | tee -a "$OUT"
  if ! timeout 300 "$@" >> "$OUT" 2>&1; then echo "plain run of $name failed: not sanitized" | tee -a "$OUT"; return 1; fi
  echo "== $name under compute-sanitizer --tool $TOOL" | tee -a "$OUT"
  timeout 1500 compute-sanitizer --tool "$TOOL" --print-limit 20 --error-exitcode 77 "$@" >> "$OUT" 2>&1
  local rc=$?
  echo "== $name: rc=$rc" | tee -a "$OUT"
  return $rc
}
rc=0
run smoke_all_kernels python tools/smoke_all_kernels.py || rc=1
run comm_emulated python tools/gpu_check_comm.py --no-dead-peer || rc=1
run mlp_engine_step python -c "import __graft_entry__ as g; g.smoke()" || rc=1
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|== .*rc=" "$OUT"
exit $rc
