#!/usr/bin/env bash
# compute-sanitizer over tools/sanitize_all.py (every kernel of the library at small ragged sizes, strided / unaligned
# inputs, a replaying CUDA graph next to eager launches).  One log per tool under $1 (default gpurun_out/); the tool's
# own refusal message is a log too.  racecheck covers the shared-memory hazards of the mbarrier / bulk-copy rings
# (pe_stream.cuh, pe_kabsch.cu, pe_addl1.cu), memcheck global/shared out-of-bounds and misaligned accesses, synccheck the
# barriers, initcheck reads of uninitialised global memory.
out=${1:-gpurun_out}
mkdir -p "$out"
export PE_SANITIZE_N=${PE_SANITIZE_N:-70001} PE_SANITIZE_REPLAYS=${PE_SANITIZE_REPLAYS:-5}
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ "$tool" = racecheck ] && extra="--racecheck-report all"
  timeout ${PE_SANITIZE_TIMEOUT:-900} compute-sanitizer --tool $tool $extra --error-exitcode 3 \
      --log-file "$out/r2_sanitizer_$tool.log" python tools/sanitize_all.py > "$out/r2_sanitizer_$tool.stdout" 2>&1
  echo "$tool: exit $?" | tee -a "$out/r2_sanitizer_summary.txt"
  tail -n 3 "$out/r2_sanitizer_$tool.log" 2>/dev/null | tee -a "$out/r2_sanitizer_summary.txt"
done
