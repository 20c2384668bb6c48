#!/usr/bin/env bash
# ONE compute-sanitizer tool per call over tools/sanitize_all.py (every kernel of the library at small ragged sizes,
# strided / unaligned inputs, a replaying CUDA graph next to eager launches).  The profiling recipe of this pool allows
# one sanitizer tool per gpurun call, so the four tools are four calls:
#     tools/run_sanitizers.sh memcheck|racecheck|synccheck|initcheck [outdir]
# racecheck covers the shared-memory hazards of the mbarrier / bulk-copy rings (pe_stream.cuh, pe_kabsch.cu,
# pe_addl1.cu), memcheck global/shared out-of-bounds and misaligned accesses, synccheck the barriers, initcheck reads of
# uninitialised global memory.  The log (or the tool's own refusal message) lands in <outdir>/r2_sanitizer_<tool>.log.
tool=${1:?tool name}
out=${2:-gpurun_out}
mkdir -p "$out"
export PE_SANITIZE_N=${PE_SANITIZE_N:-70001} PE_SANITIZE_REPLAYS=${PE_SANITIZE_REPLAYS:-5}
extra=""
[ "$tool" = racecheck ] && extra="--racecheck-report all"
timeout ${PE_SANITIZE_TIMEOUT:-900} compute-sanitizer --tool "$tool" $extra --error-exitcode 3 \
    --log-file "$out/r2_sanitizer_$tool.log" python tools/sanitize_all.py > "$out/r2_sanitizer_$tool.stdout" 2>&1
rc=$?
echo "$tool: exit $rc" | tee "$out/r2_sanitizer_${tool}_summary.txt"
tail -n 4 "$out/r2_sanitizer_$tool.log" 2>/dev/null | tee -a "$out/r2_sanitizer_${tool}_summary.txt"
tail -n 2 "$out/r2_sanitizer_$tool.stdout" | tee -a "$out/r2_sanitizer_${tool}_summary.txt"
exit 0
