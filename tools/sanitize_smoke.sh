#!/bin/bash
# compute-sanitizer passes over __graft_entry__.smoke() (one whole training iteration incl. the two copy streams,
# the pinned-flag hand-off to the host Adam thread and the DMA gradient flush) -- SURVEY.md section 5 asks for them.
# NOTE: the GPU pool of this build refuses compute-sanitizer ("closed on this pool", profiles/r02_sanitizer_refused.txt);
# the script is kept for boxes where it is allowed.
#   tools/sanitize_smoke.sh [outdir]      (run on the GPU box; writes <outdir>/sanitizer_{memcheck,racecheck,synccheck}.txt)
out=${1:-gpurun_out}
mkdir -p "$out"
cd "$(dirname "$0")/.."
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 1 \
    python -c "import __graft_entry__ as g; g.smoke()" > "$out/sanitizer_$tool.txt" 2>&1
  echo "exit code: $?" >> "$out/sanitizer_$tool.txt"
  tail -4 "$out/sanitizer_$tool.txt"
done
