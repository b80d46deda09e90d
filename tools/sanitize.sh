#!/usr/bin/env bash
# compute-sanitizer over the hand-rolled pipelines (run on the GPU box; logs -> gpurun_out/sanitizer/, summaries are
# copied to profiles/ by hand).  memcheck on every case, racecheck (shared-memory hazards) on the mbarrier / TMEM
# kernels and the Cholesky panels, initcheck on the BA workspace users.
set -u
OUT=gpurun_out/sanitizer
mkdir -p "$OUT"
CS=/usr/local/cuda/bin/compute-sanitizer
run() { # tool case
  timeout 900 $CS --tool "$1" --print-limit 20 python tools/sanitize_cases.py "$2" > "$OUT/$1_$2.log" 2>&1
  echo "$1 $2 rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$OUT/$1_$2.log" | tail -1)"
}
for c in corr_build ba_frontend ba_chains chol_chains edges lookup geom; do run memcheck $c; done
for c in corr_build ba_frontend chol_chains edges; do run racecheck $c; done
for c in ba_frontend edges; do run initcheck $c; done
