#!/usr/bin/env bash
# Run every self-test case in its own process under a timeout (a hung kernel only loses that case).
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
BIN="$ROOT/build/igemm_selftest"
mkdir -p "$ROOT/gpurun_out"
LOG="$ROOT/gpurun_out/selftest.log"
: > "$LOG"
nvidia-smi --query-gpu=name,driver_version --format=csv >> "$LOG" 2>&1
n=$("$BIN" count)
fails=0
for ((i = 0; i < n; i++)); do
  timeout 60 "$BIN" "$i" >> "$LOG" 2>&1
  rc=$?
  if [[ $rc -ne 0 ]]; then
    echo "case $i exit code $rc" >> "$LOG"
    fails=$((fails + 1))
  fi
done
echo "selftest: $fails of $n cases failed" | tee -a "$LOG"
grep -E "FAIL|exit code|CUDA error" "$LOG" | head -40
exit 0
