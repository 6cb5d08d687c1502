# one compute-sanitizer tool per gpurun call:  bash scripts/gpu_sanitize.sh memcheck|racecheck
tool=$1
for team in 1 2; do
  echo "== $tool, team $team"
  SBAYES_B200_TEAM=$team timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_step.py ${2:-120} 2>&1 | grep -v "^$" | tail -12
done
