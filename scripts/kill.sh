#!/bin/bash
# Stop the ranks started by scripts/run_exp.sh: kills exactly the PIDs recorded in ./run_exp.pids
# (the reference's kill.sh uses `pkill -f trainer.py`, which can hit unrelated processes).
set -u
[ -f run_exp.pids ] || { echo "no run_exp.pids here"; exit 0; }
while read -r pid; do
  [ -n "$pid" ] && kill "$pid" 2>/dev/null && echo "stopped $pid"
done < run_exp.pids
rm -f run_exp.pids
