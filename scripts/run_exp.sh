#!/bin/bash
# Launch an experiment of one application on the hosts listed in ./servers and ./workers (one host per
# line), one rank per line, servers first — the deployment model of the reference's run_exp.sh
# (applications/Aggregathor/run_exp.sh:18-60). With no host files every rank runs on this machine.
#
#   scripts/run_exp.sh Aggregathor --dataset cifar10 --model resnet50 --loss cross-entropy --batch 25 \
#        --gar krum --fw 1 --opt_args '{"lr":"0.2","momentum":"0.9","weight_decay":"0.0005"}'
#
# PIDs of the started ranks are written to ./run_exp.pids (scripts/kill.sh stops exactly those).
set -euo pipefail
app=${1:?usage: run_exp.sh <Aggregathor|ByzSGD|LEARN> [trainer flags...]}
shift
root=$(cd "$(dirname "$0")/.." && pwd)
trainer="$root/applications/$app/trainer.py"
servers=(); workers=()
[ -f servers ] && mapfile -t servers < servers
[ -f workers ] && mapfile -t workers < workers
[ ${#servers[@]} -eq 0 ] && servers=(127.0.0.1)
[ ${#workers[@]} -eq 0 ] && workers=(127.0.0.1 127.0.0.1 127.0.0.1 127.0.0.1 127.0.0.1 127.0.0.1)
master=${servers[0]}
: > run_exp.pids
rank=0
launch() {  # host rank
  local cmd="python3 $trainer --transport rpc --master $master --rank $2 --num_ps ${#servers[@]} --num_workers ${#workers[@]} ${extra[*]}"
  if [ "$1" = "127.0.0.1" ] || [ "$1" = "localhost" ]; then
    bash -c "$cmd" > "run_exp.rank$2.log" 2>&1 &
  else
    ssh "$1" "$cmd" > "run_exp.rank$2.log" 2>&1 &
  fi
  echo $! >> run_exp.pids
  echo "rank $2 on $1 (pid $!)"
}
extra=("$@")
if [ "$app" = "LEARN" ]; then
  nodes=("${servers[@]}" "${workers[@]}")
  for host in "${nodes[@]}"; do
    cmd="python3 $trainer --transport rpc --master $master --rank $rank --num_nodes ${#nodes[@]} ${extra[*]}"
    bash -c "$cmd" > "run_exp.rank$rank.log" 2>&1 &
    echo $! >> run_exp.pids; rank=$((rank+1))
  done
else
  for host in "${servers[@]}"; do launch "$host" $rank; rank=$((rank+1)); done
  for host in "${workers[@]}"; do launch "$host" $rank; rank=$((rank+1)); done
fi
wait
