#!/bin/bash
# A/B of the small-kernel PDL launches and the own gradient-sum kernel on one GPU's gradient pass
for cfg in "1 1" "0 1" "1 0" "0 0"; do
  set -- $cfg
  for g in 8 1; do
    echo -n "PDL_SMALL=$1 OWN_ADD=$2 : "
    GARFIELD_PDL_SMALL=$1 GARFIELD_OWN_ADD=$2 python scripts/pass_time.py --groups $g --batch 25 2>&1 | tail -1
  done
done
