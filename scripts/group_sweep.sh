#!/bin/bash
# One hc candidate batch (scripts/cand_probe.py) per number of parents of the node,
# with the candidate-group kernel on (default) and off.
for npar in 0 1 2 3 4; do
  for g in 1 0; do
    echo -n "NPAR=$npar BNL_GROUP=$g: "
    NPAR=$npar BNL_GROUP=$g timeout 120 python scripts/cand_probe.py 2>&1 | tail -1
  done
done
