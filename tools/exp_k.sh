#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for k in ${KS:-32 48 64}; do for g in ${GS:-2 3 4}; do
  echo -n "K=$k groups=$g "; SIT_ROLLOUT_GROUPS=$g python tools/prof_run.py roll K=$k iters=$((1920/k)) warm_iters=$((960/k)) 2>&1 | grep -o "'value': [0-9.]*\|'env_steps_per_s': [0-9.]*\|'ms_per_iteration': [0-9.]*\|'samples': [0-9]*\|'lane_utilisation': [0-9.]*" | tr '\n' ' '; echo
done; done
