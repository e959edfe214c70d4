#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for p in 1 0; do for g in 1 2 3 4 5; do
  echo -n "prio=$p groups=$g "; SIT_ROLLOUT_PRIO=$p SIT_ROLLOUT_GROUPS=$g python tools/prof_run.py roll 2>&1 | grep -o "'value': [0-9.]*\|'ms_rollout_only': [0-9.]*\|'ms_per_iteration': [0-9.]*\|'samples': [0-9]*" | tr '\n' ' '; echo
done; done
