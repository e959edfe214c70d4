#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for i in 1 2; do python tools/prof_run.py env32 10 2>&1 | tail -1 | grep -o "'value': [0-9.]*"; done
python tools/prof_run.py roll 2>&1 | tail -1 | grep -o "'value': [0-9.]*\|'samples': [0-9]*\|'resamples': [0-9]*" | tr '\n' ' '; echo
