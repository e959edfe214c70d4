#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
bash tools/r2_i.sh
bash tools/r2_j.sh
