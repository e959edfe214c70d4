#!/bin/bash
# the no-alloc entry point's test alone in a fresh process, then the full round script (suite with that test last, bench, ncu)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > $O/r2z_alone.log 2>&1; echo "alone rc=$?" >> $O/r2z_alone.log
tail -5 $O/r2z_alone.log
bash tools/r2_g.sh
