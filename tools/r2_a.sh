#!/bin/bash
# round 2, first GPU pass: packed K1 kernel parity + A/B speed
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_fleet.py tests/test_gpu_edges.py -x -q -m gpu -s > $O/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2a_pytest.log
python - > $O/r2a_peak.log 2>&1 <<'PY'
import ctypes as C
from sac_maritime_ast_b200 import _lib
L=_lib.lib()
for name,code in (("f32",1),("f32x2",2),("f64",0)):
    tf,ms=C.c_double(),C.c_double()
    rc=L.sit_measure_fma_peak(0,code,C.byref(tf),C.byref(ms)); print(name,rc,tf.value,ms.value)
PY
python bench.py --steps 20 --warmup 3 --no-extras > $O/r2a_bench_x2.json 2> $O/r2a_bench_x2.err
SIT_NO_X2=1 python bench.py --steps 20 --warmup 3 --no-extras > $O/r2a_bench_scalar.json 2> $O/r2a_bench_scalar.err
tail -3 $O/r2a_pytest.log; cat $O/r2a_peak.log; cat $O/r2a_bench_x2.json; cat $O/r2a_bench_scalar.json
