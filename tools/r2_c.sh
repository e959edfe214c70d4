#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_fleet.py tests/test_gpu_edges.py -q -m gpu -s > $O/r2c_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2c_pytest.log
for v in b128_m3 b128_m4 b128_m5 b64_m8 b64_m10 b256_m2; do
  SIT_LIBSHIPSIM=$PWD/sac_maritime_ast_b200/exp/lib_$v.so python bench.py --steps 20 --warmup 3 --no-extras > $O/r2c_$v.json 2> $O/r2c_$v.err
  echo "$v $(python -c "import json;d=json.load(open('$O/r2c_$v.json'));print(d['value'], d['ms_per_step'])")"
done
tail -4 $O/r2c_pytest.log
