#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_sac.py tests/test_gpu_train_loop.py -q -m gpu > $O/r2v_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2v_pytest.log
grep "passed\|failed\|rc=\|FAILED" $O/r2v_pytest.log
timeout 300 python - <<'PY'
import torch, sys
from sac_maritime_ast_b200 import sac
torch.manual_seed(0)
rows = torch.zeros(262144, 24); rows[:, :23] = torch.randn(262144, 23); rows[:, 10].clamp_(-1, 1); rows[:, 22] = 1
mem = sac.ReplayMemory(262144, device="cuda:0"); mem.extend(rows.cuda())
for b, tc in ((64, True), (256, True), (1024, True), (4096, True), (4096, False)):
    a = sac.SAC(10, None, None, device="cuda:0", hidden_size=256, batch_size=b, seed=1, tensor_cores=tc, obs_scale=None)
    a.update(mem, 50); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ls = a.update(mem, 300); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"batch {b} tc {tc}: {300 / ms * 1e3:.0f} updates/s, {ms / 300 * 1e3:.1f} us/update, losses finite {bool(torch.isfinite(ls).all())}")
PY
