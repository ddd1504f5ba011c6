"""The data-parallel gradient path on real GPUs ("1 vs 2/4/8 GPU grads equal (same global batch) within
bf16-all-reduce tolerance; bucket-order independence; overlap on/off identical"; reference: train_lm.py:151-159,
195-210, 256-271).  Needs >= 2 GPUs (gpurun --gpus 2/4); on the driver's 1-GPU test box these skip, and the same
check runs inside bench.py at every N > 1 (the "ddp_check" field of its JSON line)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("use_doc", [False, True])
def test_n_rank_gradients_equal_whole_batch_gradients(world, use_doc):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, PCV_TEST_DOC="1" if use_doc else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "ddp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    lines = [json.loads(ln[len("DDPCHECK "):]) for ln in r.stdout.splitlines() if ln.startswith("DDPCHECK ")]
    assert [ln["layers_per_bucket"] for ln in lines] == [1, 4, 8]
    for res in lines:
        print(res)
        assert res["world"] == world
        assert res["ranks_hold_identical_gradients"]
        assert res["reducer_exact_rel_l2_bf16_wire"] <= res["reducer_bf16_wire_tolerance"] <= 6.1e-3, res  # the reducer alone
        assert res["reducer_exact_rel_l2_fp32_wire"] <= 1e-6, res
        tol = max(4e-3, 2.0 * res["rel_l2_rerun_noise"])                # whole step: no tighter than its own noise
        assert res["rel_l2_bf16_wire"] <= tol and res["rel_l2_fp32_wire"] <= tol, res
        assert res["rel_l2_cuda_graph_vs_eager_bf16"] <= tol and res["cuda_graph_error"] is None, res
        assert res["overlap_off_equal"], res
        assert res["loss_abs_diff"] <= 1e-4, res
        assert res["ok"], res
