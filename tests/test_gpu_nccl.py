"""B200 x 2 (skipped on a single-GPU box): the data-parallel result over NCCL — the all-reduced encoder-weight
gradients equal the single-GPU gradient of the concatenated batch, for equal shards
and for molecule-weighted uneven shards.  One process per GPU through torch.distributed.run, rendezvous on 127.0.0.1."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("mode", ["equal", "weighted"])
def test_nccl_reduced_gradients_equal_single_gpu_gradient_of_concatenated_batch(mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(HERE, "_nccl_worker.py"), mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    out_dir = os.path.join(os.path.dirname(HERE), "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "nccl_parity_%s.json" % mode), "w") as f:
        f.write(line + "\n")
    assert res["backend"] == "nccl" and res["world"] == 2
    assert res["same_on_all_ranks"]
    if mode == "equal":
        assert res["early_group_launched_in_backward"] is True
    assert res["worst_rel_err"] <= 1e-5, res
