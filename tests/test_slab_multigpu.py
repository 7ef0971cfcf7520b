"""Real multi-GPU run of the slab decomposition: `bench.py --gpus 2` under torchrun, one process per GPU, halos
exchanged through peer memory (the product transport) and through NCCL point-to-point. bench.py steps a small fluid
through the same transport before its timed loop and compares it bit for bit with a single-device run
(`slab_parity`); it exits non-zero on a mismatch. Skipped on boxes with fewer than two GPUs."""
import json
import os
import socket
import subprocess
import sys

import pytest

import common

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("transport", ["p2p", "nccl"])
def test_two_gpu_slab_run_is_bit_exact(transport):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, SPH_SLAB_TRANSPORT=transport)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(common.ROOT, "bench.py"), "--gpus", "2", "--steps", "2",
           "--warmup", "3", "--workload", "fluid_256k", "--no-configs"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=common.ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    assert lines[0]["n_gpus"] == 2
    assert lines[0]["slab_parity"].startswith("bit-exact"), lines[0]["slab_parity"]
