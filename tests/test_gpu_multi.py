"""Multi-GPU parity through the C ABI: a valuation whose paths are sharded over two ranks must give the
one-GPU result — NPV / deltas / profile to 1e-12 relative (only the summation order of the all-reduced sums differs)
and bit-equal decision indices.  Needs two B200s: skipped on a one-GPU box (`gpurun --gpus 2` runs it)."""
import json
import os
import socket
import subprocess
import sys

import pytest

from conftest import _cuda_device_count

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_ranks(world, config, paths, tmp_path, extra=()):
    out = tmp_path / "verdict.json"
    port = free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "two_rank_worker.py"), str(out), config, str(paths), *extra],
                                      env=env))
    try:
        codes = [p.wait(timeout=600) for p in procs]
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    assert codes == [0] * world, codes
    with open(out) as f:
        return json.load(f)


@pytest.mark.skipif(_cuda_device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("config,paths", [("c3", 20_000), ("c2", 30_001)])
def test_two_ranks_equal_one_gpu(config, paths, tmp_path):
    v = run_ranks(2, config, paths, tmp_path)
    assert v["ranks_agree"]
    assert v["npv_rel"] < 1e-12 and v["backward_npv_rel"] < 1e-12
    assert v["delta_rel"] < 1e-12 and v["profile_rel"] < 1e-12
    assert v["decision_differences"] == 0
    assert v["se_rel"] < 1e-9
