"""Multi-GPU parity through the C ABI: a valuation whose paths are sharded over several ranks must give the
one-GPU result — NPV / deltas / profile to 1e-12 relative (only the summation order of the exchanged sums differs)
and the same decisions (no path decides another volume; indices may differ between decisions of equal volume).

  * one process, several contexts (lsmc_comm_init_local + lsmc_calculate_multi): the exchange goes through peer memory
    only; two or three contexts on the SAME GPU exercise every line of it on a one-GPU box;
  * one process per GPU (lsmc_comm_init: NCCL + cudaIpc-mapped exchange buffers): needs two B200s, skipped on a one-GPU
    box (`gpurun --gpus 2` runs it), once over peer memory and once with LSMC_COMM=nccl (the NCCL fallback)."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import _cuda_device_count
from storage_b200 import _abi, _native
from storage_b200._abi import Rng
from storage_b200.configs import build_config
from oracle import oracle

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    m = ~np.isnan(b)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    return float(np.max(np.abs(a[m] - b[m])) / max(np.max(np.abs(b[m])), 1e-300)) if m.any() else 0.0


def anti_normals(seed, S, F, N):
    z = np.random.default_rng(seed).standard_normal((S, F, N // 2))
    out = np.empty((S, F, N))
    out[:, :, 0::2], out[:, :, 1::2] = z, -z
    return out


@pytest.mark.parametrize("config,paths,ranks", [("c3", 6000, 2), ("c2", 9001, 3), ("c5", 4000, 2)])
def test_one_process_group_equals_one_context_with_device_rng(config, paths, ranks):
    pb = build_config(config, num_sims=paths)
    with _native.Context(0) as one:
        ref = one.calculate(pb, Rng.philox(11), Rng.philox(22), panels=("inventory", "pv"), decision_index=True)
    devices = list(range(ranks)) if _cuda_device_count() >= ranks else [0] * ranks
    with _native.Group(devices) as group:
        got = group.calculate(pb, Rng.philox(11), Rng.philox(22), panels=("inventory", "pv"), decision_index=True)
        again = group.calculate(pb, Rng.philox(11), Rng.philox(22))
    assert rel(got.npv, ref.npv) < 1e-12 and rel(got.backward_npv, ref.backward_npv) < 1e-12
    assert rel(got.deltas, ref.deltas) < 1e-12 and rel(got.profile, ref.profile) < 1e-12
    assert rel(got.npv_std_error, ref.npv_std_error) < 1e-9
    # every rank wrote its block of columns; the coefficients differ in the last bits from the one-context run (other
    # summation tree), so decisions may differ where two decisions have the same volume — never in the volume decided
    differ = got.decision_index != ref.decision_index
    assert differ.mean() < 1e-2
    assert rel(got.inventory, ref.inventory) < 1e-12 and rel(got.pv, ref.pv) < 1e-9
    assert rel(got.pv_by_sim, ref.pv_by_sim) < 1e-12
    assert rel(got.trigger_summary, ref.trigger_summary) < (1e-6 if config == "c5" else 1e-9)   # C5: cond(X) ~ 1e14 at the first steps (test_gpu_parity.py)
    assert again.npv == got.npv and np.array_equal(again.deltas, got.deltas)   # exchanges sum in rank order: reproducible


def test_one_process_group_matches_the_oracle_with_injected_normals():
    """The whole problem with injected normals through lsmc_calculate_multi (the normals are sliced per rank): NPV / deltas
    within 1e-9 of the CPU oracle and the same decisions, like the one-context parity test."""
    pb = build_config("c3", num_sims=1200)
    S, F, N = pb.n_sim_periods, pb.n_factors, 1200
    zr, zv = anti_normals(3, S, F, N), anti_normals(4, S, F, N)
    ref = oracle.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), panels=("inject_withdraw",)).results
    with _native.Group([0, 0]) as group:
        got = group.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), panels=("inject_withdraw",))
    assert rel(got.npv, ref.npv) < 1e-9 and rel(got.deltas, ref.deltas) < 1e-9 and rel(got.profile, ref.profile) < 1e-9
    scale = np.max(np.abs(ref.inject_withdraw))
    assert int(np.sum(np.abs(got.inject_withdraw - ref.inject_withdraw) > 1e-9 * scale)) == 0


def test_group_cancellation_and_reuse():
    pb = build_config("c2", num_sims=4000)
    calls = []
    with _native.Group([0, 0]) as group:
        with pytest.raises(_native.LsmcError) as e:
            group.calculate(pb, Rng.philox(1), Rng.philox(2), on_progress=lambda v: (calls.append(v), len(calls) < 5)[1])
        assert e.value.status in (_abi.LSMC_ECANCELLED, _abi.LSMC_ECUDA)
    with _native.Group([0, 0]) as group:     # a cancelled group's exchange counters are out of step: use a fresh one
        assert np.isfinite(group.calculate(pb, Rng.philox(1), Rng.philox(2)).npv)


def run_ranks(world, config, paths, tmp_path, extra=(), env_extra=None):
    out = tmp_path / "verdict.json"
    port = free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), **(env_extra or {}))
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
@pytest.mark.parametrize("comm", ["peer", "nccl"])
@pytest.mark.parametrize("config,paths", [("c3", 20_000), ("c2", 30_001)])
def test_two_ranks_equal_one_gpu(config, paths, comm, tmp_path):
    v = run_ranks(2, config, paths, tmp_path, env_extra={"LSMC_COMM": comm})
    assert v["ranks_agree"]
    assert v["npv_rel"] < 1e-12 and v["backward_npv_rel"] < 1e-12
    assert v["delta_rel"] < 1e-12 and v["profile_rel"] < 1e-12
    assert v["material_flips"] == 0 and v["decision_differences"] < 0.01 * v["decisions"]
    assert v["se_rel"] < 1e-9


def test_sharded_valuation_on_the_reference_stream_keeps_the_golden_npv():
    """LSMC_RNG_MT19937 under path sharding: one serial generator defines the draws of ALL paths, every rank generates the run
    and keeps its own columns (and with _CONTINUE walks on to the end of the set).  Three ranks, 500 paths: the reference's
    seeded golden NPV (test_multi_factor.py:226) must come out as on one GPU."""
    from test_reference_goldens import two_factor_problem, GOLDEN_NPV_2F
    pb = two_factor_problem()
    devices = list(range(3)) if _cuda_device_count() >= 3 else [0] * 3
    with _native.Group(devices) as group:
        same_seed = group.calculate(pb, Rng.mt19937(11), Rng.mt19937(11))
        cont = group.calculate(pb, Rng.mt19937(11), Rng.mt19937_continue())
    assert abs(same_seed.npv - GOLDEN_NPV_2F) < 5e-7          # the reference's own 6-decimal assertion
    with _native.Context(0) as one:
        ref = one.calculate(pb, Rng.mt19937(11), Rng.mt19937_continue())
    assert rel(cont.npv, ref.npv) < 1e-12 and cont.npv != same_seed.npv
