"""Multi-rank host logic on CPU: world_size-2 process groups over `gloo`.

What the N>1 path relies on, checked here without a GPU and with the CPU oracle as the arithmetic:
  * `shard_paths` covers every global path exactly once, on even boundaries (antithetic pairs stay on one rank);
  * a rank that simulates only its shard (sim_offset / n_sims_global) produces bit-for-bit the slice of the
    single-process simulation — the Philox counter is the GLOBAL path index;
  * the per-period regression can be assembled from per-rank normal-equation partials with one SUM all-reduce
    (what liblsmc_b200.so does over NCCL), giving the same coefficients on every rank as the single-process fit;
  * `bench.py --impl reference` under a 2-rank launch prints exactly one line (rank 0) and the other rank exits 0.
"""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_paths_partition():
    from storage_b200.sharding import shard_paths
    for n_global in (0, 1, 2, 7, 8, 1000, 1001, 1_000_000):
        for world in (1, 2, 3, 4, 8):
            shards = [shard_paths(n_global, world, r) for r in range(world)]
            assert shards[0].offset == 0
            assert sum(s.count for s in shards) == n_global
            for a, b in zip(shards, shards[1:]):
                assert a.offset + a.count == b.offset
            for s in shards:
                assert s.count == 0 or s.offset % 2 == 0
                if s.offset + s.count < n_global:
                    assert s.count % 2 == 0
            counts = [s.count for s in shards]
            assert max(counts) - min(counts) <= 3   # one pair, plus the odd tail path
    with pytest.raises(ValueError):
        shard_paths(10, 2, 2)


def _worker(rank, world, port, n_global, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from oracle import oracle
    from storage_b200._abi import Rng
    from storage_b200.configs import build_config
    from storage_b200.sharding import shard_problem

    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        oracle.set_num_threads(2)
        pb_global = build_config("c3", num_sims=n_global)
        pb, sh = shard_problem(pb_global, world, rank)
        assert pb.n_sims == sh.count and pb.n_sims_global == n_global and pb.sim_offset == sh.offset
        rng = Rng.philox(1234)
        spot, fac = oracle.simulate(pb, rng)                      # [S][n_local], [S][F][n_local]
        S, F, K = pb.n_sim_periods, pb.n_factors, pb.n_basis

        # (1) shard simulation == slice of the global simulation, bit for bit
        counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(counts, torch.tensor([sh.count]))
        width = int(max(c.item() for c in counts))
        pad = np.zeros((S, F + 1, width))
        pad[:, 0, :sh.count] = spot
        pad[:, 1:, :sh.count] = fac
        gathered = [torch.zeros(S, F + 1, width, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(pad))

        # (2) normal equations from per-rank partials + one SUM all-reduce
        s_reg = S // 2
        X = oracle.design_matrix(pb, spot[s_reg], fac[s_reg])
        y = np.maximum(spot[S - 1] - spot[s_reg], 0.0)          # any path-wise regressand does
        part = torch.from_numpy(np.concatenate([(X.T @ X).ravel(), X.T @ y]))
        dist.all_reduce(part, op=dist.ReduceOp.SUM)
        A = part[:K * K].numpy().reshape(K, K)
        b = part[K * K:].numpy()
        d = 1.0 / np.sqrt(np.diag(A))
        beta = d * np.linalg.solve(A * d[:, None] * d[None, :], b * d)
        betas = [torch.zeros(K, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(betas, torch.from_numpy(beta))
        for other in betas:                                       # every rank holds the same coefficients
            assert np.array_equal(other.numpy(), betas[0].numpy())

        if rank == 0:
            spot_g, fac_g = oracle.simulate(pb_global, rng)
            got = np.concatenate([g.numpy()[:, :, :int(c.item())] for g, c in zip(gathered, counts)], axis=2)
            ok_sim = np.array_equal(got[:, 0, :], spot_g) and np.array_equal(got[:, 1:, :], fac_g)
            Xg = oracle.design_matrix(pb_global, spot_g[s_reg], fac_g[s_reg])
            yg = np.maximum(spot_g[S - 1] - spot_g[s_reg], 0.0)
            beta_g, _ = oracle.regress(Xg, yg)
            fit_err = float(np.max(np.abs(Xg @ beta - Xg @ beta_g)) / max(1e-300, np.max(np.abs(Xg @ beta_g))))
            q.put({"ok_sim": bool(ok_sim), "fit_err": fit_err})
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_ranks_gloo_sharded_simulation_and_normal_equations(oracle_lib):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 1002, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
    for p in procs:
        assert p.exitcode == 0, "rank exited with %r" % p.exitcode
    out = q.get()
    assert out["ok_sim"], "sharded simulation differs from the single-process simulation"
    assert out["fit_err"] < 1e-9, out


@pytest.mark.timeout(600)
def test_reference_arm_two_rank_launch_prints_one_line(oracle_lib):
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
           "--warmup", "0", "--config", "c1", "--cpu-paths", "256"]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=500, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["n_gpus"] == 2
    assert line["metric"] == "lsmc_path_steps_per_sec" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
