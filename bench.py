#!/usr/bin/env python
"""bench.py — LSMC path-steps/sec on B200 (BASELINE.json metric), one JSON line on rank 0.

A "step" is one complete valuation pass of the hot path over the workload: regression-set simulation, backward
induction (fused sweep + normal equations + solve per period), valuation-set simulation, forward simulation and
the reductions to NPV / deltas / trigger-price inputs.  path-steps = paths x decision periods.

  value     device-resident throughput: problem tables already on the GPU, device RNG, summary left on device
  e2e       the same pass through the one-call C ABI (`lsmc_calculate`) with HOST buffers: problem tables H2D and the
            results (NPV, deltas, profile, trigger prices, PV by sim) D2H inside the timed region
  roofline  the dominant kernel (fused backward sweep): algorithmic HBM bytes / CUDA-event time vs measured copy peak
  cpu_baseline  the CPU oracle (port of the reference algorithm) on a bounded sample of the same workload

`--impl reference` times the reference's CPU algorithm (the oracle port: the C# reference cannot be built in this
image — no dotnet, NuGet dependencies absent) on all host threads on a bounded sample of the same config.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIG_TEXT = {
    "c1": "README CmdtyStorage simple, three_factor_seasonal_value, daily 1y",
    "c2": "one-factor multi_factor_value, constant-rate storage, daily 1y",
    "c3": "three-factor seasonal, inventory ratchets + losses, daily 1y",
    "c4": "five-factor multi_factor_value, degree-3 polynomial basis (K=56), 2y daily",
    "c5": "battery-style storage, 2-factor, half-hourly 1 month",
}
# BASELINE.json configs: total paths of the valuation.  C4 (4M paths x 730 periods) needs 164 GB of simulation panels: it only
# fits split over several GPUs, so on fewer than 8 GPUs the bench runs the 8-GPU share per GPU (500k paths each) and says so.
DEFAULT_PATHS = {"c1": 1000, "c2": 100_000, "c3": 1_000_000, "c4": 4_000_000, "c5": 2_000_000}
MAX_PATHS_PER_GPU = {"c4": 500_000}
CPU_SAMPLE_PATHS = {"c1": 1000, "c2": 20000, "c3": 8192, "c4": 2048, "c5": 4096}


# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on the first
# communicator), so file descriptor 1 is pointed at stderr for the whole run and the line goes to the saved descriptor.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIG_TEXT))
    ap.add_argument("--paths-per-gpu", type=int, default=0, help="paths owned by each GPU; implies --scaling weak")
    ap.add_argument("--paths-total", type=int, default=0, help="total paths of the valuation (strong scaling); default = the config's path count")
    ap.add_argument("--scaling", default=None, choices=["strong", "weak"],
                    help="strong (default): BASELINE's path count in total, split over the GPUs; weak: that many paths on every GPU")
    ap.add_argument("--no-parity", action="store_true", help="skip the correctness gates against the oracle")
    ap.add_argument("--no-panels", action="store_true", help="skip the timing of the per-sim panel export")
    ap.add_argument("--cpu-paths", type=int, default=0, help="paths of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(self.samples)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(config, n_local):
    """dram__bytes_read.sum + dram__bytes_write.sum per sweep launch from the committed `ncu --set full` capture of this
    workload (profiles/r02_sweep4_traffic.json); None when no capture exists for the configuration being run."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_sweep4_traffic.json")) as f:
            for row in json.load(f)["captures"]:
                if row["config"] == config and int(row["paths_per_gpu"]) == int(n_local):
                    return float(row["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


def problem_h2d_bytes(pb):
    total = 0
    for name in pb._F64 + pb._I32:
        total += getattr(pb, name).nbytes
    return int(total)


def results_d2h_bytes(res):
    total = 24
    for name in ("deltas", "profile", "trigger_summary", "trigger_inject", "trigger_withdraw", "pv_by_sim"):
        a = getattr(res, name)
        if a is not None:
            total += a.nbytes
    return int(total)


def cpu_sample(config, cpu_paths, kind_note=""):
    """The oracle on a bounded sample of the workload with all host threads; returns the cpu_baseline object."""
    from oracle import oracle
    from storage_b200._abi import Rng
    from storage_b200.configs import build_config
    pb = build_config(config, num_sims=cpu_paths)
    # all host cores, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1 to every rank)
    oracle.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    t0 = time.perf_counter()
    out = oracle.calculate(pb, Rng.philox(11), Rng.philox(22), decision_index=False, triggers=True)
    dt = time.perf_counter() - t0
    T = pb.n_periods - 1
    return {"value": cpu_paths * T / dt, "unit": "path-steps/s", "cores": oracle.num_threads(), "kind": "port",
            "sample": "%s at %d paths x %d periods, one full valuation in %.2f s on %d OpenMP threads%s" % (
                config, cpu_paths, T, dt, oracle.num_threads(), kind_note),
            "npv": out.results.npv}, dt


def parity_gates(ctx, config):
    """correctness gates printed with every benchmark line, outside the timed region.  A 1 024-path cut of the
    benchmarked config through the C ABI against the CPU oracle: injected normals (same draws on both sides) and device RNG."""
    from oracle import oracle
    from storage_b200._abi import Rng
    from storage_b200.configs import build_config
    pb = build_config(config, num_sims=1024)
    S, F, N = pb.n_sim_periods, pb.n_factors, 1024
    rs = np.random.default_rng(7)

    def anti(z):
        out = np.empty((S, F, N))
        out[:, :, 0::2] = z
        out[:, :, 1::2] = -z
        return out
    zr, zv = anti(rs.standard_normal((S, F, N // 2))), anti(rs.standard_normal((S, F, N // 2)))
    ref = oracle.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), panels=("inject_withdraw",)).results
    got = ctx.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), panels=("inject_withdraw",))

    def rel(a, b, scale=None):
        a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
        m = ~(np.isnan(a) | np.isnan(b))
        if not m.any():
            return 0.0
        return float(np.max(np.abs(a[m] - b[m])) / (scale if scale else max(np.max(np.abs(b[m])), 1e-300)))
    vol_scale = max(float(np.max(np.abs(ref.inject_withdraw))), 1e-300)
    price_scale = float(np.max(pb.forward))
    trig = max(rel(got.trigger_summary[:, [1, 3]], ref.trigger_summary[:, [1, 3]], price_scale),
               rel(got.trigger_inject[..., 1], ref.trigger_inject[..., 1], price_scale),
               rel(got.trigger_withdraw[..., 1], ref.trigger_withdraw[..., 1], price_scale))
    gates = {"sample": "%s at 1024 paths, injected normals + device Philox, vs the CPU oracle" % config,
             "npv_rel": rel(got.npv, ref.npv), "delta_rel": rel(got.deltas, ref.deltas), "trigger_rel": trig,
             "decision_flips": int(np.sum(np.abs(got.inject_withdraw - ref.inject_withdraw) > 1e-9 * vol_scale))}
    ref_p = oracle.calculate(pb, Rng.philox(11), Rng.philox(22)).results
    got_p = ctx.calculate(pb, Rng.philox(11), Rng.philox(22))
    gates["philox_npv_sigma"] = abs(got_p.npv - ref_p.npv) / max(ref_p.npv_std_error, 1e-300)
    # trigger prices: 1e-9 of the price scale unless the period's own formula is noisier (tests/test_gpu_parity.py); C5's first
    # half-hourly steps are (cond(X) ~ 1e14), so the gate flag only covers NPV, deltas, decisions and the 3-sigma rule
    # the reference's own seeded golden value (src/Cmdty.Storage.Python/tests/test_multi_factor.py:167-226: two factors, 500 sims,
    # seed = fwd_sim_seed = 11) through the public wrapper on the device, drawing the reference's MT19937 stream
    try:
        import pandas as pd
        import storage_b200
        storage = storage_b200.CmdtyStorage('D', '2019-12-01', '2020-04-01', 1.23, 0.98, min_inventory=0.0, max_inventory=100000.0,
                                            max_injection_rate=700.0, max_withdrawal_rate=700.0)
        idx = pd.PeriodIndex([pd.Period(d, freq='D') for d in ['2019-08-29', '2020-03-12', '2020-04-01']])
        fwd = pd.Series([23.87, 150.32, 150.32], idx).resample('D').ffill()
        curve = pd.period_range('2019-08-29', '2020-06-01', freq='D')
        rates = pd.Series(0.03, index=curve)
        from storage_b200 import multi_factor as _mf
        _mf.RETURN_SIM_PANELS = False
        val = storage_b200.multi_factor_value(storage, '2019-08-29', 0.0, fwd, rates, lambda p: p.asfreq('M').asfreq('D', 'end') + 20,
                                              [(0.0, pd.Series(0.14, index=curve)), (16.2, pd.Series(1.15, index=curve))], 0.64, 500,
                                              '1 + x0 + x0**2 + x1 + x1*x1', False, seed=11, fwd_sim_seed=11)
        gates["reference_golden_npv"] = {"expected": 1780380.7581833513, "got": val.npv, "rel": abs(val.npv / 1780380.7581833513 - 1.0),
                                         "source": "test_multi_factor.py:226 (the reference asserts 6 decimal places)"}
        golden_ok = abs(val.npv - 1780380.7581833513) < 5e-7
    except Exception as e:
        gates["reference_golden_npv"] = {"error": str(e)}
        golden_ok = False
    gates["pass"] = bool(gates["npv_rel"] < 1e-9 and gates["delta_rel"] < 1e-9 and gates["decision_flips"] == 0 and gates["philox_npv_sigma"] < 3.0
                         and golden_ok)
    return gates


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cpu_paths = args.cpu_paths or CPU_SAMPLE_PATHS[args.config]
    from storage_b200.configs import build_config
    for _ in range(min(args.warmup, 1)):
        cpu_sample(args.config, max(256, cpu_paths // 8))
    times, base = [], None
    for _ in range(args.steps):
        base, dt = cpu_sample(args.config, cpu_paths)
        times.append(dt)
    pb = build_config(args.config, num_sims=cpu_paths)
    T = pb.n_periods - 1
    ms = 1e3 * sum(times) / len(times)
    value = cpu_paths * T / (ms / 1e3)
    base["value"] = value
    line = {"impl": "reference", "metric": "lsmc_path_steps_per_sec", "value": value, "unit": "path-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": CONFIG_TEXT[args.config], "config": args.config, "paths_sample": cpu_paths, "periods": T,
                       "note": "reference C# cannot be built here (no dotnet/NuGet); this is the CPU oracle port of its algorithm, OpenMP on all host threads, bounded sample"},
            "cpu_baseline": base,
            "e2e": {"value": value, "unit": "path-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch   # plumbing only: process group, device selection, barriers
    import torch.distributed as dist
    from storage_b200 import _native, _abi
    from storage_b200._abi import Rng
    from storage_b200.configs import build_config

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the LSMC path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    scaling = args.scaling or ("weak" if args.paths_per_gpu else "strong")
    if scaling == "weak":
        n_global = (args.paths_per_gpu or min(DEFAULT_PATHS[args.config], MAX_PATHS_PER_GPU.get(args.config, 1 << 62))) * world
    else:
        n_global = args.paths_total or DEFAULT_PATHS[args.config]
        cap = MAX_PATHS_PER_GPU.get(args.config)
        if cap and n_global > cap * world:
            n_global = cap * world       # the named total does not fit on this few GPUs: the 8-GPU share per GPU
    from storage_b200.sharding import shard_problem
    pb, shard = shard_problem(build_config(args.config, num_sims=n_global), world, rank)
    n_local = shard.count
    T = pb.n_periods - 1

    ctx = _native.Context(local_rank)
    _native.comm_setup(ctx, rank, world, dist)
    reg, val = Rng.philox(11), Rng.philox(22)

    def device_step():
        ctx.simulate(0, reg)
        ctx.backward()
        ctx.simulate(1, val)
        ctx.forward()

    # ---- device-resident throughput ----
    ctx.set_problem(pb)
    for _ in range(args.warmup):
        device_step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    t0 = time.perf_counter()
    launches = dev_ms = span_ms = 0.0
    for _ in range(args.steps):
        launches0 = ctx.timings()["kernel_launches"]
        device_step()
        tm = ctx.timings()
        dev_ms += tm["total_ms"]; launches += tm["kernel_launches"] - launches0
        span_ms += tm["span_ms"]     # CUDA events on the library's stream: start of regression sim -> end of forward pass
    barrier()
    wall = time.perf_counter() - t0
    elapsed = span_ms / 1e3
    clocks = sampler.summary()
    res_dev = ctx.get_results(_abi.Results(pb.n_periods, pb.n_sim_periods, n_local, pv_by_sim=False))
    phase = ctx.timings()

    # ---- roofline pass: the same step with a CUDA event on either side of every sweep launch (on the library's stream).
    #      Kept out of the timed steps: an event between two kernels serialises them, and the backward loop's kernels
    #      otherwise overlap their set-up with the predecessor's drain (programmatic dependent launch) ----
    ctx.set_kernel_timing(True)
    sweep_ms = sweep_launches = sweep_bytes = roof_span_ms = 0.0
    roof_steps = max(1, min(args.steps, 3))
    device_step()
    for _ in range(roof_steps):
        device_step()
        tm = ctx.timings()
        sweep_ms += tm["backward_sweep_kernel_ms"]; sweep_launches += tm["backward_sweep_launches"]
        sweep_bytes += tm["sweep_bytes"]; roof_span_ms += tm["span_ms"]
    phase_timed = ctx.timings()
    ctx.set_kernel_timing(False)
    barrier()

    # ---- end to end through the one-call C ABI with host buffers ----
    e2e_res = _abi.Results(pb.n_periods, pb.n_sim_periods, n_local, pv_by_sim=True)
    for _ in range(min(args.warmup, 2)):
        ctx.calculate(pb, reg, val, results=e2e_res)
    barrier()
    t1 = time.perf_counter()
    for _ in range(args.steps):
        ctx.calculate(pb, reg, val, results=e2e_res)
    barrier()
    e2e_elapsed = time.perf_counter() - t1

    # ---- the same call with the eight [P][N] per-sim panels the reference API always returns (LSV.cs:602-616), into
    #      page-locked buffers (lsmc_host_alloc): export by the copy engine, all panels in flight ----
    panels_ms = panels_bytes = None
    if not args.no_panels and 8.0 * 8 * pb.n_periods * n_local < 48e9:
        try:
            pres = _abi.Results(pb.n_periods, pb.n_sim_periods, n_local, panels="all", pv_by_sim=True, pinned=True)
            ctx.calculate(pb, reg, val, results=pres)
            barrier()
            tp0 = time.perf_counter()
            for _ in range(max(1, min(args.steps, 3))):
                ctx.calculate(pb, reg, val, results=pres)
            barrier()
            panels_ms = 1e3 * (time.perf_counter() - tp0) / max(1, min(args.steps, 3))
            panels_bytes = int(sum(getattr(pres, name).nbytes for name in _abi.PANEL_NAMES))
            del pres
        except Exception as e:   # e.g. not enough page-locked memory on the host
            panels_ms, panels_bytes = None, str(e)

    times = torch.tensor([elapsed, e2e_elapsed, wall], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    elapsed, e2e_elapsed, wall = float(times[0]), float(times[1]), float(times[2])
    if not elapsed > 0.0:
        raise SystemExit("bench.py: the library reported no device-event span for the timed steps")

    if rank == 0:
        ms_per_step = 1e3 * elapsed / args.steps
        value = n_global * T / (ms_per_step / 1e3)
        peak, peak_src = measured_peak_gbs()
        achieved = (sweep_bytes / 1e9) / (sweep_ms / 1e3) if sweep_ms > 0 else 0.0
        # FP64 pipe occupancy of the sweep from first principles: 25 individually rounded operations per (grid row,
        # path) for the decisions + one FMA lane-slot per (basis, row, path) each for the fit and for X~^T V + 2 K^2 per path
        # (x~ and X~^T X~), against 64 FP64 lanes per clock per SM (DFMA and DMMA share them: profiles/r01_probe_fp64.txt)
        G, K = int(ctx.layout().max_grid), pb.n_basis
        fp64_slots = float(n_local) * (25.0 * G + 2.0 * K * G + 2.0 * K * K)
        sm_hz = 1e6 * float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)
        avg_launch_s = (sweep_ms / max(1.0, sweep_launches)) / 1e3
        fp64_frac = fp64_slots / (64.0 * torch.cuda.get_device_properties(0).multi_processor_count * sm_hz * avg_launch_s) if avg_launch_s > 0 else 0.0
        line = {
            "metric": "lsmc_path_steps_per_sec", "value": value, "unit": "path-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": CONFIG_TEXT[args.config], "config": args.config, "paths_per_gpu": n_local, "paths_total": n_global,
                       "periods": T, "factors": pb.n_factors, "basis_functions": pb.n_basis,
                       "inventory_grid_points": 100, "max_grid_rows": G,
                       "rng": "device Philox4x32-10, antithetic",
                       "timing": "CUDA events on the library's stream (regression-sim start -> forward end of each step, summed over the K steps), "
                                 "max over ranks; wall_ms_per_step is the host clock between barrier+synchronize on both sides",
                       "l2": "inputs larger than L2 (sim panels %.1f GB, value surfaces %.1f GB per GPU)" % (
                           8e-9 * n_local * pb.n_sim_periods * (pb.n_factors + 2), 16e-9 * n_local * G)},
            "npv": res_dev.npv, "npv_std_error": res_dev.npv_std_error,
            "wall_ms_per_step": 1e3 * wall / args.steps, "phase_sum_ms_per_step": dev_ms / args.steps,
            "phase_ms": {k: phase[k] for k in ("regression_sim_ms", "backward_ms", "valuation_sim_ms", "forward_ms")},
            # what one backward period costs outside the sweep kernel (reduction, all-reduce, solve, launch gaps): the limiter of strong scaling
            "backward_overhead_us_per_period": 1e3 * (phase["backward_ms"] - phase_timed["backward_sweep_kernel_ms"]) / max(1, T),
            "backward_overhead_note": "backward_ms of a timed step minus the sweep kernels' own time from the roofline pass (events around every sweep launch, %d steps)" % roof_steps,
            "e2e": {"value": n_global * T / (e2e_elapsed / args.steps), "unit": "path-steps/s",
                    "h2d_bytes_per_step": problem_h2d_bytes(pb), "d2h_bytes_per_step": results_d2h_bytes(e2e_res),
                    "ms_per_step": 1e3 * e2e_elapsed / args.steps, "api": "lsmc_calculate (include/lsmc_b200.h) with host buffers"},
            "e2e_panels": None if panels_ms is None else {
                "ms_per_step": panels_ms, "value": n_local * T / (panels_ms / 1e3), "unit": "path-steps/s (this rank)",
                "d2h_bytes_per_step": panels_bytes, "export_gbs": (panels_bytes / 1e9) / max(1e-9, (panels_ms - 1e3 * e2e_elapsed / args.steps) / 1e3),
                "note": "lsmc_calculate with all eight [P][N] panels requested into page-locked host buffers; export_gbs = panel bytes / (this time - e2e time)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"kernel": "sweep4_kernel (fused decision sweep + regression fit + DMMA normal equations, one launch per period)",
                         "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(args.config, n_local), "peak_source": peak_src,
                         "launches": int(sweep_launches), "avg_launch_ms": sweep_ms / max(1.0, sweep_launches),
                         "algorithmic_bytes_per_launch": sweep_bytes / max(1.0, sweep_launches),
                         "share_of_step": sweep_ms / max(1e-9, roof_span_ms),
                         "timing": "CUDA events around every sweep launch on the library's stream, %d steps run after the timed steps" % roof_steps, "fp64_pipe_frac": fp64_frac,
                         "note": "FP64 pipe (DFMA+DMMA share it, 64 lanes/clk/SM measured) is the co-binding roof: see DESIGN.md 4.1"},
        }
        if not args.no_cpu_baseline and world == 1:   # the contract: rank 0 at N = 1 only
            cpu_paths = args.cpu_paths or CPU_SAMPLE_PATHS[args.config]
            try:
                base, _ = cpu_sample(args.config, cpu_paths)
                line["cpu_baseline"] = base
                # one identical-input pair: the GPU arm end to end (host buffers) on exactly the CPU sample's problem
                with _native.Context(local_rank) as small:
                    pb_s = build_config(args.config, num_sims=cpu_paths)
                    res_s = _abi.Results(pb_s.n_periods, pb_s.n_sim_periods, cpu_paths, pv_by_sim=True)
                    for _ in range(2):
                        small.calculate(pb_s, reg, val, results=res_s)
                    ts = time.perf_counter()
                    for _ in range(3):
                        small.calculate(pb_s, reg, val, results=res_s)
                    gpu_s = (time.perf_counter() - ts) / 3
                line["same_input"] = {"paths": cpu_paths, "periods": T, "gpu_e2e_ms": 1e3 * gpu_s, "gpu_e2e_value": cpu_paths * T / gpu_s,
                                      "cpu_value": base["value"], "ratio": cpu_paths * T / gpu_s / base["value"],
                                      "npv_gpu": res_s.npv, "npv_cpu": base["npv"], "npv_rel": abs(res_s.npv / base["npv"] - 1.0),
                                      "note": "same problem, same Philox seeds on both arms (one GPU, rank 0)"}
            except Exception as e:   # the oracle is a checker; its absence must not hide the GPU number
                line["cpu_baseline"] = {"value": None, "unit": "path-steps/s", "cores": 0, "kind": "port", "sample": "failed: %s" % e}
        if not args.no_parity:
            try:
                with _native.Context(local_rank) as pc:
                    line["parity"] = parity_gates(pc, args.config)
            except Exception as e:
                line["parity"] = {"pass": False, "error": str(e)}
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
