"""Worker of tests/test_gpu_multi.py: one process per GPU (or per context), launched with RANK / WORLD_SIZE / MASTER_* set.

Every rank values its contiguous shard of the paths through the C ABI with the all-reduce of the normal-equation partials
between the ranks; rank 0 also values the whole problem on one GPU and compares.  Writes a JSON verdict to argv[1]."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch.distributed as dist
    from storage_b200 import _native
    from storage_b200._abi import Rng
    from storage_b200.configs import build_config
    from storage_b200.sharding import shard_problem

    out_path, config, n_global = sys.argv[1], sys.argv[2], int(sys.argv[3])
    same_device = len(sys.argv) > 4 and sys.argv[4] == "same-device"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    device = 0 if same_device else rank
    pb_global = build_config(config, num_sims=n_global)
    pb, shard = shard_problem(pb_global, world, rank)
    ctx = _native.Context(device)
    _native.comm_setup(ctx, rank, world, dist)
    res = ctx.calculate(pb, Rng.philox(11), Rng.philox(22), panels=("inject_withdraw",), decision_index=True)
    gathered = [None] * world
    dist.all_gather_object(gathered, {"npv": res.npv, "se": res.npv_std_error, "backward_npv": res.backward_npv,
                                      "deltas": res.deltas, "profile": res.profile, "offset": shard.offset,
                                      "decision_index": res.decision_index, "inject_withdraw": res.inject_withdraw,
                                      "trigger_summary": res.trigger_summary})
    ctx.close()
    verdict = {}
    if rank == 0:
        with _native.Context(device) as one:
            ref = one.calculate(pb_global, Rng.philox(11), Rng.philox(22), panels=("inject_withdraw",), decision_index=True)
        dec = np.concatenate([g["decision_index"] for g in sorted(gathered, key=lambda g: g["offset"])], axis=1)
        vol = np.concatenate([g["inject_withdraw"] for g in sorted(gathered, key=lambda g: g["offset"])], axis=1)
        vol_scale = max(float(np.max(np.abs(ref.inject_withdraw))), 1e-300)

        def rel(a, b):
            a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
            m = ~np.isnan(b)
            return float(np.max(np.abs(a[m] - b[m])) / max(np.max(np.abs(b[m])), 1e-300)) if m.any() else 0.0

        verdict = {
            "world": world,
            "ranks_agree": all(g["npv"] == gathered[0]["npv"] and np.array_equal(g["deltas"], gathered[0]["deltas"]) for g in gathered),
            "npv_rel": rel(res.npv, ref.npv), "se_rel": rel(res.npv_std_error, ref.npv_std_error),
            "backward_npv_rel": rel(res.backward_npv, ref.backward_npv),
            "delta_rel": rel(res.deltas, ref.deltas), "profile_rel": rel(res.profile, ref.profile),
            "trigger_rel": rel(res.trigger_summary, ref.trigger_summary),
            # index differences where two decisions carry the same volume are ties broken by the last bit of the coefficients
            # (the exchange sums in rank order, one GPU in CTA order); a material flip changes the volume decided
            "decision_differences": int(np.sum(dec != ref.decision_index)), "decisions": int(dec.size),
            "material_flips": int(np.sum(np.abs(vol - ref.inject_withdraw) > 1e-9 * vol_scale)),
        }
        with open(out_path, "w") as f:
            json.dump(verdict, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
