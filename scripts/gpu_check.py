"""Development check on a B200: CUDA path vs the CPU oracle with injected normals, stage by stage."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from storage_b200 import _native
from storage_b200._abi import Rng
from storage_b200.configs import build_config
from oracle import oracle


def anti_normals(rs, S, F, N):
    z = rs.standard_normal((S, F, N // 2))
    out = np.empty((S, F, N))
    out[:, :, 0::2] = z
    out[:, :, 1::2] = -z
    return out


def check(cfg, N, seed=7):
    print("=" * 30, cfg, "N =", N)
    pb = build_config(cfg, num_sims=N)
    S, F, P, K = pb.n_sim_periods, pb.n_factors, pb.n_periods, pb.n_basis
    rs = np.random.default_rng(seed)
    zr, zv = anti_normals(rs, S, F, N), anti_normals(rs, S, F, N)
    t = time.time()
    o = oracle.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), keep_values=True, keep_factors=True,
                         panels=("spot_regress", "spot_valuation", "inventory", "inject_withdraw", "pv"))
    print("oracle %.1fs npv %.6f back %.6f" % (time.time() - t, o.results.npv, o.results.backward_npv))
    ctx = _native.Context(0)
    # --- stage 1: plan + simulation
    ctx.request_outputs(panels=True, decision_index=True)
    ctx.debug_keep_values(True)
    ctx.set_problem(pb)
    lo, hi = ctx.inventory_space()
    print("inventory space equal:", np.array_equal(lo, o.inventory_space_min), np.array_equal(hi, o.inventory_space_max))
    grids_ok = all(np.array_equal(ctx.grid(p), o.grids[p, :o.grid_count[p]]) for p in range(P))
    print("grids equal:", grids_ok, "max grid", ctx.layout().max_grid)
    ctx.simulate(0, Rng.from_normals(zr))
    sp = ctx.sim_panel(0)
    print("regression spot bit-exact:", np.array_equal(sp, o.results.spot_regress), "max rel", np.max(np.abs(sp / o.results.spot_regress - 1)))
    for f in range(F):
        fp = ctx.sim_panel(0, f)
        print("  factor", f, "bit-exact:", np.array_equal(fp, o.factors_regress[:, f, :]))
    # --- stage 2: backward with oracle betas injected -> value surfaces must match bit for bit
    for p in range(P - 1):
        g = o.grid_count[p + 1]
        if p == 0 and pb.first_period_is_current:
            coeffs = np.zeros((g, K)); coeffs[:, 0] = o.current_continuation[:g]
        else:
            coeffs = o.regression_coeffs[p, :g, :]
        ctx.debug_set_regression_coeffs(p, coeffs)
    ctx.backward()
    bad = 0
    worst = 0.0
    for p in range(P - 1, -1, -1):
        v = ctx.debug_values(p)
        ov = o.values[p, :o.grid_count[p], :]
        if not np.array_equal(v, ov):
            bad += 1
            rel = np.max(np.abs(v - ov) / np.maximum(1.0, np.abs(ov)))
            worst = max(worst, rel)
            if bad <= 3:
                print("  period", p, "values differ: max rel", rel, "n diff", np.sum(v != ov), "of", v.size)
    print("injected-beta value surfaces: periods differing", bad, "of", P, "worst rel", worst)
    ctx.close()
    # --- stage 3: full run with own regression
    ctx = _native.Context(0)
    t = time.time()
    res = ctx.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), panels=("inventory", "inject_withdraw", "pv", "spot_valuation"), decision_index=True)
    print("gpu calculate %.2fs" % (time.time() - t), ctx.timings())
    r, orr = res, o.results
    print("npv gpu %.9f oracle %.9f rel %.3e | backward rel %.3e | se %.4f vs %.4f" % (
        r.npv, orr.npv, abs(r.npv / orr.npv - 1), abs(r.backward_npv / orr.backward_npv - 1), r.npv_std_error, orr.npv_std_error))
    print("valuation spot bit-exact:", np.array_equal(r.spot_valuation, orr.spot_valuation))
    mism = int(np.sum(r.decision_index != orr.decision_index))
    vol_scale = np.max(np.abs(orr.inject_withdraw))
    material = int(np.sum(np.abs(r.inject_withdraw - orr.inject_withdraw) > 1e-9 * vol_scale))
    print("forward decision index mismatches:", mism, "of", r.decision_index.size, "| material (volume differs):", material)
    if mism:
        pp, nn = np.nonzero(r.decision_index != orr.decision_index)
        print("  mismatch periods histogram (first 10):", np.unique(pp, return_counts=True)[0][:10], np.unique(pp, return_counts=True)[1][:10])
        for k in range(min(5, len(pp))):
            p_, n_ = pp[k], nn[k]
            print("   p=%d n=%d gpu idx %d vol %.6f inv %.6f | oracle idx %d vol %.6f inv %.6f" % (p_, n_, r.decision_index[p_, n_], r.inject_withdraw[p_, n_], r.inventory[p_, n_],
                  orr.decision_index[p_, n_], orr.inject_withdraw[p_, n_], orr.inventory[p_, n_]))
        print("  volume panel max abs diff", np.max(np.abs(r.inject_withdraw - orr.inject_withdraw)), "inventory panel max abs diff", np.max(np.abs(r.inventory - orr.inventory)))
    dscale = np.max(np.abs(orr.deltas))
    print("deltas max err / max|delta|: %.3e" % (np.max(np.abs(r.deltas - orr.deltas)) / dscale))
    print("profile max rel err: %.3e" % np.max(np.abs(r.profile - orr.profile) / np.maximum(1e-9, np.max(np.abs(orr.profile), axis=0))))
    m = ~np.isnan(orr.trigger_summary)
    print("trigger nan pattern equal:", np.array_equal(np.isnan(r.trigger_summary), np.isnan(orr.trigger_summary)),
          "max rel err: %.3e" % np.max(np.abs(r.trigger_summary[m] / orr.trigger_summary[m] - 1)))
    mi = ~np.isnan(orr.trigger_inject)
    print("trigger profile max rel err: %.3e" % np.max(np.abs(r.trigger_inject[mi] / orr.trigger_inject[mi] - 1)))
    # betas
    worstb = 0.0
    for p in range(1 if pb.first_period_is_current else 0, P - 1):
        b = ctx.regression_coeffs(p); ob = o.regression_coeffs[p, :b.shape[0], :]
        worstb = max(worstb, np.max(np.abs(b - ob) / np.maximum(1e-300, np.max(np.abs(ob)))))
    print("beta worst rel (to max|beta| per period):", worstb)
    ctx.close()


if __name__ == "__main__":
    cfgs = sys.argv[1:] or ["c2", "c3"]
    for c in cfgs:
        check(c, 2048)
