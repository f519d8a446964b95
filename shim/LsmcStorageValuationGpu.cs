// LsmcStorageValuationGpu.cs — the reference-side binding of liblsmc_b200.so: the body of
// LsmcStorageValuation.Calculate<T> (src/Cmdty.Storage/LsmcValuation/LsmcStorageValuation.cs:55-617) replaced by one call
// into the C ABI of include/lsmc_b200.h.  Add this file to src/Cmdty.Storage/LsmcValuation/ together with the five
// one-line internal accessors listed in shim/README.md; nothing else in the reference changes, and the Python wrapper
// (cmdty_storage/multi_factor.py:386-455) keeps working through pythonnet unchanged.
//
// DELIVERED AS SOURCE: there is no dotnet / mono in the build image, so this file has never been compiled.  Its
// flattening is mirrored statement for statement by storage_b200/flatten.py (flatten_problem), which IS tested against
// the same C ABI; the struct layouts are checked by tests/test_abi.py against sizeof/offsetof of the C header.
//
//   section 1  P/Invoke declarations                       include/lsmc_b200.h
//   section 2  GpuSimulationSpec<T>                         what LVP.cs:150-191 closes over, kept visible
//   section 3  Flatten: LsmcValuationParameters<T> -> lsmc_problem     flatten.py:42-172
//   section 4  status -> exception                          lsmc_status in the header
//   section 5  results -> LsmcStorageValuationResults<T>    LSV.cs:593-616
using System;
using System.Collections.Generic;
using System.Linq;
using System.Runtime.InteropServices;
using Cmdty.Core.Simulation;
using Cmdty.TimePeriodValueTypes;
using Cmdty.TimeSeries;

namespace Cmdty.Storage
{
    // ------------------------------------------------------------------------------------------------ 1. P/Invoke
    internal static unsafe class LsmcNative
    {
        private const string Lib = "lsmc_b200";   // liblsmc_b200.so / lsmc_b200.dll on the native probing path

        internal const int Ok = 0, EInval = 1, EConstraint = 2, ECancelled = 3, ECuda = 4, EUnsupported = 5;
        internal const int ConstraintConstant = 0, ConstraintPiecewiseLinear = 1, ConstraintStep = 2, ConstraintPolynomial = 3;
        internal const int RngPhilox = 0;
        internal const int RngMt19937 = 3, RngMt19937Continue = 4;   // the reference's seeded stream (include/lsmc_b200.h LSMC_RNG_MT19937)
        internal const int NumTriggerVolumes = 10;

        [StructLayout(LayoutKind.Sequential)]
        internal struct Problem   // lsmc_problem
        {
            public int struct_size, n_periods, n_factors, n_basis;
            public long n_sims, n_sims_global, sim_offset;
            public int extra_decisions, first_period_is_current, discount_deltas, must_be_empty_at_end;
            public double numerical_tolerance, start_inventory, grid_spacing;
            public fixed double terminal_coef[4];
            public double* sim_time, mean_reversion, factor_vol, factor_corr, forward;
            public int* spot_power, factor_power;
            public double* min_inventory, max_inventory;
            public int* constraint_id;
            public double* inject_cost, inject_cost_df, withdraw_cost, withdraw_cost_df, consumed_inject, consumed_withdraw,
                           inventory_loss, inventory_cost, inventory_cost_df, settle_df;
            public int n_constraints, n_pillars;
            public int* constraint_kind, constraint_first, constraint_count;
            public double* pillar_inventory, pillar_min_rate, pillar_max_rate;
        }

        [StructLayout(LayoutKind.Sequential)]
        internal struct Rng { public int mode, antithetic; public ulong seed; public double* normals, spot, factors; }

        [StructLayout(LayoutKind.Sequential)]
        internal struct Results   // lsmc_results
        {
            public int struct_size, reserved;
            public double npv, npv_std_error, backward_npv;
            public double* deltas, profile, trigger_summary, trigger_inject, trigger_withdraw, pv_by_sim, spot_regress, spot_valuation,
                           inventory, inject_withdraw, consumed, loss, net_volume, pv;
            public sbyte* decision_index;
        }

        [UnmanagedFunctionPointer(CallingConvention.Cdecl)]
        internal delegate int ProgressCallback(IntPtr user, double progress);

        [DllImport(Lib, CallingConvention = CallingConvention.Cdecl)] internal static extern int lsmc_create(int device, out IntPtr ctx);
        [DllImport(Lib, CallingConvention = CallingConvention.Cdecl)] internal static extern void lsmc_destroy(IntPtr ctx);
        [DllImport(Lib, CallingConvention = CallingConvention.Cdecl)] internal static extern IntPtr lsmc_last_error(IntPtr ctx);
        [DllImport(Lib, CallingConvention = CallingConvention.Cdecl)]
        internal static extern int lsmc_calculate(IntPtr ctx, Problem* problem, Rng* regressionRng, Rng* valuationRng,
                                                  ProgressCallback progress, IntPtr user, Results* results);
        [DllImport(Lib, CallingConvention = CallingConvention.Cdecl)]
        internal static extern int lsmc_comm_init_local(IntPtr* contexts, int n);
        [DllImport(Lib, CallingConvention = CallingConvention.Cdecl)]
        internal static extern int lsmc_calculate_multi(IntPtr* contexts, int n, Problem* problem, Rng* regressionRng, Rng* valuationRng,
                                                        ProgressCallback progress, IntPtr user, Results* results);
    }

    // ------------------------------------------------------------------------------------------------ 2. simulation spec
    /// <summary>What SimulateWithMultiFactorModelAndMersenneTwister (LsmcValuationParameters.cs:181-191) closes over and the
    /// native library needs in the open: the factors as (mean reversion, volatility curve by delivery period) — the arguments
    /// of Cmdty.Core's Factor&lt;T&gt; (multi_factor.py:101-105) — their correlation matrix, the number of simulations and the
    /// seeds.  Set by the new builder method LsmcValuationParameters&lt;T&gt;.Builder.SimulateOnGpu (shim/README.md).</summary>
    public sealed class GpuSimulationSpec<T> where T : ITimePeriod<T>
    {
        public IReadOnlyList<(double MeanReversion, TimeSeries<T, double> Volatility)> Factors { get; }
        public double[,] FactorCorrelations { get; }
        public int NumSims { get; }
        public ulong RegressionSeed { get; }
        public ulong ValuationSeed { get; }
        /// <summary>True (the default): seeded runs draw the reference's own stream — MathNet MT19937 + Normal.Samples in
        /// Cmdty.Core's draw order, generated on the host by the native library — so that the same seeds give the same numbers as
        /// LsmcStorageValuation.Calculate on the CPU (its seeded golden NPVs come out to 1e-15).  False: the device Philox stream,
        /// for runs with millions of paths where a serial generator would dominate.</summary>
        public bool ReferenceCompatibleStream { get; }
        public bool ValuationContinuesRegressionStream { get; }   // no valuation seed was given (LVP.cs:186-187)
        public IReadOnlyList<int> Devices { get; }   // CUDA device indices; more than one: paths are split over them

        public GpuSimulationSpec(IReadOnlyList<(double, TimeSeries<T, double>)> factors, double[,] factorCorrelations, int numSims,
                                 ulong regressionSeed, ulong? valuationSeed = null, IReadOnlyList<int> devices = null,
                                 bool referenceCompatibleStream = true)
        {
            if (factors == null || factors.Count == 0) throw new ArgumentException("factors cannot be empty.", nameof(factors));
            if (factorCorrelations.GetLength(0) != factors.Count || factorCorrelations.GetLength(1) != factors.Count)
                throw new ArgumentException("factors and factor correlations are of inconsistent sizes.");
            if (numSims < 1) throw new ArgumentException("Number of simulations must be positive.", nameof(numSims));
            Factors = factors; FactorCorrelations = factorCorrelations; NumSims = numSims;
            RegressionSeed = regressionSeed;
            // no valuation seed: the reference continues the regression generator's stream (LVP.cs:187-189); with a
            // counter-based generator that is a distinct key derived from the same seed
            ValuationSeed = valuationSeed ?? unchecked(regressionSeed + 0x9E3779B97F4A7C15UL);
            ValuationContinuesRegressionStream = valuationSeed == null;
            ReferenceCompatibleStream = referenceCompatibleStream;
            Devices = devices ?? new[] { 0 };
        }
    }

    public static class LsmcStorageValuationGpu
    {
        // ------------------------------------------------------------------------------------------------ 3. flattening
        /// <summary>Replaces LsmcStorageValuation.Calculate (LSV.cs:55).  Lines 57-84 of the original stay as they are.</summary>
        public static unsafe LsmcStorageValuationResults<T> Calculate<T>(LsmcValuationParameters<T> lsmcParams, GpuSimulationSpec<T> sim,
                                                                      IReadOnlyList<PowerMonomialBuilder> basis)
            where T : ITimePeriod<T>
        {
            // --- argument checks and trivial exits, verbatim from LSV.cs:57-84 ---
            if (lsmcParams.Inventory < 0) throw new ArgumentException("Inventory cannot be negative.", nameof(lsmcParams));
            ICmdtyStorage<T> storage = lsmcParams.Storage;
            if (lsmcParams.CurrentPeriod.CompareTo(storage.EndPeriod) > 0) return LsmcStorageValuationResults<T>.CreateExpiredResults();
            if (lsmcParams.CurrentPeriod.Equals(storage.EndPeriod))
            {
                if (storage.MustBeEmptyAtEnd)
                {
                    if (lsmcParams.Inventory > 0)
                        throw new InventoryConstraintsCannotBeFulfilledException("Storage must be empty at end, but inventory is greater than zero.");
                    return LsmcStorageValuationResults<T>.CreateExpiredResults();
                }
                double terminalNpv = storage.TerminalStorageNpv(lsmcParams.ForwardCurve[storage.EndPeriod], lsmcParams.Inventory);
                return LsmcStorageValuationResults<T>.CreateEndPeriodResults(terminalNpv);
            }

            // --- periods (SH.cs:45; LSV.cs:89,154): startActive .. EndPeriod inclusive ---
            T current = lsmcParams.CurrentPeriod;
            T startActive = storage.StartPeriod.CompareTo(current) > 0 ? storage.StartPeriod : current;
            T[] periods = startActive.EnumerateTo(storage.EndPeriod).ToArray();
            int P = periods.Length, F = sim.Factors.Count, K = basis.Count;
            bool firstIsCurrent = startActive.Equals(current);
            TimeSeries<T, double> forwardCurve = lsmcParams.ForwardCurve;
            if (forwardCurve.Start.CompareTo(startActive) > 0)        // LSV.cs:91-95
                throw new ArgumentException("Forward curve starts too late. Must start on or before the period " + startActive + ".");
            if (forwardCurve.End.CompareTo(storage.EndPeriod) < 0)
                throw new ArgumentException("Forward curve does not extend until storage end period.");

            // --- discount factors, memoised as at LSV.cs:131-143: to the settlement date of the commodity (LSV.cs:222-223)
            //     and to the payment dates of the per-unit costs ---
            Day today = current.First<Day>();
            var dfCache = new Dictionary<Day, double>();
            double Df(Day date)
            {
                if (!dfCache.TryGetValue(date, out double df)) { df = lsmcParams.DiscountFactors(today, date); dfCache[date] = df; }
                return df;
            }

            var simTime = new double[P]; var forward = new double[P]; var vol = new double[P * F]; var meanReversion = new double[F];
            var corr = new double[F * F];
            var minInv = new double[P]; var maxInv = new double[P]; var constraintId = new int[P];
            var injCost = new double[P]; var injDf = Ones(P); var wdCost = new double[P]; var wdDf = Ones(P);
            var consInj = new double[P]; var consWd = new double[P]; var loss = new double[P];
            var invCost = new double[P]; var invCostDf = Ones(P); var settleDf = Ones(P);
            for (int f = 0; f < F; f++)
            {
                meanReversion[f] = sim.Factors[f].MeanReversion;
                for (int g = 0; g < F; g++) corr[f * F + g] = sim.FactorCorrelations[f, g];
            }

            // constraint tables, de-duplicated by reference: the builder hands out one constraint object per ratchet table
            var constraintIndex = new Dictionary<IInjectWithdrawConstraint, int>();
            var kind = new List<int>(); var first = new List<int>(); var count = new List<int>();
            var pillarInv = new List<double>(); var pillarMin = new List<double>(); var pillarMax = new List<double>();

            for (int p = 0; p < P; p++)
            {
                T period = periods[p];
                simTime[p] = TimeFunctions.Act365(current.Start, period.Start);                 // LVP.cs:172-176
                forward[p] = forwardCurve[period];
                minInv[p] = storage.MinInventory(period);
                maxInv[p] = storage.MaxInventory(period);                                       // 0 at the end when must-be-empty
                if (p >= (firstIsCurrent ? 1 : 0))                                               // vols of simulated delivery periods only
                    for (int f = 0; f < F; f++) vol[p * F + f] = sim.Factors[f].Volatility[period];
                if (p == P - 1) { constraintId[p] = constraintId[p - 1]; break; }               // no decision in the end period
                settleDf[p] = Df(lsmcParams.SettleDateRule(period));

                // per-unit costs and consumption: the builder's closures are linear in the volume (CmdtyStorage.cs:327-497);
                // probe at two volumes and refuse anything else (SURVEY 7.3-3)
                double probeInventory = 0.5 * (minInv[p] + maxInv[p]);
                (injCost[p], injDf[p]) = UnitCost(v => storage.InjectionCost(period, probeInventory, v), Df, "injection cost");
                (wdCost[p], wdDf[p]) = UnitCost(v => storage.WithdrawalCost(period, probeInventory, v), Df, "withdrawal cost");
                consInj[p] = UnitRate(v => storage.CmdtyVolumeConsumedOnInject(period, probeInventory, v), "commodity consumed on injection");
                consWd[p] = UnitRate(v => storage.CmdtyVolumeConsumedOnWithdraw(period, probeInventory, v), "commodity consumed on withdrawal");
                loss[p] = storage.CmdtyInventoryPercentLoss(period);
                (invCost[p], invCostDf[p]) = UnitCost(v => storage.CmdtyInventoryCost(period, v), Df, "inventory cost");

                IInjectWithdrawConstraint constraint = ((CmdtyStorage<T>)storage).ConstraintFor(period);   // internal accessor, shim/README.md
                if (!constraintIndex.TryGetValue(constraint, out int id))
                {
                    id = kind.Count; constraintIndex[constraint] = id; first.Add(pillarInv.Count);
                    switch (constraint)
                    {
                        case ConstantInjectWithdrawConstraint c:                                 // Constant…cs:45
                            kind.Add(LsmcNative.ConstraintConstant); count.Add(1);
                            pillarInv.Add(0.0); pillarMin.Add(c.Range.MinInjectWithdrawRate); pillarMax.Add(c.Range.MaxInjectWithdrawRate);
                            break;
                        case PiecewiseLinearInjectWithdrawConstraint c:                          // PiecewiseLinear…cs:67
                            kind.Add(LsmcNative.ConstraintPiecewiseLinear); count.Add(c.Pillars.Count); AddPillars(c.Pillars);
                            break;
                        case StepInjectWithdrawConstraint c:                                     // Step…cs:72
                            kind.Add(LsmcNative.ConstraintStep); count.Add(c.Pillars.Count); AddPillars(c.Pillars);
                            break;
                        case PolynomialInjectWithdrawConstraint c:                               // Polynomial…cs:80: MathNet's own coefficients
                            kind.Add(LsmcNative.ConstraintPolynomial); count.Add(c.Inventories.Length);
                            pillarInv.AddRange(c.Inventories);
                            pillarMin.AddRange(Padded(c.MinCoefficients, c.Inventories.Length));
                            pillarMax.AddRange(Padded(c.MaxCoefficients, c.Inventories.Length));
                            break;
                        default:
                            throw new NotSupportedException("Inject/withdraw constraint of type " + constraint.GetType().Name +
                                                            " has no device descriptor (LSMC_EUNSUPPORTED).");
                    }
                }
                constraintId[p] = id;
            }
            void AddPillars(IReadOnlyList<InjectWithdrawRangeByInventory> rows)
            {
                foreach (InjectWithdrawRangeByInventory row in rows.OrderBy(r => r.Inventory))
                {
                    pillarInv.Add(row.Inventory);
                    pillarMin.Add(row.InjectWithdrawRange.MinInjectWithdrawRate);
                    pillarMax.Add(row.InjectWithdrawRange.MaxInjectWithdrawRate);
                }
            }

            // basis: the exponents the PowerMonomialBuilder delegates hide (BasisFunctionsBuilder.cs:116-130)
            var spotPower = new int[K]; var factorPower = new int[K * F];
            for (int k = 0; k < K; k++)
            {
                spotPower[k] = basis[k].SpotPower;
                foreach (KeyValuePair<int, int> pair in basis[k].MarkovFactorPowers)
                {
                    if (pair.Key >= F) throw new ArgumentException("Basis function refers to factor x" + pair.Key + " but the model has " + F + " factors.");
                    factorPower[k * F + pair.Key] = pair.Value;
                }
            }

            // terminal value: the device descriptor is c0 + c1 price + c2 inventory + c3 price inventory; verify by probing
            double[] terminal = storage.MustBeEmptyAtEnd ? new double[4] : BilinearTerminalValue(storage.TerminalStorageNpv);

            if (!(lsmcParams.GridCalc is FixedSpacingStateSpaceGridCalc gridCalc))
                throw new NotSupportedException("Only FixedSpacingStateSpaceGridCalc has a device descriptor.");

            // --- call ---
            int[] kindA = kind.ToArray(), firstA = first.ToArray(), countA = count.ToArray();
            double[] pInv = pillarInv.ToArray(), pMin = pillarMin.ToArray(), pMax = pillarMax.ToArray();
            int N = sim.NumSims;
            int S = P - (firstIsCurrent ? 1 : 0);
            var deltas = new double[P]; var profile = new double[P * 6];
            var triggerSummary = new double[(P - 1) * 4];
            var triggerInject = new double[(P - 1) * LsmcNative.NumTriggerVolumes * 2];
            var triggerWithdraw = new double[(P - 1) * LsmcNative.NumTriggerVolumes * 2];
            var pvBySim = new double[N];
            // the panels the reference returns (LSV.cs:351-356,602-603); a single array above 2 GB needs gcAllowVeryLargeObjects
            var spotRegress = new double[(long)S * N]; var spotValuation = new double[(long)S * N];
            var inventory = new double[(long)P * N]; var injectWithdraw = new double[(long)P * N]; var consumed = new double[(long)P * N];
            var lossBySim = new double[(long)P * N]; var netVolume = new double[(long)P * N]; var pv = new double[(long)P * N];

            LsmcNative.ProgressCallback callback = (user, progress) =>
            {
                lsmcParams.OnProgressUpdate?.Invoke(progress);                                   // LSV.cs:337-338,488-489,604
                return lsmcParams.CancellationToken.IsCancellationRequested ? 1 : 0;             // LSV.cs:339,490
            };

            int status; string message;
            int nDev = sim.Devices.Count;
            IntPtr* ctx = stackalloc IntPtr[nDev];
            try
            {
                for (int d = 0; d < nDev; d++) Check(LsmcNative.lsmc_create(sim.Devices[d], out ctx[d]), IntPtr.Zero);
                if (nDev > 1) Check(LsmcNative.lsmc_comm_init_local(ctx, nDev), ctx[0]);
                fixed (double* pSimTime = simTime, pMr = meanReversion, pVol = vol, pCorr = corr, pFwd = forward, pMinInv = minInv, pMaxInv = maxInv,
                       pInjCost = injCost, pInjDf = injDf, pWdCost = wdCost, pWdDf = wdDf, pConsInj = consInj, pConsWd = consWd, pLoss = loss,
                       pInvCost = invCost, pInvCostDf = invCostDf, pSettleDf = settleDf, pPillarInv = pInv, pPillarMin = pMin, pPillarMax = pMax,
                       pDeltas = deltas, pProfile = profile, pTrigSum = triggerSummary, pTrigInj = triggerInject, pTrigWd = triggerWithdraw,
                       pPvBySim = pvBySim, pSpotReg = spotRegress, pSpotVal = spotValuation, pInventory = inventory, pInjWd = injectWithdraw,
                       pConsumed = consumed, pLossBySim = lossBySim, pNetVolume = netVolume, pPv = pv)
                fixed (int* pSpotPow = spotPower, pFacPow = factorPower, pCid = constraintId, pKind = kindA, pFirst = firstA, pCount = countA)
                {
                    var problem = new LsmcNative.Problem
                    {
                        struct_size = sizeof(LsmcNative.Problem), n_periods = P, n_factors = F, n_basis = K, n_sims = N, n_sims_global = 0, sim_offset = 0,
                        extra_decisions = lsmcParams.ExtraDecisions, first_period_is_current = firstIsCurrent ? 1 : 0,
                        discount_deltas = lsmcParams.DiscountDeltas ? 1 : 0, must_be_empty_at_end = storage.MustBeEmptyAtEnd ? 1 : 0,
                        numerical_tolerance = lsmcParams.NumericalTolerance, start_inventory = lsmcParams.Inventory, grid_spacing = gridCalc.Spacing,
                        sim_time = pSimTime, mean_reversion = pMr, factor_vol = pVol, factor_corr = pCorr, forward = pFwd,
                        spot_power = pSpotPow, factor_power = pFacPow, min_inventory = pMinInv, max_inventory = pMaxInv, constraint_id = pCid,
                        inject_cost = pInjCost, inject_cost_df = pInjDf, withdraw_cost = pWdCost, withdraw_cost_df = pWdDf,
                        consumed_inject = pConsInj, consumed_withdraw = pConsWd, inventory_loss = pLoss, inventory_cost = pInvCost,
                        inventory_cost_df = pInvCostDf, settle_df = pSettleDf, n_constraints = kindA.Length, n_pillars = pInv.Length,
                        constraint_kind = pKind, constraint_first = pFirst, constraint_count = pCount,
                        pillar_inventory = pPillarInv, pillar_min_rate = pPillarMin, pillar_max_rate = pPillarMax,
                    };
                    for (int i = 0; i < 4; i++) problem.terminal_coef[i] = terminal[i];
                    // antithetic = 1: the reference always constructs MersenneTwisterGenerator(seed, antithetic: true) (LVP.cs:185-189)
                    var regressionRng = new LsmcNative.Rng { mode = sim.ReferenceCompatibleStream ? LsmcNative.RngMt19937 : LsmcNative.RngPhilox,
                                                             antithetic = 1, seed = sim.RegressionSeed };
                    var valuationRng = new LsmcNative.Rng { antithetic = 1, seed = sim.ValuationSeed,
                        mode = !sim.ReferenceCompatibleStream ? LsmcNative.RngPhilox
                             : sim.ValuationContinuesRegressionStream ? LsmcNative.RngMt19937Continue : LsmcNative.RngMt19937 };
                    var results = new LsmcNative.Results
                    {
                        struct_size = sizeof(LsmcNative.Results), deltas = pDeltas, profile = pProfile, trigger_summary = pTrigSum,
                        trigger_inject = pTrigInj, trigger_withdraw = pTrigWd, pv_by_sim = pPvBySim, spot_regress = pSpotReg, spot_valuation = pSpotVal,
                        inventory = pInventory, inject_withdraw = pInjWd, consumed = pConsumed, loss = pLossBySim, net_volume = pNetVolume, pv = pPv,
                    };
                    status = nDev > 1
                        ? LsmcNative.lsmc_calculate_multi(ctx, nDev, &problem, &regressionRng, &valuationRng, callback, IntPtr.Zero, &results)
                        : LsmcNative.lsmc_calculate(ctx[0], &problem, &regressionRng, &valuationRng, callback, IntPtr.Zero, &results);
                    message = status == LsmcNative.Ok ? null : Marshal.PtrToStringAnsi(LsmcNative.lsmc_last_error(ctx[0]));
                    GC.KeepAlive(callback);
                    if (status != LsmcNative.Ok) Throw(status, message);
                    return Wrap(periods, S, N, results.npv, deltas, profile, triggerSummary, triggerInject, triggerWithdraw, pvBySim,
                                spotRegress, spotValuation, inventory, injectWithdraw, consumed, lossBySim, netVolume, pv);
                }
            }
            finally
            {
                for (int d = 0; d < nDev; d++) if (ctx[d] != IntPtr.Zero) LsmcNative.lsmc_destroy(ctx[d]);
            }
        }

        private static double[] Ones(int n) { var a = new double[n]; for (int i = 0; i < n; i++) a[i] = 1.0; return a; }

        private static IEnumerable<double> Padded(double[] coefficients, int n) =>
            coefficients.Concat(Enumerable.Repeat(0.0, Math.Max(0, n - coefficients.Length))).Take(n);

        // A cost closure returns cash flows (amount, payment date) for a volume (CmdtyStorage.cs:327-413, 544-567): the per-unit
        // cost is the amount at volume 1 and the discount factor that of its (single) payment date; checked at volume 2.
        private static (double unitCost, double discountFactor) UnitCost(Func<double, IReadOnlyList<DomesticCashFlow>> cashFlows,
                                                                         Func<Day, double> df, string what)
        {
            IReadOnlyList<DomesticCashFlow> one = cashFlows(1.0), two = cashFlows(2.0);
            if (one.Count == 0) return (0.0, 1.0);
            if (one.Count != 1 || two.Count != 1 || !one[0].Date.Equals(two[0].Date) || Math.Abs(two[0].Amount - 2.0 * one[0].Amount) > 1e-12 * Math.Abs(one[0].Amount))
                throw new NotSupportedException("The " + what + " is not a single per-unit cash flow; no device descriptor exists for it (LSMC_EUNSUPPORTED).");
            return (one[0].Amount, df(one[0].Date));
        }

        private static double UnitRate(Func<double, double> consumed, string what)
        {
            double one = consumed(1.0), two = consumed(2.0);
            if (Math.Abs(two - 2.0 * one) > 1e-12 * Math.Abs(one))
                throw new NotSupportedException("The " + what + " is not proportional to the volume; no device descriptor exists for it (LSMC_EUNSUPPORTED).");
            return one;
        }

        private static double[] BilinearTerminalValue(Func<double, double, double> f)   // storage_b200/cmdty_storage.py terminal_coefficients
        {
            double c0 = f(0.0, 0.0), c1 = f(1.0, 0.0) - c0, c2 = f(0.0, 1.0) - c0, c3 = f(1.0, 1.0) - c0 - c1 - c2;
            foreach ((double price, double inv) in new[] { (37.25, 1234.5), (101.0, 0.75), (3.0, 52000.0) })
            {
                double expected = c0 + c1 * price + c2 * inv + c3 * price * inv;
                if (Math.Abs(f(price, inv) - expected) > 1e-9 * Math.Max(1.0, Math.Abs(expected)))
                    throw new NotSupportedException("TerminalStorageNpv is not of the form c0 + c1*price + c2*inventory + c3*price*inventory (LSMC_EUNSUPPORTED).");
            }
            return new[] { c0, c1, c2, c3 };
        }

        // ------------------------------------------------------------------------------------------------ 4. status -> exception
        private static void Check(int status, IntPtr ctx)
        {
            if (status != LsmcNative.Ok) Throw(status, Marshal.PtrToStringAnsi(LsmcNative.lsmc_last_error(ctx)));
        }

        private static void Throw(int status, string message)
        {
            switch (status)
            {
                case LsmcNative.EInval: throw new ArgumentException(message);                                   // LSV.cs:62,92,95; SH.cs:113-116
                case LsmcNative.EConstraint: throw new InventoryConstraintsCannotBeFulfilledException(message); // LSV.cs:75; SH.cs:100,130,154
                case LsmcNative.ECancelled: throw new OperationCanceledException(message);                      // LSV.cs:339,490
                case LsmcNative.EUnsupported: throw new NotSupportedException(message);
                default: throw new ApplicationException(message);                                               // CUDA / exchange failure
            }
        }

        // ------------------------------------------------------------------------------------------------ 5. results
        private static LsmcStorageValuationResults<T> Wrap<T>(T[] periods, int S, int N, double npv, double[] deltas, double[] profile,
            double[] triggerSummary, double[] triggerInject, double[] triggerWithdraw, double[] pvBySim, double[] spotRegress, double[] spotValuation,
            double[] inventory, double[] injectWithdraw, double[] consumed, double[] loss, double[] netVolume, double[] pv)
            where T : ITimePeriod<T>
        {
            int P = periods.Length, nv = LsmcNative.NumTriggerVolumes;
            var profiles = new StorageProfile[P];
            for (int p = 0; p < P; p++)
                profiles[p] = new StorageProfile(profile[p * 6], profile[p * 6 + 1], profile[p * 6 + 2], profile[p * 6 + 3], profile[p * 6 + 4], profile[p * 6 + 5]);
            double? Nullable(double v) => double.IsNaN(v) ? (double?)null : v;                    // NaN = null (TriggerPrices.cs:30-33)
            var triggerPrices = new TriggerPrices[P - 1];
            var triggerProfiles = new TriggerPriceVolumeProfiles[P - 1];
            for (int p = 0; p < P - 1; p++)
            {
                triggerPrices[p] = new TriggerPrices(Nullable(triggerSummary[p * 4]), Nullable(triggerSummary[p * 4 + 1]),
                                                     Nullable(triggerSummary[p * 4 + 2]), Nullable(triggerSummary[p * 4 + 3]));
                IEnumerable<TriggerPricePoint> Points(double[] src) =>
                    Enumerable.Range(0, nv).Where(i => !double.IsNaN(src[(p * nv + i) * 2]))
                              .Select(i => new TriggerPricePoint(src[(p * nv + i) * 2], src[(p * nv + i) * 2 + 1])).ToArray();
                triggerProfiles[p] = new TriggerPriceVolumeProfiles(Points(triggerInject), Points(triggerWithdraw));
            }
            T[] simulated = periods.Skip(P - S).ToArray();
            return new LsmcStorageValuationResults<T>(npv, new DoubleTimeSeries<T>(periods[0], deltas),
                new TimeSeries<T, StorageProfile>(periods[0], profiles),
                Panel.UseRawDataArray(spotRegress, simulated, N), Panel.UseRawDataArray(spotValuation, simulated, N),
                Panel.UseRawDataArray(inventory, periods, N), Panel.UseRawDataArray(injectWithdraw, periods, N),
                Panel.UseRawDataArray(consumed, periods, N), Panel.UseRawDataArray(loss, periods, N), Panel.UseRawDataArray(netVolume, periods, N),
                new TimeSeries<T, TriggerPrices>(periods[0], triggerPrices), new TimeSeries<T, TriggerPriceVolumeProfiles>(periods[0], triggerProfiles),
                Panel.UseRawDataArray(pv, periods, N), pvBySim);
        }
    }
}
