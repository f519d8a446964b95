"""lsmc_npv_stats_noalloc (include/lsmc_b200.h): the allocation-free variant of the NPV / standard-error reduction, whose
scratch is a static __device__ array of the library.  Called through the C ABI with device pointers and checked against
the existing entry point (lsmc_calculate's npv / npv_std_error) on two calls in a row — the second call finds the first
call's partials in the shared array.  Collected last on purpose."""
import numpy as np
import pytest

from storage_b200 import _abi, _native
from storage_b200._abi import Rng
from storage_b200.configs import build_config

pytestmark = pytest.mark.gpu


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def test_noalloc_npv_stats_match_the_existing_entry_point_on_two_calls_in_a_row():
    import torch   # device buffers for the caller's side only; the entry point itself allocates nothing
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    out = torch.empty(4, dtype=torch.float64, device="cuda")
    cases = [("c3", 40_000, True), ("c2", 10_001, True), ("c2", 24_576, False)]   # large then small: stale partials must not leak
    with _native.Context(0) as ctx:
        for config, paths, antithetic in cases:
            pb = build_config(config, num_sims=paths)
            ref = ctx.calculate(pb, Rng.philox(11, antithetic), Rng.philox(22, antithetic))
            d_pv = torch.from_numpy(ref.pv_by_sim).to("cuda")
            out.fill_(float("nan"))
            torch.cuda.synchronize()
            for _ in range(2):   # two calls in a row on the shared scratch
                _native.npv_stats_noalloc(d_pv.data_ptr(), paths, antithetic, out.data_ptr(), stream)
                got = out.cpu().numpy()
                assert rel(got[0], ref.npv) < 1e-12, (config, paths, got[0], ref.npv)            # summation order only
                assert rel(got[1], ref.npv_std_error) < 1e-9, (config, paths, got[1], ref.npv_std_error)
                assert rel(got[2], float(np.sum(ref.pv_by_sim))) < 1e-12


def test_noalloc_rejects_sizes_beyond_the_static_scratch_and_null_pointers():
    import torch
    torch.cuda.set_device(0)
    buf = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(_native.LsmcError) as e:
        _native.npv_stats_noalloc(buf.data_ptr(), (1 << 24) + 1, True, buf.data_ptr(), 0)
    assert e.value.status == _abi.LSMC_EUNSUPPORTED and "LSMC_NOALLOC_MAX_SIMS" in e.value.message
    with pytest.raises(_native.LsmcError) as e:
        _native.npv_stats_noalloc(0, 16, True, buf.data_ptr(), 0)
    assert e.value.status == _abi.LSMC_EINVAL
    with pytest.raises(_native.LsmcError) as e:
        _native.npv_stats_noalloc(buf.data_ptr(), 0, True, buf.data_ptr(), 0)
    assert e.value.status == _abi.LSMC_EINVAL
