"""storage_b200 — B200-native LSMC hot path of Cmdty.Storage behind the reference's Python API names.

Public surface mirrors src/Cmdty.Storage.Python/cmdty_storage/__init__.py for the LSMC path only.
"""
from .cmdty_storage import CmdtyStorage, RatchetInterp, InjectWithdrawRange
from .utils import FREQ_TO_PERIOD_TYPE

__version__ = "0.1.0"


def __getattr__(name):
    # multi_factor needs the native library only when a valuation is actually run; import lazily
    if name in ("multi_factor_value", "three_factor_seasonal_value", "MultiFactorSpotSim",
                "MultiFactorValuationResults", "create_3_factor_season_params"):
        from . import multi_factor
        return getattr(multi_factor, name)
    if name == "MultiFactorModel":
        from .multi_factor_model import MultiFactorModel
        return MultiFactorModel
    if name in ("intrinsic_value", "IntrinsicValuationResults"):
        from . import intrinsic
        return getattr(intrinsic, name)
    raise AttributeError(name)
