"""heat_b200 — B200-native wall-conduction time-stepping path of SIMPLE `heat`.

Host mirror of the reference interface (`ThermalModel.new/march/allocate_memory`, src/model.rs:188-428) plus the
batched entry point (`ThermalBatch`), both thin layers over the C ABI in include/heat_b200.h.
"""
from . import desc  # noqa: F401
from .model import CurrentWeather, Date, ThermalBatch, ThermalModel, ThermalModelMemory  # noqa: F401
from ._lib import HeatB200Error  # noqa: F401
