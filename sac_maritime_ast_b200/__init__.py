"""B200-native batched ship-in-transit simulator behind the reference's MultiShipRLEnv interface.

Hot path only: ``ShipFleet`` (kernel K1: fused ship-step, K steps per launch) and
``MultiShipRLEnv`` (kernel K2: fused two-ship AST env step) over ``libshipsim.so`` (C ABI in
``include/shipsim.h``).  No CPU fallback: the CUDA library must be built and a GPU present.
"""
from ._lib import ShipSimError, lib, LIB_PATH  # noqa: F401
from .config import (  # noqa: F401
    DEFAULT_MAP, EngineThrottleFromSpeedSetPoint, EnvironmentConfiguration, FuelConsumptionCoefficients,
    HeadingByRouteController, HeadingBySampledRouteController, HeadingControllerGains, LosParameters,
    MachineryMode, MachineryModeParams, MachineryModes, MachinerySystemConfiguration, PolygonObstacle,
    ShipConfiguration, ShipDraw, ShipModel, ShipModelAST, ShipModelSimplifiedPropulsion, SimplifiedMachineryModel,
    SimplifiedPropulsionMachinerySystemConfiguration, SimulationConfiguration, StaticObstacle,
    ThrottleFromSpeedSetPointSimplifiedPropulsion,
    SpecificFuelConsumptionBaudouin6M26Dot3, SpecificFuelConsumptionWartila6L26, ThrottleControllerGains,
    default_ship, ship_params, simplified_ship)
from .env import MultiShipRLEnv, ShipAssets  # noqa: F401
from .fleet import ShipFleet  # noqa: F401
from .results import LOG_KEYS, REWARD_KEYS, RewardResults, SimulationResults  # noqa: F401

__all__ = ["ShipFleet", "MultiShipRLEnv", "ShipAssets", "PolygonObstacle", "default_ship", "ShipSimError", "SimulationResults", "RewardResults"]
