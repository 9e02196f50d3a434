"""nfopp_b200 - B200-native (sm_100a) NFOPP optimisation inner loop behind the reference's
neural_field_optimal_planner API.  Host side: Python + ctypes over libnfopp.so (include/nfopp.h)."""
from .engine import EngineConfig, OnfSpec, PlannerEngine  # noqa: F401
from .onf_model import ONF  # noqa: F401
from .nerf_opt_planner import NERFOptPlanner  # noqa: F401
from .constrained_nerf_opt_planner import ConstrainedNERFOptPlanner  # noqa: F401
from .planner_factory import PlannerFactory, DEFAULT_PARAMETERS  # noqa: F401
from .test_environment_builder import TestEnvironmentBuilder  # noqa: F401
from .batched import BatchedPlanner  # noqa: F401
