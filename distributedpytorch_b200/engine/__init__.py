from .native_unet import NativeUNetEngine, ParamArena  # noqa: F401
from .program import Program, run_parallel  # noqa: F401
