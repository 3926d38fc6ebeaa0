from .evaluate import evaluate, evaluate_dice  # noqa: F401
from .fit import fit, fit_DDP, fit_DP  # noqa: F401
from .fit_mp import fit_MP  # noqa: F401
