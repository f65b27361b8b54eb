"""archs/ shim for the reference recipe: `arch=RSH_b200` in run_train.sh."""
from ssb200.RSH import *  # noqa: F401,F403
from ssb200.RSH import Collator, SepDNN, TestSet, TrainSet, compute_cv_loss, compute_loss, compute_masks  # noqa: F401
