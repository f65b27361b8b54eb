"""archs/ shim for the reference recipe: `arch=uPIT_b200` in run_train.sh."""
from ssb200.uPIT import *  # noqa: F401,F403
from ssb200.uPIT import Collator, SepDNN, TestSet, TrainSet, compute_cv_loss, compute_loss, compute_masks  # noqa: F401
