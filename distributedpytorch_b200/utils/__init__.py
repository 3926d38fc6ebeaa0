"""Framework-layer helpers of distributedpytorch_b200: loss oracles, datasets and samplers, the plateau scheduler, checkpoint
and loss-curve bookkeeping, logging, seeding, device-side timing and the (headless) plot helper.  Import the submodules
directly (`from distributedpytorch_b200.utils import data, loss, ...`); nothing is re-exported here on purpose so that
importing the package never pulls in pandas / PIL before they are needed."""
