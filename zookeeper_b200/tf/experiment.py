"""``Experiment`` field bag (reference zookeeper/tf/experiment.py:10-27); sub-classes implement
their training loop in ``run``."""
from typing import Optional, Sequence, Union

from ..core.field import ComponentField, Field
from .dataset import Dataset
from .models import Model
from .optimizers import Optimizer
from .preprocessing import Preprocessing


class Experiment:
    # Nested components
    dataset: Dataset = ComponentField()
    preprocessing: Preprocessing = ComponentField()
    model: Model = ComponentField()

    # Parameters
    epochs: int = Field()
    batch_size: int = Field()
    loss: Optional[Union[Sequence[Optional[str]], str]] = Field()
    optimizer: Union[Optimizer, str] = Field()
