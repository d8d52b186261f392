"""Optimizer descriptions (the arithmetic is ``zkb_adam_clip``)."""
from dataclasses import dataclass


@dataclass
class Adam:
    """``tf.keras.optimizers.Adam`` hyper-parameters (examples/larq_experiment.py:120-122)."""

    learning_rate: float = 1e-3
    beta_1: float = 0.9
    beta_2: float = 0.999
    epsilon: float = 1e-7


Optimizer = Adam
