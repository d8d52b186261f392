"""zookeeper-b200: the zookeeper configuration surface plus a B200-native (sm_100a) engine for the
BinaryNet train loop of ``examples/larq_experiment.py``."""
from .core import ComponentField, Field, PartialComponent, cli, component, configure, factory, task

__version__ = "0.1.0"

__all__ = ["cli", "ComponentField", "component", "configure", "factory", "Field", "PartialComponent", "task"]
