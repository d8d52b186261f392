from .cli import cli
from .component import component, configure
from .factory import factory
from .field import ComponentField, Field
from .partial_component import PartialComponent
from .task import task

__all__ = ["component", "ComponentField", "configure", "cli", "factory", "Field", "PartialComponent", "task"]
