"""Process-global map: type produced by ``build()`` -> set of @factory classes producing it
(reference zookeeper/core/factory_registry.py:8)."""
from typing import Dict, Set, Type

FACTORY_REGISTRY: Dict[Type, Set] = {}
