"""Alias package: ``from zookeeper import ...`` resolves to the B200 build, so an experiment file
written against the reference imports unchanged."""
from zookeeper_b200 import *  # noqa: F401,F403
from zookeeper_b200 import __all__, __version__  # noqa: F401
