"""Alias of zookeeper_b200.core.partial_component (import path of the reference)."""
from zookeeper_b200.core.partial_component import *  # noqa: F401,F403
import zookeeper_b200.core.partial_component as _m

globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
