"""Alias of zookeeper_b200.core.task (import path of the reference)."""
from zookeeper_b200.core.task import *  # noqa: F401,F403
import zookeeper_b200.core.task as _m

globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
