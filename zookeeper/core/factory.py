"""Alias of zookeeper_b200.core.factory (import path of the reference)."""
from zookeeper_b200.core.factory import *  # noqa: F401,F403
import zookeeper_b200.core.factory as _m

globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
