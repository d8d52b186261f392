"""Alias of zookeeper_b200.tf.preprocessing (import path of the reference)."""
import zookeeper_b200.tf.preprocessing as _m

globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
