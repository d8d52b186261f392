from zookeeper_b200.tf import *  # noqa: F401,F403
from zookeeper_b200.tf import __all__  # noqa: F401
