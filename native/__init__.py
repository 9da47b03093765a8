"""Reference-compatible top-level `native` package (alias of garfield_b200.native)."""

from garfield_b200.native import *  # noqa: F401,F403
from garfield_b200.native import krum, bulyan, median, brute, available, supports  # noqa: F401
