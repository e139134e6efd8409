"""Small helpers: Statistics (trimean etc.), GPU clock sampling for benchmarks, the in-tree build driver."""
from .statistics import Statistics  # noqa: F401
