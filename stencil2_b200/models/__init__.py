"""User-facing objects: DistributedDomain (the halo-exchange library proper) and the Jacobi3D / AstarothSim engines built on it."""
from .domain import DistributedDomain, Method, Placement  # noqa: F401
from .jacobi import Jacobi3D, jacobi_host_step  # noqa: F401
from .astaroth import AstarothSim, astaroth_host_step  # noqa: F401
