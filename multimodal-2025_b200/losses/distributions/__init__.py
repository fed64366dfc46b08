"""Mirrors of the reference `losses/distributions/` package + the `power_spherical` dependency."""
from .VonMisesFisher import VonMisesFisher, vMFLogPartition  # noqa: F401
from .ProjectedNormal import ProjectedNormal  # noqa: F401
from .power_spherical import PowerSpherical, HypersphericalUniform  # noqa: F401
