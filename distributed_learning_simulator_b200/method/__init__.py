"""Importing this package registers every algorithm name with
``CentralizedAlgorithmFactory`` (reference: simulation_lib/method/__init__.py:1-9)."""
from .algorithm_factory import CentralizedAlgorithmFactory  # noqa: F401
from . import fed_avg, fed_paq  # noqa: F401
