"""Importing this package registers every algorithm name with
``CentralizedAlgorithmFactory`` (reference: simulation_lib/method/__init__.py:1-9, plus the
four algorithms the reference's configs name without registering: SURVEY.md 2.4)."""
from .algorithm_factory import CentralizedAlgorithmFactory  # noqa: F401
from . import fed_avg, fed_dropout_avg, fed_obd, fed_paq, shapley_value, sign_sgd, smafd  # noqa: F401
