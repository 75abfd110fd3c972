"""Importing this package registers every algorithm name with
``CentralizedAlgorithmFactory`` (reference: simulation_lib/method/__init__.py:1-9, plus the
four algorithms the reference's configs name without registering: SURVEY.md 2.4)."""
from .algorithm_factory import CentralizedAlgorithmFactory  # noqa: F401
from . import (fed_aas, fed_avg, fed_dropout_avg, fed_gcn, fed_gnn, fed_obd, fed_paq, shapley_value,  # noqa: F401
               sign_sgd, smafd)
