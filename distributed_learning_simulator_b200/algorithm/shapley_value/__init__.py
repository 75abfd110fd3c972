"""Shapley-value engines.  The reference imports them from the external
``cyy_torch_algorithm.shapely_value`` package
(/root/reference/simulation_lib/method/shapley_value/GTG_shapley_value_algorithm.py:1,
multiround_shapley_value_algorithm.py:1-2); their contract (ctor ``(players,
last_round_metric)``, ``set_metric_function``, ``compute(round_number)``, attributes
``shapley_values`` / ``shapley_values_S``) is reconstructed from the call sites at
method/shapley_value/shapley_value_algorithm.py:38-52 and the GTG-Shapley paper."""
from .base import ShapleyValue  # noqa: F401
from .gtg_shapley_value import GTGShapleyValue  # noqa: F401
from .hierarchical_shapley_value import HierarchicalShapleyValue  # noqa: F401
from .multiround_shapley_value import MultiRoundShapleyValue  # noqa: F401
