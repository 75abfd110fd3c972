"""Offline post-processing of logs and session directories (capability parity with
/root/reference/simulation_lib/analysis/: analyze_log, analyze_round, session,
graph_exp_analyzer, module_diff)."""
from .module_diff import ModuleDiff  # noqa: F401
from .session import GraphSession, Session  # noqa: F401
