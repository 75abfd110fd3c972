"""Stand-in for the private ``cyy_naive_lib`` (log, time counter, storage, pipe topology,
process-pool initialisation data) with the contracts the reference uses."""
