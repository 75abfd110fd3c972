from .aggregation_worker import AggregationWorker


class GraphWorker(AggregationWorker):  # filled in by the graph milestone
    pass
