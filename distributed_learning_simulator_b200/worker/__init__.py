from .aggregation_worker import AggregationWorker  # noqa: F401
from .client import Client  # noqa: F401
from .error_feedback_worker import ErrorFeedbackWorker  # noqa: F401
from .gradient_worker import GradientWorker  # noqa: F401
from .graph_worker import GraphWorker  # noqa: F401
from .worker import Worker  # noqa: F401
