"""Star topology over multiprocessing pipes (reference: algorithm_factory.py:26-28)."""


class ProcessPipeCentralTopology:
    def __init__(self, mp_context, worker_num: int) -> None:
        self.worker_num = worker_num
        ctx = mp_context.get_ctx() if hasattr(mp_context, "get_ctx") else mp_context
        self._pipes = [ctx.Pipe(duplex=True) for _ in range(worker_num)]  # (server end, worker end)

    def server_end(self, worker_id: int):
        return self._pipes[worker_id][0]

    def worker_end(self, worker_id: int):
        return self._pipes[worker_id][1]
