"""Spawn-context process pool: one process per submitted job, all alive at once
(reference: simulation_lib/training.py:105-135,143,149)."""
import concurrent.futures

import torch.multiprocessing
from cyy_naive_lib.data_structure.process_initialization import default_initializer


class TorchProcessPool:
    def __init__(self, initargs=None, max_workers: int = 64) -> None:
        self._executor = concurrent.futures.ProcessPoolExecutor(
            max_workers=max_workers, mp_context=torch.multiprocessing.get_context("spawn"),
            initializer=default_initializer, initargs=tuple(initargs or [{}]))
        self._futures: list = []

    def submit(self, fn, *args, **kwargs):
        fut = self._executor.submit(fn, *args, **kwargs)
        self._futures.append(fut)
        return fut

    def wait_results(self, timeout=None):
        done, not_done = concurrent.futures.wait(self._futures, timeout=timeout)
        results = {}
        for fut in done:
            results[fut] = fut.result()
        self._futures = list(not_done)
        return results, not_done

    def shutdown(self, wait: bool = True) -> None:
        self._executor.shutdown(wait=wait)
