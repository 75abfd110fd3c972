"""Opportunistic block dropout: keep the blocks that changed most, under a parameter budget.

Same selection rule as /root/reference/simulation_lib/method/fed_obd/obd_algorithm.py:
default block types (:12-22), leftover parameterised sub-modules become singleton blocks
(:35-55), the blocks must partition the parameter set exactly (:65-86), block score
``||cur - prev||_2 / numel`` (:129-145), greedy by descending score under the budget
``(1 - dropout_rate) * P`` skipping blocks that would overflow (:93-127), and the log line
``partial_parameter_num %s threshold %s parameter_num %s`` (:120-125).

The reference concatenates every block, copies it to the CPU and calls
``torch.norm(...).item()`` once per block (B host synchronisations per upload).  Here all
block scores come from ONE pass over the flat arena (``ops.misc.block_sqdiff``: segmented
sum of squared differences with a tensor->block table) and one device->host read of B doubles.
"""
from typing import Any

import torch
from torch import nn

from ...engine.log import get_logger
from ...engine.model_util import ModelUtil, ParameterLayout, as_flat
from ...ops import misc as misc_ops

DEFAULT_BLOCK_TYPES = {
    ("AlbertTransformer",),
    ("AlbertEmbeddings",),
    ("Bottleneck",),
    ("TransformerEncoderLayer",),
    (nn.BatchNorm2d, nn.ReLU, nn.Conv2d),
    (nn.BatchNorm2d, nn.Conv2d),
    (nn.Conv2d, nn.BatchNorm2d),
}


def get_module_blocks(model_util: ModelUtil, block_types: set | None = None) -> list:
    return model_util.get_module_blocks(block_types=block_types or DEFAULT_BLOCK_TYPES)


class OpportunisticBlockDropoutAlgorithm:
    def __init__(self, dropout_rate: float, worker_id: int) -> None:
        self.__dropout_rate = dropout_rate
        get_logger().warning("use dropout rate %s", self.__dropout_rate)
        self.__block_names: list[list[str]] | None = None   # parameter names per block
        self.__block_sizes: list[int] = []
        self.__parameter_num: int = 0
        self.__worker_id = worker_id
        self.__tables: dict = {}

    @property
    def block_names(self) -> list[list[str]]:
        assert self.__block_names is not None
        return self.__block_names

    def find_blocks(self, model_util: ModelUtil) -> None:
        blocks = get_module_blocks(model_util)
        for submodule_name, submodule in model_util.get_modules():
            if not submodule_name or not list(submodule.parameters()):
                continue
            related = any(
                b == submodule_name or submodule_name.startswith(b + ".") or b.startswith(submodule_name + ".")
                for block in blocks for b, _ in block)
            if not related and not list(submodule.children()):
                blocks.append([(submodule_name, submodule)])
                if self.__worker_id == 0:
                    get_logger().info("identify a submodule:%s", submodule_name)
        names_per_block: list[list[str]] = []
        seen: set[str] = set()
        for block in blocks:
            names = []
            for submodule_name, submodule in block:
                for p_name, _ in submodule.named_parameters():
                    full = f"{submodule_name}.{p_name}" if submodule_name else p_name
                    if full not in seen:
                        seen.add(full)
                        names.append(full)
            if names:
                names_per_block.append(names)
        parameter_dict = model_util.get_parameter_dict()
        # parameters owned directly by a container (no leaf module) form singleton blocks too
        for name in parameter_dict:
            if name not in seen:
                seen.add(name)
                names_per_block.append([name])
        for a in seen:
            if a not in parameter_dict:
                raise RuntimeError(a + " not in model")
        if self.__worker_id == 0:
            get_logger().info("identify these blocks in model:")
            for names in names_per_block:
                get_logger().info("%s", names)
        self.__block_names = names_per_block
        self.__block_sizes = [sum(parameter_dict[n].numel() for n in names) for names in names_per_block]
        self.__parameter_num = sum(p.numel() for p in parameter_dict.values())
        assert sum(self.__block_sizes) == self.__parameter_num

    def block_scores(self, parameter_dict: dict, prev_parameter_dict: dict) -> list[float]:
        """``||cur - prev||_2 / numel`` per block."""
        cur, prev = as_flat(parameter_dict), as_flat(prev_parameter_dict)
        if cur is not None and prev is not None and cur.layout == prev.layout and cur.flat.device == prev.flat.device:
            layout: ParameterLayout = cur.layout
            key = (hash(layout), str(cur.flat.device))
            if key not in self.__tables:
                block_of = torch.full((len(layout.names),), -1, dtype=torch.int32)
                for b, names in enumerate(self.block_names):
                    for n in names:
                        block_of[layout.index[n]] = b
                self.__tables[key] = (layout.segment_offsets(cur.flat.device), block_of.to(cur.flat.device))
            offsets, block_of = self.__tables[key]
            sq = misc_ops.block_sqdiff(cur.flat, prev.flat, offsets, block_of, len(self.block_names)).cpu()
            return [float(sq[b].sqrt()) / self.__block_sizes[b] for b in range(len(self.block_names))]
        scores = []
        for names, size in zip(self.block_names, self.__block_sizes):
            sq = sum(float(((parameter_dict[n] - prev_parameter_dict[n].to(parameter_dict[n].device)).double() ** 2).sum())
                     for n in names)
            scores.append(sq ** 0.5 / size)
        return scores

    def select_blocks(self, scores: list[float]) -> list[int]:
        threshold = (1 - self.__dropout_rate) * self.__parameter_num
        partial_parameter_num = 0
        chosen: list[int] = []
        for b in sorted(range(len(scores)), key=lambda i: (-scores[i], i)):
            if partial_parameter_num > threshold:
                break
            if partial_parameter_num + self.__block_sizes[b] > threshold:
                continue
            partial_parameter_num += self.__block_sizes[b]
            chosen.append(b)
        get_logger().info("partial_parameter_num %s threshold %s parameter_num %s", partial_parameter_num, threshold,
                          self.__parameter_num)
        return chosen

    def get_block_parameter(self, parameter_dict: dict, model_util: ModelUtil, model_cache: Any) -> dict:
        """The kept blocks' parameters (a ``PartialFlat`` when the input is a flat row)."""
        if self.__block_names is None:
            self.find_blocks(model_util=model_util)
        scores = self.block_scores(parameter_dict, model_cache.parameter_dict)
        chosen = self.select_blocks(scores)
        names = [n for b in chosen for n in self.block_names[b]]
        get_logger().debug("choose blocks %s", names)
        flat = as_flat(parameter_dict)
        if flat is not None:
            from ...ops.reduce import PartialFlat

            return PartialFlat(flat.flat, flat.layout, names)
        return {n: parameter_dict[n] for n in names}
