"""Messages, flat payloads, topology hand-off."""
import copy
import threading

import torch

from distributed_learning_simulator_b200.engine.model_util import FlatParameter, ParameterLayout, as_flat
from distributed_learning_simulator_b200.message import (DeltaParameterMessage, Message, ParameterMessage,
                                                          get_message_size)
from distributed_learning_simulator_b200.topology import ClientEndpoint, ServerEndpoint, ThreadCentralTopology


def _flat():
    layout = ParameterLayout(["a", "b"], [torch.Size([2, 3]), torch.Size([4])])
    return FlatParameter(torch.arange(10, dtype=torch.float32), layout)


def test_messages_and_restore():
    old = {"a": torch.ones(2), "b": torch.zeros(3)}
    d = DeltaParameterMessage(dataset_size=5, delta_parameter={"a": torch.full((2,), 2.0)}, other_data={"x": 1},
                              in_round=True)
    msg = d.restore(old)
    assert isinstance(msg, ParameterMessage) and msg.dataset_size == 5 and msg.in_round and msg.other_data == {"x": 1}
    assert torch.equal(msg.parameter["a"], torch.full((2,), 3.0)) and torch.equal(msg.parameter["b"], old["b"])
    assert torch.equal(old["a"], torch.ones(2))  # not modified in place
    p = ParameterMessage(parameter={"a": torch.ones(2)})
    p.complete(old)
    assert set(p.parameter) == {"a", "b"}
    assert get_message_size(p) == 4 * 5
    assert get_message_size(ParameterMessage(parameter=_flat())) == 40
    assert Message().other_data == {} and not Message().end_training


def test_flat_parameter_survives_shallow_copy_and_handoff():
    fp = _flat()
    assert as_flat(fp) is fp and fp["a"].shape == (2, 3)
    assert as_flat(copy.copy(fp)) is not None
    fp2 = _flat()
    fp2["a"] = torch.zeros(2, 3)
    assert as_flat(fp2) is None  # an entry was replaced: only the dict view is valid
    topo = ThreadCentralTopology(2)
    server, c0, c1 = ServerEndpoint(topo), ClientEndpoint(topo, 0), ClientEndpoint(topo, 1)
    msg = ParameterMessage(parameter=fp, other_data={"k": 1})
    server.broadcast(msg, worker_ids={0, 1})
    got0, got1 = c0.get(), c1.get()
    assert as_flat(got0.parameter) is not None and as_flat(got1.parameter) is not None
    assert got0 is not got1 and got0.other_data is not got1.other_data
    assert got0.parameter.flat.data_ptr() == fp.flat.data_ptr()  # zero-copy hand-off
    c1.send(None)
    c0.send(DeltaParameterMessage(dataset_size=1, delta_parameter=fp))
    assert server.wait_for_data({0, 1}, timeout=1.0) == {0, 1}
    assert server.get(1) is None
    assert as_flat(server.get(0).delta_parameter) is not None
    assert server.wait_for_data({0}, timeout=0.05) == set()


def test_event_driven_wait_wakes_up():
    topo = ThreadCentralTopology(1)
    server, client = ServerEndpoint(topo), ClientEndpoint(topo, 0)
    t = threading.Timer(0.05, lambda: client.send(Message(other_data={"hello": True})))
    t.start()
    assert server.wait_for_data({0}, timeout=5.0) == {0}
    assert server.get(0).other_data["hello"]
    assert not client.has_data() and not client.wait_for_data(timeout=0.01)


def test_layout_with_channels_last_segments():
    import torch

    conv = torch.nn.Conv2d(3, 4, 3)
    fc = torch.nn.Linear(5, 2)
    model = torch.nn.Sequential(conv, torch.nn.Flatten(), fc)
    from distributed_learning_simulator_b200.engine.model_util import ModelUtil

    util = ModelUtil(model)
    ref = {k: v.clone() for k, v in util.get_parameter_dict().items()}
    flat = util.flatten_parameters(channels_last=True)
    layout = util.layout
    assert layout.formats == [1, 0, 0, 0]
    assert util.is_flat() and conv.weight.is_contiguous(memory_format=torch.channels_last)
    for k, v in util.get_parameter_dict().items():
        assert torch.equal(v, ref[k])
    # the dict view of the flat vector is format aware, in both directions
    fp = FlatParameter(flat.clone(), layout)
    assert all(torch.equal(fp[k], ref[k]) for k in ref)
    assert torch.equal(layout.flatten(ref), flat)
    plain = ParameterLayout.from_dict(ref)
    assert plain != layout and plain.formats == [0, 0, 0, 0]
    assert ParameterLayout.from_dict(util.get_parameter_dict()) == layout


def test_bf16_shadow_keeps_the_fp32_arena_as_master():
    """ModelUtil.enable_shadow: matrices become bf16 views of the shadow arena, every public
    accessor still speaks fp32, loads refresh the shadow, disable restores plain fp32 views."""
    import torch

    from distributed_learning_simulator_b200.engine.model_util import ModelUtil
    from distributed_learning_simulator_b200 import load_config_from_dict

    torch.manual_seed(0)
    config = load_config_from_dict({"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg",
                                    "worker_number": 1, "round": 1, "epoch": 1, "log_level": "WARNING",
                                    "dataset_kwargs": {"train_size": 64, "test_size": 32}})
    model = config.create_trainer().model
    util = ModelUtil(model)
    flat = util.flatten_parameters(with_grad=False)
    before = {k: v.clone() for k, v in util.get_parameter_dict().items()}
    shadow = util.enable_shadow()
    assert shadow.dtype == torch.bfloat16 and shadow.numel() == flat.numel() and util.is_flat()
    flags = dict(zip(util.layout.names, util.shadow_flags()))
    assert sum(flags.values()) >= 2
    for name, p in model.named_parameters():
        assert p.dtype == (torch.bfloat16 if flags[name] else torch.float32), name
        assert not flags[name] or (p.dim() >= 2 and p.data_ptr() % 16 == 0)
    after = util.get_parameter_dict()
    assert all(v.dtype == torch.float32 and torch.equal(v, before[k]) for k, v in after.items())
    util.load_parameter_dict({k: v + 1 for k, v in before.items()})
    assert torch.equal(shadow, flat.to(torch.bfloat16))
    assert torch.allclose(util.get_parameter_list(), torch.cat([(v + 1).reshape(-1) for v in before.values()]))
    x = torch.randn(4, 1, 28, 28)
    with torch.autocast("cpu", dtype=torch.bfloat16):
        out = model(x)
    assert out.shape == (4, 10)
    util.disable_shadow()
    assert all(p.dtype == torch.float32 for p in model.parameters()) and util.is_flat()
    assert util.enable_shadow().data_ptr() == shadow.data_ptr()  # same storage: captured graphs stay valid
