"""signSGD data plane: the fused majority-vote + SGD kernel against the torch statement of the same step, and
the arena-order contract of the exchanged gradient on a convolutional model (4-D weights are stored
channels-last inside the flat arena, so a gradient flattened in logical order would hit the wrong weights)."""
import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,k,momentum,nesterov", [(1000003, 5, 0.9, False), (4097, 1, 0.0, False), (70000, 8, 0.5, True),
                                                   (33, 3, 0.9, False)])
def test_sign_vote_sgd_kernel(n, k, momentum, nesterov):
    from distributed_learning_simulator_b200.ops import misc as M

    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(n + k)
    grads = [torch.randn(n, device=dev, generator=g) for _ in range(k)]
    rows = [M.sign_pack(x) for x in grads]
    params = torch.randn(n, device=dev, generator=g)
    mom = torch.randn(n, device=dev, generator=g)
    shadow = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    for ready in (False, True):
        p_ref, m_ref = params.clone(), mom.clone()
        vote = torch.sign(sum(torch.where(x < 0, -1.0, 1.0) for x in grads))
        d = vote
        if momentum:
            m_ref = momentum * m_ref + d if ready else d.clone()
            d = d + momentum * m_ref if nesterov else m_ref
        p_ref = p_ref - 0.05 * d
        M.sign_vote_sgd(rows, n, params, mom if momentum else None, 0.05, momentum, 0.0, nesterov, ready, shadow=shadow)
        torch.testing.assert_close(params, p_ref, rtol=1e-5, atol=1e-5)
        if momentum:
            torch.testing.assert_close(mom, m_ref, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(shadow.float(), params.bfloat16().float(), rtol=0, atol=0)
    assert torch.equal(M.sign_majority(rows, n), vote)


def test_arena_order_update_on_conv_model():
    """What GradientWorker does on CUDA: gradient laid out through ``layout.view_of`` (arena order), signs packed,
    voted and applied to the flat arena -- every parameter must move by exactly -lr * sign(grad)."""
    from distributed_learning_simulator_b200.engine.model_util import ModelUtil
    from distributed_learning_simulator_b200.ops import misc as M

    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    model = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Conv2d(8, 16, 5), nn.Flatten(), nn.LazyLinear(10)).to(dev)
    model(torch.zeros(2, 3, 12, 12, device=dev))
    util = ModelUtil(model)
    flat = util.flatten_parameters(channels_last=True)
    layout = util.layout
    assert any(layout.formats), "the arena must hold a channels-last segment for this test to mean anything"
    before = {n: p.detach().clone() for n, p in model.named_parameters()}
    grads = {n: torch.randn_like(p) for n, p in model.named_parameters()}
    gradient = torch.empty(layout.total, device=dev)
    for i, name in enumerate(layout.names):
        layout.view_of(gradient, i).copy_(grads[name])
    M.sign_vote_sgd([M.sign_pack(gradient)], layout.total, flat, None, 0.1, 0.0)
    for name, p in model.named_parameters():
        torch.testing.assert_close(p.detach(), before[name] - 0.1 * torch.sign(grads[name]), rtol=0, atol=1e-6, msg=name)


def test_sign_sgd_trains_a_conv_model():
    from distributed_learning_simulator_b200 import load_config_from_dict, train
    from distributed_learning_simulator_b200.training import get_last_training_result

    config = load_config_from_dict({
        "dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "sign_SGD", "optimizer_name": "SGD",
        "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 3, "batch_size": 64, "round": 1, "epoch": 3,
        "learning_rate": 0.002, "momentum": 0.9, "distribute_init_parameters": False, "log_level": "WARNING",
        "dataset_kwargs": {"train_size": 1920, "test_size": 500}})
    train(config)
    perf = get_last_training_result()["performance"]
    last = perf[max(perf)]
    assert last["test_accuracy"] > 0.5, perf  # the synthetic classes are separable: a wrong update order stays near 0.1
