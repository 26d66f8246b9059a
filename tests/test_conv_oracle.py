"""CPU checks of the convolutional path's oracle and host logic (no GPU): the restated input pipeline is pinned to
torchvision's own transform object (the one task.py:250-255 builds), the parameter layout to torchvision's resnet18,
and the one-process DDP emulation to real DistributedDataParallel over gloo with world_size 2."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_restated_pipeline_equals_torchvision_transforms_bit_for_bit():
    from b200step.convnet import draw_augmentation, make_synthetic_images
    from oracle.resnet_step import augment_with_draws, augment_with_torchvision

    ds = make_synthetic_images(64, seed=7)
    torch.manual_seed(123)
    draws = draw_augmentation(64)
    assert draws[:, 0].max() <= 8 and draws[:, 1].max() <= 8 and set(draws[:, 2].tolist()) == {0, 1}
    ref = augment_with_torchvision(ds.images, seed=123)
    assert torch.equal(augment_with_draws(ds.images, draws), ref)
    # and the evaluation form (no draws) is the centre crop without flip
    centre = torch.tensor([[4, 4, 0, 0]] * 64, dtype=torch.int8)
    assert torch.equal(augment_with_draws(ds.images, None), augment_with_draws(ds.images, centre))


def test_layout_follows_torchvision_parameter_order_and_round_trips():
    import torchvision

    from b200step.convnet import filter_to_gemm, flatten_state, gemm_to_filter, init_state, resnet_layout

    for arch in ("resnet18", "resnet34"):
        lay = resnet_layout(arch)
        model = torchvision.models.__dict__[arch]()
        assert [s.key for s in lay.params] == [k for k, _ in model.named_parameters()]
        assert [s.shape for s in lay.params] == [tuple(p.shape) for p in model.parameters()]
        assert all(s.offset % 4 == 0 for s in lay.params)
        bn_keys = [k[: -len(".running_mean")] for k in model.state_dict() if k.endswith("running_mean")]
        assert [n for n, _ in lay.bns] == bn_keys
    lay = resnet_layout("resnet18")
    assert lay.pool_hw == 1 and lay.stem.Kp == 152 and lay.n_params == 11689512 + 64 * 5
    sd = init_state("resnet18", 0)
    torch.manual_seed(0)
    assert torch.equal(sd["conv1.weight"], torchvision.models.resnet18().state_dict()["conv1.weight"])
    flat = flatten_state(lay, sd)
    for s in lay.params:
        if s.conv is not None:
            g = flat[s.offset: s.offset + s.numel].view(s.conv.cout, s.conv.Kp)
            assert torch.equal(gemm_to_filter(g, s.shape), sd[s.key])
            assert not g[:, s.conv.K:].any()
    w = torch.randn(8, 3, 7, 7)
    assert torch.equal(gemm_to_filter(filter_to_gemm(w, 152), w.shape), w)
    with pytest.raises(Exception, match="BasicBlock"):
        resnet_layout("resnet50")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _ddp_worker(rank, world, port, images, labels, rank_rows, rank_draws, batch, steps, out):
    import torchvision

    from oracle.resnet_step import augment_with_draws

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    torch.manual_seed(0)
    model = torch.nn.parallel.DistributedDataParallel(torchvision.models.resnet18())       # task.py:194
    crit = torch.nn.CrossEntropyLoss()
    opt = torch.optim.SGD(model.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    losses = []
    for k in range(steps):
        rows = rank_rows[rank][k * batch: (k + 1) * batch]
        x = augment_with_draws(images[rows], rank_draws[rank][k * batch: (k + 1) * batch])
        opt.zero_grad()
        loss = crit(model(x), labels[rows])
        loss.backward()
        opt.step()
        losses.append(float(loss))
    if rank == 0:
        torch.save({"losses": losses, "state": model.state_dict()}, out)
    else:
        torch.save({"losses": losses}, out + ".1")
    dist.destroy_process_group()


def test_ddp_emulation_equals_real_ddp_over_gloo(tmp_path):
    """The oracle's world=2 emulation (local BatchNorm statistics per rank, mean gradient, rank 0's buffers) against
    real DistributedDataParallel with two gloo processes."""
    from b200step.convnet import draw_augmentation, make_synthetic_images
    from b200step.sampler import rank_indices
    from oracle.resnet_step import ResNetOracle

    ds = make_synthetic_images(64, seed=4)
    rank_rows = [rank_indices(64, 2, r) for r in range(2)]
    torch.manual_seed(9)
    rank_draws = [draw_augmentation(32) for _ in range(2)]
    out = str(tmp_path / "ddp.pt")
    mp.spawn(_ddp_worker, args=(2, _free_port(), ds.images, ds.labels, rank_rows, rank_draws, 8, 3, out), nprocs=2,
             join=True)
    real = torch.load(out)
    real1 = torch.load(out + ".1")
    oracle = ResNetOracle("resnet18", ds.images, ds.labels, rank_rows, batch_size=8, lr=0.05, rank_draws=rank_draws)
    torch.set_num_threads(2)
    losses = [oracle.step() for _ in range(3)]
    for k in range(3):
        assert abs(losses[k][0] - real["losses"][k]) <= 2e-4 * max(1, abs(real["losses"][k]))
        assert abs(losses[k][1] - real1["losses"][k]) <= 2e-4 * max(1, abs(real1["losses"][k]))
    so = oracle.state_dict()
    assert list(so.keys()) == list(real["state"].keys())
    for k, v in real["state"].items():
        if v.dtype.is_floating_point:
            err = ((so[k] - v).norm() / v.norm().clamp_min(1e-12)).item()
            assert err <= 2e-3, (k, err)
        else:
            assert torch.equal(so[k], v), k


@pytest.mark.parametrize("seed,n", [(0, 1), (5, 207), (123, 1000), (7, 3000)])
def test_vectorised_draws_equal_the_per_sample_torch_calls(seed, n):
    """draw_augmentation replays torch's own mt19937 in one vectorised pass: same draws as the calls torchvision makes,
    and the global generator is left in the same state (the next torch draw is identical) — fresh after manual_seed,
    mid-stream, and across several state refills."""
    from b200step.convnet import draw_augmentation, draw_augmentation_loop

    for warm in (0, 3, 700):
        torch.manual_seed(seed)
        if warm:
            torch.rand(warm)
        a = draw_augmentation_loop(n)
        a2 = draw_augmentation_loop(5)
        tail_a = (torch.randint(0, 1 << 30, (4,)), torch.rand(3), torch.randn(2))
        torch.manual_seed(seed)
        if warm:
            torch.rand(warm)
        b = draw_augmentation(n)
        b2 = draw_augmentation(5)
        tail_b = (torch.randint(0, 1 << 30, (4,)), torch.rand(3), torch.randn(2))
        assert torch.equal(a, b) and torch.equal(a2, b2)
        assert all(torch.equal(x, y) for x, y in zip(tail_a, tail_b))


@pytest.mark.parametrize("arch", ["resnet18", "resnet34"])
def test_gradient_buckets_are_ddps_own(arch):
    """The bucket plan the all-reduces follow is the one torch's DistributedDataParallel computes for the same model
    (reducer defaults: 1 MiB first bucket, 25 MiB cap, parameters in reverse order)."""
    import torchvision

    from b200step.convnet import ddp_bucket_plan, resnet_layout

    lay = resnet_layout(arch)
    model = torchvision.models.__dict__[arch]()
    named = list(model.named_parameters())
    rev = list(reversed(named))
    idx, _ = dist._compute_bucket_assignment_by_size([p for _, p in rev], [1024 * 1024, 25 * 1024 * 1024],
                                                     [False] * len(rev), [])
    ref = [[rev[i][0] for i in b] for b in idx]
    plan = ddp_bucket_plan(lay)
    assert plan == ref
    if arch == "resnet18":
        assert [len(b) for b in plan] == [2, 12, 48]
    # every bucket is one contiguous range of the flat buffer, together they cover it
    end = lay.n_params
    for keys in plan:
        segs = sorted((lay.seg(k) for k in keys), key=lambda s: s.offset)
        assert segs[-1].offset + ((segs[-1].numel + 3) // 4) * 4 >= end - 3 and segs[-1].offset < end
        for a, b in zip(segs, segs[1:]):
            assert a.offset + ((a.numel + 3) // 4) * 4 == b.offset
        end = segs[0].offset
    assert end == 0
