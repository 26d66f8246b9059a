"""Step-level parity of the reference-literal path: `ConvTrainStep` (ResNet-18, task.py:171,303-312) on the GPU against
oracle/resnet_step.py — torchvision's resnet18 + nn.CrossEntropyLoss + torch.optim.SGD on the CPU in fp32, fed by
torchvision's own transform — on the same seeded images, sampler order and augmentation draws.

Tolerances (floating point; stated per precision):
  fp32  (3xTF32 GEMMs)      loss rel <= 1e-4 per step, parameters rel-L2 <= 1e-3 after 4 steps
  tf32  (one TF32 pass)     loss rel <= 1e-2, parameters rel-L2 <= 5e-2
  bf16  (bf16 activations)  loss rel <= 5e-2, parameters rel-L2 <= 1.5e-1   (BatchNorm over 32 samples of a 1x1 map
                            amplifies bf16 rounding; the oracle keeps fp32 activations)
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

BARS = {"fp32": (1e-4, 1e-3), "tf32": (1e-2, 5e-2), "bf16": (5e-2, 1.5e-1)}


def _setup(precision, *, world=1, rank=0, batch=32, rows=256, steps=4, augment=True, optimizer="sgd", lr=0.05,
           use_graph=True, arch="resnet18"):
    from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout
    from b200step.engine import StepConfig
    from b200step.sampler import rank_indices
    from oracle.resnet_step import ResNetOracle

    ds = make_synthetic_images(rows, seed=1)
    rank_rows = [rank_indices(rows, world, r) for r in range(world)]
    torch.manual_seed(5)
    rank_draws = [draw_augmentation(len(rr)) if augment else None for rr in rank_rows]
    oracle = ResNetOracle(arch, ds.images, ds.labels, rank_rows, batch_size=batch, lr=lr, seed=0,
                          rank_draws=rank_draws, optimizer=optimizer)
    lay = resnet_layout(arch)
    cfg = StepConfig(batch_size=batch, world=1, rank=0, precision=precision, optimizer=optimizer, lr=lr,
                     use_graph=use_graph)
    eng = ConvTrainStep(lay, init_state(arch, 0), ds, rank_rows[rank], cfg, augment=augment)
    if augment:
        eng.set_draws(rank_draws[rank])
    return ds, oracle, eng


def _rel_l2(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("precision", ["fp32", "tf32", "bf16"])
def test_resnet18_steps_match_the_torchvision_oracle(precision):
    ds, oracle, eng = _setup(precision)
    lo, lg = [], []
    for _ in range(4):
        lo.append(oracle.step()[0])
        eng.step()
        lg.append(eng.loss_value())
    eng.synchronize()
    rel = [abs(a - b) / max(1.0, abs(a)) for a, b in zip(lo, lg)]
    print(f"{precision}: oracle {lo}\n      engine {lg}\n      rel {rel}")
    loss_bar, param_bar = BARS[precision]
    assert max(rel) <= loss_bar, rel
    so, sg = oracle.state_dict(), eng.state_dict()
    assert list(so.keys()) == list(sg.keys())
    worst = ("", 0.0)
    for k in so:
        assert so[k].shape == sg[k].shape and so[k].dtype == sg[k].dtype, k
        if k.endswith("num_batches_tracked"):
            assert int(so[k]) == int(sg[k]) == 4
            continue
        e = _rel_l2(sg[k].float(), so[k].float())
        if e > worst[1]:
            worst = (k, e)
    print(f"{precision}: worst parameter / buffer rel-L2 {worst}")
    assert worst[1] <= param_bar, worst
    eng.close()


def test_initial_state_dict_is_the_reference_models():
    """Before any step, state_dict() is torchvision's seeded resnet18 bit for bit (OIHW filters restored from the GEMM
    layout), under DDP's `module.` prefix (task.py:283)."""
    import torchvision

    ds, oracle, eng = _setup("tf32")
    torch.manual_seed(0)
    ref = torchvision.models.resnet18().state_dict()
    sg = eng.state_dict()
    assert list(sg.keys()) == ["module." + k for k in ref.keys()]
    for k, v in ref.items():
        assert torch.equal(sg["module." + k], v), k
    eng.close()


def test_graph_replay_equals_eager_steps():
    _, _, a = _setup("tf32", use_graph=True)
    _, _, b = _setup("tf32", use_graph=False)
    for _ in range(4):
        a.step()
        b.step()
    la, lb = a.loss_value(), b.loss_value()
    assert la == lb
    sa, sb = a.state_dict(), b.state_dict()
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
    assert a.kernels_per_step == b.kernels_per_step > 100
    a.close(); b.close()


def test_tail_batch_and_epoch_wraparound():
    """8 steps of 32 over 256 rows = one epoch, then the cursor wraps; a 40-row rank list leaves a tail batch of 8."""
    from b200step.convnet import ConvTrainStep, init_state, make_synthetic_images, resnet_layout
    from b200step.engine import StepConfig
    from oracle.resnet_step import ResNetOracle

    ds = make_synthetic_images(40, seed=2)
    rows = torch.arange(40, dtype=torch.int64)
    oracle = ResNetOracle("resnet18", ds.images, ds.labels, [rows], batch_size=32, lr=0.05)
    eng = ConvTrainStep(resnet_layout(), init_state(), ds, rows, StepConfig(batch_size=32, precision="fp32", lr=0.05),
                        augment=False)
    lo = [oracle.step()[0], oracle.step()[0]]          # the oracle's second batch is the 8-row tail
    eng.step()
    l0 = eng.loss_value()
    eng.step_tail()
    l1 = eng.loss_value()
    assert abs(l0 - lo[0]) <= 1e-4 * max(1, abs(lo[0])) and abs(l1 - lo[1]) <= 1e-3 * max(1, abs(lo[1])), (lo, l0, l1)
    eng.close()


def test_evaluate_uses_running_statistics_like_model_eval():
    ds, oracle, eng = _setup("fp32", steps=3)
    for _ in range(3):
        oracle.step()
        eng.step()
    acc_o, pred_o = oracle.evaluate(ds.images, ds.labels)
    acc_g, pred_g = eng.evaluate(return_predictions=True)
    agree = (pred_o == pred_g).float().mean().item()
    print(f"eval accuracy oracle {acc_o:.4f} engine {acc_g:.4f}, predictions agree on {agree:.4f}")
    assert agree >= 0.98 and abs(acc_o - acc_g) <= 0.02
    eng.close()
