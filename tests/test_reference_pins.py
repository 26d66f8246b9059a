"""Fixtures produced by the reference's own functions (tests/golden/make_reference_pins.py runs
task.set_random_seeds, task.py:22-28, and task.evaluate, task.py:30-46, from /root/reference in the build
container) against this repo's restatements.  CPU part: seeding + construction order; GPU part: the untrained
state_dict bit for bit and TrainStep.evaluate() row by row."""
import dataclasses
import os

import pytest
import torch

from b200step.data import make_synthetic
from b200step.model import ARCHS, init_parameters
from tests.golden.make_reference_pins import PIN_SPECS, tensor_sha

PINS = os.path.join(os.path.dirname(__file__), "golden", "reference_pins.pt")


def _spec(name):
    return dataclasses.replace(ARCHS["tabular_mlp"], **PIN_SPECS[name][0])


@pytest.fixture(scope="module")
def pins():
    return torch.load(PINS)


@pytest.mark.parametrize("name", sorted(PIN_SPECS))
def test_seeded_init_matches_reference_seeding(pins, name):
    """task.set_random_seeds(0) followed by model construction (task.py:162 -> :171): init_parameters() and the
    oracle's build_model() must reproduce every parameter tensor bit for bit."""
    from oracle.reference_step import build_model

    spec = _spec(name)
    ref = pins[name]["init_sha256"]
    mine = init_parameters(spec, 0)
    assert sorted(mine) == sorted(ref)
    for k, v in mine.items():
        assert tensor_sha(v) == ref[k], k
    sd = dataclasses.asdict(spec)
    sd.pop("name")
    for k, v in build_model(sd, 0).state_dict().items():
        assert tensor_sha(v) == ref[k], k


@pytest.mark.skipif(not os.path.exists("/root/reference/task.py"), reason="reference tree only in the build container")
def test_pins_match_the_live_reference(pins):
    from tests.golden.make_reference_pins import load_reference, make_pins

    live = make_pins(load_reference())
    for name in ("small", "t0", "small_trained"):
        assert live[name]["accuracy"] == pins[name]["accuracy"]
        assert torch.equal(live[name]["argmax"], pins[name]["argmax"])
        if "init_sha256" in pins[name]:
            assert live[name]["init_sha256"] == pins[name]["init_sha256"]


def test_oracle_evaluate_equals_reference_evaluate(pins):
    """The oracle's eval-mode forward + argmax (what tests compare TrainStep.evaluate with elsewhere) gives exactly
    the predictions task.evaluate saw."""
    from oracle.reference_step import TabularMLP, standardise

    for name in ("small", "small_trained"):
        pin = pins[name]
        spec = _spec("small")
        ds = make_synthetic(pin["rows"], spec)
        model = TabularMLP(**pin["spec"])
        model.load_state_dict(pin["params"] if "params" in pin else init_parameters(spec, 0))
        model.eval()
        with torch.no_grad():
            out = model(standardise(ds.dense, ds.mean, ds.std), ds.cat)
        assert torch.equal(out.argmax(1).to(torch.int16), pin["argmax"])
        assert (out.argmax(1) == ds.label).float().mean().item() == pin["accuracy"]


# ------------------------------------------------------------------------------------------------ GPU
def _check_rows(pin, acc, preds):
    """Every row whose two best logits are further apart than bf16 rounding can move them must get the reference's
    prediction; the accuracy may differ by at most the share of the remaining (near-tie) rows."""
    # bf16 path: inputs, weights, hidden activations and logits are rounded to 8 significant bits; three layers
    # -> a logit moves by well under 4 * 2^-8 of the largest logit magnitude of its row
    tol = 4.0 * 2.0 ** -8 * pin["logit_abs_max"]
    decided = pin["gap"] > tol
    ref = pin["argmax"].to(torch.int64)
    assert decided.sum().item() >= 0.5 * ref.numel()
    assert torch.equal(preds[decided], ref[decided]), int((preds[decided] != ref[decided]).sum())
    flips = int((preds != ref).sum())
    assert flips <= int((~decided).sum())
    assert abs(acc - pin["accuracy"]) <= flips / ref.numel() + 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["small", "t0", "small_trained"])
def test_evaluate_row_by_row_against_reference_evaluate(pins, name):
    from b200step.engine import StepConfig, TrainStep
    from b200step.sampler import rank_indices

    pin = pins[name]
    spec = _spec("t0" if name == "t0" else "small")
    ds = make_synthetic(pin["rows"], spec)
    params = pin["params"] if "params" in pin else init_parameters(spec, 0)
    for precision in ("bf16", "fp32"):
        eng = TrainStep(spec, params, ds, rank_indices(pin["rows"], 1, 0),
                        StepConfig(batch_size=256, precision=precision))
        if "init_sha256" in pin:      # untrained: the flat buffer round-trips the seeded init bit for bit
            for k, v in eng.state_dict().items():
                assert tensor_sha(v) == pin["init_sha256"][k[len("module."):]], k
        acc, preds = eng.evaluate(return_predictions=True)
        _check_rows(pin, acc, preds)
        eng.close()
