"""BASELINE.json's full sizes (T0, 65536 rows per GPU) — a short oracle comparison plus size-independent properties:
softmax-gradient rows sum to zero, wgrad is additive over row halves, the forward GEMM is linear, preprocessing is
bit-exact at full size, identical seeds give the same loss curve."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200step import ops  # noqa: E402
from b200step.data import make_synthetic  # noqa: E402
from b200step.engine import StepConfig, TrainStep  # noqa: E402
from b200step.model import ARCHS, init_parameters  # noqa: E402
from b200step.sampler import rank_indices  # noqa: E402

B = 65536
SPEC = ARCHS["tabular_mlp"]


@pytest.fixture(scope="module")
def dataset():
    return make_synthetic(2 * B, SPEC)


def test_two_full_size_steps_against_the_oracle(dataset):
    import dataclasses

    from oracle.reference_step import OracleTrainer

    sd = dataclasses.asdict(SPEC)
    sd.pop("name")
    oracle = OracleTrainer(sd, dataset, world=1, batch_size=B, precision="bf16")
    eng = TrainStep(SPEC, init_parameters(SPEC, 0), dataset, rank_indices(2 * B, 1, 0),
                    StepConfig(batch_size=B, precision="bf16"))
    for _ in range(2):
        lo = oracle.step()[0]
        eng.step()
        assert abs(lo - eng.loss_value()) <= 2e-3 * max(1.0, abs(lo))
    so, sg = oracle.state_dict(), eng.state_dict()
    for k in so:
        rel = ((so[k] - sg[k]).norm() / so[k].norm().clamp_min(1e-12)).item()
        assert rel <= 5e-3, (k, rel)


def test_full_size_properties(dataset):
    eng = TrainStep(SPEC, init_parameters(SPEC, 0), dataset, rank_indices(2 * B, 1, 0),
                    StepConfig(batch_size=B, precision="bf16", use_graph=False, prefetch_pre=False))
    eng.step()
    eng.synchronize()
    # (1) preprocessing at full size: bit-exact against the torch restatement of the rule (sub then div)
    r = eng.row_idx[:B]
    dense_std = eng.ds.dense[r].sub(eng.ds.mean).div(eng.ds.std).to(torch.bfloat16)
    ne = SPEC.num_cat * SPEC.embed_dim
    assert torch.equal(eng.x0[:, ne: ne + SPEC.num_dense].view(torch.int16), dense_std.view(torch.int16))
    assert torch.equal(eng.y, eng.ds.label[r]) and torch.equal(eng.idx.long(), eng.ds.cat[r])
    assert torch.count_nonzero(eng.x0[:, ne + SPEC.num_dense:]).item() == 0
    # (2) every softmax-gradient row sums to zero (softmax - onehot), up to bf16 rounding of 1000 terms
    row_sums = eng.dlogits.float().sum(1) * B
    assert row_sums.abs().max().item() <= 2e-2
    # (3) wgrad is additive over the two halves of the batch
    dy, x = eng.dlogits, eng.acts[1]
    full = torch.zeros(1000, 512, device="cuda")
    halves = torch.zeros(1000, 512, device="cuda")
    ops.linear_wgrad(dy, x, full, M=B, N=1000, K=512, accumulate=True)
    ops.linear_wgrad(dy[: B // 2], x[: B // 2], halves, M=B // 2, N=1000, K=512, accumulate=True)
    ops.linear_wgrad(dy[B // 2:], x[B // 2:], halves, M=B // 2, N=1000, K=512, accumulate=True)
    torch.cuda.synchronize()
    assert (full - halves).abs().max().item() <= 1e-5 * max(1.0, full.abs().max().item())
    # (4) the forward GEMM is linear in its input: f(2x) == 2 f(x) exactly in bf16 (power-of-two scaling)
    w, xin = eng._w(1), eng.acts[0]
    y1 = torch.empty(B, 512, dtype=torch.bfloat16, device="cuda")
    y2 = torch.empty_like(y1)
    ops.linear_fwd(xin, w, None, y1, M=B, N=512, K=512, relu=False)
    ops.linear_fwd((xin.float() * 2).to(torch.bfloat16), w, None, y2, M=B, N=512, K=512, relu=False)
    torch.cuda.synchronize()
    assert torch.equal((y1.float() * 2).to(torch.bfloat16), y2)


def test_same_seed_same_loss_curve(dataset):
    def run():
        eng = TrainStep(SPEC, init_parameters(SPEC, 0), dataset, rank_indices(2 * B, 1, 0),
                        StepConfig(batch_size=B, precision="bf16"))
        out = []
        for _ in range(4):
            eng.step()
            out.append(eng.loss_value())
        return out

    a, b = run(), run()
    assert a[0] == b[0]                                       # first loss: no atomics upstream, bit-identical
    for x, y in zip(a, b):                                    # later: float-atomic summation order may differ
        assert abs(x - y) <= 1e-5 * max(1.0, abs(x))
    assert a[-1] < a[0]
