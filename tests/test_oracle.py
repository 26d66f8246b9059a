"""CPU tests of the oracle itself: it must agree with torch's own autograd / DDP / sampler (the libraries the
reference's arithmetic lives in) and with the committed golden vectors, before it is trusted as the checker."""
import dataclasses
import os

import pytest
import torch

from b200step.data import make_synthetic
from b200step.model import ARCHS, FlatLayout, init_parameters
from b200step.sampler import batch_bounds, rank_indices
from oracle.philox import dropout_keep_mask, philox4x32_10
from oracle.reference_step import OracleTrainer, rank_batches, run_ddp_gloo, standardise

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
SMALL = dataclasses.replace(ARCHS["tabular_mlp"], cardinality=50, hidden=(64, 32), num_classes=40)


def _sd(spec):
    d = dataclasses.asdict(spec)
    d.pop("name")
    return d


def test_philox_known_answers():
    # Random123 kat_vectors: philox4x32-10
    r = philox4x32_10([0], [0], [0], [0], 0, 0)
    assert [int(x[0]) for x in r] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    r = philox4x32_10([0xFFFFFFFF], [0xFFFFFFFF], [0xFFFFFFFF], [0xFFFFFFFF], 0xFFFFFFFF, 0xFFFFFFFF)
    assert [int(x[0]) for x in r] == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    r = philox4x32_10([0x243F6A88], [0x85A308D3], [0x13198A2E], [0x03707344], 0xA4093822, 0x299F31D0)
    assert [int(x[0]) for x in r] == [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]
    m = dropout_keep_mask(256, 512, 0.25, 1234, 3, 2)
    assert abs(m.mean() - 0.75) < 0.01


def test_sampler_restatement_equals_torch_classes():
    """b200step.sampler (product) == DistributedSampler + DataLoader (what task.py:260-263 instantiates)."""
    for n, world, bs in [(10000, 6, 1024), (10000, 8, 1024), (1400, 1, 256), (7, 3, 2), (3, 8, 4)]:
        for rank in range(world):
            ref = rank_batches(n, world, rank, bs)
            mine = rank_indices(n, world, rank)
            bounds = batch_bounds(mine.numel(), bs)
            assert len(ref) == len(bounds)
            for rb, bb in zip(ref, bounds):
                assert torch.equal(rb, mine[bb.start: bb.stop])
    # SURVEY.md App. D probe: N=10000, W=6 -> 1667 per rank, rank 0 starts with 6044
    assert rank_indices(10000, 6, 0).numel() == 1667 and rank_indices(10000, 6, 0)[0].item() == 6044
    with pytest.raises(ValueError):
        rank_indices(10, 2, 2)


def test_sampler_set_epoch_equals_torch():
    """`task.py --set_epoch` (additive): epoch e's order == DistributedSampler.set_epoch(e)."""
    from torch.utils.data.distributed import DistributedSampler

    class _DS(torch.utils.data.Dataset):
        def __len__(self):
            return 1003

        def __getitem__(self, i):
            return i

    for world in (1, 3, 8):
        for rank in range(world):
            smp = DistributedSampler(_DS(), num_replicas=world, rank=rank)
            for epoch in (0, 1, 7):
                smp.set_epoch(epoch)
                assert list(smp) == rank_indices(1003, world, rank, epoch=epoch).tolist()
    assert not torch.equal(rank_indices(1003, 1, 0, epoch=0), rank_indices(1003, 1, 0, epoch=1))


def test_init_matches_module_construction_order():
    ds = make_synthetic(64, SMALL)
    tr = OracleTrainer(_sd(SMALL), ds, batch_size=32, seed=3)
    mine = init_parameters(SMALL, 3)
    for k, v in tr.model.state_dict().items():
        assert torch.equal(v, mine[k]), k
    lay = FlatLayout(SMALL)
    flat = lay.flatten(mine)
    back = lay.unflatten(flat)
    assert list(back.keys()) == list(tr.state_dict().keys())
    for k, v in tr.state_dict().items():
        assert torch.equal(v, back[k]), k
    assert SMALL.in_features_padded % 64 == 0 and lay.total % 64 == 0


def test_manual_backward_equals_autograd():
    ds = make_synthetic(512, SMALL)
    a = OracleTrainer(_sd(SMALL), ds, world=2, batch_size=64, precision="fp32")
    b = OracleTrainer(_sd(SMALL), ds, world=2, batch_size=64, precision="fp32", use_autograd=False)
    for _ in range(6):
        la, lb = a.step(), b.step()
        assert max(abs(x - y) for x, y in zip(la, lb)) < 1e-5
    for k, v in a.state_dict().items():
        torch.testing.assert_close(v, b.state_dict()[k], rtol=1e-5, atol=1e-6)


def test_bf16_oracle_tracks_fp32_oracle():
    ds = make_synthetic(512, SMALL)
    a = OracleTrainer(_sd(SMALL), ds, batch_size=128, precision="fp32")
    b = OracleTrainer(_sd(SMALL), ds, batch_size=128, precision="bf16")
    for _ in range(8):
        la, lb = a.step()[0], b.step()[0]
        assert abs(la - lb) <= 2e-2 * abs(la)  # bf16 loss rel 2e-2 per step
    pa = torch.cat([v.flatten() for v in a.state_dict().values()])
    pb = torch.cat([v.flatten() for v in b.state_dict().values()])
    assert torch.nn.functional.cosine_similarity(pa, pb, dim=0).item() >= 0.999


def test_single_process_emulation_equals_ddp_gloo_world2():
    """BASELINE config 1: real DistributedDataParallel over gloo, world_size=2 (task.py:148,194)."""
    ds = make_synthetic(512, SMALL)
    emu = OracleTrainer(_sd(SMALL), ds, world=2, batch_size=64, precision="fp32")
    losses = [emu.step() for _ in range(5)]
    out = run_ddp_gloo(_sd(SMALL), ds, world=2, batch_size=64, steps=5, threads=2, port=29541)
    for r in range(2):
        for s in range(5):
            assert abs(out[r]["losses"][s] - losses[s][r]) < 1e-5
    for k, v in emu.state_dict().items():
        torch.testing.assert_close(v, out[0]["state"][k], rtol=1e-5, atol=1e-6)
        assert torch.equal(out[0]["state"][k], out[1]["state"][k])  # replicas stay identical


def test_dropout_oracle_reproducible_and_active():
    spec = dataclasses.replace(SMALL, dropout=0.25)
    ds = make_synthetic(256, spec)
    a = OracleTrainer(_sd(spec), ds, batch_size=64, precision="fp32")
    b = OracleTrainer(_sd(spec), ds, batch_size=64, precision="fp32", use_autograd=False)
    c = OracleTrainer(_sd(SMALL), ds, batch_size=64, precision="fp32")
    la, lb, lc = a.step()[0], b.step()[0], c.step()[0]
    assert abs(la - lb) < 1e-5 and abs(la - lc) > 1e-4


def test_standardise_rule():
    x = torch.tensor([[1.0, 2.0], [3.0, 5.0]])
    m, s = torch.tensor([2.0, 3.5]), torch.tensor([0.5, 3.0])
    assert torch.equal(standardise(x, m, s), (x - m) / s)


@pytest.mark.parametrize("name", ["small_fp32_sgd", "small_bf16_sgd", "small_bf16_adam"])
def test_golden_vectors(name):
    """Committed fixtures (tests/golden/make_golden.py): the oracle must keep reproducing them."""
    from tests.golden.make_golden import CASES, run_case

    blob = torch.load(os.path.join(GOLDEN, f"{name}.pt"))
    got = run_case(CASES[name])
    assert torch.equal(got["row_idx"], blob["row_idx"])
    assert torch.equal(got["x0_bits"], blob["x0_bits"])           # preprocessing: bit-exact
    torch.testing.assert_close(got["losses"], blob["losses"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(got["param_checksums"], blob["param_checksums"], rtol=1e-4, atol=1e-5)


def test_bucket_plan_covers_flat_buffer_and_respects_cap():
    """DDP-style buckets (first 1 MiB, then 25 MiB): contiguous, disjoint, cover every parameter, in backward order;
    the embedding tables stay one bucket (one large all-reduce measured faster than 17 capped ones)."""
    for name, spec in ARCHS.items():
        lay = FlatLayout(spec)
        b = lay.buckets()
        covered = sorted((s, e) for s, e, _ in b)
        assert covered[0][0] == 0 and covered[-1][1] == lay.total
        for (s0, e0), (s1, e1) in zip(covered, covered[1:]):
            assert e0 == s1
        assert all((e - s) * 4 <= (25 << 20) for s, e, r in b if r != -1)
        readies = [r for _, _, r in b]
        assert readies[0] == len(spec.layer_dims) - 1            # output layer first, like DDP's reverse order
        assert readies.count(-1) == 1


def test_bench_reference_arm_config1_line():
    """`bench.py --impl reference --workload t0_fp32_b4096_cpu_gloo2` (BASELINE config 1: the oracle under real DDP over
    gloo, 2 CPU processes) prints one JSON line with the reference arm's keys."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--workload",
                        "t0_fp32_b4096_cpu_gloo2", "--steps", "1", "--warmup", "1"], capture_output=True, text=True,
                       timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "samples/s" and line["value"] > 0
    assert line["config"]["ranks"] == 2 and line["config"]["backend"] == "gloo"
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
