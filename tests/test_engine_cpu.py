"""Host-side schedule of the engine (operand views, epilogue wiring, hand-derived backward)
checked against the live-reference goldens through the CPU emulator of the C ABI."""
import pytest
import torch

from tests import abi_emulator
from tests.engine_util import check_against_golden, run_engine_train_step
from tests.golden_util import load_case, rel_l2


@pytest.fixture()
def engine_mod(monkeypatch):
    from wav2vec2forbrain_b200 import engine
    monkeypatch.setattr(engine, "ops", abi_emulator)
    return engine


@pytest.mark.parametrize("name", ["tiny_base", "tiny_base_h0_uni", "tiny_base_stable_ln", "tiny_conformer"])
def test_engine_schedule_matches_reference(engine_mod, name):
    rec, _, _, sd, batch = load_case(name)
    logits, loss, lens, G, feats, _, _ = run_engine_train_step(engine_mod, rec, sd, batch, "cpu")
    assert rel_l2(feats.view(rec["features"].shape), rec["features"]) < 5e-5
    check_against_golden(rec, logits, loss, lens, G, tol_logits=5e-5, tol_grad=3e-4)
