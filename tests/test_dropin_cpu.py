"""Row (b) of SURVEY.md §8: the product drops into the reference's run.py path.  The reference's own argument parser
(src/args/argparsing.py:82-88), Experiment constructors (src/experiments/b2t_gru_w2v_experiment.py:78-145,
b2t_gru_w2v_conformer_experiment.py:61-123) and Trainer (src/train/train_loop.py:37-109) run UNCHANGED from
/root/reference on the product model after `compat.install()`; a checkpoint written by the reference's own model
classes loads with strict=True (and vice versa) and both then produce the same logits.  Skipped where the reference
checkout does not exist (the GPU box)."""
import json
import os
import subprocess
import sys

import pytest

REF = os.environ.get("B2T_REFERENCE", "/root/reference")
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src", "experiments")),
                                reason="reference checkout not present")


@pytest.mark.parametrize("kind", ["base", "conformer"])
def test_reference_run_py_path_on_the_product_model(kind):
    p = subprocess.run([sys.executable, os.path.join(REPO, "tests", "dropin_harness.py"), kind], capture_output=True,
                       text=True, timeout=900, cwd=REPO)
    assert p.returncode == 0, p.stdout[-2000:] + "\n" + p.stderr[-4000:]
    rec = json.loads(p.stdout.strip().splitlines()[-1])
    assert rec["kind"] == kind and rec["moved"] > 0 and rec["logits_rel_l2_vs_reference_classes"] < 2e-4
