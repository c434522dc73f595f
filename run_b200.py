#!/usr/bin/env python
"""The reference's run.py (run.py:1-5) on the B200 model: same flags, same Experiment / Trainer code.

    cd /path/to/Wav2Vec2ForBrain && python /path/to/this/repo/run_b200.py --experiment_type=b2p2t_gru+w2v ...
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 run_b200.py --experiment_type=b2p2t_gru+w2v_conformer ...

The reference checkout is the current directory (or env B2T_REFERENCE); it is imported, never modified.
B2T_SYNTHETIC_DATA=<n> replaces the private .mat dataset by n synthetic trials (offline smoke runs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from wav2vec2forbrain_b200 import compat  # noqa: E402

if __name__ == "__main__":
    compat.install()
    if os.environ.get("B2T_SYNTHETIC_DATA"):
        compat.install_synthetic_dataset(int(os.environ["B2T_SYNTHETIC_DATA"]))
    from src.args.argparsing import get_experiment_from_args
    experiment = get_experiment_from_args()
    experiment.run()
