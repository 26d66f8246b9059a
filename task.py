#!/usr/bin/env python
"""task.py — drop-in replacement for the reference's training script (reference: task.py, same file name).

Same command line, environment variables, launch topology, stdout lines and checkpoint file as the reference;
the training step behind them is the B200-native one (hand-written sm_100a CUDA behind libb200step.so).
This file is deliberately tiny so that it can be shipped alone through Vertex AI's
CustomTrainingJob(script_path=...) (pytorch-pipeline.ipynb:167-171): the package is looked up next to this file,
then in $B200STEP_HOME (an image path baked into TRAIN_IMAGE, pytorch-pipeline.ipynb:137-138).
"""
import os
import sys


def _find_package() -> str:
    here = os.path.dirname(os.path.abspath(__file__))
    cands = [os.path.join(here, "kubeflow-v2-distributed-pytorch_b200")]
    if os.environ.get("B200STEP_HOME"):
        home = os.environ["B200STEP_HOME"]
        cands += [os.path.join(home, "kubeflow-v2-distributed-pytorch_b200"), home]
    for c in cands:
        if os.path.isdir(os.path.join(c, "b200step")):
            return c
    raise ImportError("b200step package not found next to task.py or in $B200STEP_HOME: " + ", ".join(cands))


_pkg = _find_package()
if _pkg not in sys.path:
    sys.path.insert(0, _pkg)

from b200step.cli import main, main_worker  # noqa: E402,F401  (main_worker must be importable for mp.spawn)

if __name__ == "__main__":
    main()
