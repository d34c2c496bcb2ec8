"""Runs the REFERENCE's own driver scripts, unmodified, against this package (CPU; build container only).

applications/rdiff/dino_paper/fast_training.py and dino_training.py do ``import tensorflow as tf;
tf.disable_v2_behavior(); sys.path.append(os.environ['DINO_PATH']); from dino import *``, parse their single-dash
flags, fill the settings dict and call ``jacobian_training_driver(settings)``.  Here each script is executed from its
own directory with DINO_PATH = this repository and dino_b200/compat (the no-op tensorflow stand-in) on PYTHONPATH, with
the driver replaced by a recorder: the import block, the argparse -> settings mapping and the settings contract of
``jacobian_network_settings`` all run for real; only the data-dependent training call is intercepted (the PDE data
files do not exist here)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
APP = "/root/reference/applications/rdiff/dino_paper"

RUNNER = r'''
import json, os, runpy, sys
sys.path.insert(0, os.environ["DINO_PATH"])
import dino
rec = {}
def recorder(settings, *a, **k):
    rec["settings"] = {k: v for k, v in settings.items() if k != "problem_settings"}
    rec["problem_settings_type"] = type(settings["problem_settings"]).__name__
    return None
dino.jacobian_training_driver = recorder
dino.observable_training_driver = recorder
sys.argv = [sys.argv[1]] + sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
print("SETTINGS_JSON " + json.dumps(rec, default=str))
'''


def _run(script, flags):
    env = dict(os.environ, DINO_PATH=ROOT, PYTHONPATH=os.path.join(ROOT, "dino_b200", "compat"))
    out = subprocess.run([sys.executable, "-c", RUNNER, script] + flags, cwd=APP, env=env, capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("SETTINGS_JSON ")][-1]
    return json.loads(line[len("SETTINGS_JSON "):])


@pytest.mark.skipif(not os.path.isdir(APP), reason="the reference tree is only present in the build container")
def test_fast_training_script_runs_unchanged_up_to_the_driver_call():
    rec = _run("fast_training.py", ["-network_name", "cli_test", "-fixed_input_rank", "200", "-fixed_output_rank", "100",
                                    "-total_epochs", "7", "-train_data_size", "256", "-batch_rank", "10", "-target_rank", "100",
                                    "-train_full_jacobian", "0"])
    s = rec["settings"]
    assert rec["problem_settings_type"] == "dict"
    assert s["network_name"] == "cli_test" and s["fixed_input_rank"] == 200 and s["fixed_output_rank"] == 100
    assert s["reduced_input_training"] is True and s["reduced_output_training"] is False
    assert s["train_data_size"] == 256 and s["architecture"] == "rb_dense" and s["depth"] == 6
    # batch_rank != target_rank: the column-reshuffling outer loop (fast_training.py:108-115)
    assert s["shuffle_every_epoch"] is True and s["outer_epochs"] == 7 and s["opt_parameters"]["keras_epochs"] == 1
    assert s["opt_parameters"]["loss_weights"] == [1.0, 1.0] and s["opt_parameters"]["train_full_jacobian"] == 0
    assert s["data_dir"] == "../data/rdiff_nobs_50_g0.1_d1.0_nx64/"


@pytest.mark.skipif(not os.path.isdir(APP), reason="the reference tree is only present in the build container")
def test_dino_training_script_runs_unchanged_up_to_the_driver_call():
    rec = _run("dino_training.py", ["-network_name", "cli_full", "-architecture", "generic_dense", "-total_epochs", "3"])
    s = rec["settings"]
    assert s["architecture"] == "generic_dense" and s["network_name"] == "cli_full"
    assert s["opt_parameters"]["keras_epochs"] == 3 or s["outer_epochs"] == 3
