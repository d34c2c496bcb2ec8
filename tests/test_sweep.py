"""CPU tests of the sweep harness (dino_b200/sweep.py): its command list is the reference's, and the scheduler keeps
the devices busy with one job each."""
import os
import runpy
import sys
import time

import pytest

from dino_b200 import sweep

REF = "/root/reference/applications/rdiff/dino_paper/training_runs.py"


@pytest.mark.skipif(not os.path.isfile(REF), reason="the reference tree is only present in the build container")
def test_command_list_is_the_references(monkeypatch):
    """Execute the reference's training_runs.py with os.system recording instead of running: same commands, same order."""
    seen = []
    monkeypatch.setattr(os, "system", lambda c: seen.append(c) or 0)
    runpy.run_path(REF, run_name="__main__")
    assert len(seen) == 60
    assert sweep.paper_sweep_commands() == seen


def test_sweep_shape_without_the_reference():
    cmds = sweep.paper_sweep_commands()
    assert len(cmds) == 60 and sum("fast_training.py" in c for c in cmds) == 10
    assert cmds[0].startswith("python dino_training.py -total_epochs 100 -architecture rb_dense ")
    assert "-network_name DIPNet100-50_ndata16l2 " in cmds[0] and "-batch_rank 10 " in cmds[3]


def test_scheduler_one_job_per_device(tmp_path):
    py = sys.executable
    cmds = [f'{py} -c "import os,time; time.sleep(0.3); open(r\'{tmp_path}/j{i}\',\'w\').write(os.environ[\'CUDA_VISIBLE_DEVICES\'])"'
            for i in range(6)]
    t0 = time.time()
    recs = sweep.run_jobs(cmds, n_gpus=3, log_dir=str(tmp_path / "logs"), poll=0.02)
    wall = time.time() - t0
    assert [r["returncode"] for r in recs] == [0] * 6
    gpus = [open(tmp_path / f"j{i}").read() for i in range(6)]
    assert sorted(gpus) == ["0", "0", "1", "1", "2", "2"]
    assert gpus[:3] == [str(r["gpu"]) for r in recs[:3]] and len(set(gpus[:3])) == 3     # the first wave fills every device
    assert wall < 6 * 0.3 + 1.5                                                         # two waves, not six serial jobs
    bad = sweep.run_jobs([f'{py} -c "raise SystemExit(3)"'], n_gpus=1)
    assert bad[0]["returncode"] == 3
