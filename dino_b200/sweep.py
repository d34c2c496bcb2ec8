"""The paper's training sweep as concurrent jobs over the GPUs of one box.

The reference's ``applications/*/dino_paper/training_runs.py`` (lines 20-157; the three copies are identical) builds one
command line per run - 5 training-set sizes x [2 DIPNet rank pairs + generic_dense] x 4 formulations (L2, full H1,
truncated H1, truncated H1 with matrix subsampling) = 60 runs of 100 epochs - and executes them one after another
with ``os.system``.  The runs are independent, so here the same command list (``paper_sweep_commands``, pinned to the
reference's by tests/test_sweep.py) is handed to a small scheduler that keeps one job per GPU running
(``CUDA_VISIBLE_DEVICES`` per job); a training step of these small networks leaves most of a B200 idle, so
``jobs_per_gpu`` > 1 packs several on one device.

  python -m dino_b200.sweep --gpus 8 --cwd applications/rdiff/dino_paper [--jobs-per-gpu 2] [--dry-run]
"""
import argparse
import os
import subprocess
import sys
import time
from typing import Dict, List, Optional, Sequence


def default_settings() -> Dict:
    """training_runs.py:20-40, key order included (the order fixes the command line)."""
    s = {}
    s['total_epochs'] = 100
    s['architecture'] = 'rb_dense'
    s['input_basis'] = 'as'
    s['output_basis'] = 'jjt'
    s['fixed_input_rank'] = 100
    s['fixed_output_rank'] = 50
    s['train_full_jacobian'] = 1
    s['l2_weight'] = 1.0
    s['h1_weight'] = 1.0
    s['target_rank'] = 100
    s['batch_rank'] = 100
    return s


def build_command(settings: Dict, command: str = 'python dino_training.py') -> str:
    """training_runs.py:42-49."""
    return command + ' ' + ''.join(f'-{k} {v} ' for k, v in settings.items())


NDATAS = [16, 64, 256, 1024, 4096]          # training_runs.py:56
RB_DIMS = [(100, 50), (200, 100)]           # training_runs.py:62


def _four_formulations(settings: Dict, name: str, ndata: int, fast_full_h1: bool) -> List[str]:
    """The four runs of one (architecture, ndata) cell, with the reference's sticky settings: 'batch_rank' = 10 of
    the last run is set on a dict that is rebuilt for every cell (training_runs.py:64-109, 116-157)."""
    out = []
    base = name + '_ndata' + str(ndata)
    settings['h1_weight'] = 0.0
    settings['network_name'] = base + 'l2'
    out.append(build_command(settings))
    settings['h1_weight'] = 1.0
    settings['train_full_jacobian'] = 1
    settings['network_name'] = base + 'fullh1'
    out.append(build_command(settings, 'python fast_training.py') if fast_full_h1 else build_command(settings))
    settings['train_full_jacobian'] = 0
    settings['network_name'] = base + 'th1'
    out.append(build_command(settings))
    settings['batch_rank'] = 10
    settings['network_name'] = base + 'th1ms'
    out.append(build_command(settings))
    return out


def paper_sweep_commands() -> List[str]:
    """Every command training_runs.py passes to os.system, in its order."""
    cmds = []
    for ndata in NDATAS:
        for rM, rQ in RB_DIMS:
            s = default_settings()
            s['architecture'] = 'rb_dense'
            s['train_data_size'] = ndata
            s['fixed_input_rank'] = rM
            s['fixed_output_rank'] = rQ
            cmds += _four_formulations(s, 'DIPNet' + str(rM) + '-' + str(rQ), ndata, fast_full_h1=True)
    for ndata in NDATAS:
        s = default_settings()
        s['architecture'] = 'generic_dense'
        s['train_data_size'] = ndata
        cmds += _four_formulations(s, 'generic_dense', ndata, fast_full_h1=False)
    return cmds


def run_jobs(commands: Sequence[str], n_gpus: int, jobs_per_gpu: int = 1, cwd: Optional[str] = None,
             env: Optional[Dict[str, str]] = None, log_dir: Optional[str] = None, poll: float = 0.2) -> List[Dict]:
    """Run shell ``commands`` with at most ``jobs_per_gpu`` at a time on each of ``n_gpus`` devices; a job sees only its
    own device (CUDA_VISIBLE_DEVICES).  Jobs start in list order as slots free up.  Returns one record per command:
    {'command', 'gpu', 'returncode', 'seconds'} in list order."""
    assert n_gpus >= 1 and jobs_per_gpu >= 1
    slots = [g for g in range(n_gpus) for _ in range(jobs_per_gpu)]
    free = list(reversed(slots))
    running = []                                  # (proc, index, gpu, t0, logfile)
    records: List[Optional[Dict]] = [None] * len(commands)
    nxt = 0
    if log_dir:
        os.makedirs(log_dir, exist_ok=True)
    while nxt < len(commands) or running:
        while free and nxt < len(commands):
            gpu = free.pop()
            e = dict(os.environ if env is None else env)
            e['CUDA_VISIBLE_DEVICES'] = str(gpu)
            logf = open(os.path.join(log_dir, f'job_{nxt:03d}.log'), 'w') if log_dir else None
            p = subprocess.Popen(commands[nxt], shell=True, cwd=cwd, env=e, stdout=logf or subprocess.DEVNULL,
                                 stderr=subprocess.STDOUT)
            running.append((p, nxt, gpu, time.time(), logf))
            nxt += 1
        still = []
        for p, i, gpu, t0, logf in running:
            rc = p.poll()
            if rc is None:
                still.append((p, i, gpu, t0, logf))
                continue
            if logf:
                logf.close()
            records[i] = {'command': commands[i], 'gpu': gpu, 'returncode': rc, 'seconds': time.time() - t0}
            free.append(gpu)
        running = still
        if running:
            time.sleep(poll)
    return records  # type: ignore[return-value]


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split('\n')[0])
    ap.add_argument('--gpus', type=int, default=8)
    ap.add_argument('--jobs-per-gpu', type=int, default=1)
    ap.add_argument('--cwd', default='.', help="an applications/<problem>/dino_paper directory")
    ap.add_argument('--log-dir', default='sweep_logs')
    ap.add_argument('--dry-run', action='store_true')
    args = ap.parse_args(argv)
    cmds = paper_sweep_commands()
    if args.dry_run:
        print('\n'.join(cmds))
        return 0
    recs = run_jobs(cmds, args.gpus, args.jobs_per_gpu, cwd=args.cwd, log_dir=args.log_dir)
    bad = [r for r in recs if r['returncode'] != 0]
    print(f'{len(recs) - len(bad)} of {len(recs)} runs finished; wall time of the longest {max(r["seconds"] for r in recs):.1f} s')
    for r in bad:
        print('FAILED', r['returncode'], r['command'])
    return 1 if bad else 0


if __name__ == '__main__':
    sys.exit(main())
