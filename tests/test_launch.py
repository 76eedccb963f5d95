"""CPU: the per-rank launcher (dace_b200/launch.py) - environment contract, waiting, failure handling."""
import json
import os
import subprocess
import sys
import time

from dace_b200 import launch as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rank_environment_contract():
    env = L.rank_environment(2, 4, [4, 5, 6, 7], '/tmp/x.id', 29999, base={'PATH': '/bin', 'CUDA_VISIBLE_DEVICES': '0,1'})
    assert env['RANK'] == '2' and env['WORLD_SIZE'] == '4' and env['LOCAL_RANK'] == '2'
    assert env['CUDA_VISIBLE_DEVICES'] == '6'          # the rank's GPU becomes device 0 of its process
    assert env['DACE_B200_RENDEZVOUS'] == '/tmp/x.id' and env['MASTER_ADDR'] == '127.0.0.1' and env['MASTER_PORT'] == '29999'
    assert env['PATH'] == '/bin'
    keep = L.rank_environment(0, 2, None, '/tmp/x.id', 1, base={'CUDA_VISIBLE_DEVICES': '0,1'})
    assert keep['CUDA_VISIBLE_DEVICES'] == '0,1'       # --no-isolate


def test_launches_one_process_per_rank(tmp_path):
    script = tmp_path / 'w.py'
    script.write_text(
        'import os, json, sys\n'
        'keys = ["RANK", "WORLD_SIZE", "LOCAL_RANK", "CUDA_VISIBLE_DEVICES", "DACE_B200_RENDEZVOUS", "MASTER_PORT"]\n'
        'json.dump({k: os.environ.get(k) for k in keys}, open(sys.argv[1] + "/rank%s.json" % os.environ["RANK"], "w"))\n')
    r = subprocess.run([sys.executable, '-m', 'dace_b200.launch', '--nproc', '3', '--gpus', '5,6,7', str(script), str(tmp_path)],
                       cwd=ROOT, timeout=120)
    assert r.returncode == 0
    envs = [json.load(open(tmp_path / ('rank%d.json' % i))) for i in range(3)]
    assert [e['RANK'] for e in envs] == ['0', '1', '2'] and {e['WORLD_SIZE'] for e in envs} == {'3'}
    assert [e['CUDA_VISIBLE_DEVICES'] for e in envs] == ['5', '6', '7']
    assert len({e['DACE_B200_RENDEZVOUS'] for e in envs}) == 1 and len({e['MASTER_PORT'] for e in envs}) == 1
    assert not os.path.exists(envs[0]['DACE_B200_RENDEZVOUS'])


def test_first_failure_stops_the_other_ranks(tmp_path):
    script = tmp_path / 'f.py'
    script.write_text('import os, sys, time\n'
                      'if os.environ["RANK"] == "1":\n    sys.exit(3)\n'
                      'time.sleep(60)\n')
    t0 = time.monotonic()
    rc = L.launch([sys.executable, str(script)], 3, isolate=False)
    assert rc == 3 and time.monotonic() - t0 < 30
    slow = tmp_path / 's.py'
    slow.write_text('import time\ntime.sleep(60)\n')
    t0 = time.monotonic()
    assert L.launch([sys.executable, str(slow)], 2, isolate=False, timeout=1.0) == 124 and time.monotonic() - t0 < 30
