"""bench.py's stdout contract: one JSON line, whatever native libraries print on file descriptor 1 during the run
(NCCL prints its version banner there under torchrun)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_stdout_is_one_json_line_despite_library_noise():
    code = ("import os, bench\n"
            "bench.claim_stdout()\n"
            "print('python noise')\n"
            "os.write(1, b'native noise\\n')\n"
            "os.system('echo child noise')\n"
            "bench.emit({'metric': 'm', 'value': 1.5})\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert r.stdout.count("\n") == 1 and json.loads(r.stdout) == {"metric": "m", "value": 1.5}
    for noise in ("python noise", "native noise", "child noise"):
        assert noise in r.stderr


def test_emit_without_claim_prints_to_stdout():
    code = "import bench\nbench.emit({'impl': 'reference'})\n"
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and json.loads(r.stdout) == {"impl": "reference"}
