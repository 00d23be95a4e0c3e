"""N>1 host logic on CPU: two `gloo` ranks.  The sharded occupancy update
(each rank evaluates its slice, decays the whole sample, MAX all-reduce) must equal the
single-process update bit for bit; bench.py's aggregation is MAX over times, SUM over units."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_rank_update_and_aggregation(tmp_path):
    out = tmp_path / "r0.json"
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NFB_NO_BUILD="1",
                   OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "_dist_worker.py"), str(out)],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    logs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            p.kill()
            o, _ = p.communicate()
        logs.append(o.decode(errors="replace"))
    assert all(p.returncode == 0 for p in procs), "\n".join(logs)
    r = json.loads(out.read_text())
    assert r == {"update": True, "estimator": True, "parts": True, "ms": 11.0, "units": 3000, "distinct": True, "world": 2}
