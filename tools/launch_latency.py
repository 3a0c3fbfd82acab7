"""Submit-to-first-step latency of the launcher (BASELINE.json's third headline metric).

Runs the SAME job script (rendezvous over gloo on 127.0.0.1, one all-reduce, "first step" mark)
  (a) through this repo's launcher (client -> ApplicationMaster -> TaskExecutors, pytorch runtime)
  (b) through ``python -m torch.distributed.run`` (what a stock PyTorch user would type)
and prints one JSON line per arm: wall time from submit to the first completed step.  The reference's
own floor for this path is structural: 1 s client poll + 1 s AM-RM heartbeat + 3 s gang-barrier poll
on top of JVM starts; it cannot be run here (no JDK / YARN).

``python tools/launch_latency.py [--workers 2] [--repeat 3]`` — CPU only, no GPU needed.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

JOB = r'''
import os, sys, time
sys.path.insert(0, os.environ["TONY_REPO"])
import torch, torch.distributed as dist
dist.init_process_group("gloo")
t = torch.ones(4)
dist.all_reduce(t)
assert t[0].item() == dist.get_world_size()
stamp = os.environ.get("LATENCY_STAMP_DIR")
if stamp:
    with open(os.path.join(stamp, f"first.{dist.get_rank()}"), "w") as f:
        f.write(f"{time.time():.6f}")
from tony_b200 import runtime
runtime.mark_first_step()
dist.destroy_process_group()
'''


def ours(workers: int, script: str, work: str) -> float:
    from tony_b200.cluster.mini import MiniCluster

    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.setdefault("TONY_B200_FAKE_GPUS", "0")
    with MiniCluster(num_node_managers=workers) as cluster:
        client = cluster.new_client()
        assert client.init(["--python_binary_path", sys.executable, "--executes", script,
                            "--conf", "tony.application.framework=pytorch",
                            "--conf", f"tony.worker.instances={workers}",
                            "--conf", "tony.am.wait-client-signal-stop-timeout-sec=1",
                            "--shell_env", f"TONY_REPO={REPO}"])
        rc = client.start()
        assert rc == 0, "launcher job failed"
        try:
            with open(os.path.join(client.app_dir, "timeline.json")) as f:
                ours.timeline = [(e["event"], e["t_ms"]) for e in json.load(f)]
        except OSError:
            ours.timeline = []
        return float(client.submit_to_first_step_ms)


def torchrun(workers: int, script: str, work: str) -> float:
    stamp = tempfile.mkdtemp(dir=work)
    env = dict(os.environ, TONY_REPO=REPO, LATENCY_STAMP_DIR=stamp)
    t0 = time.time()
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={workers}",
                    "--master-addr", "127.0.0.1", "--master-port", "29731", script], env=env, check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    firsts = [float(open(os.path.join(stamp, f)).read()) for f in os.listdir(stamp)]
    return 1000.0 * (min(firsts) - t0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--arms", default="ours,torchrun", help="comma list of: ours, torchrun")
    a = ap.parse_args()
    arms = set(a.arms.split(","))
    work = tempfile.mkdtemp(prefix="tony_latency_")
    script = os.path.join(work, "job.py")
    with open(script, "w") as f:
        f.write(JOB)
    for name, fn in (("tony_b200 launcher (pytorch runtime)", ours), ("torch.distributed.run", torchrun)):
        if fn.__name__ not in arms:
            continue
        vals = [fn(a.workers, script, work) for _ in range(a.repeat)]
        print(json.dumps({"arm": name, "workers": a.workers, "submit_to_first_step_ms": [round(v) for v in vals],
                          "median_ms": round(statistics.median(vals))}), flush=True)
        if fn is ours and getattr(ours, "timeline", None):
            print(json.dumps({"launcher_timeline_ms_since_submit": ours.timeline}), flush=True)


if __name__ == "__main__":
    main()
