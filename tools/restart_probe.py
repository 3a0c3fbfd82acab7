#!/usr/bin/env python
"""BASELINE config 5 at full size: GPT-2 medium DDP on every GPU of the box through the launcher, the chief
dies after step 6, the ApplicationMaster retries the whole session and the job resumes from the sharded
checkpoint.  Prints one JSON line: wall time of the job, time from the chief's death to the first step of
the resumed session ("time to recover"), and the launcher timeline."""
import json
import os
import re
import sys
import tempfile
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch

from tony_b200.cluster.mini import MiniCluster


def main():
    n = torch.cuda.device_count()
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    ckpt = tempfile.mkdtemp(prefix="tony_restart_")
    with MiniCluster() as cluster:
        client = cluster.new_client()
        argv = ["--executes", os.path.join(REPO, "examples", "gpt2-medium-ddp", "gpt2_train.py"),
                "--python_binary_path", sys.executable,
                "--task_params", f"--steps 14 --batch 8 --seq 1024 --ckpt_every 4 --crash_at 6 --ckpt_dir {ckpt}",
                "--shell_env", f"TONY_REPO={REPO}", "--shell_env", "TONY_B200_COMM_TIMEOUT_MS=60000",
                "--conf", "tony.am.wait-client-signal-stop-timeout-sec=2",
                "--conf", "tony.application.framework=pytorch", "--conf", f"tony.worker.instances={n}",
                "--conf", "tony.worker.gpus=1", "--conf", "tony.am.retry-count=1"]
        assert client.init(argv)
        t0 = time.time()
        rc = client.start()
        wall = time.time() - t0
        text = ""
        cdir = os.path.join(client.app_dir, "containers")
        stamps = {}
        for c in sorted(os.listdir(cdir)) if os.path.isdir(cdir) else []:
            for f in ("stdout", "stderr"):
                p = os.path.join(cdir, c, f)
                if os.path.exists(p):
                    body = open(p, errors="replace").read()
                    text += body
                    if "simulating a crash" in body:
                        stamps["crash_mtime"] = os.path.getmtime(p)
                    if "resumed from step" in body:
                        stamps.setdefault("resume_files", []).append(p)
        timeline = []
        try:
            timeline = json.load(open(os.path.join(client.app_dir, "timeline.json")))
        except (OSError, ValueError):
            pass
        firsts = [e["t_ms"] for e in timeline if e.get("event") == "first_step"]
        resets = [e["t_ms"] for e in timeline if e.get("event") == "session_reset"]
        fails = [e["t_ms"] for e in timeline if e.get("event") == "task_failed"]
        recover = None
        if fails and resets:
            after = [t for t in firsts if t > resets[0]]
            if after:
                recover = round((min(after) - min(fails)) / 1000.0, 2)
        m = re.search(r"done: ([0-9.]+) samples/s", text)
        out = {"config": f"GPT-2 medium DDP, {n} x B200, seq 1024, 8 samples/GPU, chief exits at step 6, resume from step 4",
               "rc": rc, "job_wall_s": round(wall, 1), "resumed": "resumed from step 4" in text,
               "crashed": "simulating a crash" in text, "samples_per_sec_after_resume": float(m.group(1)) if m else None,
               "time_to_recover_s": recover, "chief_failure_ms": fails[:1], "session_reset_ms": resets,
               "first_step_ms_since_submit": [round(x) for x in firsts],
               "timeline_events": [(e.get("event"), e.get("t_ms")) for e in timeline][-40:]}
        print(json.dumps(out))


if __name__ == "__main__":
    main()
