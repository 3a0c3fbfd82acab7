"""Launcher -> GPU job, end to end: tony.xml-style submission through the mini cluster places one
task per GPU (CUDA_VISIBLE_DEVICES), the tasks build symmetric memory across processes, train
with the fused data plane, report the first-step marker; plus BASELINE config #5's path: the
chief dies mid-training, the whole session is retried and resumes from the sharded checkpoint."""
import os
import sys

import pytest
import torch

from tony_b200.am.session import TaskStatus
from tony_b200.cluster.mini import MiniCluster

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EX = os.path.join(REPO, "examples")


def _submit(cluster, script, params, confs):
    client = cluster.new_client()
    argv = ["--executes", os.path.join(EX, script), "--python_binary_path", sys.executable,
            "--task_params", params, "--shell_env", f"TONY_REPO={REPO}",
            "--shell_env", "TONY_B200_COMM_TIMEOUT_MS=30000",
            "--conf", "tony.am.wait-client-signal-stop-timeout-sec=2"]
    for c in confs:
        argv += ["--conf", c]
    assert client.init(argv)
    rc = client.start()
    return rc, client


def _logs(client):
    out = []
    cdir = os.path.join(client.app_dir, "containers")
    for c in sorted(os.listdir(cdir)) if os.path.isdir(cdir) else []:
        for f in ("stdout", "stderr"):
            p = os.path.join(cdir, c, f)
            if os.path.exists(p):
                out.append(f"--- {c}/{f}\n" + open(p, errors="replace").read()[-1500:])
    return "\n".join(out)


def test_horovod_bert_tiny_through_launcher():
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    n = min(2, torch.cuda.device_count())
    with MiniCluster() as cluster:
        rc, client = _submit(cluster, "bert-large-hvd/bert_pretrain.py", "--tiny --steps 6 --batch 4",
                             ["tony.application.framework=horovod", f"tony.worker.instances={n}",
                              "tony.worker.gpus=1"])
        assert rc == 0, _logs(client)
        st = {f"{t.name}:{t.index}": t.status for t in client.task_infos}
        assert all(st[f"worker:{i}"] == TaskStatus.SUCCEEDED for i in range(n)), st
        assert client.submit_to_first_step_ms is not None
        print(f"submit_to_first_step_ms={client.submit_to_first_step_ms:.0f} (n={n})")
        text = _logs(client)
        assert "samples/s" in text


def test_gpt2_chief_failure_restart_resumes_from_checkpoint(tmp_path):
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    n = min(2, torch.cuda.device_count())
    with MiniCluster() as cluster:
        rc, client = _submit(cluster, "gpt2-medium-ddp/gpt2_train.py",
                             f"--tiny --steps 12 --batch 2 --seq 64 --ckpt_every 4 --crash_at 6 --ckpt_dir {tmp_path}",
                             ["tony.application.framework=pytorch", f"tony.worker.instances={n}",
                              "tony.worker.gpus=1", "tony.am.retry-count=1"])
        assert rc == 0, _logs(client)
        text = _logs(client)
        assert "simulating a crash" in text and "resumed from step 4" in text, text[-3000:]


def test_horovod_shaped_mnist_job_through_launcher():
    """examples/horovod-on-tony: a job written against the hvd-shaped API (init / broadcast_parameters /
    DistributedOptimizer / fp32 metric all-reduce), launched with framework=horovod."""
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    n = min(2, torch.cuda.device_count())
    with MiniCluster() as cluster:
        rc, client = _submit(cluster, "horovod-on-tony/torch_mnist_hvd.py", "--steps 30 --batch 128",
                             ["tony.application.framework=horovod", f"tony.worker.instances={n}",
                              "tony.worker.gpus=1"])
        assert rc == 0, _logs(client)
        text = _logs(client)
        assert "done: loss" in text, text[-2000:]


def test_horovod_keras_shaped_job_with_callbacks_through_launcher():
    """examples/horovod-on-tony/torch_mnist_hvd_callbacks.py: fit() driven by broadcast / metric-average /
    learning-rate warm-up / rank-0 checkpoint callbacks (the reference's tensorflow2_keras_mnist.py shape)."""
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    n = min(2, torch.cuda.device_count())
    with MiniCluster() as cluster:
        rc, client = _submit(cluster, "horovod-on-tony/torch_mnist_hvd_callbacks.py", "--epochs 3 --batches 8 --batch 128",
                             ["tony.application.framework=horovod", f"tony.worker.instances={n}",
                              "tony.worker.gpus=1"])
        assert rc == 0, _logs(client)
        text = _logs(client)
        assert "done: loss" in text and "epoch 2: mean loss" in text, text[-2000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs (1 ps + 1 worker at least)")
def test_widedeep_dedicated_parameter_servers_through_launcher():
    """BASELINE config 4's role split: ps tasks on their own GPUs own the shards and run only the fused
    reduce + update; workers train; the untracked ps tasks are torn down by the AM when the workers end.
    2 ps + 6 workers on an 8-GPU box, 1 + 1 on a 2-GPU box."""
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    gpus = torch.cuda.device_count()
    n_ps, n_w = (2, 6) if gpus >= 8 else (1, gpus - 1 if gpus < 4 else 3)
    with MiniCluster() as cluster:
        rc, client = _submit(cluster, "widedeep-ps/widedeep_train.py", "--tiny --steps 12 --batch 512",
                             ["tony.application.framework=pytorch", f"tony.ps.instances={n_ps}", "tony.ps.gpus=1",
                              f"tony.worker.instances={n_w}", "tony.worker.gpus=1"])
        assert rc == 0, _logs(client)
        st = {f"{t.name}:{t.index}": t.status for t in client.task_infos}
        assert all(st[f"worker:{i}"] == TaskStatus.SUCCEEDED for i in range(n_w)), st
        text = _logs(client)
        assert f"done ({n_ps} ps + {n_w} workers)" in text, text[-3000:]
        assert "serving done, waiting for teardown" in text
        import re
        m = re.search(r"loss ([0-9.]+) -> ([0-9.]+)", text)
        assert m and float(m.group(2)) < float(m.group(1)), text[-1500:]


def test_nvml_sampler_reports_every_gpu():
    """The in-process NVML path the TaskMonitor uses instead of forking nvidia-smi."""
    from tony_b200.cluster.topology import GpuDiscoverer

    os.environ.pop("TONY_B200_FAKE_GPUS", None)
    d = GpuDiscoverer()
    assert d.nvml is not None, "libnvidia-ml.so.1 did not load on a GPU box"
    info = d.sample()
    assert info is not None and len(info.gpus) == torch.cuda.device_count()
    g = info.gpus[0]
    assert g.fb_total_mib > 100_000 and 0 <= g.gpu_util_pct <= 100 and g.bar1_total_mib > 0
