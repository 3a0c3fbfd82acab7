"""More unit tests mirroring reference classes that the first unit file did not reach:
TestUtils (poll, shell, job-type helpers), TestTaskExecutor, TestTaskStatus, TestEventHandler
(queue drain / missing job dir), TestTonyClient (task merge / sort), framework plug-ins,
the reusable-port helper and the docker env file."""
import os
import subprocess
import sys
import time

import pytest

from tony_b200.am.session import TaskInfo, TaskStatus
from tony_b200.conf import keys as K
from tony_b200.conf.configuration import TonyConfiguration
from tony_b200.utils import utils as U


def test_poll_and_poll_till_non_null():
    calls = {"n": 0}

    def third():
        calls["n"] += 1
        return calls["n"] >= 3

    assert U.poll(third, 0.01, 1.0) is True and calls["n"] == 3
    t0 = time.monotonic()
    assert U.poll(lambda: False, 0.01, 0.1) is False
    assert 0.09 <= time.monotonic() - t0 < 0.5
    box = []
    assert U.poll_till_non_null(lambda: (box.append(1), "x" if len(box) == 2 else None)[1], 0.01, 1.0) == "x"
    assert U.poll_till_non_null(lambda: None, 0.01, 0.05) is None


def test_execute_shell_exit_code_env_and_timeout(tmp_path):
    assert U.execute_shell("exit 7") == 7
    out = tmp_path / "o.txt"
    with open(out, "wb") as f:
        rc = U.execute_shell("echo $FOO-$MALLOC_ARENA_MAX", env={"FOO": "bar"}, stdout=f)
    assert rc == 0 and out.read_text().strip() == "bar-"      # MALLOC_ARENA_MAX is dropped
    t0 = time.monotonic()
    assert U.execute_shell("sleep 30", timeout_ms=200) == -1  # whole process group killed
    assert time.monotonic() - t0 < 5


def test_job_type_helpers():
    conf = TonyConfiguration()
    for k, v in (("tony.ps.instances", "2"), ("tony.worker.instances", "3"), ("tony.evaluator.instances", "1"),
                 ("tony.application.untracked.jobtypes", "ps,evaluator")):
        conf.set(k, v)
    assert sorted(conf.job_types()) == ["evaluator", "ps", "worker"]
    assert sorted(conf.untracked_job_types()) == ["evaluator", "ps"]
    reqs = U.parse_container_requests(conf)
    assert sum(r.num_instances for r in reqs.values()) == 6
    assert U.task_id("worker", 2) == "worker:2"


def test_task_status_ordinals_and_rows():
    # ordinals are wire-significant (they travel in getTaskInfos)
    assert [s.value for s in (TaskStatus.NEW, TaskStatus.READY, TaskStatus.RUNNING, TaskStatus.FAILED,
                              TaskStatus.SUCCEEDED, TaskStatus.FINISHED)] == [0, 1, 2, 3, 4, 5]
    t = TaskInfo("worker", 1, "http://h/logs", TaskStatus.RUNNING)
    again = TaskInfo.from_dict(t.to_dict())
    assert (again.name, again.index, again.url, again.status) == ("worker", 1, "http://h/logs", TaskStatus.RUNNING)
    rows = [TaskInfo("worker", 1, "", TaskStatus.RUNNING), TaskInfo("worker", 0, "", TaskStatus.FAILED),
            TaskInfo("ps", 0, "", TaskStatus.SUCCEEDED)]
    assert [r.status for r in sorted(rows, key=lambda r: r.sort_key())][0] == TaskStatus.FAILED


def test_task_executor_needs_its_environment(tmp_path):
    """TestTaskExecutor.testTaskExecutorConfShouldThrowException: a container without the
    launcher's environment contract must fail at construction, not later."""
    from tony_b200.executor.task_executor import TaskExecutor

    with pytest.raises(KeyError):
        TaskExecutor(env={})
    conf = TonyConfiguration()
    conf.set("tony.worker.instances", "1")
    final = tmp_path / "tony-final.xml"
    conf.write_xml(str(final))
    ex = TaskExecutor(env={K.JOB_NAME: "worker", K.TASK_INDEX: "0", K.AM_HOST: "127.0.0.1", K.AM_PORT: "1",
                           K.TONY_CONF_PATH: str(final)})
    assert ex.task_id == "worker:0" and ex.framework == "tensorflow" and ex.secret is None


def test_event_handler_drains_queue_and_survives_missing_dir(tmp_path):
    from tony_b200.events import history as H

    job_dir = tmp_path / "job"
    job_dir.mkdir()
    eh = H.EventHandler(str(job_dir), "application_1_0001", int(time.time() * 1000), "alice")
    eh.start()
    from tony_b200.events import schema

    for i in range(50):
        eh.emit(schema.task_started("worker", i, "h"))
    final = eh.stop("SUCCEEDED")
    assert os.path.exists(final) and final.endswith(".jhist")
    events = H.parse_events(str(job_dir))
    assert sum(1 for e in events if e.type == "TASK_STARTED") == 50       # nothing lost at stop()
    # an event the schema cannot encode must not take the writer thread (or the job) down
    eh2 = H.EventHandler(str(tmp_path / "job2"), "application_1_0002", 1, "alice")
    eh2.start()
    eh2.emit({"type": "TASK_STARTED", "event": {"bogus": object()}, "timestamp": 1})
    eh2.emit(schema.task_started("worker", 0, "h"))
    final2 = eh2.stop("FAILED")
    assert "FAILED" in os.path.basename(final2)
    assert any(e.type == "TASK_STARTED" for e in H.parse_events(str(tmp_path / "job2")))


def test_framework_plugin_registration(monkeypatch, tmp_path):
    from tony_b200.runtimes import frameworks as fw

    mod = tmp_path / "my_runtime_plugin.py"
    mod.write_text(
        "from tony_b200.runtimes import frameworks as fw\n"
        "class AM(fw.AMAdapter):\n    pass\n"
        "class Task(fw.TaskAdapter):\n    pass\n"
        "fw.register_framework('myfw', AM, Task)\n")
    monkeypatch.syspath_prepend(str(tmp_path))
    monkeypatch.setenv("TONY_B200_FRAMEWORK_PLUGINS", "my_runtime_plugin")
    assert type(fw.get_am_adapter("myfw")).__name__ == "AM"


def test_reusable_port_helper_process(tmp_path):
    """`python -m tony_b200.executor.ports -p PORT` holds the port and drops <PORT>___PORT___."""
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    proc = subprocess.Popen([sys.executable, "-m", "tony_b200.executor.ports", "-p", str(port), "-d", "30"],
                            cwd=str(tmp_path), env={**os.environ, "PYTHONPATH": os.getcwd()})
    try:
        marker = tmp_path / f"{port}___PORT___"
        assert U.poll(marker.exists, 0.05, 10.0)
        # the reservation uses SO_REUSEPORT: a cooperating process can bind the same port
        s2 = socket.socket()
        s2.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEPORT, 1)
        s2.bind(("", port))
        s2.close()
    finally:
        proc.terminate()
        proc.wait(timeout=10)


def test_docker_env_file(tmp_path):
    from tony_b200.utils import docker as D

    conf = TonyConfiguration()
    assert D.write_env_file(conf, {"A": "1"}, str(tmp_path)) is None          # docker disabled
    conf.set(K.DOCKER_ENABLED, "true")
    path = D.write_env_file(conf, {"A": "1", "B": "x y", "BAD": "line1\nline2"}, str(tmp_path))
    lines = open(path).read().splitlines()
    assert "A=1" in lines and "B=x y" in lines and not any(l.startswith("BAD=") for l in lines)


def test_batchnorm_act_cpu_fallback_matches_torch():
    """BatchNormAct off the GPU (or in eval mode) must be plain BatchNorm2d (+ residual) (+ ReLU)."""
    import torch

    from tony_b200.models.nn import BatchNormAct

    torch.manual_seed(0)
    bn = BatchNormAct(6, relu=True)
    ref = torch.nn.BatchNorm2d(6)
    with torch.no_grad():
        bn.weight.copy_(torch.rand(6) + 0.5)
        bn.bias.copy_(torch.randn(6) * 0.1)
        ref.weight.copy_(bn.weight.float())
        ref.bias.copy_(bn.bias.float())
    x = torch.randn(4, 6, 5, 5).to(torch.bfloat16)
    res = torch.randn(4, 6, 5, 5).to(torch.bfloat16)
    y = bn(x, residual=res)
    want = torch.relu(ref(x.float()) + res.float())
    assert (y.float() - want).abs().max().item() < 0.1
    assert (bn.running_mean - ref.running_mean).abs().max().item() < 2e-2     # running statistics are updated
    bn.eval()
    ref.eval()
    y2 = bn(x)
    assert (y2.float() - torch.relu(ref(x.float()))).abs().max().item() < 0.1
