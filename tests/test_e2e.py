"""Mini-cluster end-to-end scenarios on CPU (SURVEY.md §4 tier 3).

Each test drives a real ``TonyClient.init() + start()`` against a private MiniCluster — real
ApplicationMaster and TaskExecutor processes, real RPC, tiny Python payloads — mirroring the
scenario list of the reference's TestTonyE2E.java (tony-core/src/test/java/com/linkedin/tony/
TestTonyE2E.java:128-677) including its fault-injection hooks (same env-var names).
"""
import json
import os
import sys
import time

import pytest

from tony_b200.am.session import TaskStatus
from tony_b200.client.tony_client import CallbackHandler, TaskUpdateListener
from tony_b200.cluster.mini import MiniCluster
from tony_b200.conf import keys as K
from tony_b200.events import history

SCRIPTS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scripts")
PY = sys.executable


@pytest.fixture(scope="module")
def cluster():
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.setdefault("TONY_B200_FAKE_GPUS", "0")
    with MiniCluster(num_node_managers=3) as c:
        yield c


def run(cluster, script, confs, extra=(), env=None, listener=None, handler=None):
    client = cluster.new_client(handler)
    argv = ["--python_binary_path", PY] + list(extra)
    if script:
        argv += ["--executes", os.path.join(SCRIPTS, script)]
    for c in confs:
        argv += ["--conf", c]
    argv += ["--conf", "tony.task.heartbeat-interval-ms=100", "--conf", "tony.am.wait-client-signal-stop-timeout-sec=2"]
    old = {}
    for k, v in (env or {}).items():
        old[k] = os.environ.get(k)
        os.environ[k] = v
    try:
        assert client.init(argv)
        if listener:
            client.add_listener(listener)
        rc = client.start()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return rc, client


def statuses(client):
    return {f"{t.name}:{t.index}": t.status for t in client.task_infos}


def test_single_node_pass_and_fail(cluster):
    assert run(cluster, "exit_0.py", ["tony.application.security.enabled=false"])[0] == 0
    assert run(cluster, "exit_1.py", [])[0] == -1


def test_ps_worker_training_should_pass(cluster):
    rc, client = run(cluster, "exit_0_check_env.py", ["tony.ps.instances=1", "tony.worker.instances=2"],
                     extra=["--shell_env", "ENV_CHECK=ENV_CHECK"])
    assert rc == 0
    st = statuses(client)
    assert st["worker:0"] == st["worker:1"] == TaskStatus.SUCCEEDED
    assert st["ps:0"] in (TaskStatus.SUCCEEDED, TaskStatus.FINISHED)


def test_ps_worker_training_should_fail(cluster):
    rc, _ = run(cluster, "exit_1.py", ["tony.ps.instances=1", "tony.worker.instances=2"])
    assert rc == -1


def test_untracked_ps_keeps_running_and_is_finished_by_am(cluster):
    rc, client = run(cluster, None, ["tony.ps.instances=1", "tony.worker.instances=1",
                                     f"tony.ps.command={PY} {SCRIPTS}/sleep_30.py",
                                     f"tony.worker.command={PY} {SCRIPTS}/exit_0.py"])
    assert rc == 0
    st = statuses(client)
    assert st["worker:0"] == TaskStatus.SUCCEEDED and st["ps:0"] == TaskStatus.FINISHED


def test_untracked_ps_crash_fails_the_app(cluster):
    rc, client = run(cluster, None, ["tony.ps.instances=1", "tony.worker.instances=1",
                                     f"tony.ps.command={PY} {SCRIPTS}/exit_1.py",
                                     f"tony.worker.command={PY} {SCRIPTS}/sleep_30.py"])
    assert rc == -1
    st = statuses(client)
    assert st["ps:0"] == TaskStatus.FAILED and st["worker:0"] == TaskStatus.FINISHED


def test_sidecar_crash_is_ignored(cluster):
    rc, _ = run(cluster, None, ["tony.worker.instances=1", "tony.sidecar.instances=1",
                                "tony.application.sidecar.jobtypes=sidecar",
                                f"tony.sidecar.command={PY} {SCRIPTS}/exit_1.py",
                                f"tony.worker.command={PY} {SCRIPTS}/exit_0.py"])
    assert rc == 0


def test_pytorch_env_contract(cluster):
    rc, _ = run(cluster, "exit_0_check_pytorchenv.py",
                ["tony.application.framework=pytorch", "tony.worker.instances=2", "tony.ps.instances=0"])
    assert rc == 0


def test_tensorflow_env_contract(cluster):
    rc, _ = run(cluster, "check_tf_env.py", ["tony.worker.instances=2", "tony.ps.instances=1", "tony.evaluator.instances=1"])
    assert rc == 0


def test_horovod_env_contract_and_driver(cluster):
    rc, client = run(cluster, "check_horovod_env.py",
                     ["tony.application.framework=horovod", "tony.worker.instances=2",
                      "tony.horovod.mode.test=true"])
    assert rc == 0
    assert statuses(client)["driver:0"] == TaskStatus.FINISHED


def test_horovod_real_plan(cluster):
    rc, _ = run(cluster, "check_horovod_env.py", ["tony.application.framework=horovod", "tony.worker.instances=3"])
    assert rc == 0


def test_horovod_driver_crash_fails(cluster):
    rc, _ = run(cluster, "check_horovod_env.py",
                ["tony.application.framework=horovod", "tony.worker.instances=2",
                 "tony.horovod.mode.test.fast.fail=true"])
    assert rc == -1


def test_horovod_rejects_user_driver_conf(cluster):
    rc, _ = run(cluster, "exit_0.py", ["tony.application.framework=horovod", "tony.worker.instances=1",
                                       "tony.driver.instances=2"])
    assert rc == -1


def test_standalone_ok_and_rejects_two_instances(cluster):
    assert run(cluster, "exit_0.py", ["tony.application.framework=standalone", "tony.worker.instances=1"])[0] == 0
    assert run(cluster, "exit_0.py", ["tony.application.framework=standalone", "tony.worker.instances=2"])[0] == -1


def test_staged_dag(cluster):
    rc, client = run(cluster, "exit_0.py",
                     ["tony.db.instances=1", "tony.dbloader.instances=1", "tony.ps.instances=1",
                      "tony.worker.instances=2", "tony.application.prepare-stage=db,dbloader",
                      "tony.application.training-stage=ps,worker", "tony.application.untracked.jobtypes=db,ps",
                      f"tony.db.command={PY} {SCRIPTS}/sleep_30.py"])
    assert rc == 0
    st = statuses(client)
    assert st["dbloader:0"] == TaskStatus.SUCCEEDED and st["worker:1"] == TaskStatus.SUCCEEDED


def test_non_chief_worker_failure_still_succeeds(cluster):
    rc, _ = run(cluster, None, ["tony.chief.instances=1", "tony.worker.instances=1",
                                f"tony.chief.command={PY} {SCRIPTS}/exit_0.py",
                                f"tony.worker.command={PY} {SCRIPTS}/exit_1.py"])
    assert rc == 0


def test_fail_on_worker_failure_flag(cluster):
    rc, _ = run(cluster, None, ["tony.chief.instances=1", "tony.worker.instances=1",
                                "tony.application.fail-on-worker-failure-enabled=true",
                                f"tony.chief.command={PY} {SCRIPTS}/sleep_30.py",
                                f"tony.worker.command={PY} {SCRIPTS}/exit_1.py"])
    assert rc == -1


def test_tb_port_only_on_chief(cluster):
    rc, _ = run(cluster, "check_tb_port_set_in_chief_only.py", ["tony.worker.instances=2", "tony.ps.instances=1"])
    assert rc == 0


def test_sidecar_tensorboard(cluster, tmp_path):
    rc, client = run(cluster, "exit_0.py", ["tony.worker.instances=1"],
                     extra=["--sidecar_tensorboard_log_dir", str(tmp_path)], env={"SIDECAR_TB_TEST": "1"})
    assert rc == 0
    assert "tensorboard:0" in statuses(client)


def test_resources_src_dir_and_venv(cluster, tmp_path):
    import zipfile

    venv = tmp_path / "venv.zip"
    with zipfile.ZipFile(venv, "w") as z:
        z.writestr("123.xml", "<x/>")
    common = tmp_path / "common.zip"
    with zipfile.ZipFile(common, "w") as z:
        z.writestr("a.txt", "a")
    test2 = tmp_path / "test2.zip"
    with zipfile.ZipFile(test2, "w") as z:
        z.writestr("b.txt", "b")
    plain = tmp_path / "plain.txt"
    plain.write_text("p")
    rc, _ = run(cluster, "check_env_and_venv.py", ["tony.worker.instances=1"], extra=["--python_venv", str(venv)])
    assert rc == 0
    rc, _ = run(cluster, "check_archive_file_localization.py",
                ["tony.worker.instances=1",
                 f"tony.worker.resources={common}#archive,{test2}::renamed.zip#archive,{plain}"])
    assert rc == 0


def test_missed_heartbeats_fail_the_app(cluster):
    t0 = time.time()
    rc, _ = run(cluster, "sleep_30.py", ["tony.worker.instances=1", "tony.task.max-missed-heartbeats=5"],
                env={"TEST_TASK_EXECUTOR_NUM_HB_MISS": "20"})
    assert rc == -1
    assert time.time() - t0 < 20


def test_task_skew_still_passes(cluster):
    rc, _ = run(cluster, "exit_0.py", ["tony.worker.instances=2"], env={"TEST_TASK_EXECUTOR_SKEW": "worker#0#1500"})
    assert rc == 0


def test_am_crash_fails_without_retry_and_recovers_with_retry(cluster):
    assert run(cluster, "exit_0.py", ["tony.worker.instances=1"], env={"TEST_AM_CRASH": "1"})[0] == -1
    assert run(cluster, "exit_0.py", ["tony.worker.instances=1", "tony.am.retry-count=1"],
               env={"TEST_AM_CRASH": "1"})[0] == 0


def test_worker_termination_then_session_retry(cluster):
    """BASELINE config #5 plumbing: chief killed, whole session restarted, job succeeds."""
    rc, client = run(cluster, "exit_0_if_retry.py", ["tony.worker.instances=2", "tony.am.retry-count=1"])
    assert rc == 0
    rc, _ = run(cluster, "sleep_30.py", ["tony.worker.instances=2"], env={"TEST_WORKER_TERMINATION": "1"})
    assert rc == -1


def test_late_completion_notification(cluster):
    rc, _ = run(cluster, "exit_0.py", ["tony.worker.instances=1", "tony.task.max-missed-heartbeats=3"],
                env={"TEST_TASK_COMPLETION_NOTIFICATION_DELAYED": "1"})
    assert rc == 0


def test_preprocess_forwards_model_params(cluster):
    rc, _ = run(cluster, "print_model_params.py", ["tony.worker.instances=1", "tony.application.enable-preprocess=true"])
    assert rc == 0


def test_application_timeout(cluster):
    t0 = time.time()
    rc, _ = run(cluster, "sleep_30.py", ["tony.worker.instances=1", "tony.application.timeout=1500"])
    assert rc == -1 and time.time() - t0 < 15


def test_worker_timeout_kills_user_process(cluster):
    rc, _ = run(cluster, "sleep_30.py", ["tony.worker.instances=1", "tony.worker.timeout=500"])
    assert rc == -1


def test_callbacks_listener_history_and_first_step(cluster):
    got = {"id": None, "updates": 0}

    class H(CallbackHandler):
        def on_application_id_received(self, app_id):
            got["id"] = app_id

    class L(TaskUpdateListener):
        def on_task_infos_updated(self, infos):
            got["updates"] += 1

    rc, client = run(cluster, "mark_first_step.py", ["tony.worker.instances=2", "tony.ps.instances=1"],
                     extra=["--shell_env", f"TONY_REPO={os.path.dirname(os.path.dirname(SCRIPTS))}"],
                     listener=L(), handler=H())
    assert rc == 0
    assert got["id"] == client.app_id and got["updates"] >= 1
    assert client.submit_to_first_step_ms is not None and 0 < client.submit_to_first_step_ms < 30000
    job_dir = os.path.join(cluster.history_dir, "intermediate", client.app_id)
    md = history.parse_metadata(job_dir)
    assert md.status == "SUCCEEDED" and md.id == client.app_id
    events = history.parse_events(job_dir)
    types = [e.type for e in events]
    assert types[0] == "APPLICATION_INITED" and types[-1] == "APPLICATION_FINISHED"
    assert types.count("TASK_STARTED") == 3 and types.count("TASK_FINISHED") == 3
    # training metrics reported by the user process + the launcher's own latency reach TaskFinished.metrics
    fin = [e for e in events if e.type == "TASK_FINISHED" and e.event["taskType"] == "worker"]
    for e in fin:
        got_m = {m["name"]: m["value"] for m in e.event["metrics"]}
        assert got_m.get("STEP_TIME_MS") == 12.5 and got_m.get("SAMPLES_PER_SEC") == 2560.0, got_m
        assert 0 < got_m.get("SUBMIT_TO_FIRST_STEP_MS", -1) < 30000, got_m
    assert any(c.name == "tony.worker.instances" and c.value == "2" for c in history.parse_config(job_dir))
    assert os.path.exists(os.path.join(client.app_dir, K.TONY_FINAL_XML))


def test_limits_are_enforced_at_init(cluster):
    client = cluster.new_client()
    with pytest.raises(RuntimeError):
        client.init(["--executes", "x.py", "--conf", "tony.worker.instances=3", "--conf", "tony.worker.max-instances=2"])
    client = cluster.new_client()
    with pytest.raises(RuntimeError):
        client.init(["--executes", "x.py", "--conf", "tony.worker.instances=3", "--conf", "tony.worker.gpus=2",
                     "--conf", "tony.task.max-total-gpus=4"])


def test_mnist_pytorch_two_workers_gloo(cluster):
    """BASELINE config #1: mnist-pytorch distributed, 2 workers on CPU/gloo, launched from tony.xml."""
    repo = os.path.dirname(os.path.dirname(SCRIPTS))
    ex = os.path.join(repo, "examples", "mnist-pytorch")
    client = cluster.new_client()
    assert client.init(["--conf_file", os.path.join(ex, "tony.xml"), "--executes", os.path.join(ex, "mnist_distributed.py"),
                        "--python_binary_path", PY, "--task_params", "--epochs 2",
                        "--shell_env", f"TONY_REPO={repo}", "--conf", "tony.am.wait-client-signal-stop-timeout-sec=2"])
    assert client.start() == 0
    assert statuses(client) == {"worker:0": TaskStatus.SUCCEEDED, "worker:1": TaskStatus.SUCCEEDED}
    assert client.submit_to_first_step_ms is not None


def _example(cluster, rel, script, extra_conf=(), conf="tony.xml"):
    repo = os.path.dirname(os.path.dirname(SCRIPTS))
    ex = os.path.join(repo, "examples", rel)
    client = cluster.new_client()
    argv = ["--conf_file", os.path.join(ex, conf), "--executes", os.path.join(ex, script),
            "--python_binary_path", PY, "--conf", "tony.am.wait-client-signal-stop-timeout-sec=2"]
    for c in extra_conf:
        argv += ["--conf", c]
    assert client.init(argv)
    return client.start(), client


def test_example_tensorflow_ps_worker(cluster):
    rc, client = _example(cluster, "mnist-tensorflow", "mnist_ps_worker.py")
    assert rc == 0
    st = statuses(client)
    assert st["worker:0"] == st["worker:1"] == TaskStatus.SUCCEEDED and st["ps:0"] == TaskStatus.FINISHED


def test_example_tensorflow_multiworker(cluster):
    """MultiWorkerMirrored-shaped job: workers only, gradients all-reduced over the TF_CONFIG addresses."""
    rc, client = _example(cluster, "mnist-tensorflow", "mnist_multiworker.py", conf="tony-multiworker.xml")
    assert rc == 0
    assert statuses(client) == {"worker:0": TaskStatus.SUCCEEDED, "worker:1": TaskStatus.SUCCEEDED}


def test_example_tensorflow_estimator_chief_worker_evaluator(cluster, tmp_path):
    """train_and_evaluate-shaped job: chief + worker train; the untracked evaluator is outside their TF_CONFIG."""
    client = cluster.new_client()
    repo = os.path.dirname(os.path.dirname(SCRIPTS))
    ex = os.path.join(repo, "examples", "mnist-tensorflow")
    assert client.init(["--conf_file", os.path.join(ex, "tony-estimator.xml"),
                        "--executes", os.path.join(ex, "mnist_estimator_style.py"), "--python_binary_path", PY,
                        "--task_params", f"--model_dir {tmp_path}/model",
                        "--conf", "tony.am.wait-client-signal-stop-timeout-sec=2"])
    assert client.start() == 0
    st = statuses(client)
    assert st["chief:0"] == st["worker:0"] == TaskStatus.SUCCEEDED
    assert st["evaluator:0"] in (TaskStatus.FINISHED, TaskStatus.SUCCEEDED)
    assert os.path.exists(os.path.join(str(tmp_path), "model", "DONE"))


def test_example_ray_shaped_program_on_head_and_workers(cluster):
    rc, client = _example(cluster, "ray-on-tony", "example.py")
    assert rc == 0
    assert statuses(client)["head:0"] == TaskStatus.SUCCEEDED


def test_example_mxnet_dmlc_roles(cluster):
    rc, client = _example(cluster, "linearregression-mxnet", "dmlc_linear_regression.py")
    assert rc == 0


def test_example_custom_job_types_head_worker(cluster):
    rc, client = _example(cluster, "ray-on-tony", "discovery.py")
    assert rc == 0
    assert statuses(client)["head:0"] == TaskStatus.SUCCEEDED


def test_timeline_is_recorded(cluster):
    rc, client = run(cluster, "exit_0.py", ["tony.worker.instances=2"])
    assert rc == 0
    tl = json.load(open(os.path.join(client.app_dir, "timeline.json")))
    events = [e["event"] for e in tl]
    assert events[0] == "am_started" and "gang_barrier_open" in events and events.count("task_registered") == 2
    assert all(a["t_ms"] <= b["t_ms"] for a, b in zip(tl, tl[1:]))


def test_am_throwing_exception_fails_the_app(cluster):
    """TestTonyE2E.testAMThrowExceptionCrashTonYShouldFail: an exception inside the AM's monitor
    loop (not a process crash) must still end the application as FAILED."""
    rc, _ = run(cluster, "exit_0.py", ["tony.worker.instances=1"], env={"TEST_AM_THROW_EXCEPTION_CRASH": "1"})
    assert rc == -1


def test_horovod_debug_mode_runs_user_driver(cluster):
    """TestTonyE2E.testTonyHorovodWithDebugModeShouldPass: in driver debug mode the user's own
    driver command runs instead of the built-in rendezvous driver (same port-file contract)."""
    import sys

    from tony_b200.runtimes import horovod_driver as hd

    driver_cmd = f"{sys.executable} {os.path.abspath(hd.__file__)} -t -p 9999 -w $CLUSTER_WORKER_LIST -o $DRIVER_OUTPUT_PATH"
    rc, client = run(cluster, "check_horovod_env.py",
                     ["tony.application.framework=horovod", "tony.worker.instances=2",
                      "tony.horovod.driver.mode.debug=true", "tony.driver.command=" + driver_cmd])
    assert rc == 0
    assert statuses(client)["driver:0"] == TaskStatus.FINISHED
    # debug mode without a driver command is rejected up front
    rc, _ = run(cluster, "check_horovod_env.py",
                ["tony.application.framework=horovod", "tony.worker.instances=2",
                 "tony.horovod.driver.mode.debug=true"])
    assert rc == -1


def test_final_conf_is_frozen_for_am_and_tasks(cluster):
    """TestTonyE2E.testTonyFinalConf: everything the client resolved (CLI --conf included) is in
    the tony-final.xml of the application directory, which is what the AM and the tasks read."""
    from tony_b200.conf.configuration import TonyConfiguration

    rc, client = run(cluster, "exit_0.py", ["tony.worker.instances=2", "tony.application.name=final-conf-check",
                                            "tony.worker.gpus=0"])
    assert rc == 0
    final = TonyConfiguration.from_final_xml(os.path.join(client.app_dir, "tony-final.xml"))
    assert final.get("tony.application.name") == "final-conf-check"
    assert final.get_int("tony.worker.instances", 0) == 2
    assert final.get("tony.task.heartbeat-interval-ms") == "100"
    # a hadoop-classpath style entry that does not exist is recorded, not fatal
    # (testTonyFinalConfWithNonExistentHDFSClasspath)
    rc, client = run(cluster, "exit_0.py", ["tony.worker.instances=1",
                                            "tony.application.hadoop.location=/no/such/hadoop"])
    assert rc == 0
    final = TonyConfiguration.from_final_xml(os.path.join(client.app_dir, "tony-final.xml"))
    assert final.get("tony.application.hadoop.location") == "/no/such/hadoop"


def test_client_survives_am_that_never_comes_up(cluster, tmp_path):
    """TestTonyE2E.testNullAMRpcClient: with no AM endpoint the client's monitor loop must end in
    failure instead of raising or spinning forever."""
    client = cluster.new_client()
    assert client.init(["--python_binary_path", PY, "--executes", os.path.join(SCRIPTS, "exit_0.py"),
                        "--conf", "tony.worker.instances=1"])
    os.environ["TEST_AM_NEVER_STARTS"] = "1"
    try:
        rc = client.start()
    finally:
        os.environ.pop("TEST_AM_NEVER_STARTS", None)
    assert rc == -1
