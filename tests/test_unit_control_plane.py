"""Unit tier (SURVEY.md §4 tier 1-2): pure functions and small classes of the control plane,
mirroring TestUtils / TestTaskScheduler / TestTonySession / TestApplicationMaster /
TestLocalizableResource / TestTonyClient / TestHorovodRuntime / TestMLGenericRuntime /
TestHistoryFileUtils / TestParserUtils / TestTaskMonitor / TestTonyConfigurationFields /
TestPortAllocation / TestHorovodDriver / TestGpuDeviceInformationParser of the reference."""
import json
import os
import socket
import time
import zipfile

import pytest

from tony_b200.am.scheduler import TaskScheduler, is_dag
from tony_b200.am.session import FinalStatus, TaskInfo, TaskStatus, TonySession
from tony_b200.am.liveliness import LivelinessMonitor
from tony_b200.client import tony_client as tc
from tony_b200.cluster import topology
from tony_b200.conf import keys as K
from tony_b200.conf.configuration import (DEFAULT_XML_PATH, TonyConfiguration, build_configuration,
                                          parse_memory_string, read_xml, write_xml)
from tony_b200.events import avro, history, schema
from tony_b200.executor import ports
from tony_b200.executor.task_monitor import (AVG_GPU_UTILIZATION, AVG_MEMORY_BYTES, MAX_GPU_FB_MEMORY_USAGE,
                                             MAX_MEMORY_BYTES, TaskMonitor, process_tree_rss_bytes)
from tony_b200.rpc.server import RpcClient, RpcServer
from tony_b200.rpc.wire import RpcError
from tony_b200.runtimes import frameworks as fw
from tony_b200.runtimes import horovod_driver as hd
from tony_b200.utils import utils as U
from tony_b200.utils.docker import container_runtime_env
from tony_b200.utils.localizable_resource import LocalizableResource, ResourceParseError

REF = "/root/reference"


def conf_with(**kv):
    c = TonyConfiguration()
    for k, v in kv.items():
        c.set(k.replace("__", "-").replace("_", "."), v)
    return c


# ---- configuration ---------------------------------------------------------------------------
def test_parse_memory_string():
    assert parse_memory_string("2g") == 2048
    assert parse_memory_string("2G") == 2048
    assert parse_memory_string("512m") == 512
    assert parse_memory_string("1024") == 1024
    with pytest.raises(ValueError):
        parse_memory_string("2gb")


def test_default_xml_matches_keys():
    """XML <-> keys parity (TestTonyConfigurationFields.java:20-74)."""
    xml = {n: v for n, v, _ in read_xml(DEFAULT_XML_PATH)}
    assert xml == K.DEFAULTS_MAP
    declared = {v for k, v in vars(K).items() if isinstance(v, str) and v.startswith("tony.") and k.isupper()}
    missing = {k for k in declared if k not in xml and k not in K.KEYS_WITHOUT_DEFAULT and not k.endswith(".") and not k.endswith(".xml")}  # noqa
    assert not missing, missing
    # same defaults as the reference's tony-default.xml for every key it defines
    ref = os.path.join(REF, "tony-core/src/main/resources/tony-default.xml")
    if os.path.exists(ref):
        for n, v, _ in read_xml(ref):
            assert xml.get(n) == v, (n, v, xml.get(n))


def test_conf_precedence_and_multivalue(tmp_path):
    f = tmp_path / "tony.xml"
    write_xml(f, [("tony.worker.instances", "3", {}), ("tony.containers.envs", "A=1", {}),
                  ("tony.application.name", "fromfile", {})])
    site_dir = tmp_path / "site"
    site_dir.mkdir()
    write_xml(site_dir / "tony-site.xml", [("tony.application.name", "fromsite", {})])
    c = build_configuration(str(f), ["tony.worker.instances=5", "tony.containers.envs=B=2",
                                     "tony.application.name=fromcli"], env={"TONY_CONF_DIR": str(site_dir)})
    assert c.get_int("tony.worker.instances") == 5           # cli > file
    assert c.get("tony.containers.envs") == "A=1,B=2"        # multi-value appends
    assert c.get("tony.application.name") == "fromsite"      # site wins last
    assert c.get("tony.yarn.queue") == "default"             # defaults underneath
    c2 = build_configuration(None, [], cwd=str(tmp_path))    # ./tony.xml picked up
    assert c2.get_int("tony.worker.instances") == 3
    out = tmp_path / "final.xml"
    c.write_xml(out)
    assert TonyConfiguration.from_final_xml(out).get("tony.containers.envs") == "A=1,B=2"


def test_job_types_regex():
    c = conf_with()
    c.set("tony.worker.instances", "2")
    c.set("tony.evaluator.instances", "1")
    c.set("tony.Bad9.instances", "1")
    assert c.job_types() == ["worker", "evaluator"]
    assert not c.is_job_type_monitored("ps") and not c.is_job_type_monitored("tensorboard")
    assert c.is_job_type_monitored("worker")


# ---- utils -------------------------------------------------------------------------------------
def test_parse_container_requests_and_stages():
    c = TonyConfiguration()
    for j, n in (("worker", 3), ("evaluator", 1), ("ps", 2), ("db", 1), ("dbloader", 1)):
        c.set(K.instances_key(j), str(n))
    c.set("tony.worker.gpus", "1")
    c.set("tony.worker.memory", "3g")
    c.set("tony.evaluator.vcores", "2")
    c.set(K.PREPARE_STAGE, "dbloader,db")
    c.set(K.TRAINING_STAGE, "evaluator,worker,ps")
    c.set(K.UNTRACKED_JOBTYPES, "db,ps")
    r = U.parse_container_requests(c, gpus_available=8)
    assert r["worker"].num_instances == 3 and r["worker"].gpus == 1 and r["worker"].memory_mb == 3072
    assert r["evaluator"].vcores == 2 and r["ps"].memory_mb == 2048
    assert len({x.priority for x in r.values()}) == 5
    assert r["worker"].depends_on == ["dbloader"] and r["ps"].depends_on == ["dbloader"]
    assert r["db"].depends_on == [] and r["dbloader"].depends_on == []
    with pytest.raises(RuntimeError):
        U.parse_container_requests(c, gpus_available=0)
    c.set(K.TRAINING_STAGE, "worker")
    with pytest.raises(ValueError):
        U.parse_container_requests(c)


def test_stage_autofill():
    c = TonyConfiguration()
    for j in ("a", "b", "worker"):
        c.set(K.instances_key(j), "1")
    c.set(K.PREPARE_STAGE, "a,b")
    r = U.parse_container_requests(c)
    assert r["worker"].depends_on == ["a", "b"]


def test_construct_tf_config():
    spec = json.dumps({"worker": ["w0:1", "w1:1"], "ps": ["p0:1"], "evaluator": ["e0:1"], "tensorboard": ["t:1"]})
    cfg = json.loads(U.construct_tf_config(spec, "worker", 1))
    assert cfg["task"] == {"type": "worker", "index": 1}
    assert set(cfg["cluster"]) == {"worker", "ps"}
    cfg = json.loads(U.construct_tf_config(spec, "evaluator", 0))
    assert set(cfg["cluster"]) == {"worker", "ps", "evaluator"}


def test_build_task_command_truth_table():
    b = U.build_task_command
    assert b(None, None, None, None) is None
    assert b(None, None, "ls", None) == "ls"
    assert b(None, None, "ls", "-l") == "ls -l"
    assert b(None, "python", "a.py", None) == "python a.py"
    assert b("venv.zip", "bin/python", "a.py", "--x 1") == "venv/bin/python a.py --x 1"
    assert b("venv.zip", "/usr/bin/python", "a.py", None) == "/usr/bin/python a.py"
    assert b(None, "bin/python", "a.py", None) == "bin/python a.py"


def test_parse_key_value_and_cluster_spec_parsers():
    assert U.parse_key_value(["a=b", "c=d=e", "f"]) == {"a": "b", "c": "d=e", "f": ""}
    assert U.parse_key_value(None) == {}
    assert U.parse_cluster_spec_for_pytorch('{"worker":["h:1","g:2"]}') == "tcp://h:1"
    assert U.parse_cluster_spec_for_mxnet('{"scheduler":["localhost:99"]}') == ["127.0.0.1", "99"]


def test_is_archive_and_zip_roundtrip(tmp_path):
    d = tmp_path / "src"
    (d / "sub").mkdir(parents=True)
    (d / "a.py").write_text("x")
    (d / "sub" / "b.py").write_text("y")
    z = tmp_path / "s.zip"
    U.zip_folder(str(d), str(z))
    assert U.is_archive(str(z)) and not U.is_archive(str(d / "a.py"))
    U.unzip_archive(str(z), str(tmp_path / "out"))
    assert (tmp_path / "out" / "sub" / "b.py").read_text() == "y"
    import tarfile

    t = tmp_path / "t.tar.gz"
    with tarfile.open(t, "w:gz") as tf:
        tf.add(d / "a.py", arcname="a.py")
    assert U.is_archive(str(t))


def test_localizable_resource_grammar(tmp_path):
    f = tmp_path / "file.txt"
    f.write_text("x")
    r = LocalizableResource.parse(str(f))
    assert r.local_name == "file.txt" and not r.is_archive and not r.is_directory
    r = LocalizableResource.parse(f"{f}::renamed.txt")
    assert r.local_name == "renamed.txt"
    z = tmp_path / "a.zip"
    with zipfile.ZipFile(z, "w") as zz:
        zz.writestr("k", "v")
    r = LocalizableResource.parse(f"{z}::data#archive")
    assert r.is_archive and r.local_name == "data"
    dst = r.localize(str(tmp_path / "c"))
    assert os.path.isfile(os.path.join(dst, "k"))
    with pytest.raises(ResourceParseError):
        LocalizableResource.parse(f"{f}::a::b")
    with pytest.raises(FileNotFoundError):
        LocalizableResource.parse(str(tmp_path / "missing"))
    assert LocalizableResource.parse(str(tmp_path)).is_directory


def test_docker_env():
    c = TonyConfiguration()
    assert container_runtime_env(c, "worker") == {}
    c.set(K.DOCKER_ENABLED, "true")
    c.set(K.DOCKER_CONTAINERS_IMAGE, "img:1")
    c.set(K.docker_image_key("ps"), "psimg:2")
    c.set(K.DOCKER_CONTAINERS_MOUNT, "/a:/b:ro")
    e = container_runtime_env(c, "worker")
    assert e == {"YARN_CONTAINER_RUNTIME_TYPE": "docker", "YARN_CONTAINER_RUNTIME_DOCKER_IMAGE": "img:1",
                 "YARN_CONTAINER_RUNTIME_DOCKER_MOUNTS": "/a:/b:ro"}
    assert container_runtime_env(c, "ps")["YARN_CONTAINER_RUNTIME_DOCKER_IMAGE"] == "psimg:2"


# ---- session / scheduler -----------------------------------------------------------------------
def make_session(extra=None, **instances):
    c = TonyConfiguration()
    for j, n in instances.items():
        c.set(K.instances_key(j), str(n))
    for k, v in (extra or {}).items():
        c.set(k, v)
    return TonySession(c, U.parse_container_requests(c)), c


def fill(session):
    for job, req in session.requests.items():
        for _ in range(req.num_instances):
            t = session.init_task_by_priority(req.priority)
            t.set_host_port(f"h:{1000 + t.index}")
    return session


def test_session_chief_and_task_matching():
    s, _ = make_session(worker=2, ps=1)
    assert s.is_chief("worker", 0) and not s.is_chief("worker", 1) and not s.is_chief("ps", 0)
    s2, _ = make_session(worker=2, chief=1)
    assert s2.is_chief("chief", 0) and not s2.is_chief("worker", 0)
    assert s.task_infos() == []                     # empty until everything is scheduled
    fill(s)
    assert len(s.task_infos()) == 3
    assert s.init_task_by_priority(s.requests["worker"].priority) is None
    assert json.loads(s.cluster_spec_json()) == {"worker": ["h:1000", "h:1001"], "ps": ["h:1000"]}
    assert s.total_tracked_tasks() == 2 and s.total_tasks() == 3


def test_session_completion_policy():
    s, _ = make_session(worker=3, ps=1)
    fill(s)
    s.on_task_completed("worker", 1, 1)            # non-chief failure: keep going
    assert not s.training_finished
    s.on_task_completed("worker", 0, 0)
    s.on_task_completed("worker", 2, 0)
    s.update_session_status()
    assert s.final_status == FinalStatus.SUCCEEDED and "failedCnt=1" in s.final_message
    s, _ = make_session(worker=2)
    fill(s)
    s.on_task_completed("worker", 0, 2, "boom")    # chief failure short-circuits
    assert s.training_finished and s.final_status == FinalStatus.FAILED and "boom" in s.final_message
    s, _ = make_session(worker=2)
    fill(s)
    s.on_task_completed("worker", 0, K.KILLED_BY_APPMASTER)   # killed by AM is not a failure
    assert not s.training_finished and s.get_task("worker", 0).status == TaskStatus.FINISHED
    s, _ = make_session(worker=2)
    fill(s)
    s.on_task_completed("worker", 1, 1)
    s.on_task_completed("worker", 0, 0)
    s.fail_on_worker_failure = True
    s.update_session_status()
    assert s.final_status == FinalStatus.FAILED
    s, _ = make_session({K.STOP_ON_FAILURE_JOBTYPES: "evaluator"}, worker=2, evaluator=1)
    fill(s)
    s.on_task_completed("evaluator", 0, 3)
    assert s.training_finished
    s, _ = make_session(worker=2)
    fill(s)
    s.on_task_completed("worker", 0, 0)
    s.update_session_status()                       # worker:1 never finished
    assert s.final_status == FinalStatus.FAILED
    t = s.get_task("worker", 0)
    t.set_exit_status(5)                            # first terminal state wins
    assert t.status == TaskStatus.SUCCEEDED


def test_task_info_attention_order():
    infos = [TaskInfo("worker", 0, "", TaskStatus.RUNNING), TaskInfo("worker", 1, "", TaskStatus.FAILED),
             TaskInfo("ps", 0, "", TaskStatus.NEW), TaskInfo("worker", 2, "", TaskStatus.SUCCEEDED),
             TaskInfo("ps", 1, "", TaskStatus.FINISHED), TaskInfo("a", 0, "", TaskStatus.READY)]
    order = [t.status for t in sorted(infos, key=lambda t: t.sort_key())]
    assert order == [TaskStatus.FAILED, TaskStatus.SUCCEEDED, TaskStatus.FINISHED, TaskStatus.RUNNING,
                     TaskStatus.NEW, TaskStatus.READY]
    assert [int(s) for s in TaskStatus] == [0, 1, 2, 3, 4, 5]
    assert TaskInfo.from_dict(infos[1].to_dict()) == infos[1]


def test_scheduler_dag_release_sequence():
    s, c = make_session({K.PREPARE_STAGE: "a,b", K.TRAINING_STAGE: "worker,ps"}, a=2, b=1, worker=2, ps=1)
    asked = []
    sch = TaskScheduler(s, lambda r: asked.append(r.job_name))
    sch.schedule_tasks()
    assert sorted(asked) == ["a", "b"] and s.num_expected_tasks == 3
    sch.register_dependency_completed("a")
    assert sorted(asked) == ["a", "b"]
    sch.register_dependency_completed("a")
    sch.register_dependency_completed("b")
    assert sorted(asked) == ["a", "b", "ps", "worker"] and s.num_expected_tasks == 6
    s, c = make_session({"tony.a.depends-on": "b", "tony.b.depends-on": "a"}, a=1, b=1)
    assert not is_dag(s.requests)
    sch = TaskScheduler(s, lambda r: None)
    sch.schedule_tasks()
    assert not sch.dependency_check_passed


def test_liveliness_monitor_expires_and_ping_keeps_alive():
    dead = []
    m = LivelinessMonitor(20, 3, dead.append)   # expire after 60 ms
    m.start()
    m.register("worker:0")
    m.register("worker:1")
    for _ in range(10):
        time.sleep(0.02)
        m.received_ping("worker:0")
    assert dead == ["worker:1"]
    m.unregister("worker:0")
    time.sleep(0.15)
    assert dead == ["worker:1"]
    m.stop()


# ---- client -----------------------------------------------------------------------------------
def test_cli_parser_forms():
    a = tc.parse_cli(["-executes", "a.py", "--task_params", "--x 1", "--conf=tony.a.instances=1",
                      "--conf", "k=v", "-shell_env", "A=1", "--shell_env", "B=2", "--src_dir=src"])
    assert a["executes"] == "a.py" and a["task_params"] == "--x 1" and a["src_dir"] == "src"
    assert a["conf"] == ["tony.a.instances=1", "k=v"] and a["shell_env"] == ["A=1", "B=2"]
    with pytest.raises(tc.UsageError):
        tc.parse_cli(["--nope", "1"])
    with pytest.raises(tc.UsageError):
        tc.parse_cli(["--executes"])


def test_client_init_writes_expected_conf(tmp_path):
    c = tc.TonyClient()
    assert c.init(["--executes", "train.py", "--task_params", "--epochs 2", "--python_binary_path", "python3",
                   "--shell_env", "A=1", "--container_env", "B=2", "--conf", "tony.worker.instances=2",
                   "--conf", "tony.containers.envs=C=3"])
    conf = c.get_tony_conf()
    assert conf.get(K.CONTAINERS_COMMAND) == "python3 train.py --epochs 2"
    assert conf.get(K.EXECUTION_ENVS) == "A=1"
    assert conf.get(K.CONTAINERS_ENVS) == "C=3,B=2"
    assert conf.get("tony.version-info.version")
    with pytest.raises(RuntimeError):
        tc.validate_tony_conf(conf_with(tony_worker_instances="5", tony_task_max__total__instances="4"))
    with pytest.raises(RuntimeError):
        tc.validate_tony_conf(conf_with(tony_worker_instances="2", tony_worker_memory="3g",
                                        tony_task_max__total__memory="4096"))


# ---- runtimes -----------------------------------------------------------------------------------
def test_horovod_worker_list_and_slots():
    s, _ = make_session(worker=4, driver=1)
    fill(s)
    hosts = ["h1", "h1", "h2", "h1"]
    for i, h in enumerate(hosts):
        s.get_task("worker", i).host = h
    a = fw.HorovodAMAdapter()
    a.set_session(s)
    same = []
    assert a.build_worker_list("h1", same) == "h1:3,h2:1" and sorted(same) == [0, 1, 3]
    slots = hd.host_assignments(hd.parse_worker_list("h1:3,h2:1"))
    assert [x["rank"] for x in slots] == [0, 1, 2, 3]
    assert [x["localRank"] for x in slots] == [0, 1, 2, 0]
    assert slots[3]["crossRank"] == 1 and slots[0]["crossSize"] == 2 and slots[1]["crossSize"] == 1
    assert a.construct_cluster_spec("worker:0") is None            # driver not ready yet
    assert a.construct_cluster_spec("driver:0") == "h1:3,h2:1"
    assert not a.receive_task_callback_info("worker:0", "{}")
    assert a.receive_task_callback_info("driver:0", json.dumps({"port": "77", "host": "h1", "slotInfos": slots}))
    spec = json.loads(a.construct_cluster_spec("worker:3"))
    assert spec["port"] == "77" and spec["sameHostTaskIndexList"] == [0, 1, 3]
    env = {}
    fw.set_horovod_run_env(env, spec, 3, "h1")                     # 3rd task on h1 -> local rank 2
    assert env["HOROVOD_LOCAL_RANK"] == "2" and env["HOROVOD_RANK"] == "2" and env["HOROVOD_SIZE"] == "4"
    assert env["HOROVOD_GLOO_RENDEZVOUS_PORT"] == "77" and env["HOROVOD_CONTROLLER"] == "gloo"


def test_horovod_config_rules_and_gang_only():
    a = fw.HorovodAMAdapter()
    c = conf_with(tony_worker_instances="2")
    assert a.validate_and_update_config(c)
    assert c.get_int("tony.driver.instances") == 1 and "driver" in c.untracked_job_types()
    a2 = fw.HorovodAMAdapter()
    assert not a2.validate_and_update_config(conf_with(tony_worker_instances="2", tony_driver_vcores="3"))
    s, _ = make_session(worker=1, driver=1)
    a.set_session(s)
    assert not a.can_start_task("FCFS", "worker:0") and "FCFS" in a.app_failed_reason


def test_generic_barrier_and_standalone():
    s, _ = make_session(worker=2)
    a = fw.AMAdapter()
    a.set_session(s)
    s.num_expected_tasks = 2
    s.registered.add("worker:0")
    assert not a.can_start_task("GANG", "worker:0") and a.can_start_task("FCFS", "worker:0")
    s.registered.add("worker:1")
    assert a.can_start_task("GANG", "worker:0")
    assert fw.StandaloneAMAdapter().validate_and_update_config(conf_with(tony_worker_instances="1"))
    assert not fw.StandaloneAMAdapter().validate_and_update_config(conf_with(tony_worker_instances="2"))
    with pytest.raises(ValueError):
        fw.get_am_adapter("caffe")


class _FakeExecutor:
    def __init__(self, conf, job, chief):
        self.conf, self.job_name, self.is_chief = conf, job, chief


def test_tb_port_rule():
    c = conf_with(tony_worker_instances="2")
    assert fw.TaskAdapter(_FakeExecutor(c, "worker", True)).need_reserve_tb_port()
    assert not fw.TaskAdapter(_FakeExecutor(c, "worker", False)).need_reserve_tb_port()
    c.set(K.TENSORBOARD_LOG_DIR, "/tmp/x")
    assert not fw.TaskAdapter(_FakeExecutor(c, "worker", True)).need_reserve_tb_port()
    assert fw.TaskAdapter(_FakeExecutor(c, "tensorboard", False)).need_reserve_tb_port()


def test_horovod_driver_process_fake_and_fail_modes(tmp_path):
    drv = fw.HorovodDriver.create("127.0.0.1:2", {}, TonyConfiguration(), str(tmp_path / "d"), test_mode=True)
    try:
        assert drv.port == 9999 and len(drv.slot_infos) == 2
        info = json.loads(drv.callback_info())
        assert info["port"] == "9999" and info["slotInfos"][1]["localRank"] == 1
    finally:
        drv.close()
    with pytest.raises(RuntimeError):
        fw.HorovodDriver.create("127.0.0.1:2", {}, TonyConfiguration(), str(tmp_path / "f"), fail=True)
    drv = fw.HorovodDriver.create("127.0.0.1:2", {}, TonyConfiguration(), str(tmp_path / "r"))
    try:
        import urllib.request

        url = f"http://127.0.0.1:{drv.port}/rendezvous/rank_and_size/127.0.0.1:1"
        assert json.loads(urllib.request.urlopen(url, timeout=5).read())["rank"] == 1
    finally:
        drv.close()


# ---- ports / monitor / gpu ------------------------------------------------------------------------
def test_port_reservation():
    p = ports.EphemeralPort.create()
    assert not ports.is_port_available(p.port)
    p.close()
    assert ports.is_port_available(p.port)
    r = ports.ReusablePort.create()
    s = socket.socket()
    s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEPORT, 1)
    s.bind(("", r.port))             # the user process can bind while the reservation is held
    s.close()
    plain = socket.socket()
    with pytest.raises(OSError):
        plain.bind(("", r.port))     # ... but nobody without SO_REUSEPORT can steal it
    plain.close()
    r.close()


def test_task_monitor_avg_max():
    pushed = []
    m = TaskMonitor(pushed.append, interval_ms=1000, gpu_enabled=False)
    info = topology.GpuDeviceInformation(gpus=[topology.PerGpuInfo(gpu_util_pct=10, fb_total_mib=100, fb_used_mib=50,
                                                                  bar1_total_mib=10, bar1_used_mib=1)])
    m.refresh(info, rss=100)
    info.gpus[0].gpu_util_pct = 30
    info.gpus[0].fb_used_mib = 20
    m.refresh(info, rss=300)
    assert m.metrics[MAX_MEMORY_BYTES] == 300 and m.metrics[AVG_MEMORY_BYTES] == 200
    assert m.metrics[AVG_GPU_UTILIZATION] == 20 and m.metrics[MAX_GPU_FB_MEMORY_USAGE] == 50
    assert process_tree_rss_bytes(os.getpid()) > 0


def test_gpu_xml_parser_on_reference_sample():
    sample = os.path.join(REF, "tony-core/src/test/resources/nvidia-smi-sample-xml-output")
    if not os.path.exists(sample):
        pytest.skip("reference fixture not mounted")
    info = topology.parse_smi_xml(open(sample).read())
    assert info.driver_version and len(info.gpus) >= 1
    g = info.gpus[0]
    assert g.fb_total_mib > 0 and 0 <= g.fb_usage_pct <= 100 and g.product_name
    assert topology.GpuDiscoverer("/nonexistent").binary is None


# ---- rpc ------------------------------------------------------------------------------------------
def test_rpc_auth_acl_and_errors():
    class H:
        def getClusterSpec(self):
            return "{}"

        def finishApplication(self):
            raise ValueError("nope")

        def secretMethod(self):
            return 1

    srv = RpcServer(H(), ("getClusterSpec", "finishApplication"), secret=b"k" * 32).start()
    try:
        good = RpcClient("127.0.0.1", srv.port, b"k" * 32, retries=1)
        assert good.call("getClusterSpec") == "{}"
        with pytest.raises(RpcError, match="nope"):
            good.call("finishApplication")
        with pytest.raises(RpcError, match="not part of this protocol"):
            good.call("secretMethod")
        bad = RpcClient("127.0.0.1", srv.port, b"x" * 32, retries=1)
        with pytest.raises(RpcError):
            bad.call("getClusterSpec")
    finally:
        srv.stop()


# ---- history ----------------------------------------------------------------------------------------
def test_hist_file_names():
    assert history.inprogress_name("application_1_2", 5, "bob") == "application_1_2-5-bob.jhist.inprogress"
    assert history.finished_name("application_1_2", 5, 9, "bob", "FAILED") == "application_1_2-5-9-bob-FAILED.jhist"
    ok = history.is_valid_hist_file_name
    assert ok("application_1_2-5-bob.jhist.inprogress") and ok("application_1_2-5-9-bob-SUCCEEDED.jhist")
    assert not ok("application_1_2-5-Bob.jhist.inprogress")          # user must be lower case
    assert not ok("application_1_2-5-9-bob-succeeded.jhist")         # status must be upper case
    assert not ok("job_1_2-5-9-bob-SUCCEEDED.jhist") and not ok("application_1_2-x-9-bob-SUCCEEDED.jhist")
    assert not ok("application_1_2-5-9-bob.jhist") and not ok("application_1_2.jhist")
    md = history.JobMetadata.from_file_name("application_1_2-5-9-bob-KILLED.jhist")
    assert (md.id, md.started, md.completed, md.user, md.status) == ("application_1_2", 5, 9, "bob", "KILLED")


def test_avro_roundtrip_and_reference_sample(tmp_path):
    evs = [schema.application_inited("application_1_0001", 3, "h", "c1"), schema.task_started("worker", 0, "h"),
           schema.task_finished("worker", 0, "SUCCEEDED", {"MAX_MEMORY_BYTES": 1.5}, None),
           schema.application_finished("application_1_0001", 1, 0)]
    p = tmp_path / "x.jhist"
    w = avro.DataFileWriter(str(p), schema.EVENT)
    for e in evs:
        w.append(e)
        w.flush()
    w.close()
    raw = p.read_bytes()
    assert raw[:4] == b"Obj\x01" and b"avro.schema" in raw[:40] and b"avro.codec" not in raw
    sch, objs = avro.read_container(str(p))
    assert objs == evs and sch["name"] == "Event"
    assert [f["name"] for f in sch["fields"]] == ["type", "event", "timestamp"]
    for n in (0, 1, -1, 63, 64, -65, 2 ** 40, -(2 ** 40)):
        import io

        assert avro.decode_long(io.BytesIO(avro.encode_long(n))) == n
    assert avro.encode_long(-1) == b"\x01" and avro.encode_long(64) == b"\x80\x01"
    sample_dir = os.path.join(REF, "tony-portal/example/tony-history/finished/2019/01/08/application_123456_0001")
    if os.path.isdir(sample_dir):
        events = history.parse_events(sample_dir)
        assert [e.type for e in events] == ["APPLICATION_INITED", "APPLICATION_FINISHED"]
        assert history.parse_metadata(sample_dir).user == "testuser"
        assert len(history.parse_config(sample_dir)) > 0


def test_history_mover_and_purger(tmp_path):
    inter, fin = tmp_path / "intermediate", tmp_path / "finished"
    done = inter / "application_1_0001"
    running = inter / "application_1_0002"
    killed = inter / "application_1_0003"
    for d in (done, running, killed):
        d.mkdir(parents=True)
    t = 1546910517303
    (done / history.finished_name("application_1_0001", t - 10, t, "bob", "SUCCEEDED")).write_bytes(b"x")
    (running / history.inprogress_name("application_1_0002", t, "bob")).write_bytes(b"x")
    (killed / history.inprogress_name("application_1_0003", t, "bob")).write_bytes(b"x")
    mover = history.HistoryFileMover(str(inter), str(fin), killed_apps=lambda: {"application_1_0003"})
    moved = mover.run_once()
    assert (fin / "2019" / "01" / "08" / "application_1_0001").is_dir()
    assert running.is_dir() and len(moved) == 2
    assert any(n.endswith("-bob-KILLED.jhist") for p in moved if "0003" in p for n in os.listdir(p))
    assert history.find_job_dir(str(fin), str(inter), "application_1_0001").endswith("application_1_0001")
    purger = history.HistoryFilePurger(str(fin), retention_sec=30 * 86400)
    removed = purger.run_once(now=t / 1000 + 31 * 86400)
    assert any("2019/01/08" in r for r in removed) and not (fin / "2019").exists()
