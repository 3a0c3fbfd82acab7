"""Portal routes/caches and the workflow job-type adapter (reference tiers: tony-portal/test/**,
tony-azkaban/src/test/.../TestTonyJob.java)."""
import json
import os
import sys
import urllib.request

from tony_b200.azkaban import tony_job
from tony_b200.cluster.mini import MiniCluster
from tony_b200.portal.server import Portal
from tony_b200.proxy.proxy_server import ProxyServer

SCRIPTS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scripts")


def _get(port, path):
    with urllib.request.urlopen(f"http://127.0.0.1:{port}{path}", timeout=10) as r:
        return r.status, r.read().decode()


def test_portal_serves_history_of_a_real_job():
    os.environ["TONY_B200_QUIET"] = "1"
    os.environ.setdefault("TONY_B200_FAKE_GPUS", "0")
    with MiniCluster() as c:
        client = c.new_client()
        assert client.init(["--executes", os.path.join(SCRIPTS, "exit_0.py"), "--python_binary_path", sys.executable,
                            "--conf", "tony.worker.instances=2", "--conf", "tony.am.wait-client-signal-stop-timeout-sec=1"])
        assert client.start() == 0
        portal = Portal(c.base_conf())
        port = portal.start(0)
        try:
            # the mover ran at start-up: the finished job now lives under finished/yyyy/MM/dd
            assert not os.listdir(os.path.join(c.history_dir, "intermediate"))
            code, body = _get(port, "/")
            assert code == 200 and client.app_id in body and "SUCCEEDED" in body
            jobs = json.loads(_get(port, "/api/")[1])
            assert jobs[0]["id"] == client.app_id and jobs[0]["status"] == "SUCCEEDED"
            events = json.loads(_get(port, f"/api/jobs/{client.app_id}")[1])
            assert events[0]["type"] == "APPLICATION_INITED" and events[-1]["type"] == "APPLICATION_FINISHED"
            cfg = json.loads(_get(port, f"/api/config/{client.app_id}")[1])
            assert any(x["name"] == "tony.worker.instances" and x["value"] == "2" for x in cfg)
            logs = json.loads(_get(port, f"/api/logs/{client.app_id}")[1])
            assert len(logs) == 2 and logs[0]["task_type"] == "worker"
            assert "Events of" in _get(port, f"/jobs/{client.app_id}")[1]
            try:
                _get(port, "/jobs/application_9_9999")
                assert False
            except urllib.error.HTTPError as e:
                assert e.code == 404
            # proxy relays to the portal
            px = ProxyServer("127.0.0.1", port).start()
            try:
                assert _get(px.local_port, "/api/")[0] == 200
            finally:
                px.stop()
        finally:
            portal.stop()


def test_azkaban_job_arguments_and_conf(tmp_path):
    props = {"tony.worker.instances": "2", "tony.application.framework": "pytorch", "executes": "train.py",
             "task_params": "--epochs 2", "python_binary_path": "python3", "python_venv": "venv.zip",
             "worker_env.FOO": "bar", "src_dir": "mysrc", "azkaban.flow.execid": "77",
             "azkaban.flow.flowid": "flow", "azkaban.input.dataset": "a,b", "unrelated": "x"}
    conf = {k: v for k, v, _ in tony_job.job_configuration(props)}
    assert conf["tony.worker.instances"] == "2" and "unrelated" not in conf
    assert conf["tony.application.tags"] == "execid:77,flowid:flow"
    args = tony_job.main_arguments(props, "tony.xml")
    assert args[:4] == ["-conf_file", "tony.xml", "-src_dir", "mysrc"]
    for flag, val in (("-executes", "train.py"), ("-task_params", "--epochs 2"), ("-python_venv", "venv.zip")):
        assert args[args.index(flag) + 1] == val
    assert "FOO=bar" in args and "AZKABAN_INPUT_DATASET=a;b" in args
    job = tmp_path / "x.job"
    job.write_text("# comment\ntony.worker.instances=1\nexecutes = a.py\n")
    assert tony_job.read_properties(str(job)) == {"tony.worker.instances": "1", "executes": "a.py"}
