"""Framework runtime SPI: an ApplicationMaster half and a TaskExecutor half per framework.

Reference SPI: tony-core/src/main/java/com/linkedin/tony/Framework.java:33-67, loaded through
FrameworkRuntimeProvider.java:29-46.  Implementations mirrored here with the same semantics:
generic GANG/FCFS barrier + TB-port rule (runtime/MLGenericRuntime.java:79-97,161-177),
TensorFlow (TFRuntime.java:45-59), PyTorch (PyTorchRuntime.java:45-56), MXNet
(MXNetRuntime.java:44-66), Horovod (HorovodRuntime.java:57-59,100-232,245-350) and
Standalone (StandaloneRuntime.java:69-75).

Differences by design: the PyTorch runtime exports the SUPERSET contract — the reference's
``INIT_METHOD``/``RANK``/``WORLD`` plus torch ``env://`` names (``MASTER_ADDR``,
``MASTER_PORT``, ``WORLD_SIZE``, ``LOCAL_RANK``, ``LOCAL_WORLD_SIZE``) with a global rank that
is unique across all tracked job types (the reference's ``RANK = TASK_INDEX`` collides as
soon as two tracked types exist).
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time
from typing import Dict, List, Optional

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration
from ..utils import utils as U
from . import horovod_driver as hd

SIDECAR_TB_ROLE_NAME = K.TENSORBOARD_JOB_NAME


# ======================================================================================
# ApplicationMaster side
# ======================================================================================
class AMAdapter:
    """Generic ML runtime: gang barrier or first-come-first-served."""

    illegal_conf_regexes: List[str] = []

    def __init__(self):
        self.session = None
        self.app_failed_reason: Optional[str] = None

    def set_session(self, session) -> None:
        self.session = session

    def construct_cluster_spec(self, task_id: str) -> Optional[str]:
        return self.session.cluster_spec_json()

    def destroy(self) -> None:
        pass

    def receive_task_callback_info(self, task_id: str, callback_info: str) -> bool:
        return True

    def can_start_task(self, mode: str, task_id: str) -> bool:
        if mode == "GANG":
            return self.session.num_registered() == self.session.num_expected_tasks
        return mode == "FCFS"

    def validate_and_update_config(self, conf: TonyConfiguration) -> bool:
        import re

        for rx in self.illegal_conf_regexes:
            pat = re.compile(rx)
            bad = [k for k, _ in conf.items() if pat.match(k)]
            if bad:
                self.app_failed_reason = f"Not allowed to configure illegal conf in runtime: {bad}"
                return False
        return True


class HorovodAMAdapter(AMAdapter):
    illegal_conf_regexes = [r"tony\.driver\.(?!command$).*"]

    def __init__(self):
        super().__init__()
        self.driver_ready = False
        self.slot_infos: List[dict] = []
        self.rendezvous_port: Optional[str] = None
        self.rendezvous_host: Optional[str] = None

    def validate_and_update_config(self, conf: TonyConfiguration) -> bool:
        debug = conf.get_bool(K.HOROVOD_DRIVER_DEBUG, False)
        if not super().validate_and_update_config(conf):
            return False
        if debug and not conf.get(K.command_key(K.DRIVER_JOB_NAME)):
            self.app_failed_reason = "horovod driver debug mode needs tony.driver.command"
            return False
        # inject the driver role: 1 instance, 1 vcore, untracked (HorovodRuntime.java:210-232)
        conf.set(K.instances_key(K.DRIVER_JOB_NAME), "1", "horovod-runtime")
        conf.set(K.vcores_key(K.DRIVER_JOB_NAME), "1", "horovod-runtime")
        untracked = conf.get_strings(K.UNTRACKED_JOBTYPES)
        if K.DRIVER_JOB_NAME not in untracked:
            untracked.append(K.DRIVER_JOB_NAME)
            conf.set(K.UNTRACKED_JOBTYPES, ",".join(untracked), "horovod-runtime")
        return True

    def _is_driver(self, task_id: str) -> bool:
        return task_id.split(":")[0] == K.DRIVER_JOB_NAME

    def build_worker_list(self, task_host: Optional[str], same_host_indices: List[int]) -> str:
        per_host: Dict[str, int] = {}
        for t in self.session.all_tasks():
            if t.job_name == K.DRIVER_JOB_NAME or t.host is None:
                continue
            per_host[t.host] = per_host.get(t.host, 0) + 1
            if t.host == task_host:
                same_host_indices.append(t.index)
        return ",".join(f"{h}:{n}" for h, n in per_host.items())

    def construct_cluster_spec(self, task_id: str) -> Optional[str]:
        task = self.session.get_task_by_id(task_id)
        same: List[int] = []
        worker_list = self.build_worker_list(task.host if task else None, same)
        same.sort()
        if self._is_driver(task_id):
            return worker_list
        if not self.driver_ready:
            return None
        # workerAddrs is an addition to the reference's HorovodClusterSpec: the registered
        # host:port of every worker, so the torch env:// rendezvous can be exported as well
        addrs = [t.host_port for t in self.session.job_tasks.get(K.WORKER_JOB_NAME, []) if t is not None]
        return json.dumps({"slotInfos": self.slot_infos, "port": self.rendezvous_port,
                           "amHost": self.rendezvous_host, "sameHostTaskIndexList": same,
                           "workerAddrs": addrs}, separators=(",", ":"))

    def receive_task_callback_info(self, task_id: str, callback_info: str) -> bool:
        if not self._is_driver(task_id):
            return False
        info = json.loads(callback_info)
        self.slot_infos = info.get("slotInfos", [])
        self.rendezvous_port = str(info.get("port"))
        self.rendezvous_host = info.get("host")
        self.driver_ready = True
        return True

    def can_start_task(self, mode: str, task_id: str) -> bool:
        if mode != "GANG":
            self.app_failed_reason = f"Horovod don't support {mode} distributed mode."
            return False
        if self.session.num_registered() != self.session.num_expected_tasks:
            return False
        return self._is_driver(task_id) or self.driver_ready


class StandaloneAMAdapter(AMAdapter):
    def validate_and_update_config(self, conf: TonyConfiguration) -> bool:
        if conf.total_instances() != 1:
            self.app_failed_reason = "standalone framework runs exactly one task instance"
            return False
        return True

    def construct_cluster_spec(self, task_id: str) -> Optional[str]:
        return self.session.cluster_spec_json()


# ======================================================================================
# TaskExecutor side
# ======================================================================================
class TaskAdapter:
    def __init__(self, executor):
        self.executor = executor

    def need_reserve_tb_port(self) -> bool:
        ex = self.executor
        sidecar_tb = bool(ex.conf.get(K.TENSORBOARD_LOG_DIR)) or ex.conf.num_instances(SIDECAR_TB_ROLE_NAME) > 0
        if not sidecar_tb:
            return ex.is_chief
        return ex.job_name == SIDECAR_TB_ROLE_NAME

    def build_task_env(self) -> None:
        raise NotImplementedError

    def run(self) -> int:
        self.build_task_env()
        return self.executor.execute_user_command()

    def _basic_env(self) -> Dict[str, str]:
        ex = self.executor
        env = ex.shell_env
        env[K.JOB_NAME] = ex.job_name
        env[K.TASK_INDEX] = str(ex.task_index)
        env[K.TASK_NUM] = str(ex.num_tasks)
        env[K.DISTRIBUTED_MODE_NAME] = ex.distributed_mode
        return env


class TFTaskAdapter(TaskAdapter):
    def build_task_env(self) -> None:
        ex = self.executor
        env = self._basic_env()
        if ex.distributed_mode == "GANG" and ex.cluster_spec:
            env[K.CLUSTER_SPEC] = ex.cluster_spec
            env[K.TF_CONFIG] = U.construct_tf_config(ex.cluster_spec, ex.job_name, ex.task_index)


class PyTorchTaskAdapter(TaskAdapter):
    def build_task_env(self) -> None:
        ex = self.executor
        env = self._basic_env()
        env[K.CLUSTER_SPEC] = ex.cluster_spec or "{}"
        init = U.parse_cluster_spec_for_pytorch(ex.cluster_spec) if ex.cluster_spec else None
        if init is None:
            raise RuntimeError("Errors on getting init method.")
        # reference contract
        env[K.INIT_METHOD] = init
        env[K.WORLD] = str(ex.num_tasks)
        # superset contract: globally unique rank over tracked job types, env:// names
        spec = json.loads(ex.cluster_spec)
        tracked = [j for j in spec if ex.conf.is_job_type_monitored(j)]
        # worker first so worker:0 (whose address is the rendezvous) is global rank 0
        tracked.sort(key=lambda j: (j != K.WORKER_JOB_NAME, j != K.CHIEF_JOB_NAME, j))
        rank, world = 0, 0
        for j in tracked:
            if j == ex.job_name:
                rank = world + ex.task_index
            world += len(spec[j])
        if ex.job_name not in tracked:
            rank = ex.task_index
        env[K.RANK] = str(rank)
        # numbering over EVERY job type (tracked first, then untracked such as ps): a job whose ps tasks
        # hold GPUs joins them into the same process group with runtime.init(include_ps=True)
        everyone = tracked + sorted(j for j in spec if j not in tracked)
        all_rank, all_world, ps_ranks = 0, 0, []
        for j in everyone:
            if j == ex.job_name:
                all_rank = all_world + ex.task_index
            if j == K.PS_JOB_NAME:
                ps_ranks = list(range(all_world, all_world + len(spec[j])))
            all_world += len(spec[j])
        env["TONY_B200_ALL_RANK"] = str(all_rank)
        env["TONY_B200_ALL_WORLD"] = str(all_world)
        env["TONY_B200_PS_RANKS"] = ",".join(str(r) for r in ps_ranks)
        host, port = init[len(K.COMMUNICATION_BACKEND):].rsplit(":", 1)
        env["MASTER_ADDR"] = host
        env["MASTER_PORT"] = port
        env["WORLD_SIZE"] = str(world if world else ex.num_tasks)
        env["LOCAL_RANK"] = "0" if os.environ.get("CUDA_VISIBLE_DEVICES") else str(rank)
        my_host = U.local_hostname()
        env["LOCAL_WORLD_SIZE"] = str(sum(1 for j in tracked for a in spec[j] if a.rsplit(":", 1)[0] == my_host))


class MXNetTaskAdapter(TaskAdapter):
    def build_task_env(self) -> None:
        ex = self.executor
        env = self._basic_env()
        env[K.CLUSTER_SPEC] = ex.cluster_spec or "{}"
        parts = U.parse_cluster_spec_for_mxnet(ex.cluster_spec)
        if parts is None:
            raise RuntimeError("Errors on getting dmlc server address.")
        spec = json.loads(ex.cluster_spec)
        env["DMLC_ROLE"] = ex.job_name
        env["DMLC_PS_ROOT_URI"] = parts[0]
        env["DMLC_PS_ROOT_PORT"] = parts[1]
        env["DMLC_LOCAL"] = "0"
        env["DMLC_NUM_SERVER"] = str(len(spec.get(K.SERVER_JOB_NAME, [])))
        env["DMLC_NUM_WORKER"] = str(len(spec.get(K.WORKER_JOB_NAME, [])))
        env["PS_VERBOSE"] = "1"


class StandaloneTaskAdapter(TaskAdapter):
    def build_task_env(self) -> None:
        env = self._basic_env()
        env[K.CLUSTER_SPEC] = self.executor.cluster_spec or "{}"


class HorovodDriver:
    """Spawn the rendezvous driver and wait for its port file (HorovodDriver.java:97-216)."""

    def __init__(self, proc: subprocess.Popen, port: int, host: str, slot_infos: List[dict], out_dir: str):
        self.proc, self.port, self.host, self.slot_infos, self.out_dir = proc, port, host, slot_infos, out_dir

    @staticmethod
    def create(worker_list: str, env: Dict[str, str], conf: TonyConfiguration, out_dir: str,
               debug_command: Optional[str] = None, test_mode: bool = False, fail: bool = False,
               wait_s: float = 10.0) -> "HorovodDriver":
        os.makedirs(out_dir, exist_ok=True)
        full_env = dict(os.environ)
        full_env.update(env)
        full_env["CLUSTER_WORKER_LIST"] = worker_list
        full_env["DRIVER_OUTPUT_PATH"] = out_dir
        if debug_command:
            cmd = ["bash", "-c", debug_command]
        else:
            cmd = [sys.executable, os.path.abspath(hd.__file__), "-w", worker_list, "-o", out_dir]
            if test_mode or fail:
                cmd += ["-t", "-p", "9999"]
            if fail:
                cmd += ["-f"]
        proc = subprocess.Popen(cmd, env=full_env, start_new_session=True)
        deadline = time.monotonic() + wait_s
        while time.monotonic() < deadline:
            if proc.poll() is not None:
                raise RuntimeError(f"Horovod driver exited with {proc.returncode} before publishing its port")
            for name in os.listdir(out_dir):
                if name.endswith(hd.PORT_FILE_SUFFIX):
                    port = int(name[: -len(hd.PORT_FILE_SUFFIX)])
                    with open(os.path.join(out_dir, name)) as f:
                        slots = json.load(f)
                    return HorovodDriver(proc, port, U.local_hostname(), slots, out_dir)
            time.sleep(0.05)
        U.kill_process_group(proc)
        raise RuntimeError("Horovod driver did not publish its port file in time")

    def callback_info(self) -> str:
        return json.dumps({"port": str(self.port), "host": self.host, "slotInfos": self.slot_infos})

    def wait(self) -> int:
        return self.proc.wait()

    def close(self) -> None:
        U.kill_process_group(self.proc)


class HorovodTaskAdapter(TaskAdapter):
    def build_task_env(self) -> None:
        ex = self.executor
        env = self._basic_env()
        env[K.CLUSTER_SPEC] = ex.cluster_spec or ""
        if ex.job_name == K.DRIVER_JOB_NAME:
            return
        spec = json.loads(ex.cluster_spec)
        set_horovod_run_env(env, spec, ex.task_index, U.local_hostname())

    def run(self) -> int:
        ex = self.executor
        self.build_task_env()
        if ex.job_name != K.DRIVER_JOB_NAME:
            return ex.execute_user_command()
        conf = ex.conf
        debug_cmd = conf.get(K.command_key(K.DRIVER_JOB_NAME)) if conf.get_bool(K.HOROVOD_DRIVER_DEBUG, False) else None
        driver = HorovodDriver.create(
            ex.cluster_spec, ex.shell_env, conf, os.path.join(os.getcwd(), "horovod_driver"),
            debug_command=debug_cmd, test_mode=conf.get_bool(K.HOROVOD_TEST_MODE, False),
            fail=conf.get_bool(K.HOROVOD_TEST_FAST_FAIL, False))
        ex.register_child(driver.proc)
        ex.callback_info_to_am(driver.callback_info())
        return driver.wait()


def set_horovod_run_env(env: Dict[str, str], spec: dict, task_index: int, host: str) -> None:
    """Pick this task's slot: the k-th local-rank slot on this host, k = position of the task
    index among the task indices placed on this host (HorovodRuntime.java:312-350)."""
    env["HOROVOD_CONTROLLER"] = "gloo"
    env["HOROVOD_CPU_OPERATIONS"] = "gloo"
    env["HOROVOD_GLOO_TIMEOUT_SECONDS"] = "2000"
    env["HOROVOD_GLOO_RENDEZVOUS_PORT"] = str(spec["port"])
    env["HOROVOD_GLOO_RENDEZVOUS_ADDR"] = str(spec["amHost"])
    local = sorted((s for s in spec["slotInfos"] if s["hostname"] == host), key=lambda s: s["localRank"])
    seq = spec["sameHostTaskIndexList"].index(task_index)
    slot = local[seq]
    env["HOROVOD_CROSS_RANK"] = str(slot["crossRank"])
    env["HOROVOD_CROSS_SIZE"] = str(slot["crossSize"])
    env["HOROVOD_LOCAL_RANK"] = str(slot["localRank"])
    env["HOROVOD_LOCAL_SIZE"] = str(slot["localSize"])
    env["HOROVOD_SIZE"] = str(slot["size"])
    env["HOROVOD_RANK"] = str(slot["rank"])
    env["HOROVOD_HOSTNAME"] = str(slot["hostname"])
    # torch env:// names too, so a job using our hvd shim over torch.distributed just works
    addrs = spec.get("workerAddrs") or []
    if addrs and addrs[0]:
        host0, port0 = addrs[0].rsplit(":", 1)
        env.setdefault("MASTER_ADDR", host0)
        env.setdefault("MASTER_PORT", port0)
    env.setdefault("RANK", str(slot["rank"]))
    env.setdefault("WORLD_SIZE", str(slot["size"]))
    env.setdefault("LOCAL_RANK", "0" if os.environ.get("CUDA_VISIBLE_DEVICES") else str(slot["localRank"]))


# ======================================================================================
_AM = {"tensorflow": AMAdapter, "pytorch": AMAdapter, "mxnet": AMAdapter,
       "horovod": HorovodAMAdapter, "standalone": StandaloneAMAdapter}
_TASK = {"tensorflow": TFTaskAdapter, "pytorch": PyTorchTaskAdapter, "mxnet": MXNetTaskAdapter,
         "horovod": HorovodTaskAdapter, "standalone": StandaloneTaskAdapter}


def register_framework(name: str, am_adapter_cls, task_adapter_cls) -> None:
    """Plug in a runtime for ``tony.application.framework=<name>`` (the reference discovers
    AbstractFrameworkRuntime implementations through java.util.ServiceLoader:
    FrameworkRuntimeProvider.java:29-46).  Also loaded from the entry-point style env var
    ``TONY_B200_FRAMEWORK_PLUGINS=module1,module2``: each module is imported and may call this."""
    _AM[name.lower()] = am_adapter_cls
    _TASK[name.lower()] = task_adapter_cls


def _load_plugins() -> None:
    import importlib

    for mod in filter(None, os.environ.get("TONY_B200_FRAMEWORK_PLUGINS", "").split(",")):
        importlib.import_module(mod.strip())


def get_am_adapter(framework: str) -> AMAdapter:
    _load_plugins()
    try:
        return _AM[framework.lower()]()
    except KeyError:
        raise ValueError(f"Unsupported framework {framework!r}; expected one of {sorted(_AM)}") from None


def get_task_adapter(framework: str, executor) -> TaskAdapter:
    _load_plugins()
    try:
        return _TASK[framework.lower()](executor)
    except KeyError:
        raise ValueError(f"Unsupported framework {framework!r}; expected one of {sorted(_TASK)}") from None
