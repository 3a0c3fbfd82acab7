"""Configuration keys, defaults and environment-variable names of the TonY contract.

Same key strings, defaults and per-job-type templates as the reference
(tony-core/src/main/java/com/linkedin/tony/TonyConfigurationKeys.java:14-336,
tony-core/src/main/resources/tony-default.xml, Constants.java:15-196) so an existing
``tony.xml`` submits unchanged.  Where the reference's class default and XML default
disagree (``tony.application.name``, ``tony.container.allocation.timeout``) the XML value
is what a running job sees there, so it is the default here.  Keys that only make sense on
YARN/HDFS (queue, node labels, keytabs) are accepted and recorded in ``tony-final.xml`` but
have no effect on the local B200 cluster; new ``tony.b200.*`` keys configure the data plane.
"""
from __future__ import annotations

import re

TONY_PREFIX = "tony."

# ---- static keys --------------------------------------------------------------------------
OTHER_NAMENODES = "tony.other.namenodes"
YARN_QUEUE = "tony.yarn.queue"
APPLICATION_NAME = "tony.application.name"
APPLICATION_TYPE = "tony.application.type"
APPLICATION_TAGS = "tony.application.tags"
APPLICATION_NODE_LABEL = "tony.application.node-label"
ENABLE_PREPROCESS = "tony.application.enable-preprocess"
APPLICATION_TIMEOUT = "tony.application.timeout"
FRAMEWORK = "tony.application.framework"
NUM_CLIENT_RM_CONNECT_RETRIES = "tony.application.num-client-rm-connect-retries"
DISTRIBUTED_MODE = "tony.application.distributed-mode"
PREPARE_STAGE = "tony.application.prepare-stage"
TRAINING_STAGE = "tony.application.training-stage"
UNTRACKED_JOBTYPES = "tony.application.untracked.jobtypes"
SIDECAR_JOBTYPES = "tony.application.sidecar.jobtypes"
STOP_ON_FAILURE_JOBTYPES = "tony.application.stop-on-failure-jobtypes"
FAIL_ON_WORKER_FAILURE = "tony.application.fail-on-worker-failure-enabled"
SECURITY_ENABLED = "tony.application.security.enabled"
HDFS_CONF_PATH = "tony.application.hdfs-conf-path"
YARN_CONF_PATH = "tony.application.yarn-conf-path"
MAPRED_CONF_PATH = "tony.application.mapred-conf-path"
HADOOP_LOCATION = "tony.application.hadoop.location"
HADOOP_CLASSPATH = "tony.application.hadoop.classpath"
TENSORBOARD_LOG_DIR = "tony.application.tensorboard-log-dir"

MAX_TOTAL_INSTANCES = "tony.task.max-total-instances"
MAX_TOTAL_GPUS = "tony.task.max-total-gpus"
AM_JVM_OPTS = "tony.task.am.jvm.opts"
EXECUTOR_JVM_OPTS = "tony.task.executor.jvm.opts"
HEARTBEAT_INTERVAL_MS = "tony.task.heartbeat-interval-ms"
MAX_MISSED_HEARTBEATS = "tony.task.max-missed-heartbeats"
METRICS_INTERVAL_MS = "tony.task.metrics-interval-ms"
GPU_METRICS_ENABLED = "tony.task.gpu-metrics.enabled"

AM_RETRY_COUNT = "tony.am.retry-count"
AM_MEMORY = "tony.am.memory"
AM_VCORES = "tony.am.vcores"
AM_GPUS = "tony.am.gpus"
AM_COMMAND = "tony.am.command"
AM_WAIT_CLIENT_STOP_TIMEOUT_SEC = "tony.am.wait-client-signal-stop-timeout-sec"

CONTAINER_ALLOCATION_TIMEOUT = "tony.container.allocation.timeout"
WORKER_TIMEOUT = "tony.worker.timeout"

CONTAINERS_RESOURCES = "tony.containers.resources"
CONTAINERS_COMMAND = "tony.containers.command"
CONTAINERS_ENVS = "tony.containers.envs"
EXECUTION_ENVS = "tony.execution.envs"
MULTI_VALUE_KEYS = (CONTAINERS_RESOURCES, CONTAINERS_ENVS, EXECUTION_ENVS)

DOCKER_ENABLED = "tony.docker.enabled"
DOCKER_CONTAINERS_IMAGE = "tony.docker.containers.image"
DOCKER_CONTAINERS_MOUNT = "tony.docker.containers.mount"

GPU_EXEC_PATH = "tony.gpu-exec-path"
PYTHON_EXEC_PATH = "tony.python-exec-path"

HOROVOD_TEST_MODE = "tony.horovod.mode.test"
HOROVOD_TEST_FAST_FAIL = "tony.horovod.mode.test.fast.fail"
HOROVOD_DRIVER_DEBUG = "tony.horovod.driver.mode.debug"

PORTAL_URL = "tony.portal.url"
HISTORY_LOCATION = "tony.history.location"
HISTORY_INTERMEDIATE = "tony.history.intermediate"
HISTORY_FINISHED = "tony.history.finished"
HISTORY_MOVER_INTERVAL_MS = "tony.history.mover-interval-ms"
HISTORY_RETENTION_SEC = "tony.history.retention-sec"
HISTORY_PURGER_INTERVAL_MS = "tony.history.purger-interval-ms"
HISTORY_FINISHED_DIR_TIMEZONE = "tony.history.finished-dir-timezone"
PORTAL_CACHE_MAX_ENTRIES = "tony.portal.cache.max-entries"
KEYTAB_USER = "tony.keytab.user"
KEYTAB_LOCATION = "tony.keytab.location"
HTTPS_PORT = "tony.https.port"
HTTPS_KEYSTORE_PATH = "tony.https.keystore.path"
HTTPS_KEYSTORE_TYPE = "tony.https.keystore.type"
HTTPS_KEYSTORE_PASSWORD = "tony.https.keystore.password"
HTTPS_KEYSTORE_ALGORITHM = "tony.https.keystore.algorithm"
HTTP_PORT = "tony.http.port"
SECRET_KEY = "tony.secret.key"
VERSION_INFO_PREFIX = "tony.version-info."

# ---- B200 data-plane / local-cluster extensions (new; no reference counterpart) ------------
B200_WORK_DIR = "tony.b200.work-dir"
B200_NUM_GPUS = "tony.b200.cluster.gpus"          # -1 = discover
B200_CLUSTER_MEMORY = "tony.b200.cluster.memory"
B200_CLUSTER_VCORES = "tony.b200.cluster.vcores"
B200_KERNELS = "tony.b200.kernels"                # fused | stock
B200_BUCKET_MB = "tony.b200.bucket-mb"
B200_COMM_CTAS = "tony.b200.comm-ctas"
B200_PIN_NUMA = "tony.b200.pin-numa"

# ---- per-job-type key templates (TonyConfigurationKeys.java:189-257) -----------------------
INSTANCES_REGEX = re.compile(r"^tony\.([a-z]+)\.instances$")
MAX_TOTAL_RESOURCES_REGEX = re.compile(r"^tony\.task\.max-total-([a-z]+)$")
RESERVED_NON_JOB_PREFIXES = ("application", "task", "am", "containers", "docker", "history",
                             "portal", "keytab", "https", "http", "secret", "horovod", "yarn",
                             "other", "container", "execution", "b200", "version-info",
                             "gpu-exec-path", "python-exec-path")


def instances_key(job: str) -> str:
    return f"tony.{job}.instances"


def max_instances_key(job: str) -> str:
    return f"tony.{job}.max-instances"


def memory_key(job: str) -> str:
    return f"tony.{job}.memory"


def vcores_key(job: str) -> str:
    return f"tony.{job}.vcores"


def gpus_key(job: str) -> str:
    return f"tony.{job}.gpus"


def resource_key(job: str, resource: str) -> str:
    return f"tony.{job}.{resource}"


def resources_key(job: str) -> str:
    return f"tony.{job}.resources"


def command_key(job: str) -> str:
    return f"tony.{job}.command"


def node_label_key(job: str) -> str:
    return f"tony.{job}.node-label"


def depends_on_key(job: str) -> str:
    return f"tony.{job}.depends-on"


def docker_image_key(job: str) -> str:
    return f"tony.docker.{job}.image"


def max_total_resource_key(resource: str) -> str:
    return f"tony.task.max-total-{resource}"


DEFAULT_MEMORY = "2g"
DEFAULT_VCORES = 1
DEFAULT_GPUS = 0

FRAMEWORKS = ("tensorflow", "pytorch", "horovod", "mxnet", "standalone")
DISTRIBUTED_MODES = ("GANG", "FCFS")

# (key, default, description) — the single source for tony-default.xml
DEFAULTS = [
    (OTHER_NAMENODES, "", "Extra file-system URIs a job needs tokens for (recorded only on the local cluster)."),
    (YARN_QUEUE, "default", "Scheduler queue name; recorded in the history file."),
    (APPLICATION_NAME, "TensorFlowApplication", "Display name of the application."),
    (APPLICATION_TYPE, "TONY", "Application type tag."),
    (APPLICATION_TAGS, "", "Comma-separated application tags."),
    (APPLICATION_NODE_LABEL, "", "Node label for the whole application (recorded only on the local cluster)."),
    (ENABLE_PREPROCESS, "false", "Run the AM command first as a preprocessing job."),
    (APPLICATION_TIMEOUT, "0", "Whole-application timeout in ms; 0 disables it."),
    (FRAMEWORK, "tensorflow", "tensorflow | pytorch | horovod | mxnet | standalone."),
    (NUM_CLIENT_RM_CONNECT_RETRIES, "3", "Client retries when the resource manager is unreachable."),
    (DISTRIBUTED_MODE, "GANG", "GANG: start user processes only once every task registered; FCFS: start immediately."),
    (MAX_TOTAL_INSTANCES, "-1", "Upper bound on the total number of task instances; -1 = unlimited."),
    (MAX_TOTAL_GPUS, "-1", "Upper bound on the total number of GPUs requested; -1 = unlimited."),
    (AM_JVM_OPTS, "", "Kept for compatibility; the ApplicationMaster is not a JVM here."),
    (EXECUTOR_JVM_OPTS, "-Xmx1536m", "Kept for compatibility; the TaskExecutor is not a JVM here."),
    (HEARTBEAT_INTERVAL_MS, "1000", "TaskExecutor heartbeat period."),
    (MAX_MISSED_HEARTBEATS, "25", "Missed heartbeats before a task is declared dead."),
    (METRICS_INTERVAL_MS, "5000", "TaskMonitor sampling period."),
    (GPU_METRICS_ENABLED, "true", "Collect GPU utilisation / memory metrics."),
    (AM_RETRY_COUNT, "0", "Whole-session retries after a failed attempt."),
    (AM_MEMORY, "2g", "ApplicationMaster memory."),
    (AM_VCORES, "1", "ApplicationMaster vcores."),
    (AM_GPUS, "0", "ApplicationMaster GPUs."),
    (AM_WAIT_CLIENT_STOP_TIMEOUT_SEC, "15", "How long the AM waits for the client's finishApplication."),
    ("tony.ps.memory", "2g", "Parameter-server memory."),
    ("tony.ps.vcores", "1", "Parameter-server vcores."),
    (CONTAINER_ALLOCATION_TIMEOUT, "-1", "Task-registration timeout in ms; -1 disables it."),
    (WORKER_TIMEOUT, "0", "User-process timeout in ms for every task type; 0 disables it."),
    ("tony.worker.memory", "2g", "Worker memory."),
    ("tony.worker.vcores", "1", "Worker vcores."),
    ("tony.worker.gpus", "0", "GPUs per worker (one task is pinned to each GPU it is given)."),
    (UNTRACKED_JOBTYPES, "ps", "Job types whose completion the application does not wait for."),
    (SIDECAR_JOBTYPES, "tensorboard", "Job types whose failure is ignored."),
    (STOP_ON_FAILURE_JOBTYPES, "ps", "Job types whose failure fails the application at once."),
    (FAIL_ON_WORKER_FAILURE, "false", "Fail the application on the first failing tracked task."),
    (DOCKER_ENABLED, "false", "Launch tasks through the docker container runtime."),
    (SECURITY_ENABLED, "true", "Authenticate RPC peers with the per-application secret."),
    (HDFS_CONF_PATH, "", "Extra hdfs-site.xml (recorded only: the local cluster has no HDFS)."),
    (YARN_CONF_PATH, "", "Extra yarn-site.xml (recorded only: the local cluster has no YARN)."),
    (MAPRED_CONF_PATH, "", "Extra mapred-site.xml (recorded only)."),
    (PORTAL_URL, "https://localhost:19886", "Base URL of the history portal."),
    (HISTORY_LOCATION, "/path/to/tony-history", "Root of the job-history tree."),
    (HISTORY_INTERMEDIATE, "/path/to/tony-history/intermediate", "Where running jobs write history."),
    (HISTORY_FINISHED, "/path/to/tony-history/finished", "Where finished jobs are moved (yyyy/MM/dd)."),
    (HISTORY_MOVER_INTERVAL_MS, "300000", "History mover period."),
    (HISTORY_RETENTION_SEC, "2592000", "How long finished history is kept."),
    (HISTORY_PURGER_INTERVAL_MS, "21600000", "History purger period."),
    (HISTORY_FINISHED_DIR_TIMEZONE, "UTC", "Time zone of the yyyy/MM/dd finished directories."),
    (PORTAL_CACHE_MAX_ENTRIES, "10000", "Entries per portal cache."),
    (KEYTAB_USER, "user1", "Portal keytab principal (recorded only)."),
    (KEYTAB_LOCATION, "/location/to/tony.keytab", "Portal keytab file (recorded only)."),
    (HTTPS_PORT, "19886", "Portal HTTPS port."),
    (HTTPS_KEYSTORE_PATH, "/location/to/tony.jks", "Portal keystore path."),
    (HTTPS_KEYSTORE_TYPE, "JKS", "Portal keystore type."),
    (HTTPS_KEYSTORE_PASSWORD, "password", "Portal keystore password."),
    (HTTPS_KEYSTORE_ALGORITHM, "SunX509", "Portal keystore algorithm."),
    (HTTP_PORT, "disabled", "Portal HTTP port, or 'disabled'."),
    (SECRET_KEY, "changeme", "Portal application secret."),
    (HOROVOD_TEST_FAST_FAIL, "false", "Test hook: the rendezvous driver fails at once."),
    (HOROVOD_TEST_MODE, "false", "Test hook: the rendezvous driver serves a fixed fake plan."),
    (HOROVOD_DRIVER_DEBUG, "false", "Run a user-supplied driver command instead of the built-in driver."),
    (B200_WORK_DIR, "", "Where the local cluster keeps per-application working directories."),
    (B200_NUM_GPUS, "-1", "GPUs offered by the local cluster; -1 = discover."),
    (B200_CLUSTER_MEMORY, "", "Memory offered by the local cluster; empty = physical memory."),
    (B200_CLUSTER_VCORES, "-1", "vcores offered by the local cluster; -1 = CPU count (oversubscribable)."),
    (B200_KERNELS, "fused", "fused: hand-written sm_100a data plane; stock: torch/NCCL baseline."),
    (B200_BUCKET_MB, "64", "Gradient bucket size for the fused reduce+optimizer kernel."),
    (B200_COMM_CTAS, "32", "CTAs the fused collective may occupy while backward is running."),
    (B200_PIN_NUMA, "true", "Pin each task to the CPUs nearest its GPU."),
]

DEFAULTS_MAP = {k: v for k, v, _ in DEFAULTS}

# keys that exist without a default in tony-default.xml (TestTonyConfigurationFields skip list)
KEYS_WITHOUT_DEFAULT = (
    PREPARE_STAGE, TRAINING_STAGE, HADOOP_LOCATION, HADOOP_CLASSPATH, TENSORBOARD_LOG_DIR, AM_COMMAND,
    CONTAINERS_RESOURCES, CONTAINERS_COMMAND, CONTAINERS_ENVS, EXECUTION_ENVS,
    DOCKER_CONTAINERS_IMAGE, DOCKER_CONTAINERS_MOUNT, GPU_EXEC_PATH, PYTHON_EXEC_PATH,
)

# ---- environment variables / constants (Constants.java) ------------------------------------
JOB_NAME = "JOB_NAME"
TASK_INDEX = "TASK_INDEX"
TASK_NUM = "TASK_NUM"
IS_CHIEF = "IS_CHIEF"
JOB_ID = "JOB_ID"
APP_ID = "appid"
SESSION_ID = "SESSION_ID"
DISTRIBUTED_MODE_NAME = "DISTRIBUTED_MODE"
AM_HOST = "AM_HOST"
AM_PORT = "AM_PORT"
METRICS_RPC_PORT = "METRICS_RPC_PORT"
NUM_AM_RETRIES = "NUM_AM_RETRIES"
ATTEMPT_NUMBER = "ATTEMPT_NUMBER"
MODEL_PARAMS = "MODEL_PARAMS"
PREPROCESSING_JOB = "PREPROCESSING_JOB"
TB_PORT = "TB_PORT"
TB_LOG_DIR = "TB_LOG_DIR"
CLUSTER_SPEC = "CLUSTER_SPEC"
TF_CONFIG = "TF_CONFIG"
INIT_METHOD = "INIT_METHOD"
RANK = "RANK"
WORLD = "WORLD"
COMMUNICATION_BACKEND = "tcp://"
CONTAINER_ID = "CONTAINER_ID"
TONY_SECRET = "TONY_B200_SECRET"
TONY_CONF_PATH = "TONY_B200_FINAL_CONF"
TONY_TOPOLOGY = "TONY_TOPOLOGY"

CHIEF_JOB_NAME = "chief"
WORKER_JOB_NAME = "worker"
PS_JOB_NAME = "ps"
EVALUATOR_JOB_NAME = "evaluator"
DRIVER_JOB_NAME = "driver"
SCHEDULER_JOB_NAME = "scheduler"
SERVER_JOB_NAME = "server"
NOTEBOOK_JOB_NAME = "notebook"
TENSORBOARD_JOB_NAME = "tensorboard"
AM_NAME = "am"

ARCHIVE_SUFFIX = "#archive"
RESOURCE_DIVIDER = "::"
TONY_FINAL_XML = "tony-final.xml"
TONY_DEFAULT_XML = "tony-default.xml"
TONY_XML = "tony.xml"
TONY_SITE_XML = "tony-site.xml"
TONY_CONF_DIR_ENV = "TONY_CONF_DIR"
TONY_SRC_ZIP_PREFIX = "tony_src_"
PYTHON_VENV_ZIP = "venv.zip"
PYTHON_VENV_DIR = "venv"
KILLED_BY_APPMASTER = -105  # YARN ContainerExitStatus.KILLED_BY_APPMASTER

# fault-injection hooks, same names as the reference (Constants.java:124-130)
TEST_AM_CRASH = "TEST_AM_CRASH"
TEST_AM_THROW_EXCEPTION_CRASH = "TEST_AM_THROW_EXCEPTION_CRASH"
TEST_WORKER_TERMINATED = "TEST_WORKER_TERMINATION"
TEST_TASK_COMPLETION_NOTIFICATION_DELAYED = "TEST_TASK_COMPLETION_NOTIFICATION_DELAYED"
TEST_TASK_EXECUTOR_NUM_HB_MISS = "TEST_TASK_EXECUTOR_NUM_HB_MISS"
TEST_TASK_EXECUTOR_SKEW = "TEST_TASK_EXECUTOR_SKEW"
