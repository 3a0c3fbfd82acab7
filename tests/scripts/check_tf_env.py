import json, os, sys
spec = json.loads(os.environ["CLUSTER_SPEC"])
cfg = json.loads(os.environ["TF_CONFIG"])
assert cfg["task"]["type"] == os.environ["JOB_NAME"]
assert cfg["task"]["index"] == int(os.environ["TASK_INDEX"])
assert set(cfg["cluster"]) <= set(spec)
sys.exit(0)
