import os, sys
for k in ("RANK", "WORLD", "INIT_METHOD", "MASTER_ADDR", "MASTER_PORT", "WORLD_SIZE", "LOCAL_RANK"):
    assert os.environ.get(k), k
assert os.environ["INIT_METHOD"].startswith("tcp://")
assert int(os.environ["WORLD"]) == 2 and 0 <= int(os.environ["RANK"]) < 2
sys.exit(0)
