import os, sys
assert os.environ.get("ENV_CHECK") == "ENV_CHECK", os.environ.get("ENV_CHECK")
sys.exit(0)
