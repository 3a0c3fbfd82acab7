import os, sys
is_chief = os.environ.get("IS_CHIEF") == "true"
has_tb = "TB_PORT" in os.environ
sys.exit(0 if is_chief == has_tb else 1)
