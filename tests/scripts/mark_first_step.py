import sys
sys.path.insert(0, __import__("os").environ.get("TONY_REPO", "."))
from tony_b200 import runtime
runtime.mark_first_step()
import time; time.sleep(0.3)
