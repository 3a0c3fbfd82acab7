import sys
sys.path.insert(0, __import__("os").environ.get("TONY_REPO", "."))
from tony_b200 import runtime
runtime.mark_first_step()
runtime.report_metrics({"STEP_TIME_MS": 12.5, "SAMPLES_PER_SEC": 2560.0})
import time; time.sleep(0.3)
