import os, sys, time
# first session: worker 0 fails; the retried session (NUM_AM_RETRIES set) succeeds
if os.environ.get("NUM_AM_RETRIES"):
    sys.exit(0)
time.sleep(0.2)
sys.exit(1 if os.environ["TASK_INDEX"] == "0" else 0)
