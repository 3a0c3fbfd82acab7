import os, sys
if os.environ.get("PREPROCESSING_JOB") == "true":
    print("Model parameters: --lr=0.5")
    sys.exit(0)
assert os.environ.get("MODEL_PARAMS") == "--lr=0.5", os.environ.get("MODEL_PARAMS")
sys.exit(0)
