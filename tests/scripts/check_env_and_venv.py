import os, sys
assert os.path.isfile(os.path.join("venv", "123.xml")), os.listdir(".")
sys.exit(0)
