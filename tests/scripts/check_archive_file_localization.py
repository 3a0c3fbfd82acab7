import os, sys
# tony.worker.resources = common.zip#archive , test2.zip::renamed.zip#archive, plain.txt
assert os.path.isdir("common.zip"), os.listdir(".")
assert os.path.isdir("renamed.zip"), os.listdir(".")
assert os.path.isfile("plain.txt"), os.listdir(".")
sys.exit(0)
