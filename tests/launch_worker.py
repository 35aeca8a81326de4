"""Child script of tests/test_launch.py: joins the world the launcher described and checks what it was told."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from paddlebox_b200 import fleet                                   # noqa: E402
from paddlebox_b200.role_maker import PaddleCloudRoleMaker          # noqa: E402

mode = sys.argv[1]
if mode == "fail" and os.environ["PADDLE_TRAINER_ID"] == "1":
    sys.exit(7)
if mode == "fail":
    time.sleep(60)                                                  # the launcher must stop this one
    sys.exit(0)
if mode == "rolemaker":
    fleet.init(PaddleCloudRoleMaker(is_collective=True))
else:
    fleet.init()
rank, world = fleet.worker_index(), fleet.worker_num()
assert rank == int(os.environ["PADDLE_TRAINER_ID"]) == int(os.environ["RANK"]) and world == int(os.environ["PADDLE_TRAINERS_NUM"])
assert os.environ["PADDLE_CURRENT_ENDPOINT"] == os.environ["PADDLE_TRAINER_ENDPOINTS"].split(",")[rank]
total = fleet.util.all_reduce(np.array([rank + 1.0]), "sum")
print(f"rank {rank} of {world} sum {float(total[0])} args {sys.argv[2:]}", flush=True)
fleet.stop_worker()
