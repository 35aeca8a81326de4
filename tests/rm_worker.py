import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from paddlebox_b200 import fleet
from paddlebox_b200.role_maker import GeneralRoleMaker, PaddleCloudRoleMaker
kind = sys.argv[1]
rm = GeneralRoleMaker(path="/tmp/ignored", init_timeout_seconds=120) if kind == "general" else PaddleCloudRoleMaker(is_collective=(kind == "collective"))
fleet.init(rm)
if fleet.is_server():
    fleet.run_server()
    print(json.dumps({"server": rm.server_index()}))
    sys.exit(0)
out = np.zeros(2)
rm.all_reduce_worker(np.array([fleet.worker_index() + 1.0, 10.0]), out)
gathered = rm.all_gather(fleet.worker_index())
rm.barrier_worker()
print(json.dumps({"rank": fleet.worker_index(), "world": fleet.worker_num(), "first": fleet.is_first_worker(), "sum": out.tolist(),
                  "gathered": gathered, "eps": rm.get_trainer_endpoints(), "servers": rm.get_pserver_endpoints()}))
fleet.stop_worker()
