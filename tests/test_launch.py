"""The job launcher: per-process environment, logs, failure handling."""
import os
import socket
import sys

import pytest

from paddlebox_b200 import launch as L

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_plan_single_and_multi_node():
    a = L._parse(["--gpus", "0,1,2,3", "--start_port", "7000", "train.py", "--x", "1"])
    p = L.plan(a)
    assert p["world"] == 4 and a.training_script_args == ["--x", "1"]
    e = p["envs"][2]
    assert e["PADDLE_TRAINER_ID"] == "2" and e["FLAGS_selected_gpus"] == "2" and e["LOCAL_RANK"] == "2"
    assert e["PADDLE_TRAINER_ENDPOINTS"].split(",") == [f"127.0.0.1:{7000 + i}" for i in range(4)]
    assert e["PADDLE_CURRENT_ENDPOINT"] == "127.0.0.1:7002" and e["MASTER_PORT"] == "6999" and e["WORLD_SIZE"] == "4"
    a = L._parse(["--ips", "10.0.0.1,10.0.0.2", "--node_rank", "1", "--nproc_per_node", "2", "--gpus", "4,5", "t.py"])
    p = L.plan(a)
    assert p["world"] == 4 and [e["RANK"] for e in p["envs"]] == ["2", "3"] and p["envs"][0]["MASTER_ADDR"] == "10.0.0.1"
    assert p["envs"][1]["PADDLE_CURRENT_ENDPOINT"] == "10.0.0.2:6171" and p["envs"][1]["FLAGS_selected_gpus"] == "5"
    with pytest.raises(SystemExit):
        L.plan(L._parse(["--ips", "10.9.9.1,10.9.9.2", "t.py"]))             # which node is this?
    with pytest.raises(SystemExit):
        L.plan(L._parse(["--gpus", "0", "--nproc_per_node", "2", "t.py"]))


@pytest.mark.parametrize("mode", ["plain", "rolemaker"])
def test_two_trainers_join_one_world(tmp_path, mode, capfd):
    port = _free_port()
    rc = L.launch(["--nproc_per_node", "2", "--start_port", str(port), "--master_port", str(_free_port()), "--backend", "gloo",
                   "--log_dir", str(tmp_path / "log"), "--server_num", "1", os.path.join(HERE, "launch_worker.py"), mode, "--lr", "0.1"])
    out = capfd.readouterr()
    assert rc == 0, out
    logs = [(tmp_path / "log" / f"workerlog.{i}").read_text() for i in range(2)]
    assert "rank 0 of 2 sum 3.0 args ['--lr', '0.1']" in logs[0] and "rank 1 of 2 sum 3.0" in logs[1]
    assert "rank 0 of 2" in out.out and "no parameter-server processes" in out.err


def test_a_failing_trainer_stops_the_job(tmp_path, capfd):
    import time
    t0 = time.time()
    rc = L.launch(["--nproc_per_node", "2", "--start_port", str(_free_port()), "--master_port", str(_free_port()),
                   "--log_dir", str(tmp_path / "log"), os.path.join(HERE, "launch_worker.py"), "fail"])
    assert rc == 7 and time.time() - t0 < 45
    assert "exited with code 7" in capfd.readouterr().err


def test_ported_job_script_runs_and_resumes_under_the_launcher(tmp_path, capfd):
    script = os.path.join(os.path.dirname(HERE), "examples", "ported_job.py")
    common = ["--nproc_per_node", "2", "--backend", "gloo", "--log_dir", str(tmp_path / "log")]
    job = ["--cpu", "--out", str(tmp_path / "job"), "--lines", "1500"]
    rc = L.launch(common + ["--start_port", str(_free_port()), "--master_port", str(_free_port()), script] + job + ["--epochs", "3"])
    out = capfd.readouterr().out
    assert rc == 0, out
    assert "epoch 0: 3000 instances" in out and "epochs run: [0, 1, 2]" in out
    auc = float(out.split("epoch 2: 3000 instances, auc ")[1].split()[0])
    assert auc > 0.58, out                                                     # it learns something
    rc = L.launch(common + ["--start_port", str(_free_port()), "--master_port", str(_free_port()), script] + job + ["--epochs", "4"])
    out = capfd.readouterr().out
    assert rc == 0 and "epochs run: [3]" in out and "epoch 0" not in out        # resumed after the saved epochs
    assert float(out.split("epoch 3: 3000 instances, auc ")[1].split()[0]) > 0.58, out   # ... with the trained state
