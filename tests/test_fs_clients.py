"""LocalFS / HDFSClient / AFSClient (reference fleet/utils/fs.py) and the hadoop passthrough of BoxFileMgr and
FleetUtil, against a stand-in ``hadoop`` executable (tests/fake_hadoop.py) that serves a local directory."""
import json
import os
import stat
import sys

import pytest

from paddlebox_b200.utils import fs as FS


@pytest.fixture
def hadoop(tmp_path, monkeypatch):
    home = tmp_path / "hadoop_home"
    (home / "bin").mkdir(parents=True)
    exe = home / "bin" / "hadoop"
    exe.write_text(f"#!/bin/sh\nexec {sys.executable} {os.path.join(os.path.dirname(__file__), 'fake_hadoop.py')} \"$@\"\n")
    exe.chmod(exe.stat().st_mode | stat.S_IEXEC)
    root = tmp_path / "hdfs_root"
    root.mkdir()
    monkeypatch.setenv("FAKE_HDFS_ROOT", str(root))
    monkeypatch.setenv("FAKE_HDFS_LOG", str(tmp_path / "hadoop.log"))
    monkeypatch.setenv("PATH", str(home / "bin") + os.pathsep + os.environ["PATH"])
    return str(home), str(root), str(tmp_path / "hadoop.log")


def _exercise(client, base, tmp_path):
    """The behaviours shared by every client."""
    d = f"{base}/job/day1"
    assert client.ls_dir(d) == ([], []) and not client.is_exist(d)
    client.mkdirs(d)
    client.mkdirs(d)                                          # idempotent
    assert client.is_dir(d) and not client.is_file(d)
    client.touch(f"{d}/_SUCCESS")
    client.touch(f"{d}/_SUCCESS")                             # exist_ok
    with pytest.raises(FS.FSFileExistsError):
        client.touch(f"{d}/_SUCCESS", exist_ok=False)
    local = tmp_path / "part-0"
    local.write_text("hello\nworld\n")
    client.upload(str(local), f"{d}/part-0")
    with pytest.raises(FS.FSFileExistsError):
        client.upload(str(local), f"{d}/part-0")
    client.upload(str(local), f"{d}/part-0", overwrite=True)
    with pytest.raises(FS.FSFileNotExistsError):
        client.upload(str(tmp_path / "nope"), f"{d}/x")
    client.mkdirs(f"{d}/sub")
    dirs, files = client.ls_dir(d)
    assert dirs == ["sub"] and sorted(files) == ["_SUCCESS", "part-0"] and client.list_dirs(d) == ["sub"]
    assert client.cat(f"{d}/part-0") == "hello\nworld" and client.cat(f"{d}/missing") == ""
    assert client.is_file(f"{d}/part-0") and client.list_files_info([f"{d}/part-0", f"{d}/zz"]) == \
        [{"path": f"{d}/part-0", "size": 12}]
    client.mv(f"{d}/part-0", f"{d}/part-1")
    with pytest.raises(FS.FSFileNotExistsError):
        client.mv(f"{d}/part-0", f"{d}/part-2", test_exists=True)
    client.touch(f"{d}/part-2")
    with pytest.raises(FS.FSFileExistsError):
        client.mv(f"{d}/part-1", f"{d}/part-2", test_exists=True)
    client.mv(f"{d}/part-1", f"{d}/part-2", overwrite=True, test_exists=True)
    got = tmp_path / "back"
    client.download(f"{d}/part-2", str(got))
    assert got.read_text() == "hello\nworld\n"
    with pytest.raises(FS.FSFileExistsError):
        client.download(f"{d}/part-2", str(got))
    client.delete(f"{d}/sub")
    client.delete(f"{d}/sub")                                 # idempotent
    client.delete(d)
    assert not client.is_exist(d)
    assert FS.FS._split_files(list("abcdefg"), 0, 3) == ["a", "b", "c"] and FS.FS._split_files(list("abcdefg"), 2, 3) == ["f", "g"]


def test_local_fs(tmp_path):
    c = FS.LocalFS()
    assert not c.need_upload_download()
    _exercise(c, str(tmp_path / "lfs"), tmp_path)


def test_hdfs_client_over_the_hadoop_shell(hadoop, tmp_path):
    home, root, log = hadoop
    c = FS.HDFSClient(home, {"fs.default.name": "hdfs://ns1:54310", "hadoop.job.ugi": "u,p"}, time_out=20000, sleep_inter=50)
    assert c.need_upload_download()
    _exercise(c, "hdfs:/user/test", tmp_path)
    first = open(log).readline()
    assert first.startswith("fs -Dfs.default.name=hdfs://ns1:54310 -Dhadoop.job.ugi=u,p -test -e hdfs:/user/test/job/day1")
    # a client whose hadoop binary is missing says so
    with pytest.raises(FS.ExecuteError):
        FS.HDFSClient(str(tmp_path / "no_such_home"), {}).is_exist("hdfs:/x")


def test_afs_client_and_timeout(hadoop, tmp_path, monkeypatch):
    home, root, log = hadoop
    monkeypatch.setenv("HADOOP_HOME", home)
    c = FS.AFSClient(time_out=20000, sleep_inter=50)
    c.init("afs://xx.afs.com:9902", "user", "pw", "./afs.conf")
    _exercise(c, "afs:/user/afs_test", tmp_path)
    assert "-conf ./afs.conf -Dfs.default.name=afs://xx.afs.com:9902 -Dhadoop.job.ugi=user,pw" in open(log).readline()
    # a command that keeps failing gives up at the time-out instead of retrying for ever
    slow = FS.HDFSClient(home, {}, time_out=300, sleep_inter=200)
    with pytest.raises((FS.FSTimeOut, FS.ExecuteError)):
        slow.mv("hdfs:/nope/a", "hdfs:/nope/b", test_exists=False)


def test_box_file_mgr_and_fleet_util_on_remote_paths(hadoop, tmp_path):
    """The hadoop passthrough of BoxFileMgr and the donefile logic of FleetUtil on ``hdfs:`` paths."""
    import torch
    from paddlebox_b200 import fleet_util as FU
    from paddlebox_b200.side_tables import BoxFileMgr
    home, root, log = hadoop
    mgr = BoxFileMgr()
    mgr.init("hdfs://ns1", "u,p")
    assert mgr.makedir("hdfs:/out/m") and mgr.exists("hdfs:/out/m") and not mgr.exists("hdfs:/out/zz")
    src = tmp_path / "f.txt"
    src.write_text("abc")
    assert mgr.upload(str(src), "hdfs:/out/m/f.txt") and mgr.list_dir("hdfs:/out/m") == ["hdfs:/out/m/f.txt"]
    assert mgr.list_info("hdfs:/out/m") == [("hdfs:/out/m/f.txt", 3)] and mgr.file_size("hdfs:/out/m") == 3
    assert mgr.download("hdfs:/out/m/f.txt", str(tmp_path / "g.txt")) and (tmp_path / "g.txt").read_text() == "abc"
    assert mgr.remove("hdfs:/out/m") and not mgr.exists("hdfs:/out/m")
    u = FU.FleetUtil(fs=mgr)
    out = "hdfs:/out/job"
    assert u.get_last_save_xbox(out, "hdfs://ns1", "u,p", home)[:2] == [-1, -1]
    assert u.write_xbox_donefile(out, 20240102, 1, 77, "/data", "hdfs://ns1", "u,p", hadoop_home=home)
    assert u.write_xbox_donefile(out, 20240102, 2, 77, "/data", "hdfs://ns1", "u,p", hadoop_home=home)
    assert not u.write_xbox_donefile(out, 20240102, 2, 77, "/data", "hdfs://ns1", "u,p", hadoop_home=home)
    lines = open(os.path.join(root, "out/job/xbox_patch_done.txt")).read().strip().split("\n")
    assert len(lines) == 2 and json.loads(lines[1])["input"] == "hdfs://ns1/out/job/20240102/delta-2/000"
    assert u.get_last_save_xbox(out, "hdfs://ns1", "u,p", home) == [20240102, 2, "hdfs://ns1/out/job/20240102/delta-2", 77]
    assert u.write_model_donefile(out, 20240102, 2, 77, "hdfs://ns1", "u,p", hadoop_home=home)
    assert u.get_last_save_model(out, "hdfs://ns1", "u,p", home) == [20240102, 2, "hdfs:/out/job/20240102/2/", 77]
